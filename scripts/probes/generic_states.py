"""Where does a generic alphabet (17..64 states) go non-finite?  Prints, per state count and kernel choice, the
log-likelihood, the library's error text and which buffers hold non-finite values."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package  # noqa: E402

pkg = load_product_package()
for states in (25, 28, 31, 32, 33, 40, 61):
    for mode in (0, 1):
        pkg.set_option("generic_dmma", mode)
        for attrs in (None, 0):
            kw = dict(tips=8, sites=40, states=states, rate_cats=2)
            if attrs is not None:
                kw["attributes"] = attrs
            with Harness("b200", **kw) as h:
                ll = h.loglh()
                bad = []
                for m in range(2 * 8 - 3):
                    if not np.isfinite(h.pmatrix(m)).all():
                        bad.append("pmat%d" % m)
                for i in range(8 - 2):
                    v = h.clv(8 + i)
                    if not np.isfinite(v).all():
                        bad.append("clv%d(%d)" % (8 + i, int((~np.isfinite(v)).sum())))
                err = h.lib.h_errmsg().decode() if hasattr(h.lib, "h_errmsg") else ""
                print(states, "dmma" if mode else "scalar", "tipclv" if attrs == 0 else "pattern", ll, bad[:6], err[:100], flush=True)
pkg.set_option("generic_dmma", 1)
