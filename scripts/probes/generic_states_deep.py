"""Deep trees on a generic alphabet: where do the product and the reference part?"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package  # noqa: E402

pkg = load_product_package()
for kw in (dict(tips=40, sites=130, states=32, brlen_scale=14.0), dict(tips=20, sites=64, states=61, rate_cats=2, brlen_scale=12.0),
           dict(tips=40, sites=130, states=20, brlen_scale=14.0, rate_cats=3)):
    tips = kw["tips"]
    with Harness("ref", **kw) as r:
        lr = r.loglh()
        rc = [r.clv(tips + i).copy() for i in range(tips - 2)]
        rs = [r.scaler(i).copy() for i in range(tips - 2)]
    for mode in (0, 1):
        pkg.set_option("generic_dmma", mode)
        with Harness("b200", **kw) as h:
            ll = h.loglh()
            print(kw, "dmma" if mode else "scalar", ll, lr, flush=True)
            for i in range(tips - 2):
                v, s = h.clv(tips + i), h.scaler(i)
                nf = int((~np.isfinite(v)).sum())
                seq = np.array_equal(s, rs[i])
                err = float(np.nanmax(np.abs(v - rc[i]) / (np.abs(rc[i]) + 1e-300))) if nf < v.size else float("nan")
                if nf or not seq or err > 1e-9:
                    print("   clv", tips + i, "nonfinite", nf, "scaler equal", seq, "scaler sums", int(s.sum()), int(rs[i].sum()), "relerr", err, flush=True)
pkg.set_option("generic_dmma", 1)
