import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
from pllharness import *
pkg = load_product_package()
def check(kw, variant=None):
    if variant is not None: pkg.set_option("trav_variant", variant)
    with Harness("ref", **kw) as r, Harness("b200", **kw) as d:
        lr, ld = r.loglh(), d.loglh()
        print(kw, "variant", variant, "logL", lr, ld, "rel", abs(lr-ld)/abs(lr))
        tips = kw["tips"]
        for i in range(tips-2):
            a, b = r.clv(tips+i), d.clv(tips+i)
            sa, sb = r.scaler(i), d.scaler(i)
            if not np.allclose(a, b, rtol=1e-9, atol=0) or not np.array_equal(sa, sb):
                bad = np.argwhere(~np.isclose(a, b, rtol=1e-9, atol=0))
                print("  clv", tips+i, "mismatch count", len(bad), "first", bad[:5].tolist(), "sites", a.shape[0],
                      "scaler diff", int((sa != sb).sum()))
                break
        else:
            print("  all CLVs equal")
check(dict(tips=16, sites=777, rate_cats=1))
check(dict(tips=16, sites=768, rate_cats=1))
check(dict(tips=16, sites=1000, rate_cats=4), 0)
check(dict(tips=16, sites=1000, rate_cats=4), 2)
check(dict(tips=16, sites=1000, rate_cats=4), 1)
check(dict(tips=16, sites=1000, rate_cats=2), 0)
check(dict(tips=16, sites=1000, rate_cats=8), 0)
