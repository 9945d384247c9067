import os, sys, time
sys.path.insert(0, "tests")
from pllharness import Harness, load_product_package
pkg = load_product_package()
tips, nparts, sites = 200, 100, 2500
states = [4 if p % 2 == 0 else 20 for p in range(nparts)]
alphas = [0.3 + 0.01 * p for p in range(nparts)]
with Harness("b200", tips=tips, sites=[sites] * nparts, states=states, alpha=alphas) as h:
    h.loglh(0); h.loglh(0)
    t0 = time.perf_counter(); h.loglh(0); print("wall ms", (time.perf_counter() - t0) * 1e3)
