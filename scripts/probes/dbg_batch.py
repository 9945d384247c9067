"""Debug aid for batch mode (device_ctx.cu): the same smoothing pass with the collector off (0), on (1), and
with single kernel kinds kept out of it (bit k+1 set = kind k (internal.h: 0 P-matrix, 1 DNA CLV update, 2 protein CLV update, 3 DNA sumtable) launches per partition on the lead stream)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package
pkg = load_product_package()
cases = [("mixed", dict(tips=18, sites=[900, 300, 1300, 77, 520, 33], states=[4, 20, 4, 20, 4, 4],
                        alpha=[0.3, 0.6, 1.0, 2.0, 0.5, 0.8], pinv=0.15, linkage=0)),
         ("dna-ragged", dict(tips=18, sites=[900, 1300, 520], states=[4, 4, 4], alpha=[0.3, 1.0, 0.5], linkage=0)),
         ("aa", dict(tips=18, sites=[300, 77], states=[20, 20], alpha=[0.6, 2.0], linkage=0))]
for name, kw in cases:
    for mask in (0, 1, 31 - 2, 31 - 4, 31 - 8, 31 - 16):
        pkg.set_option("batch", mask)
        with Harness("b200", **kw) as h:
            h.loglh()
            pkg.reset_stats()
            v = h.optimize_branch_lengths(method=0, radius=-1, max_iters=2, lh_epsilon=0.01)
            print(name, "mask", mask, "nll %.9f" % v, "launches", pkg.stats()["kernel_launches"], flush=True)
