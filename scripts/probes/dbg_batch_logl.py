"""Debug aid: full-traversal logL of a multi-partition treeinfo with single kernel kinds collected."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package
pkg = load_product_package()
cases = [("dna", dict(tips=18, sites=[900, 1300, 520], states=[4, 4, 4], alpha=[0.3, 1.0, 0.5], linkage=0)),
         ("aa", dict(tips=18, sites=[300, 77], states=[20, 20], alpha=[0.6, 2.0], linkage=0))]
for name, kw in cases:
    for mask in (0, 1, 31 - 2, 31 - 4, 31 - 8, 31):
        pkg.set_option("batch", mask)
        with Harness("b200", **kw) as h:
            pkg.reset_stats()
            v = h.loglh(0)
            print(name, "mask", mask, "loglh %.9f" % v, "launches", pkg.stats()["kernel_launches"], flush=True)
