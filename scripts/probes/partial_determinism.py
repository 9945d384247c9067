"""Repeat a fixed sequence of re-rootings + incremental logL evaluations under different kernel options and
report the first evaluation whose logL differs from the reference sequence (trav_variant=3, trav_forward=0)."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np
from pllharness import Harness, load_product_package
pkg = load_product_package()
tips, sites, n = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
def run(opts):
    for k, v in (("trav_short", 3), ("trav_variant", 0), ("trav_stash", 2), ("trav_forward", 1), ("trav_reorder", 1)):
        pkg.set_option(k, opts.get(k, v))
    out = []
    with Harness("b200", tips=tips, sites=sites) as h:
        h.loglh(0)
        rng = np.random.RandomState(5)
        nodes = rng.randint(tips, 4 * tips - 6, size=n)
        for nd in nodes:
            h.perturb_edge(int(nd), 0.05 + 0.1 * rng.rand())
            h.set_root(int(nd))
            c0 = h.ops_counter()
            out.append((h.loglh(1), h.ops_counter() - c0))
        out.append((h.loglh(0), 0))
    return out
ref = run({"trav_variant": 3, "trav_forward": 0})
for name, opts in (("ref_again", {"trav_variant": 3, "trav_forward": 0}), ("old_fwd", {"trav_variant": 3}), ("old_fwd_again", {"trav_variant": 3}),
                   ("old_fwd_noreorder", {"trav_variant": 3, "trav_reorder": 0}),
                   ("lean", {"trav_short": 0}), ("lean_again", {"trav_short": 0}), ("lean_nofwd", {"trav_short": 0, "trav_forward": 0, "trav_stash": 0})):
    got = run(opts)
    bad = [i for i, (a, b) in enumerate(zip(ref, got)) if a[0] != b[0]]
    print(name, "mismatches", len(bad), "first", bad[:3], [(ref[i], got[i]) for i in bad[:2]], flush=True)
