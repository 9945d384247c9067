"""Does an SPR round depend on which traversal kernel runs the short operation lists?  (It must not.)"""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package
pkg = load_product_package()
tips = int(sys.argv[1]) if len(sys.argv) > 1 else 120
sites = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
res = {}
for name, opts in (("default", {}), ("lean_all", {"trav_short": 0}), ("old_all", {"trav_variant": 3}),
                   ("lean_all_stash0", {"trav_short": 0, "trav_stash": 0})):
    for k, v in (("trav_short", 3), ("trav_variant", 0), ("trav_stash", 2)):
        pkg.set_option(k, opts.get(k, v))
    with Harness("b200", tips=tips, sites=sites, seed=17) as h:
        start = h.loglh(0)
        lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=5, ntopol_keep=10, smoothings=16, epsilon=1.0, bl_min=1e-6)
        hs, ln = h.topology_signature()
        res[name] = dict(start=start, lh=lh, ops=h.ops_counter(), topo=hash(tuple(int(x) for x in hs)), cutoff=co.tolist())
    print(name, json.dumps(res[name]), flush=True)
