"""Run-to-run determinism of an SPR round at a size where kernels overlap on the stream."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package
pkg = load_product_package()
tips, sites = int(sys.argv[1]), int(sys.argv[2])
for rep in range(int(sys.argv[3]) if len(sys.argv) > 3 else 2):
    for k in sys.argv[4:]:
        name, v = k.split("=")
        pkg.set_option(name, int(v))
    with Harness("b200", tips=tips, sites=sites) as h:
        start = h.loglh(0)
        lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=5, ntopol_keep=20, smoothings=32, epsilon=1.0, bl_min=1e-6)
        print(rep, json.dumps(dict(start=start, lh=lh, ops=h.ops_counter(), cutoff=co.tolist())), flush=True)
