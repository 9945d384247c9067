"""C4 smoothing pass only (for launch lists): python scripts/probes/c4_blo_only.py [sites_per_partition]"""
import sys, os, time, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness, load_product_package
pkg = load_product_package()
for _o in ("fused_newton", "small_stage", "batch"):
    if os.environ.get("PLLB200_" + _o.upper()):
        pkg.set_option(_o, int(os.environ["PLLB200_" + _o.upper()]))
sites = int(sys.argv[1]) if len(sys.argv) > 1 else 2500
tips, nparts = 200, 100
states = [4 if p % 2 == 0 else 20 for p in range(nparts)]
alphas = [0.3 + 0.01 * p for p in range(nparts)]
with Harness("b200", tips=tips, sites=[sites] * nparts, states=states, alpha=alphas) as h:
    h.loglh(0)
    for rep in range(2):
        pkg.reset_stats()
        t0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        t = time.perf_counter() - t0
        st = pkg.stats()
        print(json.dumps({"pass": rep, "seconds": t, "us_per_branch": t / (2 * tips - 3) * 1e6, "neg_loglh": nll,
                          "launches_per_branch": st["kernel_launches"] / (2 * tips - 3),
                          "newton_iterations_per_branch": st["newton_iterations"] / (2 * tips - 3)}), flush=True)
