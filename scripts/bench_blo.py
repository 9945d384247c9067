"""One Newton-Raphson smoothing pass (pllmod_opt_optimize_branch_lengths_local_multi, all branches):
wall time, branches/s and kernel launches.  usage: python scripts/bench_blo.py [tips] [sites] [states]"""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from pllharness import Harness, load_product_package
tips = int(sys.argv[1]) if len(sys.argv) > 1 else 200
sites = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
states = int(sys.argv[3]) if len(sys.argv) > 3 else 4
pkg = load_product_package()
for opt in ("fused_newton", "queued_newton", "small_stage", "trav_variant", "trav_stash", "trav_short"):
    if os.environ.get("PLLB200_" + opt.upper()):
        pkg.set_option(opt, int(os.environ["PLLB200_" + opt.upper()]))
with Harness("b200", tips=tips, sites=sites, states=states) as h:
    ll = h.loglh(0)
    h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)   # warm-up pass
    pkg.reset_stats()
    t0 = time.perf_counter()
    nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
    dt = time.perf_counter() - t0
    st = pkg.stats()
    nb = 2 * tips - 3
    print(json.dumps({"workload": "BLO smoothing pass, %d taxa x %d sites, %d states" % (tips, sites, states),
                      "seconds": dt, "branches": nb, "us_per_branch": dt / nb * 1e6, "neg_loglh": nll,
                      "kernel_launches": st["kernel_launches"], "launches_per_branch": st["kernel_launches"] / nb,
                      "newton_iterations_per_branch": st["newton_iterations"] / nb, "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"]}))
