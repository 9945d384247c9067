"""One SPR round at BASELINE configs[4] (C5: DNA GTR+G4, 2000 taxa x 500000 sites, fast mode, radius 1..5,
ntopol_keep 20), wall clock, with the library's phase trace (PLLB200_TRACE_SPR=1 prints it to stderr).
usage: python scripts/bench_spr_c5.py [taxa] [sites]"""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
os.environ.setdefault("PLLB200_TRACE_SPR", "1")
from pllharness import Harness, load_product_package  # noqa: E402

pkg = load_product_package()
tips = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
sites = int(sys.argv[2]) if len(sys.argv) > 2 else 500000
t0 = time.perf_counter()
with Harness("b200", tips=tips, sites=sites) as h:
    start = h.loglh(0)
    setup = time.perf_counter() - t0
    pkg.reset_stats()
    t0 = time.perf_counter()
    lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=5, ntopol_keep=20, smoothings=32, epsilon=1.0, bl_min=1e-6)
    secs = time.perf_counter() - t0
    st = pkg.stats()
    print(json.dumps(dict(workload="SPR round, %d taxa x %d sites, fast mode, radius 1..5, ntopol_keep 20" % (tips, sites),
                          setup_s=setup, start_loglh=start, loglh=lh, seconds=secs, clv_ops=h.ops_counter(),
                          kernel_launches=st["kernel_launches"], h2d_bytes=st["h2d_bytes"], d2h_bytes=st["d2h_bytes"])), flush=True)
