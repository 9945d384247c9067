"""Secondary measurements on the other BASELINE.json configs (the headline line stays bench.py):
  C1  DNA 16 taxa x 10k sites, full traversal logL (also timed on the CPU reference when present)
  C3  protein LG+G4 500 x 200k (see scripts/bench_aa.py for the kernel-only number)
  C4  100 partitions (50 DNA + 50 protein) x 20k sites, 200 taxa: this GPU's 1/8 site share (2500 sites
      per partition) by default, `--c4-sites N` for another share
  C5' DNA 2000 taxa x 500k sites: partial traversals per second (perturb one edge the way an SPR regraft
      evaluation does, then incremental pllmod_treeinfo_compute_loglh), and one full NR smoothing pass
usage: python scripts/bench_configs.py [c1] [c3] [c4] [c5] [--c4-sites N] [--c5-sites N]
Each result is one JSON line; wall clock through the C ABI (launch + sync included)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
from pllharness import Harness, have_reference, load_product_package

args = sys.argv[1:]
which = [a for a in args if not a.startswith("--")] or ["c1", "c3", "c4", "c5"]
def opt(name, default):
    return int(args[args.index(name) + 1]) if name in args else default
pkg = load_product_package()
for _o in ("fused_newton", "queued_newton", "small_stage", "batch", "aa_variant", "aa_tiptab"):
    if os.environ.get("PLLB200_" + _o.upper()):
        pkg.set_option(_o, int(os.environ["PLLB200_" + _o.upper()]))


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps


if "c1" in which:
    out = {"config": "C1 DNA GTR+G4 16 taxa x 10000 sites, pllmod_treeinfo_compute_loglh full traversal"}
    with Harness("b200", tips=16, sites=10000) as h:
        out["loglh"] = h.loglh(0)
        out["b200_ms"] = timed(lambda: h.loglh(0), 50) * 1e3
    if have_reference():
        with Harness("ref", tips=16, sites=10000) as r:
            out["ref_loglh"] = r.loglh(0)
            out["ref_cpu_1core_ms"] = timed(lambda: r.loglh(0), 20) * 1e3
    print(json.dumps(out), flush=True)

if "c3" in which:
    tips, sites = 500, opt("--c3-sites", 200000)
    with Harness("b200", tips=tips, sites=sites, states=20) as h:
        ll = h.loglh(0)
        t = timed(lambda: h.loglh(0), 3)
        c = h.op_type_counts()
        root = h.root()
        tl = timed(lambda: h.edge_loglh(root), 5)
        h.derivatives(root, [0.1])
        t1 = timed(lambda: h.derivatives(root, [0.1]), 3)
        t9 = timed(lambda: h.derivatives(root, list(np.linspace(0.01, 0.5, 9))), 3)
        td = (t9 - t1) / 8
        tb0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        tb = time.perf_counter() - tb0
        print(json.dumps({"config": "C3 protein LG+G4 %d taxa x %d sites" % (tips, sites), "loglh": ll, "ops": c,
                          "blo_one_smoothing_s": tb, "blo_branches": 2 * tips - 3, "blo_neg_loglh": nll,
                          "full_loglh_ms": t * 1e3, "site_updates_per_s": (tips - 2) * sites / t,
                          "TFLOPs_fp64": (tips - 2) * sites * 6480.0 / t / 1e12,
                          "edge_logl_ms": tl * 1e3, "edge_logl_GBps": 1292.0 * sites / tl / 1e9,
                          "deriv_ms": td * 1e3, "deriv_GBps": 644.0 * sites / td / 1e9,
                          "sumtable_ms": (t1 - td) * 1e3}), flush=True)

if "c4" in which:
    tips, nparts, sites = 200, 100, opt("--c4-sites", 2500)
    states = [4 if p % 2 == 0 else 20 for p in range(nparts)]
    alphas = [0.3 + 0.01 * p for p in range(nparts)]
    with Harness("b200", tips=tips, sites=[sites] * nparts, states=states, alpha=alphas) as h:
        ll = h.loglh(0)
        pkg.reset_stats()
        t = timed(lambda: h.loglh(0), 5)
        launches = pkg.stats()["kernel_launches"] / 6
        h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)   # first pass: allocates the sumtables
        pkg.reset_stats()
        tb0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        tb = time.perf_counter() - tb0
        blo_launches = pkg.stats()["kernel_launches"] / (2 * tips - 3)
        upd = (tips - 2) * sites * nparts
        print(json.dumps({"config": "C4 %d partitions (DNA/protein alternating, per-partition alpha) x %d sites, %d taxa"
                                    % (nparts, sites, tips), "loglh": ll, "full_loglh_ms": t * 1e3,
                          "site_updates_per_s": upd / t, "kernel_launches_per_eval": launches,
                          "blo_one_smoothing_s": tb, "blo_launches_per_branch": blo_launches,
                          "blo_neg_loglh": nll}), flush=True)

if "c5" in which:
    tips, sites = 2000, opt("--c5-sites", 500000)
    with Harness("b200", tips=tips, sites=sites) as h:
        ll = h.loglh(0)
        tfull = timed(lambda: h.loglh(0), 2)
        rng = np.random.RandomState(5)
        nodes = rng.randint(tips, 4 * tips - 6, size=200)   # inner subnodes
        c0 = h.ops_counter()
        t0 = time.perf_counter()
        for nd in nodes:
            h.perturb_edge(int(nd), 0.05 + 0.1 * rng.rand())
            h.set_root(int(nd))     # as algo_search.c does before scoring a regraft
            h.loglh(1)
        tpart = (time.perf_counter() - t0) / len(nodes)
        ops_per_eval = (h.ops_counter() - c0) / len(nodes)
        ll_inc = h.loglh(1)
        ll_full = h.loglh(0)
        tb0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        tb = time.perf_counter() - tb0
        print(json.dumps({"config": "C5' DNA GTR+G4 %d taxa x %d sites" % (tips, sites), "loglh": ll,
                          "full_loglh_ms": tfull * 1e3, "site_updates_per_s_full": (tips - 2) * sites / tfull,
                          "partial_traversal_evals_per_s": 1.0 / tpart, "partial_eval_ms": tpart * 1e3, "clv_ops_per_partial_eval": ops_per_eval,
                          "incremental_equals_full": abs(ll_inc - ll_full) <= 1e-9 * abs(ll_full),
                          "blo_one_smoothing_s": tb, "blo_branches": 2 * tips - 3, "blo_neg_loglh": nll}), flush=True)

if "c5spr" in which:
    # one full pllmod_algo_spr_round (fast mode, cutoff on) -- BASELINE configs[4]
    tips, sites, rmax = opt("--c5-tips", 2000), opt("--c5-sites", 500000), opt("--c5-radius", 5)
    with Harness("b200", tips=tips, sites=sites) as h:
        ll = h.loglh(0)
        pkg.reset_stats()
        c0 = h.ops_counter()
        t0 = time.perf_counter()
        lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=rmax, ntopol_keep=20, smoothings=32,
                             epsilon=1.0, bl_min=1e-6)
        dt = time.perf_counter() - t0
        st = pkg.stats()
        print(json.dumps({"config": "C5 pllmod_algo_spr_round fast, radius 1..%d, DNA GTR+G4 %d taxa x %d sites" % (rmax, tips, sites),
                          "start_loglh": ll, "loglh": lh, "seconds": dt, "clv_ops": h.ops_counter() - c0,
                          "clv_site_updates_per_s": (h.ops_counter() - c0) * sites / dt,
                          "kernel_launches": st["kernel_launches"], "launches_per_s": st["kernel_launches"] / dt,
                          "cutoff": co.tolist()}), flush=True)
    if have_reference() and "--no-ref" not in args:
        rs = opt("--c5-ref-sites", 2000)
        with Harness("ref", tips=tips, sites=rs) as r:
            r.loglh(0)
            c0 = r.ops_counter()
            t0 = time.perf_counter()
            lh, co = r.spr_round(thorough=False, radius_min=1, radius_max=rmax, ntopol_keep=20, smoothings=32,
                                 epsilon=1.0, bl_min=1e-6)
            dt = time.perf_counter() - t0
            print(json.dumps({"config": "C5 reference CPU (1 core, AVX2) same round at %d sites" % rs, "loglh": lh,
                              "seconds": dt, "clv_ops": r.ops_counter() - c0,
                              "clv_site_updates_per_s": (r.ops_counter() - c0) * rs / dt}), flush=True)
