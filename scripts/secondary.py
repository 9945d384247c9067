"""Secondary measurements on BASELINE.json configs[2..4], importable by bench.py (one dict per config) and
runnable on their own (`python scripts/secondary.py [c3] [c4] [c5]`).  Wall clock through the C ABI unless a key
says otherwise; every harness is closed before the next one is created (C3 and C5' need ~70 and ~130 GB).

  C3  protein LG+G4, 500 taxa x 200k sites: full traversal, tensor-pipe fraction, one NR smoothing pass
  C4  100 partitions (DNA / protein alternating, per-partition alpha) x 200 taxa, ONE GPU's 1/8 site share of the
      20k-site partitions (2500 sites each; the whole config is 152 GB of CLVs): logL and one smoothing pass
  C5' DNA 2000 taxa x 500k sites: full logL, partial-traversal evaluations/s (what an SPR regraft costs), one
      smoothing pass; and one complete pllmod_algo_spr_round on a bounded shape (taxa x sites in the key)
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
from pllharness import Harness, load_product_package

DMMA_PEAK = 37.15e12  # measured register-only mma.sync.m8n8k4.f64 peak (profiles/r01_dmma_peak.jsonl)


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps


def c3(sites=200000, tips=500):
    with Harness("b200", tips=tips, sites=sites, states=20) as h:
        ll = h.loglh(0)
        t = timed(lambda: h.loglh(0), 3)
        tk = h.time_update_partials(3)   # pll_update_partials over the whole op list + synchronise
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
    # executed DMMA work: a tip child is a table lookup; an inner child is 15 m8n8k4 (512 flop each) per 8 sites
    # and category (3 n-tiles x 5 k-steps, the third n-tile padded from 4 to 8 columns)
    inner_children = 2 * c["ii"] + c["ti"]
    dmma_flops = inner_children * (sites / 8.0) * 4 * 15 * 512
    return {"workload": "protein LG+G4 %d taxa x %d sites (BASELINE configs[2])" % (tips, sites), "loglh": ll,
            "traversal_ms": tk * 1e3, "full_loglh_ms": t * 1e3, "site_updates_per_s": (tips - 2) * sites / tk,
            "dense_equiv_TFLOPs_fp64": (tips - 2) * sites * 6480.0 / tk / 1e12,
            "dmma_frac": dmma_flops / tk / DMMA_PEAK, "dmma_peak_TFLOPs": DMMA_PEAK / 1e12,
            "edge_logl_ms": tl * 1e3, "edge_logl_GBps": 1292.0 * sites / tl / 1e9,
            "deriv_ms": td * 1e3, "deriv_GBps": 644.0 * sites / td / 1e9, "sumtable_ms": (t1 - td) * 1e3,
            "blo_one_smoothing_s": tb, "blo_branches": 2 * tips - 3, "blo_neg_loglh": nll}


def n1(sites=20000, tips=100, states=61):
    """Codon-sized alphabet (61 states, Gamma-4): full traversal on partial_dmma_kernel (one launch per operation)."""
    with Harness("b200", tips=tips, sites=sites, states=states) as h:
        ll = h.loglh(0)
        tk = h.time_update_partials(3)
        t = timed(lambda: h.loglh(0), 3)
        c = h.op_type_counts()
        h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)   # allocates the sumtable
        tb0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        tb = time.perf_counter() - tb0
    sp = (states + 3) // 4 * 4
    nt = (states + 7) // 8
    # executed DMMA work: an inner child is ceil(s/8) n-tiles x sp/4 k-steps of m8n8k4 (512 flop) per 8 sites and category
    inner_children = 2 * c["ii"] + c["ti"]
    dmma_flops = inner_children * (sites / 8.0) * 4 * nt * (sp // 4) * 512
    clv_bytes = 8.0 * sites * 4 * sp
    return {"workload": "%d states, Gamma-4, %d taxa x %d sites (row N1: codon-sized alphabet)" % (states, tips, sites),
            "loglh": ll, "traversal_ms": tk * 1e3, "full_loglh_ms": t * 1e3, "site_updates_per_s": (tips - 2) * sites / tk,
            "dense_equiv_TFLOPs_fp64": (tips - 2) * sites * 4 * (4.0 * states * states + states) / tk / 1e12,
            "dmma_frac": dmma_flops / tk / DMMA_PEAK, "dmma_peak_TFLOPs": DMMA_PEAK / 1e12,
            "clv_GBps": (inner_children + tips - 2) * clv_bytes / tk / 1e9,
            "blo_one_smoothing_s": tb, "blo_branches": 2 * tips - 3, "blo_neg_loglh": nll}


def c4(sites=2500, tips=200, nparts=100):
    pkg = load_product_package()
    states = [4 if p % 2 == 0 else 20 for p in range(nparts)]
    alphas = [0.3 + 0.01 * p for p in range(nparts)]
    with Harness("b200", tips=tips, sites=[sites] * nparts, states=states, alpha=alphas) as h:
        ll = h.loglh(0)
        pkg.reset_stats()
        t = timed(lambda: h.loglh(0), 5)
        launches = pkg.stats()["kernel_launches"] / 6
        h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)   # allocates the sumtables
        pkg.reset_stats()
        tb0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        tb = time.perf_counter() - tb0
        blo_launches = pkg.stats()["kernel_launches"] / (2 * tips - 3)
    return {"workload": "%d partitions (DNA/protein alternating, per-partition alpha) x %d sites each, %d taxa: one "
                        "GPU's 1/8 site share of BASELINE configs[3] (20k sites per partition)" % (nparts, sites, tips),
            "loglh": ll, "full_loglh_ms": t * 1e3, "site_updates_per_s": (tips - 2) * sites * nparts / t,
            "kernel_launches_per_eval": launches, "blo_one_smoothing_s": tb,
            "blo_launches_per_branch": blo_launches, "blo_neg_loglh": nll}


def c5(sites=500000, tips=2000, spr_tips=500, spr_sites=100000, spr_radius=5, spr=True):
    pkg = load_product_package()
    out = {"workload": "DNA GTR+G4 %d taxa x %d sites (BASELINE configs[4])" % (tips, sites)}
    with Harness("b200", tips=tips, sites=sites) as h:
        out["loglh"] = h.loglh(0)
        out["full_loglh_ms"] = timed(lambda: h.loglh(0), 2) * 1e3
        rng = np.random.RandomState(5)
        nodes = rng.randint(tips, 4 * tips - 6, size=100)   # inner subnodes
        c0 = h.ops_counter()
        t0 = time.perf_counter()
        for nd in nodes:
            h.perturb_edge(int(nd), 0.05 + 0.1 * rng.rand())
            h.set_root(int(nd))     # as algo_search.c does before scoring a regraft
            h.loglh(1)
        tpart = (time.perf_counter() - t0) / len(nodes)
        out["partial_traversal_evals_per_s"] = 1.0 / tpart
        out["clv_ops_per_partial_eval"] = (h.ops_counter() - c0) / len(nodes)
        h.loglh(0)
        tb0 = time.perf_counter()
        out["blo_neg_loglh"] = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        out["blo_one_smoothing_s"] = time.perf_counter() - tb0
        out["blo_branches"] = 2 * tips - 3
    if spr:
        with Harness("b200", tips=spr_tips, sites=spr_sites) as h:
            h.loglh(0)
            pkg.reset_stats()
            c0 = h.ops_counter()
            t0 = time.perf_counter()
            lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=spr_radius, ntopol_keep=20, smoothings=32,
                                 epsilon=1.0, bl_min=1e-6)
            dt = time.perf_counter() - t0
            st = pkg.stats()
            out["spr_round"] = {"shape": "%d taxa x %d sites, fast mode, radius 1..%d, ntopol_keep 20"
                                         % (spr_tips, spr_sites, spr_radius),
                                "seconds": dt, "loglh": lh, "clv_ops": h.ops_counter() - c0,
                                "kernel_launches": st["kernel_launches"]}
    return out


if __name__ == "__main__":
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c3", "c4", "c5"]
    for name in which:
        if name == "c5full":
            print(json.dumps({"c5full": c5(spr_tips=2000, spr_sites=500000)}), flush=True)
        else:
            print(json.dumps({name: {"c3": c3, "c4": c4, "c5": c5, "n1": n1}[name]()}), flush=True)
