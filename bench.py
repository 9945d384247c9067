#!/usr/bin/env python
"""bench.py — CLV site-updates/s of the phylogenetic-likelihood hot path on B200 (BASELINE.json metric).

Workload (config.workload): BASELINE.json configs[1] — DNA GTR+G4, 1000 taxa x 1 M sites, full
post-order traversal (998 CLV updates per site) + P-matrix rebuild + edge log-likelihood, i.e. one
`pllmod_treeinfo_compute_loglh(treeinfo, 0)` per step; synthetic alignment/tree per SURVEY.md §8d.
With --gpus N (torchrun) every rank owns its own contiguous 1 M-site slice of an N M-site alignment
(weak scaling, the line's `value`); the only cross-GPU traffic is the all-reduce of the per-partition logL
scalar, done inside the library (pllb200_exchange_reduce_cb: one small kernel over NVLink peer memory).
The same run also times STRONG scaling (BASELINE's "6.5x at 8 GPUs": the 1 M-site alignment split N ways,
`strong_*` keys) and checks a sharded 64-taxon case against the same alignment on one GPU (`multi_gpu_check`).
At N = 1 the line carries `secondary`: configs[2..4] measured in the same process (scripts/secondary.py).

  value   = CLV site-updates/s (whole job), device-timed with CUDA events on the library's stream,
            inputs resident in HBM (tip codes, P-matrix inputs staged before the events)
  e2e     = the same metric through the public C-ABI call from host data: new branch lengths are set
            on the host tree, pllmod_treeinfo_compute_loglh uploads op list + branch lengths and the
            logL comes back to the host; wall-clock, max over ranks
  roofline= the traversal kernel alone, DRAM reading: the bytes a fused traversal must move (every parent CLV
            and scaler written once = 132 B per site and op, tip codes read once) / CUDA-event launch duration
            vs MEASURED_PEAKS.json hbm_gbs.  SURVEY.md 8(d)'s per-op model (396/265/134 B per site for ii/ti/tt:
            no reuse between ops) is kept as `algorithmic_equiv_GBps`; it is not a bound for a fused traversal
  cpu_baseline / --impl reference = the UNMODIFIED reference (oracle/_ref, AVX2, one site slice per
            host thread — the reference's own parallel_reduce_cb model) on a bounded site sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

BYTES = {"ii": 396, "ti": 265, "tt": 134}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tips", type=int, default=1000)
    ap.add_argument("--sites", type=int, default=1000000)
    ap.add_argument("--cpu-sites-per-core", type=int, default=4000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--blo", action="store_true", help="also time one NR smoothing pass (extra keys)")
    ap.add_argument("--no-secondary", action="store_true", help="skip configs[2..4] (N = 1 only)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling arm and the sharding check (N > 1)")
    ap.add_argument("--reduce", default="native", choices=["native", "nccl"],
                    help="cross-rank reduction: the library's peer-memory kernel, or torch.distributed NCCL in a Python callback")
    return ap.parse_args()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        # in-process NVML first (a query takes tens of microseconds, so even a 100 ms timed region is
        # sampled; eight ranks each spawning nvidia-smi got no sample at all), nvidia-smi as the fallback
        self.stop_flag = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
        bits = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                try:
                    r = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.rows.append([str(sm), str(mx), "0"] + ["Active" if r & b else "Not Active" for b in (0x8, 0x40, 0x20, 0x4)])
            except Exception:
                pass
            time.sleep(0.01)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        self.stop_flag = True
        if getattr(self, "nvml", None):
            self.thread.join(timeout=1.0)
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # under load = upper half of the samples (the sampler also sees idle gaps)
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": load[len(load) // 2] if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_run(tips, sites_per_core, reps, max_cores=None):
    """Times the unmodified reference (AVX2) on a bounded sample: one site slice per host thread."""
    from pllharness import Harness, have_reference
    if not have_reference():
        return None
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    if max_cores:
        cores = min(cores, max_cores)
    try:
        import psutil
        avail = psutil.virtual_memory().available
        per_core = (tips - 2) * sites_per_core * 128 * 1.2
        cores = max(1, min(cores, int(avail * 0.5 / per_core)))
    except Exception:
        pass
    total_sites = sites_per_core * cores
    hs = [None] * cores
    times = [0.0] * cores

    def setup(i):
        hs[i] = Harness("ref", tips=tips, sites=total_sites, rank=i, nranks=cores)

    def run(i):
        times[i] = hs[i].time_update_partials(reps)

    def fan(fn):
        th = [threading.Thread(target=fn, args=(i,)) for i in range(cores)]
        [t.start() for t in th]
        [t.join() for t in th]

    fan(setup)
    fan(run)            # warm-up (includes the first compute_loglh)
    t0 = time.perf_counter()
    fan(run)
    wall = (time.perf_counter() - t0) / reps
    for h in hs:
        h.close()
    updates = (tips - 2) * total_sites
    return {"value": updates / wall, "unit": "CLV site-updates/s", "cores": cores, "kind": "reference",
            "sample": "%d taxa x %d sites (%d sites/thread), full traversal via pll_update_partials, "
                      "libpll AVX2 kernels, best-effort %d reps" % (tips, total_sites, sites_per_core, reps),
            "seconds_per_traversal": wall}


def attach_reduce(h, pkg, a, rank, world):
    """Cross-rank reduction of the logL / derivative sums for a site-sharded harness."""
    if world == 1:
        return
    import numpy as np
    import torch
    import torch.distributed as dist

    def allgather_bytes(b):
        mine = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
        every = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        return [bytes(t.cpu().numpy()) for t in every]

    if a.reduce == "native":
        # the library's own collective (CUDA-IPC inboxes over NVLink); also serves the fused Newton kernel
        h.set_reduce_native(pkg, rank, world, allgather_bytes)
        return

    def reduce_cb(view, op):
        n = len(view)
        t = torch.from_numpy(np.array(view, copy=True)).to("cuda", non_blocking=False)
        dist.all_reduce(t, op={0: dist.ReduceOp.SUM, 1: dist.ReduceOp.MAX, 2: dist.ReduceOp.MIN}[op])
        view[:] = t.cpu().numpy()[:n]

    h.set_reduce_cb(reduce_cb)
    if os.environ.get("PLLB200_EXCHANGE", "1") != "0":
        part = h.partition_ptr(0)
        pkg.exchange_connect(part, rank, world, b"".join(allgather_bytes(pkg.exchange_create(part))))


def multi_gpu_check(pkg, a, rank, world, barrier):
    """A 64-taxon x (world * 40k)-site alignment sharded over the ranks against the same alignment on rank 0
    alone: logL, -logL after one Newton-Raphson smoothing pass (exercises the in-kernel exchange), and one SPR round."""
    import numpy as np
    from pllharness import Harness
    tips, S = 64, 40000 * world
    with Harness("b200", tips=tips, sites=S, rank=rank, nranks=world) as hs:
        attach_reduce(hs, pkg, a, rank, world)
        ll_s = hs.loglh(0)
        blo_s = hs.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        barrier()
    # one SPR round (batched regraft scoring: the per-candidate scores are summed over the ranks before the
    # reference's cutoff logic is replayed) on a smaller alignment
    stips, sS = 40, 6000 * world
    spr_kw = dict(thorough=False, radius_min=1, radius_max=4, ntopol_keep=5, smoothings=4, epsilon=0.1, bl_min=1e-6)
    with Harness("b200", tips=stips, sites=sS, rank=rank, nranks=world) as hs:
        attach_reduce(hs, pkg, a, rank, world)
        hs.loglh(0)
        spr_s, _ = hs.spr_round(**spr_kw)
        topo_s = hs.topology_signature()[0]
        barrier()
    out = None
    if rank == 0:
        with Harness("b200", tips=tips, sites=S) as h1:
            ll_1 = h1.loglh(0)
            blo_1 = h1.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        with Harness("b200", tips=stips, sites=sS) as h1:
            h1.loglh(0)
            spr_1, _ = h1.spr_round(**spr_kw)
            topo_1 = h1.topology_signature()[0]
        out = {"shape": "%d taxa x %d sites over %d ranks vs one GPU" % (tips, S, world),
               "logl_sharded": ll_s, "logl_one_gpu": ll_1, "logl_rel_err": abs(ll_s - ll_1) / abs(ll_1),
               "blo_sharded": blo_s, "blo_one_gpu": blo_1, "blo_rel_err": abs(blo_s - blo_1) / abs(blo_1),
               "tolerance": 1e-10, "reduce": a.reduce}
        out["spr_round"] = {"shape": "%d taxa x %d sites, radius 1..4, ntopol_keep 5" % (stips, sS), "loglh_sharded": spr_s,
                            "loglh_one_gpu": spr_1, "rel_err": abs(spr_s - spr_1) / abs(spr_1),
                            "same_topology": bool(np.array_equal(np.asarray(topo_s), np.asarray(topo_1)))}
        out["ok"] = bool(out["logl_rel_err"] <= 1e-10 and out["blo_rel_err"] <= 1e-10 and
                         out["spr_round"]["rel_err"] <= 1e-9 and out["spr_round"]["same_topology"])
    barrier()
    return out


def strong_scaling_run(pkg, a, rank, world, barrier, torch, dist):
    """BASELINE's strong-scaling target: the 1000 x 1 M alignment split `world` ways (S / world sites per GPU)."""
    from pllharness import Harness
    with Harness("b200", tips=a.tips, sites=a.sites, rank=rank, nranks=world) as h:
        stream = torch.cuda.Stream()
        pkg.load().pllb200_partition_set_stream(h.partition_ptr(0), stream.cuda_stream)
        attach_reduce(h, pkg, a, rank, world)
        ll = h.loglh(0)
        for _ in range(a.warmup):
            h.loglh(0)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(a.steps):
            ll = h.loglh(0)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1) / a.steps
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pkg.load().pllb200_partition_set_stream(h.partition_ptr(0), None)
    return float(t.cpu()[0]), ll


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if a.impl == "reference":
        if rank != 0:
            return 0
        res = cpu_reference_run(a.tips, a.cpu_sites_per_core, max(1, a.steps))
        if res is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built on this box"}))
            return 0
        line = {"impl": "reference", "metric": "CLV site-updates/sec", "value": res["value"],
                "unit": "CLV site-updates/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": res["seconds_per_traversal"] * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "DNA GTR+G4, %d taxa x %d sites per GPU, full traversal (BASELINE configs[1]); "
                                       "reference CPU arm on a bounded site sample, throughput is per site" % (a.tips, a.sites),
                           "taxa": a.tips, "l2": "inputs larger than L2"},
                "cpu_baseline": res,
                "e2e": {"value": res["value"], "unit": "CLV site-updates/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries exactly one JSON line
    import numpy as np
    import torch
    import torch.distributed as dist
    from pllharness import Harness, load_product_package

    pkg = load_product_package()
    if not torch.cuda.is_available() or pkg.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: libpll_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    pkg.load().pllb200_set_device(local_rank)
    for opt in ("trav_variant", "trav_reorder", "trav_ctas_per_sm", "trav_forward", "trav_stash", "trav_debug", "trav_prefetch", "trav_split", "trav_split_rounds", "fused_newton", "queued_newton"):      # tuning knobs (experiments)
        if os.environ.get("PLLB200_" + opt.upper()):
            pkg.set_option(opt, int(os.environ["PLLB200_" + opt.upper()]))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    tips, S = a.tips, a.sites

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    check = strong_ms = strong_ll = None
    if world > 1 and not a.no_strong:
        check = multi_gpu_check(pkg, a, rank, world, barrier)
        strong_ms, strong_ll = strong_scaling_run(pkg, a, rank, world, barrier, torch, dist)
    t_setup = time.perf_counter()
    # weak scaling: a global alignment of world*S sites, this rank owns slice `rank`
    h = Harness("b200", tips=tips, sites=S * world, rank=rank, nranks=world)
    stream = torch.cuda.Stream()
    pkg.load().pllb200_partition_set_stream(h.partition_ptr(0), stream.cuda_stream)

    attach_reduce(h, pkg, a, rank, world)

    ll0 = h.loglh(0)            # first full evaluation: uploads tips, builds everything
    setup_s = time.perf_counter() - t_setup
    counts = h.op_type_counts()
    alg_bytes = sum(BYTES[k] * counts[k] for k in counts) * S
    updates_per_step = (tips - 2) * S

    # ---- device-timed value: K steps of compute_loglh(0) --------------------------------------
    for _ in range(a.warmup):
        h.loglh(0)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    pkg.reset_stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(a.steps):
        ll = h.loglh(0)
    ev1.record(stream)
    barrier()
    dev_ms = ev0.elapsed_time(ev1) / a.steps
    st = pkg.stats()
    launches = st["kernel_launches"]

    # ---- traversal kernel alone (roofline) -------------------------------------------------------
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    for e0, e1 in kev:
        e0.record(stream)
        h.queue_update_partials()
        e1.record(stream)
    torch.cuda.synchronize()
    kern_ms = sorted(e0.elapsed_time(e1) for e0, e1 in kev)
    kern_avg = sum(kern_ms) / len(kern_ms)

    # ---- e2e: host branch lengths -> public API -> logL on host, wall clock ----------------------
    brl = h.branch_lengths()
    rng = np.random.RandomState(1234)
    pkg.reset_stats()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        h.set_branch_lengths(brl * rng.uniform(0.9, 1.1, size=brl.shape))
        ll_e2e = h.loglh(0)
    barrier()
    e2e_s = (time.perf_counter() - t0) / a.steps
    st2 = pkg.stats()
    clocks = sampler.stop()

    extra = {}
    # ---- logL and derivative evaluations/s at fixed S (second half of BASELINE's metric) -----------
    root = h.root()
    nrep = max(20, a.steps)
    h.edge_loglh(root)
    barrier()
    logl_s = h.time_edge_loglh(root, nrep)            # C loop: launch + kernel + host wait per call
    barrier()
    h.derivatives(root, [0.1])
    barrier()
    t0 = time.perf_counter()
    h.derivatives(root, [0.1])
    barrier()
    t1 = time.perf_counter() - t0
    deriv_s = max(1e-9, h.time_derivatives(root, nrep))
    barrier()
    # the kernels alone, by CUDA events the library records around them on its stream
    pkg.set_option("eval_timing", 1)
    lk_us, dv_us = [], []
    for i in range(nrep):
        h.edge_loglh(root)
        lk_us.append(pkg.last_eval_kernel_us())
    for i in range(nrep):
        h.derivatives(root, [0.01 + 0.02 * i])
        dv_us.append(pkg.last_eval_kernel_us())
    pkg.set_option("eval_timing", 0)
    lk_us, dv_us = sorted(lk_us), sorted(dv_us)
    root_is_ti = h.node_info(root)["tip"] or h.node_info(h.node_info(root)["back"])["tip"]
    extra["logl_evals_per_s"] = world / logl_s if world == 1 else 1.0 / logl_s
    extra["logl_GBps_algorithmic"] = (137 if root_is_ti else 268) * S / logl_s / 1e9
    extra["deriv_evals_per_s"] = 1.0 / deriv_s
    extra["deriv_GBps_algorithmic"] = 132 * S / deriv_s / 1e9
    extra["sumtable_builds_per_s"] = 1.0 / max(1e-9, t1 - deriv_s)
    per_site_logl = 137 if root_is_ti else 268
    extra["logl_kernel_us"] = lk_us[len(lk_us) // 2]
    extra["logl_kernel_GBps"] = per_site_logl * S / (lk_us[len(lk_us) // 2] * 1e-6) / 1e9
    extra["deriv_kernel_us"] = dv_us[len(dv_us) // 2]
    extra["deriv_kernel_GBps"] = 132 * S / (dv_us[len(dv_us) // 2] * 1e-6) / 1e9
    extra["eval_kernel_note"] = ("median of %d launches, CUDA events around the kernel on its stream; the derivative pass re-reads "
                                 "a %.0f MB table, part of which the previous pass left in the 126 MB L2" % (nrep, 128e-6 * S))
    extra["evals_note"] = ("host wall clock of a C loop over the C ABI incl. launch + host wait per call, %d sites per GPU, root edge is %s"
                           % (S, "tip-inner" if root_is_ti else "inner-inner"))
    if a.blo:
        h.set_branch_lengths(brl)
        h.loglh(0)
        barrier()
        t0 = time.perf_counter()
        nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
        barrier()
        extra["blo_one_smoothing_s"] = time.perf_counter() - t0
        extra["blo_neg_loglh"] = nll

    # max over ranks
    if world > 1:
        t = torch.tensor([dev_ms, e2e_s, kern_avg], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, kern_avg = [float(x) for x in t.cpu()]
    peak, peak_kind = measured_peak()
    # DRAM reading of the roofline: what a fused traversal has to move through HBM — every parent CLV (128 B)
    # and scaler count (4 B) written once, every tip code read once; children come from registers / shared memory
    fused_min_bytes = (tips - 2) * S * 132 + tips * S
    achieved = fused_min_bytes / (kern_avg * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of the kernel in one `ncu --set full` capture at 200k sites
    # (profiles/r02zz_trav_dna4_ncu_full_summary.csv): 0.745 GB read + 26.309 GB written; linear in sites
    traffic = 27.054e9 * S / 200000.0 * (tips - 2) / 998.0
    # the kernel's DRAM traffic is writes, and a write-only stream does not reach the copy figure of
    # MEASURED_PEAKS.json: time one here (4 GiB fill, best of 5)
    write_peak = None
    try:
        wbuf = torch.empty(1 << 30, dtype=torch.float32, device="cuda")
        wts = []
        for _ in range(6):
            w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            w0.record()
            wbuf.fill_(1.0)
            w1.record()
            torch.cuda.synchronize()
            wts.append(w0.elapsed_time(w1))
        write_peak = wbuf.numel() * 4 / (min(wts[1:]) * 1e-3) / 1e9
        del wbuf
    except Exception:
        pass

    if rank == 0:
        line = {
            "metric": "CLV site-updates/sec", "value": updates_per_step * world / (dev_ms * 1e-3),
            "unit": "CLV site-updates/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "DNA GTR+G4, %d taxa x %d sites per GPU, full traversal + P-matrices + edge logL "
                                   "(pllmod_treeinfo_compute_loglh, BASELINE configs[1])" % (tips, S),
                       "sites_per_gpu": S, "taxa": tips, "ops": counts, "parallelism": "site-sharded x%d" % world,
                       "l2": "inputs larger than L2 (%.1f GB of CLVs per step)" % ((tips - 2) * S * 128 / 1e9),
                       "setup_s": round(setup_s, 2)},
            "e2e": {"value": updates_per_step * world / e2e_s, "unit": "CLV site-updates/s",
                    "h2d_bytes_per_step": st2["h2d_bytes"] // a.steps, "d2h_bytes_per_step": st2["d2h_bytes"] // a.steps,
                    "ms_per_step": e2e_s * 1e3,
                    "what": "host tree with new branch lengths -> pllmod_treeinfo_compute_loglh -> logL on host"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic,
                         "traffic_note": "constant from one ncu --set full capture of this kernel at 200k sites "
                                         "(profiles/r02zz_trav_dna4_ncu_full_summary.csv: dram read 0.745 GB + write 26.309 GB), "
                                         "scaled linearly in sites and operations; not measured by this run "
                                         "(a 1M-site launch cannot be replayed by ncu)",
                         "achieved_is": "fused-minimum bytes / kernel time: (taxa-2) x sites x 132 B written + taxa x sites "
                                        "tip codes read; children of an operation come from registers / shared memory",
                         "fused_min_bytes_per_launch": fused_min_bytes,
                         "write_stream_GBps": write_peak,
                         "frac_of_write_stream": achieved / write_peak if write_peak else None,
                         "algorithmic_equiv_GBps": alg_bytes / (kern_avg * 1e-3) / 1e9,
                         "algorithmic_bytes_per_launch": alg_bytes,
                         "peak_kind": peak_kind, "kernel": "trav_dna4_kernel(LeanArgs)",
                         "kernel_ms": kern_avg, "kernel_share_of_step": kern_avg / dev_ms},
            "clocks": clocks, "loglh": ll, "loglh_first": ll0,
        }
        line.update(extra)
        if world > 1:
            line["reduce"] = ("library kernel over CUDA-IPC peer memory (pllb200_exchange_reduce_cb)" if a.reduce == "native"
                              else "torch.distributed NCCL all-reduce in a Python parallel_reduce_cb")
        if strong_ms is not None:
            # N = 1 time of the same alignment = this run's per-GPU step (1 M sites per GPU) less nothing:
            # the weak arm IS the one-GPU workload plus one 8-byte all-reduce
            line["strong_scaling"] = {"workload": "the %d x %d alignment split over %d GPUs (%d sites per GPU)"
                                                  % (tips, S, world, S // world),
                                      "strong_ms_per_step": strong_ms, "n1_ms_per_step": dev_ms,
                                      "n1_is": "this run's per-GPU step on %d sites (the weak arm)" % S,
                                      "strong_speedup_vs_n1": dev_ms / strong_ms,
                                      "value": updates_per_step / (strong_ms * 1e-3), "loglh": strong_ll}
            line["strong_ms_per_step"] = strong_ms
            line["strong_speedup_vs_n1"] = dev_ms / strong_ms
        if check is not None:
            line["multi_gpu_check"] = check
        if not a.no_cpu_baseline and world == 1:
            cb = cpu_reference_run(tips, a.cpu_sites_per_core, 2)
            if cb:
                line["cpu_baseline"] = cb
    h.close()
    rc = 0
    if rank == 0:
        if world == 1 and not a.no_secondary:
            # configs[2..4] under the same clock (each frees its HBM before the next starts)
            sys.path.insert(0, os.path.join(ROOT, "scripts"))
            import secondary
            sec = {}
            for name, fn in (("C3", secondary.c3), ("C4", secondary.c4), ("C5", secondary.c5), ("N1", secondary.n1)):
                t0 = time.perf_counter()
                try:
                    sec[name] = fn()
                except Exception as e:  # a secondary config must not cost the headline line
                    sec[name] = {"error": repr(e)}
                sec[name]["wall_s"] = round(time.perf_counter() - t0, 2)
            line["secondary"] = sec
        print(json.dumps(line))
        if check is not None and not check["ok"]:
            sys.stderr.write("multi_gpu_check failed: %r\n" % (check,))
            rc = 1
    if world > 1:
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
