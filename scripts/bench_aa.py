"""Protein (LG+G4) traversal timing: BASELINE configs[2] shape (500 taxa x 200k sites by default)."""
import os, sys, time, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import torch
from pllharness import Harness, load_product_package
tips = int(sys.argv[1]) if len(sys.argv) > 1 else 500
sites = int(sys.argv[2]) if len(sys.argv) > 2 else 200000
pkg = load_product_package()
if os.environ.get("PLLB200_AA_TIPTAB"):
    pkg.set_option("aa_tiptab", int(os.environ["PLLB200_AA_TIPTAB"]))
if os.environ.get("PLLB200_TRAV_PREFETCH"):
    pkg.set_option("trav_prefetch", int(os.environ["PLLB200_TRAV_PREFETCH"]))
if os.environ.get("PLLB200_AA_VARIANT"):
    pkg.set_option("aa_variant", int(os.environ["PLLB200_AA_VARIANT"]))
h = Harness("b200", tips=tips, sites=sites, states=20)
stream = torch.cuda.Stream()
pkg.load().pllb200_partition_set_stream(h.partition_ptr(0), stream.cuda_stream)
ll = h.loglh(0)
c = h.op_type_counts()
alg = (c["ii"] * 1932 + c["ti"] * 1289 + c["tt"] * 646) * sites
flops = (tips - 2) * sites * 6480.0
ts = []
for _ in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); h.queue_update_partials(); e1.record(stream); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[1]
print(json.dumps({"workload": "protein LG+G4 %d taxa x %d sites" % (tips, sites), "loglh": ll, "ops": c, "kernel_ms": ms,
                  "site_updates_per_s": (tips - 2) * sites / ms * 1e3, "algorithmic_GBps": alg / ms / 1e6,
                  "TFLOPs_fp64": flops / ms / 1e9,
                  # what the tensor pipe executes: 15 m8n8k4 DMMAs (512 flop) per 8 sites, category and inner
                  # child (n padded 20 -> 24); tip children are table lookups.  Peak: scripts/probes/dmma_peak.cu
                  "dmma_TFLOPs_issued": (c["ii"] * 2 + c["ti"]) * sites * 4 * 15 * 512 / 8 / ms / 1e9,
                  "dmma_frac_of_37.15_peak": (c["ii"] * 2 + c["ti"]) * sites * 4 * 15 * 512 / 8 / ms / 1e9 / 37.15}))
