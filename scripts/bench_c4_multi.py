"""C4 on N GPUs (BASELINE configs[3]: 100 partitions, DNA/protein alternating, site-sharded): logL and one
Newton smoothing pass.  torchrun --nproc-per-node N scripts/bench_c4_multi.py [sites_per_partition_per_gpu]
PLLB200_EXCHANGE=0 keeps the host callback (one NCCL all-reduce per Newton iteration) for comparison."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import numpy as np
import torch
import torch.distributed as dist
from pllharness import Harness, load_product_package

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
sites = int(sys.argv[1]) if len(sys.argv) > 1 else 2500
pkg = load_product_package()
torch.cuda.set_device(local_rank)
pkg.load().pllb200_set_device(local_rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
tips, nparts = 200, 100
states = [4 if p % 2 == 0 else 20 for p in range(nparts)]
alphas = [0.3 + 0.01 * p for p in range(nparts)]
h = Harness("b200", tips=tips, sites=[sites * world] * nparts, states=states, alpha=alphas, rank=rank, nranks=world)


def reduce_cb(view, op):
    n = len(view)
    t = torch.from_numpy(np.array(view, copy=True)).to("cuda", non_blocking=False)
    dist.all_reduce(t, op={0: dist.ReduceOp.SUM, 1: dist.ReduceOp.MAX, 2: dist.ReduceOp.MIN}[op])
    view[:] = t.cpu().numpy()[:n]


exchange = world > 1 and os.environ.get("PLLB200_EXCHANGE", "1") != "0"
if world > 1:
    h.set_reduce_cb(reduce_cb)
    if exchange:
        mine = torch.frombuffer(bytearray(pkg.exchange_create(h.partition_ptr(0))), dtype=torch.uint8).cuda()
        every = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        pkg.exchange_connect(h.partition_ptr(0), rank, world, b"".join(bytes(t.cpu().numpy()) for t in every))


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


ll = h.loglh(0)
h.loglh(0)
barrier()
t0 = time.perf_counter()
for _ in range(3):
    ll = h.loglh(0)
barrier()
t_ll = (time.perf_counter() - t0) / 3
h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)   # allocates the sumtables
pkg.reset_stats()
barrier()
t0 = time.perf_counter()
nll = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
barrier()
tb = time.perf_counter() - t0
st = pkg.stats()
tm = torch.tensor([t_ll, tb], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"config": "C4 %d partitions x %d sites per GPU, %d taxa, %d GPU(s)" % (nparts, sites, tips, world),
                      "newton_exchange": "device (NVLink peer memory)" if exchange else ("host callback + NCCL" if world > 1 else "none"),
                      "loglh": ll, "full_loglh_ms": float(tm[0]) * 1e3,
                      "site_updates_per_s": (tips - 2) * sites * nparts * world / float(tm[0]),
                      "blo_one_smoothing_s": float(tm[1]), "blo_launches_per_branch": st["kernel_launches"] / (2 * tips - 3),
                      "blo_neg_loglh": nll}), flush=True)
h.close() if hasattr(h, "close") else None
if world > 1:
    dist.destroy_process_group()
