"""One SPR round at BASELINE configs[4] (2000 taxa x 500000 sites, fast mode, radius 1..5, ntopol_keep 20) with the
alignment site-sharded over the ranks of a torchrun launch (one process per GPU; per-candidate scores, logL and
Newton sums cross NVLink through the library's peer-memory exchange).
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
       scripts/bench_spr_c5_multi.py [taxa] [sites]"""
import json, os, sys, time, types
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import torch
import torch.distributed as dist
from pllharness import Harness, load_product_package  # noqa: E402
import bench  # noqa: E402  (attach_reduce)

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
tips = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
sites = int(sys.argv[2]) if len(sys.argv) > 2 else 500000
pkg = load_product_package()
torch.cuda.set_device(local)
pkg.load().pllb200_set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


t0 = time.perf_counter()
with Harness("b200", tips=tips, sites=sites, rank=rank, nranks=world) as h:
    bench.attach_reduce(h, pkg, types.SimpleNamespace(reduce="native"), rank, world)
    start = h.loglh(0)
    setup = time.perf_counter() - t0
    barrier()
    pkg.reset_stats()
    t0 = time.perf_counter()
    lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=5, ntopol_keep=20, smoothings=32, epsilon=1.0, bl_min=1e-6)
    barrier()
    secs = time.perf_counter() - t0
    st = pkg.stats()
    if rank == 0:
        print(json.dumps(dict(workload="SPR round, %d taxa x %d sites over %d GPUs (%d sites per GPU), fast mode, radius 1..5, ntopol_keep 20"
                                       % (tips, sites, world, sites // world),
                              n_gpus=world, setup_s=setup, start_loglh=start, loglh=lh, seconds=secs, clv_ops=h.ops_counter(),
                              kernel_launches_rank0=st["kernel_launches"])), flush=True)
    barrier()
if world > 1:
    dist.destroy_process_group()
