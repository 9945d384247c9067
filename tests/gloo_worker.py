"""Worker for test_multi_rank_site_sharding_with_gloo (launched by torch.distributed.run, world_size 2)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from pllharness import Harness  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
kw = dict(tips=20, sites=[600, 350], states=[4, 20], alpha=[0.5, 1.2])
OPS = {0: dist.ReduceOp.SUM, 1: dist.ReduceOp.MAX, 2: dist.ReduceOp.MIN}


def reduce_cb(view, op):
    t = torch.from_numpy(np.array(view, copy=True))
    dist.all_reduce(t, op=OPS[op])
    view[:] = t.numpy()


with Harness("ref", rank=rank, nranks=world, **kw) as h:
    h.set_reduce_cb(reduce_cb)
    sharded = h.loglh()
    blo = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
single = None
if rank == 0:
    with Harness("ref", **kw) as h:
        single = h.loglh()
        blo1 = h.optimize_branch_lengths(method=0, radius=-1, max_iters=1, lh_epsilon=1.0)
    assert abs(blo - blo1) <= 1e-9 * abs(blo1), (blo, blo1)
    print("RESULT %.12f %.12f" % (sharded, single))
dist.barrier()
dist.destroy_process_group()
