"""CPU, world_size 2 over gloo: the host logic of the sharded outer sync.

The product's N>1 path is bucketed reduce-scatter -> update of the owned slice -> all-gather
(nanodiloco_b200/csrc/nd_sync.cu).  Its kernels need a GPU, but the partitioning arithmetic
(`shard_index_map`, padded sizes, `SyncContext.gather_shards`' reassembly) is host logic shared
with the GPU path.  Here two gloo ranks follow exactly that schedule with the ORACLE doing the
arithmetic, and must land on the reference's own results (tests/golden/outer_w2.npz)."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nanodiloco_b200.flat import FlatLayout
from nanodiloco_b200.sync import shard_index_map

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _worker(rank, world, store, bucket, result_path):
    from oracle import oracle as O
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=rank, world_size=world)
    z = np.load(os.path.join(GOLDEN, f"outer_w{world}.npz"))
    n_real = z["init"].size
    n = (n_real + 4 * world - 1) // (4 * world) * (4 * world)      # nd_sync_padded_elems
    pad = lambda a: np.concatenate([a, np.zeros(n - n_real, np.float32)])
    master = pad(z["init"])
    mine = shard_index_map(n, world, rank, bucket).numpy()
    mom_shard = np.zeros(mine.size, np.float32)
    ok = True
    for r in range(3):
        local = pad(z[f"local_r{r}_w{rank}"])
        pg = torch.from_numpy(O.pseudo_grad(master, local))             # K1, all ranks, full space
        dist.all_reduce(pg, op=dist.ReduceOp.SUM)                       # RS == all-reduce + own slices
        g = (pg.numpy()[mine] / np.float32(world)).astype(np.float32)   # K2 on the owned shard
        p_shard = master[mine].copy()
        O.sgd_step(p_shard, mom_shard, g, lr=0.7, momentum=0.9, first_step=(r == 0))
        parts = [torch.empty(mine.size) for _ in range(world)]          # AG
        dist.all_gather(parts, torch.from_numpy(p_shard))
        new = np.empty(n, np.float32)
        for q, part in enumerate(parts):
            new[shard_index_map(n, world, q, bucket).numpy()] = part.numpy()
        master = new                                                    # master refresh
        ok &= np.array_equal(master[:n_real], z[f"params_r{r}"])
        # momentum: gather shards the same way SyncContext.gather_shards does
        mparts = [torch.empty(mine.size) for _ in range(world)]
        dist.all_gather(mparts, torch.from_numpy(mom_shard))
        mom = np.empty(n, np.float32)
        for q, part in enumerate(mparts):
            mom[shard_index_map(n, world, q, bucket).numpy()] = part.numpy()
        ok &= np.array_equal(mom[:n_real], z[f"mom_r{r}"])
    if rank == 0:
        with open(result_path, "w") as f:
            f.write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("bucket", [8 * 64, 8 * 1000, 1 << 20])
def test_sharded_schedule_matches_reference_w2(bucket):
    with tempfile.TemporaryDirectory() as td:
        res = os.path.join(td, "res")
        mp.spawn(_worker, args=(2, os.path.join(td, "store"), bucket, res), nprocs=2, join=True)
        assert open(res).read() == "ok"


def test_layout_quantum_gives_even_slices():
    ts = [torch.zeros(5), torch.zeros(130), torch.zeros(3, 7)]
    for W in (1, 2, 4, 8):
        lay = FlatLayout.plan(ts, quantum=4 * W)
        assert lay.total % (4 * W) == 0
