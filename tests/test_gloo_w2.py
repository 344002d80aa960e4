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


# ---- the peer-memory schedule: rank r owns the contiguous shard [r*n/W, (r+1)*n/W) -----------------------------
def _peer_worker(rank, world, store, result_path):
    """What nd_outer_sync_p2p does, with gloo standing in for the mapped peer memory and the oracle for the kernel:
    the owner gathers every worker's copy of ITS shard, sums (master - local_w) in worker order, updates its
    master/momentum shard and hands the new parameters back to every worker.  On BASELINE config 1's fixture
    (tiny Llama, produced by the reference itself) the result must be the reference's, bit for bit."""
    from oracle import oracle as O
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=rank, world_size=world)
    z = np.load(os.path.join(GOLDEN, "config1_tiny_llama_w2.npz"))
    n = z["init"].size
    assert n % (4 * world) == 0
    lo, hi = rank * n // world, (rank + 1) * n // world
    master_shard, mom_shard = z["init"][lo:hi].copy(), np.zeros(hi - lo, np.float32)
    ok = True
    for r in range(3):
        local = torch.from_numpy(z[f"local_r{r}_w{rank}"].copy())
        copies = []
        for q in range(world):                                   # "reads": owner q gets everybody's copy of shard q
            qlo, qhi = q * n // world, (q + 1) * n // world
            got = [torch.empty(qhi - qlo) for _ in range(world)] if rank == q else None
            dist.gather(local[qlo:qhi].contiguous(), got, dst=q)
            if rank == q:
                copies = [g.numpy() for g in got]
        O.outer_step(master_shard, copies, mom_shard, lr=0.7, momentum=0.9, first_step=(r == 0))
        parts = [torch.empty(n // world) for _ in range(world)]   # "writes": every worker's copy gets every shard
        dist.all_gather(parts, torch.from_numpy(master_shard))
        new_local = torch.cat(parts).numpy()
        ok &= np.array_equal(new_local, z[f"params_r{r}"])
        ok &= np.array_equal(mom_shard, z[f"mom_r{r}"][lo:hi])
    open(f"{result_path}.{rank}", "w").write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


def test_peer_shard_schedule_matches_the_reference_on_config1():
    with tempfile.TemporaryDirectory() as td:
        res = os.path.join(td, "res")
        mp.spawn(_peer_worker, args=(2, os.path.join(td, "store"), res), nprocs=2, join=True)
        assert [open(f"{res}.{r}").read() for r in range(2)] == ["ok", "ok"]


# ---- the inner step of a worker that spans two GPUs (nd_inner_step_replicas) ----------------------------------
def _replica_worker(rank, world, store, result_path):
    """The schedule of nd_inner_step_replicas with gloo for the peer memory and the oracle for the kernels: replica j
    averages and steps the j-th half, the clip coefficient comes from the two partial sums of squares, the halves
    are exchanged.  Must equal all_reduce(AVG) of the gradient followed by clip + AdamW over the whole space."""
    from oracle import oracle as O
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=rank, world_size=world)
    n = 40_000
    rng = np.random.default_rng(5)
    p0 = (rng.standard_normal(n) * 0.02).astype(np.float32)
    ok = True
    p, m, v = p0.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)          # the replica schedule
    rp, rm, rv = p0.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)       # the plain way
    part = n // world
    lo, hi = rank * part, (rank + 1) * part
    for step in range(1, 4):
        g = (np.random.default_rng(100 * step + rank).standard_normal(n) * (0.05 if step % 2 else 1e-4)).astype(np.float32)
        # plain: all_reduce(AVG), one norm, one AdamW over everything
        ga = torch.from_numpy(g.copy())
        dist.all_reduce(ga, op=dist.ReduceOp.SUM)
        ga = (ga.numpy() / np.float32(world)).astype(np.float32)
        O.adamw_step(rp, ga, rm, rv, lr=1e-3, step=step, clip=O.clip_coef(float(np.sqrt(np.float32(O.sumsq(ga))))),
                     recipe="cuda")
        # replicas: K_a on the owned part
        mine = [torch.empty(part) for _ in range(world)]
        for q in range(world):
            got = [torch.empty(part) for _ in range(world)] if rank == q else None
            dist.gather(torch.from_numpy(g[q * part:(q + 1) * part].copy()), got, dst=q)
            if rank == q:
                mine = got
        acc = mine[0].numpy().copy()
        for t in mine[1:]:
            acc = (acc + t.numpy()).astype(np.float32)
        avg = (acc / np.float32(world)).astype(np.float32)
        partial = torch.tensor([np.float32(O.sumsq(avg))])
        parts = [torch.empty(1) for _ in range(world)]
        dist.all_gather(parts, partial)                                       # the partial norms in the flag pages
        total = np.float32(0)
        for k, t in enumerate(parts):
            total = np.float32(t.item()) if k == 0 else np.float32(total + np.float32(t.item()))
        # K_b on the owned part, then the new parameters go to every replica
        ps, ms, vs = p[lo:hi].copy(), m[lo:hi].copy(), v[lo:hi].copy()
        O.adamw_step(ps, avg, ms, vs, lr=1e-3, step=step, clip=O.clip_coef(float(np.sqrt(total))), recipe="cuda")
        m[lo:hi], v[lo:hi] = ms, vs
        halves = [torch.empty(part) for _ in range(world)]
        dist.all_gather(halves, torch.from_numpy(ps))
        p = torch.cat(halves).numpy()
        ok &= np.allclose(p, rp, rtol=2e-6, atol=1e-9)          # the norm is summed in two parts: 1 ulp on the coefficient
        ok &= np.allclose(m[lo:hi], rm[lo:hi], rtol=2e-6, atol=1e-12) and np.allclose(v[lo:hi], rv[lo:hi], rtol=4e-6, atol=1e-16)
    open(f"{result_path}.{rank}", "w").write("ok" if ok else "mismatch")
    dist.barrier()
    dist.destroy_process_group()


def test_replica_inner_step_schedule_equals_allreduce_then_adamw():
    with tempfile.TemporaryDirectory() as td:
        res = os.path.join(td, "res")
        mp.spawn(_replica_worker, args=(2, os.path.join(td, "store"), res), nprocs=2, join=True)
        assert [open(f"{res}.{r}").read() for r in range(2)] == ["ok", "ok"]
