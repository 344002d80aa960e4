"""GPU, world size >= 2 (skipped where fewer GPUs are visible): the multi-GPU outer sync through the drop-in
wrapper, every way the worker copies can meet, against (a) the reference's golden fixtures and (b) the
CPU oracle at sizes where every CTA runs many tiles and stages:

  p2p          nd_outer_sync_p2p, pull: one fused kernel per GPU reading and writing NVLink peer memory, the two
               rendezvous of the ranks folded into the kernel
  p2p_nofold   the same with the rendezvous as barrier launches of their own
  p2p_push     nd_outer_sync_p2p, push: workers write their slices into the owners' staging, owners reduce locally
  p2p_bf16     bf16 working copy, fp32 master and momentum (extension, BASELINE config 2)
  p2p_offload  the 1/W master shard in pinned host memory (nd_outer_sync_p2p_offloaded)
  p2p_rep2     workers of 2 GPUs each (extension, BASELINE config 5): world/2 workers
  nccl         nd_outer_sync: K1 -> reduce-scatter -> K2 on the owned slice -> all-gather, bucketed
  nvls/hybrid  in-switch reduction (multimem), alone or sharing the space with the peer kernel

All cases of one world size run inside ONE group of processes (a spawn costs ~30 s of N GPUs); the
parametrised tests below look their case up in that run's results.  Nothing is skipped for not having
been seen green: a case skips only for a stated hardware reason (too few GPUs, no multicast support).
"""
import json
import os
import tempfile
import traceback

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(1500)]
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

MODES = ["p2p", "p2p_nofold", "p2p_push", "p2p_bf16", "p2p_offload", "p2p_rep2", "nccl", "nvls", "hybrid"]
GOLDEN_CASES = [("golden", m, b) for m, b in [("nccl", 0.002), ("nccl", 64.0), ("p2p", 64.0), ("p2p_nofold", 64.0), ("p2p_push", 64.0),
                                              ("p2p_offload", 64.0), ("p2p_bf16", 64.0), ("p2p_rep2", 64.0),
                                              ("nvls", 64.0), ("nvls", 0.002), ("hybrid", 0.002)]]
# tiny Llama (57 tensors, 9,766,528 parameters: BASELINE configs[0]) and the 150M-class Llama (111 tensors,
# 214,983,680 parameters: configs[1]) -- at 8 GPUs that is 1.2 M and 26.9 M elements per shard, i.e. 1.2 K and
# 26 K tiles of the W=8 kernel
LARGE_CASES = [("large", m, wl) for wl in ("tiny", "llama150m") for m in MODES]
# BASELINE configs[0]: what the reference computed around the real tiny Llama with 2 workers (sampled elements,
# tests/golden/config1_tiny_llama_w2.npz); runs where the job has exactly two workers
CONFIG1_CASES = [("config1", m, 1.0) for m in ("p2p", "p2p_push", "p2p_offload", "p2p_rep2", "nccl", "nvls", "hybrid")]
# the inner step of workers that span 2 GPUs (nd_inner_step_replicas) against all_reduce(AVG) + the flat kernels
REPLICA_CASES = [("replica_inner", "p2p_rep2", wl) for wl in ("tiny", "llama150m")]
CASES = GOLDEN_CASES + CONFIG1_CASES + REPLICA_CASES + LARGE_CASES


def _case_id(c):
    return "-".join(str(x) for x in c)


class Skip(Exception):
    pass


_FAILED = []          # names of the checks that failed in the case that is running (this rank)


def chk(label, cond) -> bool:
    cond = bool(cond)
    if not cond:
        _FAILED.append(label)
    return cond


def _diloco_kwargs(mode, bucket_mb, offload_chunk_mb=0.001):
    sync = "p2p" if mode.startswith("p2p") else mode
    kw = dict(bucket_mb=bucket_mb, n_streams=2, sync=sync)
    if mode == "p2p_offload":
        kw.update(offload_master=True, offload_chunk_mb=offload_chunk_mb)
    if mode == "p2p_rep2":
        kw.update(replicas_per_worker=2)
    if mode == "p2p_push":
        kw.update(p2p_exchange="push")
    elif sync == "p2p":
        kw.update(p2p_exchange="pull")
    if mode == "p2p_nofold":
        kw.update(p2p_fold=False)
    return kw


# ---- (a) the reference's golden fixtures (5,397 elements) ---------------------------------------------------
def _golden_case(mode, bucket_mb, world, rank):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_gpu_diloco import Bag, _flat
    from nanodiloco_b200 import Diloco
    from oracle import oracle as O
    rep = 2 if mode == "p2p_rep2" else 1
    workers = world // rep
    z = np.load(os.path.join(GOLDEN, f"outer_w{workers}.npz"))
    model = Bag(z["init"]).cuda()
    bf16 = mode == "p2p_bf16"
    if bf16:                                        # extension: bf16 working copy, fp32 master/momentum
        model = model.bfloat16()
    if rank != 0:                                   # diloco.py:21-22: rank 0's parameters win
        with torch.no_grad():
            for p in model.parameters():
                p.add_(1.0)
    inner = torch.optim.AdamW(model.parameters(), lr=4e-4)
    outer = torch.optim.SGD(model.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(model, inner, outer, 2, 100, **_diloco_kwargs(mode, bucket_mb))
    params = list(model.parameters())
    widen = (lambda a: torch.from_numpy(a).to(torch.bfloat16).float().numpy()) if bf16 else (lambda a: a)
    f32 = lambda ts: _flat(ts).float().cpu().numpy()
    ok = np.array_equal(f32(params), widen(z["init"]))
    om, ob = widen(z["init"]).copy(), np.zeros_like(z["init"])
    exact_vs_reference = workers <= 2 and not bf16
    for r in range(3):
        loc, off = z[f"local_r{r}_w{rank // rep}"], 0         # replicas of a worker hold the same weights
        with torch.no_grad():
            for p in params:
                p.copy_(torch.from_numpy(loc[off:off + p.numel()]).reshape(p.shape))
                off += p.numel()
        d.outer_step()
        got = f32(params)
        master = f32(d.offloaded_last_sync_parameters)
        mom = f32([outer.state_dict()["state"][i]["momentum_buffer"] for i in range(len(params))])   # the hook gathers
        O.outer_step(om, [widen(z[f"local_r{r}_w{w}"]).copy() for w in range(workers)], ob, lr=0.7, momentum=0.9,
                     first_step=(r == 0))
        ok &= chk("np.array_equal(got, widen(master))", np.array_equal(got, widen(master)))               # working copy == cast(theta_sync)
        if mode.startswith("p2p"):     # fixed summation order: bit-exact against the oracle for ANY world size
            ok &= chk("np.array_equal(master, om) and np.array_equal(mom, ob)", np.array_equal(master, om) and np.array_equal(mom, ob))
        if exact_vs_reference:         # order-free average: bit-exact against the reference itself
            ok &= chk("np.array_equal(master, z[f'params_r{r}']) and np.array_equal(mom, z[f'", np.array_equal(master, z[f"params_r{r}"]) and np.array_equal(mom, z[f"mom_r{r}"]))
        elif not bf16:                 # north-star tolerance
            ok &= chk("np.allclose(master, z[f'params_r{r}'], rtol=1e-6, atol=1e-6 * np.abs(z", np.allclose(master, z[f"params_r{r}"], rtol=1e-6, atol=1e-6 * np.abs(z[f"params_r{r}"]).max()))
            ok &= chk("np.allclose(mom, z[f'mom_r{r}'], rtol=1e-6, atol=1e-6 * np.abs(z[f'mom", np.allclose(mom, z[f"mom_r{r}"], rtol=1e-6, atol=1e-6 * np.abs(z[f"mom_r{r}"]).max()))
    ok &= chk("d.avg_sync_time > 0", d.avg_sync_time > 0)
    if mode in ("p2p", "p2p_nofold", "p2p_push"):
        w = d.avg_sync_wait_time
        ok &= chk("avg_sync_wait_time is a non-negative number", w is not None and w >= 0.0)
    c = d.sync_counters()
    ok &= chk("sync_counters: ms and NVLink GB/s of the last sync", c["ms"] > 0 and c["nvlink_gbs_per_dir"] > 0)
    if mode in ("p2p", "p2p_nofold", "p2p_push"):
        ok &= chk("sync_counters: wait + work = total", abs(c["wait_ms"] + c["work_ms"] - c["total_ms"]) < 1e-3)
    sd = d.state_dict()                       # collective: full master and momentum on every rank
    saved = torch.cat([v.reshape(-1) for v in d.layout.views(sd["master"])]).numpy()
    ok &= chk("np.array_equal(widen(saved), f32(params))", np.array_equal(widen(saved), f32(params)))                 # just synced: theta_sync == params
    d.load_state_dict(sd)                     # scatter back into this rank's shard: nothing changes
    ok &= chk("np.array_equal(d.state_dict()['momentum'].numpy(), sd['momentum'].nump", np.array_equal(d.state_dict()["momentum"].numpy(), sd["momentum"].numpy()))
    if mode == "p2p_rep2":
        # inner step: replicas average their gradients and stay identical
        for p in params:
            p.grad.copy_(torch.full_like(p, 0.01 * (rank + 1)))
        d.inner_step(); d.inner_step()
        mine = _flat(params)
        both = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(both, mine)
        ok &= chk("torch.equal(both[rank], both[rank ^ 1])", torch.equal(both[rank], both[rank ^ 1]))
    d.close()
    return bool(ok)


def _config1_case(mode, world, rank):
    from nanodiloco_b200 import Diloco
    rep = 2 if mode == "p2p_rep2" else 1
    if world // rep != 2:
        raise Skip("config 1 has two workers")
    z = np.load(os.path.join(GOLDEN, "config1_tiny_llama_w2.npz"))
    p = torch.nn.Parameter(torch.from_numpy(z["init"].copy()).cuda() + (0.0 if rank == 0 else 1.0))
    holder = torch.nn.ParameterList([p])
    inner = torch.optim.AdamW(holder.parameters(), lr=4e-4)
    outer = torch.optim.SGD(holder.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(holder, inner, outer, 5, 30, **_diloco_kwargs(mode, 0.05, offload_chunk_mb=0.01))
    ok = np.array_equal(p.detach().cpu().numpy(), z["init"])
    for r in range(3):
        with torch.no_grad():
            p.copy_(torch.from_numpy(z[f"local_r{r}_w{rank // rep}"]))
        d.outer_step()
        mom = outer.state_dict()["state"][0]["momentum_buffer"]          # the hook gathers the shards
        ok &= chk("np.array_equal(p.detach().cpu().numpy(), z[f'params_r{r}'])", np.array_equal(p.detach().cpu().numpy(), z[f"params_r{r}"]))         # two workers: the mean is order-free,
        ok &= chk("np.array_equal(mom.cpu().numpy(), z[f'mom_r{r}'])", np.array_equal(mom.cpu().numpy(), z[f"mom_r{r}"]))                    # so every transport is bit-exact
    d.close()
    return bool(ok)


def _replica_inner_case(workload, world, rank):
    """Workers of 2 GPUs: three inner steps through the wrapper (gradient average + norm + AdamW as two peer-memory
    kernels per GPU, moments sharded over the replicas) against the same steps done the plain way on every rank:
    NCCL all_reduce(AVG) of the gradient, nd_sumsq, nd_adamw over the whole space."""
    from nanodiloco_b200 import Diloco, _lib
    from nanodiloco_b200.workloads import workload_shapes
    if world % 2:
        raise Skip("needs an even number of GPUs")
    dev = torch.device("cuda", rank)
    model = _Shapes(workload_shapes(workload), dev, torch.float32)
    params = list(model.parameters())
    inner = torch.optim.AdamW(params, lr=4e-4)
    outer = torch.optim.SGD(params, lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(model, inner, outer, 1, 100, sync="p2p", replicas_per_worker=2, p2p_exchange="pull")
    n = d.layout.total
    ok = d.grad_buffer == "flat" and d._gpeer is not None
    ref_p = d._local[:n].clone()
    ref_m, ref_v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    ss, sc = torch.zeros(1, device=dev), _lib.scratch(dev)
    grp = d._intra_group
    for step in range(1, 4):
        gen = torch.Generator(device=dev).manual_seed(500 + 10 * rank + step)      # every replica its own gradient
        g = torch.zeros(n, device=dev)
        for gv, p in zip(d.layout.views(g), params):
            gv.copy_(torch.randn(p.shape, generator=gen, device=dev) * (0.02 if step % 2 else 1e-4))
            p.grad.copy_(gv)
        lr = inner.param_groups[0]["lr"]
        dist.all_reduce(g, op=dist.ReduceOp.AVG, group=grp)
        _lib.sumsq(g, ss, sc)
        _lib.adamw(ref_p, g, ref_m, ref_v, lr=lr, step=step, sumsq_dev=ss, max_norm=1.0)
        d.inner_step()
        torch.cuda.synchronize()
        ok &= chk("float(d.last_grad_norm) == pytest.approx(float(ss.sqrt()), rel=2e-6)", float(d.last_grad_norm) == pytest.approx(float(ss.sqrt()), rel=2e-6))
        ok &= chk("bool(torch.allclose(d._local[:n], ref_p, rtol=2e-6, atol=1e-8))", bool(torch.allclose(d._local[:n], ref_p, rtol=2e-6, atol=1e-8)))
        ok &= chk("int(torch.count_nonzero(d._grad)) == 0", int(torch.count_nonzero(d._grad)) == 0)                              # zero_grad (diloco.py:60)
        mine = d._local[:n].clone()
        other = torch.empty_like(mine)
        pair = [torch.empty_like(mine) for _ in range(2)]
        dist.all_gather(pair, mine, group=grp)
        ok &= chk("torch.equal(pair[0], pair[1])", torch.equal(pair[0], pair[1]))                                       # replicas stay bit-identical
        del other, pair, mine
    sd = inner.state_dict()                                                        # the hook completes the sharded moments
    m_full = torch.cat([sd["state"][i]["exp_avg"].reshape(-1) for i in range(len(params))])
    v_full = torch.cat([sd["state"][i]["exp_avg_sq"].reshape(-1) for i in range(len(params))])
    want_m = torch.cat([t.reshape(-1) for t in d.layout.views(ref_m)])
    want_v = torch.cat([t.reshape(-1) for t in d.layout.views(ref_v)])
    ok &= chk("bool(torch.allclose(m_full, want_m, rtol=2e-6, atol=1e-10)) and bool(t", bool(torch.allclose(m_full, want_m, rtol=2e-6, atol=1e-10)) and bool(torch.allclose(v_full, want_v, rtol=4e-6, atol=1e-14)))
    d.close()
    del d, model, params
    torch.cuda.empty_cache()
    return bool(ok)


# ---- (b) model-sized parameter spaces against the oracle ----------------------------------------------------
class _Shapes(torch.nn.Module):
    def __init__(self, shapes, device, dtype):
        super().__init__()
        g = torch.Generator(device=device).manual_seed(1337)            # main.py:43: the same model on every rank
        self.ps = torch.nn.ParameterList(
            [torch.nn.Parameter((torch.randn(s, generator=g, device=device) * 0.02).to(dtype)) for s in shapes])


def _large_case(mode, workload, world, rank):
    from nanodiloco_b200 import Diloco
    from nanodiloco_b200.workloads import workload_shapes
    from oracle import oracle as O
    dev = torch.device("cuda", rank)
    rep = 2 if mode == "p2p_rep2" else 1
    workers = world // rep
    bf16 = mode == "p2p_bf16"
    wd = torch.bfloat16 if bf16 else torch.float32
    model = _Shapes(workload_shapes(workload), dev, wd)
    params = list(model.parameters())
    init = torch.cat([p.detach().reshape(-1) for p in params]).clone()
    if rank != 0:
        with torch.no_grad():
            for p in params:
                p.add_(0.5)
    inner = torch.optim.AdamW(params, lr=4e-4)
    outer = torch.optim.SGD(params, lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(model, inner, outer, 2, 100, **_diloco_kwargs(mode, 16.0 if workload == "llama150m" else 1.0, offload_chunk_mb=4.0))
    flat = lambda: d._local[: d.layout.total]
    ok = True
    # after the constructor's broadcast every rank holds rank 0's parameters (padding between tensors is zero)
    got0 = torch.cat([p.detach().reshape(-1) for p in params])
    ok &= chk("torch.equal(got0, init)", torch.equal(got0, init))
    idx = d._shard_indices()                                   # global indices of the elements this rank owns
    exact = mode.startswith("p2p") or workers <= 2
    for rnd in range(2):
        gen = torch.Generator(device=dev).manual_seed(100 + 10 * (rank // rep) + rnd)    # per WORKER drift
        with torch.no_grad():
            for p in params:
                p.add_((torch.randn(p.shape, generator=gen, device=dev) * 1e-3).to(wd))
        torch.cuda.synchronize()
        master = d._master if d._master.numel() != d.layout.total else d._master[idx.to(d._master.device)]
        m_in = master.float().cpu().numpy().copy()
        b_in = d._mom.float().cpu().numpy().copy()
        all_idx = [torch.empty_like(idx) for _ in range(world)]
        dist.all_gather(all_idx, idx)
        locs = None
        for q in range(world):                                 # every worker's copy of rank q's elements -> rank q
            piece = flat()[all_idx[q]].float()
            got = [torch.empty_like(piece) for _ in range(world)] if rank == q else None
            dist.gather(piece, got, dst=q)
            if rank == q:
                locs = [got[k * rep].cpu().numpy() for k in range(workers)]
        del all_idx
        d.outer_step()
        torch.cuda.synchronize()
        master = d._master if d._master.numel() != d.layout.total else d._master[idx.to(d._master.device)]
        got_m, got_b = master.float().cpu().numpy(), d._mom.float().cpu().numpy()
        got_l = flat()[idx].float().cpu().numpy()
        om, ob = m_in.copy(), b_in.copy()
        pg_sq = float(np.sum(O.average([O.pseudo_grad(m_in, l) for l in locs]).astype(np.float64) ** 2))
        O.outer_step(om, locs, ob, lr=0.7, momentum=0.9, first_step=(rnd == 0))      # rewinds `locs` in place
        want_l = torch.from_numpy(om).to(wd).float().numpy()
        if exact:
            ok &= chk("np.array_equal(got_m, om) and np.array_equal(got_b, ob) and np.array_e", np.array_equal(got_m, om) and np.array_equal(got_b, ob) and np.array_equal(got_l, want_l))
        else:
            ok &= chk("np.allclose(got_m, om, rtol=1e-6, atol=1e-6 * float(np.abs(om).max()))", np.allclose(got_m, om, rtol=1e-6, atol=1e-6 * float(np.abs(om).max())))
            ok &= chk("np.allclose(got_b, ob, rtol=1e-6, atol=1e-6 * float(np.abs(ob).max()))", np.allclose(got_b, ob, rtol=1e-6, atol=1e-6 * float(np.abs(ob).max())))
            ok &= chk("np.array_equal(got_l, torch.from_numpy(got_m).to(wd).float().numpy())", np.array_equal(got_l, torch.from_numpy(got_m).to(wd).float().numpy()))
        # every rank's whole working copy is the same bits
        words = flat().view(torch.int16 if bf16 else torch.int32).sum(dtype=torch.int64).reshape(1)
        lo, hi = words.clone(), words.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        ok &= chk("int(lo) == int(hi)", int(lo) == int(hi))
        if mode in ("p2p", "p2p_nofold", "p2p_push", "p2p_rep2"):            # the pseudo-gradient norm the kernel computes on the way
            sq = torch.tensor([pg_sq], device=dev, dtype=torch.float64)
            dist.all_reduce(sq)
            got_norm, want_norm = float(d.last_outer_grad_norm), float(sq.sqrt())
            ok &= chk(f"last_outer_grad_norm {got_norm!r} vs oracle {want_norm!r}", got_norm == pytest.approx(want_norm, rel=1e-5))
    d.close()
    del d, model, params
    torch.cuda.empty_cache()
    return bool(ok)


def _worker(rank, world, store, cases, result_path):
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"file://{store}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    results = {}
    for case in cases:
        kind, mode = case[0], case[1]
        del _FAILED[:]
        try:
            if mode == "p2p_rep2" and world % 2:
                raise Skip("needs an even number of GPUs")
            if mode in ("nvls", "hybrid"):
                import torch.distributed._symmetric_memory as symm   # noqa: F401
            if kind == "golden":
                ok = _golden_case(mode, case[2], world, rank)
            elif kind == "config1":
                ok = _config1_case(mode, world, rank)
            elif kind == "replica_inner":
                ok = _replica_inner_case(case[2], world, rank)
            else:
                ok = _large_case(mode, case[2], world, rank)
            verdict = "ok" if ok else f"mismatch on rank {rank}: " + "; ".join(_FAILED[:6])
        except Skip as e:
            verdict = f"skip: {e}"
        except Exception as e:  # noqa: BLE001
            msg = repr(e)
            verdict = ("skip: no NVLS multicast on this system" if "multicast" in msg.lower() and mode in ("nvls", "hybrid")
                       else "error: " + msg + " | " + traceback.format_exc(limit=3).replace("\n", " / "))
        # a case passes only if it passed on every rank
        code = torch.tensor([0 if verdict == "ok" else (1 if verdict.startswith("skip") else 2)], device="cuda")
        try:
            dist.all_reduce(code, op=dist.ReduceOp.MAX)
        except Exception:  # noqa: BLE001
            pass
        if int(code) == 2:       # collect every rank's account on rank 0
            box = [None] * world
            dist.all_gather_object(box, verdict)
            bad = [v for v in box if v != "ok"]
            verdict = " | ".join(bad) if bad else verdict
        results[_case_id(case)] = verdict
        if rank == 0:
            with open(result_path, "w") as f:
                json.dump(results, f)
        if verdict.startswith("error"):
            break                                   # a CUDA error poisons the context: stop, report what ran
    dist.barrier()
    dist.destroy_process_group()


_RESULTS = {}
# ND_SYNC_CASES=substring[,substring] runs only the matching cases (debugging on a charged box)
_ONLY = [x for x in os.environ.get("ND_SYNC_CASES", "").split(",") if x]


def _run_world(world):
    if world not in _RESULTS:
        with tempfile.TemporaryDirectory() as td:
            res = os.path.join(td, "res.json")
            try:
                cases = [c for c in CASES if not _ONLY or any(o in _case_id(c) for o in _ONLY)]
                mp.spawn(_worker, args=(world, os.path.join(td, "store"), cases, res), nprocs=world, join=True)
            except Exception as e:  # noqa: BLE001
                _RESULTS[world] = {"__spawn__": repr(e)}
            got = json.load(open(res)) if os.path.exists(res) else {}
            _RESULTS[world] = {**_RESULTS.get(world, {}), **got}
        keep = os.environ.get("ND_SYNC_TEST_LOG")
        if keep:
            with open(keep.replace("%W", str(world)), "w") as f:
                json.dump(_RESULTS[world], f, indent=1)
    return _RESULTS[world]


# ND_SYNC_WORLDS=8 restricts a run to some world sizes (a visit to an 8-GPU box is charged 8x)
WORLDS = [int(x) for x in os.environ.get("ND_SYNC_WORLDS", "2,4,8").split(",")]


@pytest.mark.parametrize("world", WORLDS)
@pytest.mark.parametrize("case", CASES, ids=_case_id)
def test_sharded_outer_sync(world, case):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = _run_world(world)
    verdict = res.get(_case_id(case))
    if verdict is None and _ONLY:
        pytest.skip("not selected by ND_SYNC_CASES")
    if verdict is None:
        pytest.fail(f"case did not run: {res.get('__spawn__', 'an earlier case ended the run')}")
    if verdict.startswith("skip"):
        pytest.skip(verdict)
    assert verdict == "ok", verdict
