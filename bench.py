#!/usr/bin/env python
"""bench.py -- DiLoCo outer-sync benchmark (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one DiLoCo outer synchronisation (pseudo-gradient -> cross-worker average ->
Nesterov outer SGD -> copy-back; reference nanodiloco/diloco/diloco.py:34-54) over the flat
fp32 parameter space of the 1B-class Llama of BASELINE.json configs[3] (201 tensors,
1,261,529,088 parameters), every GPU holding one DiLoCo worker (weak scaling).

Prints ONE JSON line (rank 0).  `value` = device milliseconds per outer sync with everything
resident in HBM (max over ranks); `e2e` = the same step with the last-sync copy in HOST memory,
where the reference keeps it (diloco.py:25-32,47,54), host<->device copies inside the timed
region; `roofline` = the dominant kernel against the measured HBM copy bandwidth;
`cpu_baseline` = the reference's CPU path (oracle/torch_reference.py port) on a bounded sample.

`--impl reference` times only that CPU path, with all host threads, on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

OUTER_LR, OUTER_MU = 0.7, 0.9          # reference main.py:51, main.py:101
METRIC, UNIT = "diloco_outer_sync_ms", "ms"
WORKLOADS = {"llama1b": "llama1b", "llama150m": "llama150m", "tiny": "tiny"}


def workload_shapes(name: str):
    from nanodiloco_b200.workloads import workload_shapes as _ws   # shape tables only
    return _ws(name)


def numel(shape) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


# ---------------------------------------------------------------------------------------------
# clocks: sampled with NVML from a thread while the timed region runs
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int, period_s: float = 0.005):
        self.samples, self.reason_bits, self.sm_max = [], 0, None
        self._stop, self._active = threading.Event(), threading.Event()
        self._thread, self._h, self._nv = None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._period = period_s
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        except Exception as e:  # noqa: BLE001
            self.error = str(e)

    def _run(self):
        nv = self._nv
        while not self._stop.is_set():
            if self._active.is_set():
                try:
                    self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                    self.reason_bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:  # noqa: BLE001
                    pass
            time.sleep(self._period)

    def start(self):
        self._active.set()

    def stop(self):
        self._active.clear()

    def close(self):
        self._stop.set()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0,
                    "note": getattr(self, "error", "timed region shorter than one sample")}
        reasons = [n for b, n in self.REASONS.items() if self.reason_bits & b and n != "gpu_idle"]
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.sm_max, "reasons": reasons,
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------
# the reference's CPU path (port), timed on the host cores
# ---------------------------------------------------------------------------------------------
def _cpu_worker(rank, world, store, shapes, threads, steps, warmup, out_path):
    import torch
    import torch.distributed as dist
    from oracle.torch_reference import ReferenceDilocoPort
    torch.set_num_threads(max(1, threads))
    if world > 1:
        dist.init_process_group("gloo", init_method=f"file://{store}", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(1337)
    params = [torch.nn.Parameter(torch.randn(s, generator=g) * 0.02) for s in shapes]
    inner = torch.optim.AdamW(params, lr=4e-4)
    outer = torch.optim.SGD(params, lr=OUTER_LR, momentum=OUTER_MU, nesterov=True)
    d = ReferenceDilocoPort(params, inner, outer, 100, 10_000)
    gd = torch.Generator().manual_seed(100 + rank)
    with torch.no_grad():
        for p in params:
            p.add_(torch.randn(p.shape, generator=gd) * 1e-3)
    times = []
    for i in range(warmup + steps):
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        d.outer_step()
        if world > 1:
            dist.barrier()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump(times, f)
    if world > 1:
        dist.destroy_process_group()


def time_reference_cpu(shapes, world: int, steps: int, warmup: int, threads_total: int):
    """Seconds per outer step of the reference's CPU path on `shapes`, W=world gloo workers."""
    import tempfile
    import torch.multiprocessing as mp
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "times.json")
        if world == 1:
            _cpu_worker(0, 1, None, shapes, threads_total, steps, warmup, out)
        else:
            mp.spawn(_cpu_worker, args=(world, os.path.join(td, "store"), shapes, max(1, threads_total // world),
                                        steps, warmup, out), nprocs=world, join=True)
        return json.load(open(out))


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


def cpu_sample_shapes(all_shapes, budget_elems: int):
    """A prefix of the model's tensor list (embedding + whole layers) within budget_elems."""
    out, tot = [], 0
    for s in all_shapes:
        if out and tot + numel(s) > budget_elems:
            break
        out.append(s)
        tot += numel(s)
    return out, tot


def mem_available_bytes() -> int:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except Exception:  # noqa: BLE001
        pass
    return 0


# host bytes one CPU worker of the reference's path holds per parameter: parameters, the last-sync
# snapshot, the pseudo-gradient, the momentum buffer and the clone of diloco.py:54 in flight
CPU_BYTES_PER_PARAM = 5 * 4 + 2


def run_cpu_baseline(all_shapes, n_total: int, world: int, steps: int, warmup: int, max_elems=None):
    """The reference's CPU path (the port in oracle/torch_reference.py) on the WHOLE tensor list when the
    host has the memory for `world` workers of it, else on the largest prefix that fits -- and then the
    value reported is the prefix's own time, never a scaled one (`extrapolated` says so)."""
    cores = host_cores()
    avail = mem_available_bytes()
    fit_elems = int(0.8 * avail / (CPU_BYTES_PER_PARAM * max(1, world))) if avail else n_total
    budget = min(n_total, fit_elems, max_elems or n_total)
    shapes, sample_n = (all_shapes, n_total) if budget >= n_total else cpu_sample_shapes(all_shapes, budget)
    times = time_reference_cpu(shapes, world, steps, warmup, cores)
    sec = statistics.median(times)
    whole = sample_n == n_total
    what = (f"oracle/torch_reference.py (statement-by-statement torch-CPU port of the reference's Diloco.outer_step, "
            f"{'gloo W=%d' % world if world > 1 else 'W=1'}) on "
            + (f"the whole model ({len(shapes)} tensors, {sample_n} params per worker)" if whole else
               f"the first {len(shapes)} tensors = {sample_n} of {n_total} params per worker (host memory: "
               f"{avail / 2**30:.0f} GiB available for {world} workers x {CPU_BYTES_PER_PARAM} B/param); the value is "
               f"this sample's own time, NOT scaled to the full model")
            + f"; median of {steps} steps after {warmup} warm-up; torch threads = {cores} in total")
    return {
        "value": sec * 1e3, "unit": UNIT, "cores": cores, "kind": "port", "sample": what,
        "sample_elems": sample_n, "full_elems": n_total, "extrapolated": False, "whole_workload": whole,
        "best_ms": min(times) * 1e3, "steps_timed": len(times),
    }


def _extras_n1(torch, _lib, dev, n, g, local, shapes, peak):
    """N=1 side measurements: the inner step (ours vs torch foreach) and the reference's outer step on this GPU."""
    extras = {}
    # (3) inner step the way the wrapper runs it by default: gradients are separate tensors (as autograd
    # allocates them), p / exp_avg / exp_avg_sq are views of flat buffers; nd_sumsq_mt + nd_adamw_mt over a
    # device table, no zeroing store (`.grad = None` afterwards, diloco.py:60): 32 B/elem algorithmic
    from nanodiloco_b200.flat import FlatLayout
    metas = [torch.empty(sh, device="meta") for sh in shapes]
    lay = FlatLayout.plan(metas, quantum=4, align=4)
    m1, v1 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    grads = [torch.empty(sh, device=dev).normal_(0, 1e-3, generator=g) for sh in shapes]
    rows = list(zip(lay.views(local[:lay.total]), grads, lay.views(m1[:lay.total]), lay.views(v1[:lay.total])))
    tab, T, chunks = _lib.make_table(rows, dev)
    gtab, gT, gchunks = _lib.make_table([(gr,) for gr in grads], dev)
    ss, sc = torch.zeros(1, device=dev), _lib.scratch(dev)
    L, st = _lib.load(), _lib._stream()
    Ki = 10

    def inner(step_no):
        _lib.check(L.nd_sumsq_mt(gtab.data_ptr(), gT, gchunks, _lib.ND_F32, ss.data_ptr(), sc.data_ptr(), st))
        _lib.check(L.nd_adamw_mt(tab.data_ptr(), T, chunks, _lib.ND_F32, 4e-4, 0.9, 0.999, 1e-8, 1e-2, step_no,
                                 ss.data_ptr(), 1.0, 0, st))

    def timed(fn):
        for i in range(3):
            fn(i + 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for i in range(Ki):
            fn(i + 4)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / Ki

    i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    inner_ms = timed(inner)
    extras["inner_step"] = {"ms": inner_ms, "algorithmic_gbs": 32.0 * n / (inner_ms * 1e-3) / 1e9,
                            "frac_of_hbm_peak": 32.0 * n / (inner_ms * 1e-3) / 1e9 / peak,
                            "what": "nd_sumsq_mt + nd_adamw_mt over the model's %d gradient tensors with the clip "
                                    "coefficient read on the device (diloco.py:57-60), no zeroing store: 32 B/elem "
                                    "algorithmic" % len(shapes)}
    del grads, rows, tab, gtab
    # the flat-gradient-buffer mode (grad_buffer='flat'): one range, AdamW zeroes the gradients as it goes (36 B/elem)
    grad = torch.empty(n, device=dev).normal_(0, 1e-3, generator=g)

    def inner_flat(step_no):
        _lib.sumsq(grad, ss, sc)
        _lib.adamw(local, grad, m1, v1, lr=4e-4, step=step_no, sumsq_dev=ss, max_norm=1.0, zero_grad=True)

    extras["inner_step"]["flat_zero_grad_ms"] = timed(inner_flat)
    del grad, m1, v1
    # the same inner step the way the reference runs it on a GPU: clip_grad_norm_ + AdamW.step()
    # (torch foreach kernels) over the model's 201 separate tensors
    ps = [torch.nn.Parameter(torch.empty(sh, device=dev).normal_(0, 0.02, generator=g)) for sh in shapes]
    opt = torch.optim.AdamW(ps, lr=4e-4)
    gs = [torch.empty_like(p_).normal_(0, 1e-3, generator=g) for p_ in ps]

    def ref_inner():
        for p_, g_ in zip(ps, gs):
            p_.grad = g_
        torch.nn.utils.clip_grad_norm_(ps, max_norm=1.0)
        opt.step()
        opt.zero_grad()

    for _ in range(3):
        ref_inner()
    torch.cuda.synchronize()
    i0.record()
    for _ in range(Ki):
        ref_inner()
    i1.record()
    torch.cuda.synchronize()
    extras["inner_step"]["torch_foreach_ms"] = i0.elapsed_time(i1) / Ki
    del opt, gs
    # the reference's outer step on this GPU (per-tensor pageable H2D, sub, foreach SGD, clone + D2H;
    # diloco.py:46-54) -- the path to beat on the same box
    from oracle.torch_reference import ReferenceDilocoPort
    inner_o = torch.optim.AdamW(ps, lr=4e-4)
    outer_o = torch.optim.SGD(ps, lr=OUTER_LR, momentum=OUTER_MU, nesterov=True)
    port = ReferenceDilocoPort(ps, inner_o, outer_o, 100, 10_000)
    port.outer_step()
    torch.cuda.synchronize()
    t_ref = []
    for _ in range(2):
        t_a = time.perf_counter()
        port.outer_step()
        torch.cuda.synchronize()
        t_ref.append((time.perf_counter() - t_a) * 1e3)
    extras["reference_gpu_outer_step_ms"] = {
        "value": min(t_ref), "what": "oracle/torch_reference.py port of the reference's outer_step run with CUDA "
        "parameters on this GPU (host-offloaded snapshot, per-tensor copies, torch foreach SGD), wall ms, best of 2"}
    del port, inner_o, outer_o, ps
    torch.cuda.empty_cache()

    return extras


def _reference_gpu_outer_ms(torch, dist, dev, shapes, world):
    """Wall ms of the reference's outer step (oracle/torch_reference.py port) with CUDA parameters and the
    NCCL process group, max over ranks, best of 2."""
    from oracle.torch_reference import ReferenceDilocoPort
    g = torch.Generator(device=dev).manual_seed(1)
    ps = [torch.nn.Parameter(torch.empty(sh, device=dev).normal_(0, 0.02, generator=g)) for sh in shapes]
    inner_o = torch.optim.AdamW(ps, lr=4e-4)
    outer_o = torch.optim.SGD(ps, lr=OUTER_LR, momentum=OUTER_MU, nesterov=True)
    port = ReferenceDilocoPort(ps, inner_o, outer_o, 100, 10_000)
    port.outer_step()
    best = None
    for _ in range(2):
        dist.barrier()
        torch.cuda.synchronize()
        t_a = time.perf_counter()
        port.outer_step()
        torch.cuda.synchronize()
        t = torch.tensor([(time.perf_counter() - t_a) * 1e3], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = float(t) if best is None else min(best, float(t))
    del port, inner_o, outer_o, ps
    torch.cuda.empty_cache()
    return {"value": best, "what": f"the reference's outer_step (port) on {world} GPUs: per-tensor pageable H2D, sub, "
            "NCCL all_reduce(AVG) per tensor, foreach SGD, clone + D2H; wall ms, max over ranks, best of 2"}


def _replica_inner_ms(torch, dist, _lib, dev, pctx, n, world, rank, replicas, peak, K=10):
    """One inner step of workers that span `replicas` GPUs (BASELINE configs[4]): nd_inner_step_replicas (gradient
    average + norm + AdamW as two peer-memory kernels per GPU) against the plain way on the same buffers (NCCL
    all_reduce(AVG) inside the worker, nd_sumsq, nd_adamw with fused zero-grad).  Device ms, max over ranks."""
    from nanodiloco_b200.sync import PeerContext
    gctx = PeerContext(world, rank, n, elem_bytes=4, device=dev, replicas=replicas, exchange="pull")
    grad = gctx.local()[:n]
    g = torch.Generator(device=dev).manual_seed(900 + rank)
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    ss, sc = torch.zeros(1, device=dev), _lib.scratch(dev)
    local = pctx.local()[:n]
    grp = None
    for k in range(world // replicas):                       # every rank creates every group
        ranks = list(range(k * replicas, (k + 1) * replicas))
        gk = dist.new_group(ranks)
        if rank in ranks:
            grp = gk

    def fused(step):
        gctx.inner_step_replicas(pctx, m, v, lr=4e-4, step=step, max_norm=1.0, sumsq_total_out=ss)

    def plain(step):
        dist.all_reduce(grad, op=dist.ReduceOp.AVG, group=grp)
        _lib.sumsq(grad, ss, sc)
        _lib.adamw(local, grad, m, v, lr=4e-4, step=step, sumsq_dev=ss, max_norm=1.0, zero_grad=True)

    out = {}
    for name, fn in (("fused_peer_ms", fused), ("allreduce_then_flat_kernels_ms", plain)):
        for i in range(3):
            grad.normal_(0, 1e-3, generator=g)
            fn(i + 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for i in range(K):
            grad.normal_(0, 1e-3, generator=g)               # what the backward pass leaves behind (not timed)
            dist.barrier()
            torch.cuda.synchronize()
            a.record()
            fn(i + 4)
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        t = torch.tensor([tot / K], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name] = float(t)
    R = replicas
    out["nvlink_bytes_per_dir"] = 2 * 4.0 * n * (R - 1) / R          # phase 1: gradient reads; phase 2: parameter writes
    out["hbm_bytes_local"] = (4 + 4 + 16 + 16) * n / R + 4.0 * n * (R - 1) / R + 4.0 * n * (R - 1) / R
    out["nvlink_gbs_per_dir"] = out["nvlink_bytes_per_dir"] / (out["fused_peer_ms"] * 1e-3) / 1e9
    out["what"] = ("per inner step and GPU: each replica averages and steps 1/R of the flat space; NVLink per direction = "
                   "(R-1)/R x 4 B/elem of gradient reads served/received plus (R-1)/R x 4 B/elem of parameter writes, in two "
                   "phases; both timings include their flag barriers / NCCL launch")
    del grad, m, v
    gctx.close()
    return out


def _tokens_per_s(torch, dist, dev, world, args):
    """tokens/s of the drop-in training loop (BASELINE configs[1] model size) on synthetic tokens."""
    tokens = None
    from nanodiloco_b200.main import train_model
    cfg150 = {"architectures": ["LlamaForCausalLM"], "hidden_size": 1024, "intermediate_size": 2688,
              "num_attention_heads": 16, "num_hidden_layers": 12, "rms_norm_eps": 1e-05, "use_cache": False}
    H, rounds = args.tokens_inner_steps, 3
    kw = dict(batch_size=8, per_device_batch_size=8, seq_length=1024, warmup_steps=10, total_steps=1000 * H,
              inner_steps=H, max_steps=H * rounds, llama_config=cfg150)
    st = train_model(**kw)
    tk = torch.tensor([st["tokens_per_s_per_worker"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tk, op=dist.ReduceOp.MIN)
    tokens = {"value": float(tk) * world, "unit": "tokens/s", "per_worker_min": float(tk),
              "outer_sync_ms": st["outer_sync_ms"], "avg_sync_ms": st.get("avg_sync_ms"),
              "avg_sync_wait_ms_rank0": st.get("avg_sync_wait_ms"), "params": st["params"],
              "what": f"nanodiloco_b200.main.train_model, 150M-class Llama (h1024 i2688 a16 L12, fp32 like the "
                      f"reference), per-device batch 8 x seq 1024, H={H} inner steps per outer sync ({rounds} rounds, "
                      f"first excluded), synthetic tokens; fwd/bwd is stock HF Llama (out of scope)"}
    if world == 1:
        from oracle.torch_reference import ReferenceDilocoModelPort
        st_ref = train_model(diloco_factory=ReferenceDilocoModelPort, **kw)
        tokens["reference_port_on_gpu"] = {"value": st_ref["tokens_per_s"], "outer_sync_ms": st_ref["outer_sync_ms"]}

    return tokens


# ---------------------------------------------------------------------------------------------
def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    shapes = workload_shapes(args.workload)
    n_total = sum(numel(s) for s in shapes)
    steps, warmup = args.steps, max(1, args.warmup)
    cb = run_cpu_baseline(shapes, n_total, args.gpus, steps, warmup,
                          max_elems=args.cpu_max_elems if args.cpu_max_elems > 0 else None)
    cfg = workload_config(args, n_total, len(shapes), args.gpus)
    if not cb["whole_workload"]:
        cfg["workload"] = (f"SAMPLE of {args.workload}: first {cb['sample_elems']} of {n_total} params per worker "
                           f"(capped by host memory or --cpu-max-elems); " + cfg["workload"])
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": cb["value"], "higher_is_better": False,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg, "cpu_baseline": cb, "partial_workload": not cb["whole_workload"],
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, n_total, n_tensors, world):
    rep = max(1, getattr(args, "replicas", 1))
    workers = world // rep
    copy = "bf16 working copy, fp32 master and momentum" if getattr(args, "dtype", "f32") == "bf16" else "fp32"
    return {
        "workload": (f"{args.workload} outer sync: {n_tensors} tensors, {n_total} params per worker ({copy}), "
                     f"outer SGD lr {OUTER_LR} momentum {OUTER_MU} nesterov, steady-state (momentum warmed), "
                     f"{workers} worker(s), {rep} GPU(s) per worker"),
        "l2": "each step streams several GB per GPU (>> 126 MB L2): inputs larger than L2, no flush needed",
        "parallelism": f"diloco-dp{workers}" + (f"x{rep}" if rep > 1 else ""),
    }


def host_link_ceiling_ms(torch, dist, dev, nbytes: int, world: int, reps: int = 3):
    """What the host link allows for the e2e step on this box: `nbytes` host->device and `nbytes`
    device->host from page-locked memory, both directions at once on two streams, all ranks at once;
    device ms per repetition, max over ranks.  No kernel, no dependency between the two copies -- the
    e2e step cannot beat this, whatever the kernels do."""
    elems = max(1, nbytes // 4)
    h_in = torch.empty(elems, dtype=torch.float32, pin_memory=True)
    h_out = torch.empty(elems, dtype=torch.float32, pin_memory=True)
    d_in, d_out = torch.empty(elems, device=dev), torch.empty(elems, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    best = None
    for _ in range(reps + 1):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_stream(torch.cuda.current_stream(dev))
        s2.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream(dev).wait_stream(s1)
        torch.cuda.current_stream(dev).wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = float(t) if best is None else min(best, float(t))       # first repetition warms the path
    return best


# ---------------------------------------------------------------------------------------------
# NVLink byte counters (NVML), sampled around the timed region: the wire evidence for the fused
# compute+exchange kernel, which ncu cannot wrap (one process per GPU)
# ---------------------------------------------------------------------------------------------
class NvlinkCounters:
    # nvml.h: NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX/RX (payload) and _RAW_TX/RX (payload + protocol), KiB,
    # scope UINT_MAX = sum over the device's links
    FIELDS = {"data_tx": 138, "data_rx": 139, "raw_tx": 140, "raw_rx": 141}

    def __init__(self, index: int):
        self.ok, self.err = False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.read()
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def read(self):
        nv = self.nv
        if getattr(self, "per_link", None) is None:
            # some drivers answer the all-links scope, some only per link: settle it on the first read
            vals = nv.nvmlDeviceGetFieldValues(self.h, [(fid, 0xFFFFFFFF) for fid in self.FIELDS.values()])
            self.per_link = any(v.nvmlReturn != 0 for v in vals)
        if not self.per_link:
            vals = nv.nvmlDeviceGetFieldValues(self.h, [(fid, 0xFFFFFFFF) for fid in self.FIELDS.values()])
            bad = [n for n, v in zip(self.FIELDS, vals) if v.nvmlReturn != 0]
            if bad:
                raise RuntimeError(f"NVML NVLink throughput fields {bad}: not available")
            return {n: int(v.value.ullVal) * 1024 for n, v in zip(self.FIELDS, vals)}          # KiB -> bytes
        out = {n: 0 for n in self.FIELDS}
        seen = 0
        for link in range(18):
            vals = nv.nvmlDeviceGetFieldValues(self.h, [(fid, link) for fid in self.FIELDS.values()])
            if any(v.nvmlReturn != 0 for v in vals):
                continue
            seen += 1
            for n, v in zip(self.FIELDS, vals):
                out[n] += int(v.value.ullVal) * 1024
        if not seen:
            raise RuntimeError("NVML NVLink throughput counters: return code 3 (NOT_SUPPORTED) for the all-links scope "
                               "and for every link -- this driver/VM does not expose them")
        return out


def parity_check(torch, dist, O, *, mode, world, rank, workers, replicas, dev, step_fn, local, master, mom,
                 idx, master_pos, mom_pos, bf16, elems):
    """ONE extra outer step from saved inputs, outside the timed region, checked against the CPU oracle:
    every rank compares `elems` elements of its own shard (new master, momentum and working copy) with
    oracle.outer_step run on the inputs gathered from all workers, and all ranks' whole working copies
    are compared with each other through a 64-bit checksum.  Bit-exact where the summation order is ours
    (one GPU, peer-memory sync); rtol 1e-6 + atol 1e-6*max|x| where NCCL or the switch fixes the order."""
    import numpy as np
    exact = mode in ("none", "p2p") or workers <= 2
    # --- inputs of the elements this rank checks (its own shard: it holds their master and momentum)
    m_in = master[master_pos].float().cpu().numpy().copy()
    b_in = mom[mom_pos].float().cpu().numpy().copy()
    mine = local[idx].float()
    if world > 1:
        # worker k's copy of rank q's elements travels to rank q (first replica of each worker sends)
        counts = torch.tensor([idx.numel()], device=dev)
        all_idx = [torch.empty_like(idx) for _ in range(world)]
        dist.all_gather(all_idx, idx)
        locs = None
        for q in range(world):
            piece = local[all_idx[q]].float()
            got = [torch.empty_like(piece) for _ in range(world)] if rank == q else None
            dist.gather(piece, got, dst=q)
            if rank == q:
                locs = [got[k * replicas].cpu().numpy() for k in range(workers)]
        del counts
    else:
        locs = [mine.cpu().numpy()]
    # --- the step under test
    step_fn()
    torch.cuda.synchronize()
    got_m = master[master_pos].float().cpu().numpy()
    got_b = mom[mom_pos].float().cpu().numpy()
    got_l = local[idx].float().cpu().numpy()
    # --- oracle
    om, ob = m_in.copy(), b_in.copy()
    O.outer_step(om, [np.ascontiguousarray(l) for l in locs], ob, lr=OUTER_LR, momentum=OUTER_MU, first_step=False)
    want_l = om
    if bf16:
        want_l = torch.from_numpy(om).to(torch.bfloat16).float().numpy()

    def ulps(a, b):
        ai = a.view(np.int32).astype(np.int64)
        bi = b.view(np.int32).astype(np.int64)
        ai = np.where(ai < 0, np.int64(-(2 ** 31)) - ai, ai)
        bi = np.where(bi < 0, np.int64(-(2 ** 31)) - bi, bi)
        return int(np.abs(ai - bi).max()) if a.size else 0

    max_ulp = max(ulps(got_m, om), ulps(got_b, ob))
    if exact:
        ok = np.array_equal(got_m, om) and np.array_equal(got_b, ob) and np.array_equal(got_l, want_l)
    else:
        ok = all(np.allclose(g, w, rtol=1e-6, atol=1e-6 * float(np.abs(w).max()))
                 for g, w in ((got_m, om), (got_b, ob)))
        ok = ok and np.array_equal(got_l, torch.from_numpy(got_m).to(local.dtype).float().numpy())
    # --- every rank's whole working copy is the same bits
    words = local.view(torch.int16 if bf16 else torch.int32)
    csum = words.sum(dtype=torch.int64).reshape(1)
    identical = True
    if world > 1:
        lo_, hi_ = csum.clone(), csum.clone()
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
        identical = bool(int(lo_) == int(hi_))
        flag = torch.tensor([1 if ok else 0, max_ulp], device=dev, dtype=torch.int64)
        dist.all_reduce(flag[0:1], op=dist.ReduceOp.MIN)
        dist.all_reduce(flag[1:2], op=dist.ReduceOp.MAX)
        ok, max_ulp = bool(int(flag[0])), int(flag[1])
    return {"mode": mode, "elems": int(elems) * world, "elems_per_rank": int(elems), "ranks_checked": world,
            "max_ulp": max_ulp, "ok": bool(ok and identical), "oracle_match": bool(ok),
            "cross_rank_identical": identical, "checksum": int(csum),
            "criterion": ("bit-exact (master, momentum, working copy) vs oracle/nd_oracle.c outer_step" if exact else
                          "rtol 1e-6 + atol 1e-6*max|x| vs oracle/nd_oracle.c outer_step (summation order fixed by "
                          "NCCL / the switch); working copy == cast(master) bit-exact"),
            "what": "one extra outer step after the timed region; inputs gathered from every worker"}


# ---------------------------------------------------------------------------------------------
def native_arm(args):
    import torch
    import torch.distributed as dist
    from nanodiloco_b200 import _lib
    from nanodiloco_b200.flat import FlatLayout
    from nanodiloco_b200.sync import SyncContext

    # a benchmark's ranks reach every sync together: a rank that waits two minutes for a peer has lost it, and an
    # error then beats spinning until the driver's own time limit (training jobs keep the library's 1800 s default)
    os.environ.setdefault("ND_PEER_TIMEOUT_S", "120")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()

    shapes = workload_shapes(args.workload)
    metas = [torch.empty(s, device="meta") for s in shapes]
    bf16 = args.dtype == "bf16"
    wdtype = torch.bfloat16 if bf16 else torch.float32
    gran = 8 if bf16 else 4
    layout = FlatLayout.plan(metas, quantum=gran * world, align=gran)
    n_real = sum(numel(s) for s in shapes)
    n = layout.total
    K, Wm = args.steps, max(3, args.warmup)

    from nanodiloco_b200.sync import PeerContext
    mode = "none" if world == 1 else args.sync
    if (bf16 or args.replicas > 1) and mode not in ("none", "p2p"):
        raise SystemExit("--dtype bf16 / --replicas need the peer-memory sync (--sync p2p)")
    workers = world // max(1, args.replicas)
    g = torch.Generator(device=dev).manual_seed(0)
    master = torch.empty(n, device=dev).normal_(0, 0.02, generator=g)      # same on every rank (seed 0)
    gl = torch.Generator(device=dev).manual_seed(100 + rank)
    hyper = dict(lr=OUTER_LR, momentum=OUTER_MU, dampening=0.0, weight_decay=0.0, nesterov=True)
    ctx = pctx = nctx = None
    pctx_exchange = None
    if mode == "p2p":
        # one fused kernel per GPU over NVLink peer memory; master and momentum sharded 1/W
        pctx = PeerContext(world, rank, n, elem_bytes=2 if bf16 else 4, group=None, device=dev,
                           replicas=args.replicas, exchange=args.p2p_exchange,
                           fold=False if args.p2p_no_fold else None)
        pctx_exchange = pctx.exchange
        local = pctx.local(wdtype)[:n]
        # replicas of one worker hold identical parameters: seed the drift per worker, not per rank
        gl = torch.Generator(device=dev).manual_seed(100 + rank // max(1, args.replicas))
        local.copy_(master + torch.empty(n, device=dev).normal_(0, 1e-3, generator=gl))
        lo, hi = pctx.shard_range()
        master = master[lo:hi].clone()
        mom = torch.zeros(hi - lo, device=dev)
        pg = None
        torch.cuda.empty_cache()

        def step(first=False):
            pctx.outer_sync(master, mom, dtype=wdtype, first_step=first, **hyper)
    elif mode in ("nvls", "hybrid"):
        # in-switch reduction (multimem) over torch symmetric memory; full master replica
        from nanodiloco_b200.sync import NvlsContext
        nctx = NvlsContext(world, rank, n, bucket_elems=int(args.bucket_mb * (1 << 20)) // 4, device=dev,
                           nvls_fraction=1.0 if mode == "nvls" else args.nvls_fraction)
        local = nctx.local()
        local.copy_(master)
        local.add_(torch.empty(n, device=dev).normal_(0, 1e-3, generator=gl))
        lo, hi = nctx.shard_range()
        mom = torch.zeros(hi - lo, device=dev)
        pg = None

        def step(first=False):
            nctx.outer_sync(master, mom, first_step=first, **hyper)
    else:
        ctx = SyncContext(world, rank, bucket_elems=int(args.bucket_mb * (1 << 20)) // 4, n_streams=args.streams,
                          device=dev, ag_comm=not args.nccl_single_comm)
        if world > 1 and not args.nccl_no_register:
            local, pg = ctx.alloc(n), ctx.alloc(n)       # NCCL-allocated, registered: collectives may run in-switch
        else:
            local = torch.empty(n, device=dev, dtype=wdtype)
            pg = torch.empty(n, device=dev) if world > 1 else None
        local.copy_((master + torch.empty(n, device=dev).normal_(0, 1e-3, generator=gl)).to(wdtype))
        mom = torch.zeros(ctx.shard_elems(n), device=dev)

        def step(first=False):
            if world == 1:
                _lib.outer_fused(master, mom, [local], first_step=first, **hyper)
            else:
                ctx.outer_sync(local, master, mom, pg, first_step=first, **hyper)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step(first=True)                                    # first-step branch, also warms the momentum
    local.add_(torch.empty(n, device=dev).normal_(0, 1e-3, generator=gl).to(wdtype))   # drift again
    for _ in range(Wm):
        step()
    sampler = ClockSampler(local_rank) if rank == 0 else None

    # ---- timed region: exactly K steps ---------------------------------------------------------
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    nvl = NvlinkCounters(local_rank) if world > 1 else None
    barrier()
    nvl0 = nvl.read() if nvl is not None and nvl.ok else None
    if sampler:
        sampler.start()
    t0.record()
    for _ in range(K):
        step()
    t1.record()
    barrier()
    if sampler:
        sampler.stop()
    nvl1 = nvl.read() if nvl0 is not None else None
    ms_total = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms = float(ms_total) / K

    # ---- the dominant kernel alone, CUDA events on its launching stream -------------------------
    # N=1: the step IS one launch of the fused kernel (24 B/elem).  N>1: kernel (1) pseudo-gradient
    # over the whole space (12 B/elem) is the largest HBM mover; time it on its own.
    lb = 2.0 if bf16 else 4.0                                  # bytes per element of the working copy
    if world == 1:
        kern_ms, kern_bytes, kern_name = ms, (16.0 + 2 * lb) * n, "outer_fused<W=1> (tma::tile_kernel)"
    elif mode in ("nvls", "hybrid"):
        kern_ms, kern_bytes = ms, (12.0 + 8.0 + 8.0 * (world - 1) / world + 16.0 / world) * n
        kern_name = f"K1 + nvls_shard_kernel (multimem.ld_reduce / multimem.st) + master refresh, W={world}"
    elif mode == "p2p":
        # the step IS one launch of the fused kernel over this rank's shard (+ two flag barriers);
        # HBM bytes this GPU moves for it: its copy is read once and written once by the W owners
        # (8 B/elem) and it reads+writes its master and momentum shards (16/W B/elem)
        kern_ms, kern_bytes = ms, (2 * lb + 16.0 / world) * n
        kern_name = f"outer_fused<W={workers}> over peer memory (tma::tile_kernel), NVLink-bound"
    else:
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        k0.record()
        for _ in range(K):
            _lib.pseudo_grad(master, local, pg)
        k1.record()
        torch.cuda.synchronize()
        kern_ms, kern_bytes, kern_name = k0.elapsed_time(k1) / K, 12.0 * n, "pseudo_grad (flat_kernel)"
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "of measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "of fallback (B200_PROFILING.md)"
    achieved = kern_bytes / (kern_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            if world == 1 and not bf16 and tj.get("kernel") == "outer_fused_w1" and tj.get("elems"):
                traffic = tj["dram_bytes_per_launch"] * (n / tj["elems"])
        except Exception:  # noqa: BLE001
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": kern_name, "peak_source": peak_src,
                "traffic_source": ("dram__bytes_read.sum + dram__bytes_write.sum of ONE `ncu --set full` capture of this "
                                   "kernel on the 150M-class space (profiles/traffic.json, 214,983,680 elements), scaled "
                                   "by element count to this launch -- not measured in this run") if traffic else None,
                "algorithmic_bytes_per_launch": kern_bytes, "kernel_ms": kern_ms}

    # ---- whole-step roofline (HBM and NVLink), for the record ---------------------------------
    hbm_bytes = ((16.0 + 2 * lb) if world == 1 else (2 * lb + 16.0 / world if mode == "p2p" else 12.0 + 24.0 / world)) * n
    # reduce-scatter + all-gather bytes of the working copy, per direction per GPU (the roofline)
    # Worst direction of the worst rank.  One rank per worker: every rank sends (W-1)/W of the copy as read
    # responses and (W-1)/W as writes, and receives the same.  Workers of R replicas (configs[4]): only each
    # worker's first replica is read, so ITS transmit side still carries (W-1) shards of read responses plus
    # (W-1) shards of writes -- the same 2(W-1)/W -- while its receive side carries (W/R-1) + (W-1) shards and
    # the other replicas less; the bound is unchanged, the mean load is lower.
    nvl_bytes = 0.0 if world == 1 else 2 * lb * (world - 1) / world * n
    nvl_bytes_rx_first_replica = 0.0 if world == 1 else lb * ((workers - 1) + (world - 1)) / world * n
    step_roof = {"hbm_algorithmic_gbs": hbm_bytes / (ms * 1e-3) / 1e9,
                 "nvlink_algorithmic_gbs_per_dir": nvl_bytes / (ms * 1e-3) / 1e9,
                 "t_hbm_ms": hbm_bytes / (peak * 1e9) * 1e3, "t_nvlink_ms_at_900": nvl_bytes / 900e9 * 1e3}
    bound_ms = max(step_roof["t_hbm_ms"], step_roof["t_nvlink_ms_at_900"])
    step_roof["frac_of_overlapped_bound"] = bound_ms / ms
    step_roof["t_nvlink_ms_at_770_measured_peer_copy"] = nvl_bytes / 770e9 * 1e3
    step_roof["frac_of_bound_at_770"] = max(step_roof["t_hbm_ms"], nvl_bytes / 770e9 * 1e3) / ms
    step_roof["sync"] = mode
    if args.replicas > 1:
        step_roof["nvlink_bytes_per_dir"] = {"tx_first_replica": nvl_bytes, "rx_first_replica": nvl_bytes_rx_first_replica,
                                             "tx_other_replicas": lb * (world - 1) / world * n,
                                             "rx_other_replicas": lb * (workers + world - 1) / world * n}
    if world > 1 and mode in ("p2p", "nvls", "hybrid"):
        # A fused compute+exchange step is bounded by the slower of its HBM time and its NVLink time; at
        # W >= 2 that is NVLink.  Denominator: the measured peer-copy bandwidth of this pool, 770 GB/s per
        # direction per GPU (B200_PROFILING.md; 900 nominal -- that fraction is in step_roofline).
        nvl_gbs = nvl_bytes / (ms * 1e-3) / 1e9
        roofline = {"bound": "nvlink", "achieved": nvl_gbs, "peak": 770.0, "unit": "GB/s", "frac": nvl_gbs / 770.0,
                    "traffic": None, "kernel": kern_name,
                    "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); 900 GB/s nominal",
                    "algorithmic_bytes_per_launch": nvl_bytes, "kernel_ms": ms,
                    "hbm": {"achieved": achieved, "peak": peak, "frac": achieved / peak, "peak_source": peak_src,
                            "algorithmic_bytes_per_launch": kern_bytes}}

    # ---- NCCL pipeline: per-bucket stage timeline and bus bandwidth of one extra (profiled) step -------
    nccl_timeline = None
    if mode == "nccl" and world > 1:
        ctx.set_profiling(True)
        step()
        torch.cuda.synchronize()
        rows = ctx.timeline()
        ctx.set_profiling(False)
        if rows:
            blen = min(ctx.bucket, n)
            wire = 4.0 * blen * (world - 1) / world               # bytes each rank sends per collective per bucket
            dur = lambda k: [r[k][1] - r[k][0] for r in rows]
            med = lambda xs: statistics.median(xs) if xs else None
            span = lambda k: (min(r[k][0] for r in rows), max(r[k][1] for r in rows))
            busy = lambda k: sum(dur(k))
            total_ms = max(r["refresh"][1] for r in rows)
            nccl_timeline = {
                "buckets": len(rows), "bucket_elems": blen, "total_ms": total_ms,
                "median_ms": {k: med(dur(k)) for k in ("k1", "rs", "k2", "ag", "refresh")},
                "sum_ms": {k: busy(k) for k in ("k1", "rs", "k2", "ag", "refresh")},
                "span_ms": {k: span(k) for k in ("k1", "rs", "k2", "ag", "refresh")},
                "rs_bus_gbs_median": wire / (med(dur("rs")) * 1e-3) / 1e9 if med(dur("rs")) else None,
                "ag_bus_gbs_median": wire / (med(dur("ag")) * 1e-3) / 1e9 if med(dur("ag")) else None,
                "note": "event times on the issuing streams; a stage's duration includes waiting for the stream's "
                        "previous op and, for NCCL kernels, for the communicator's previous collective",
                "first_buckets": rows[:6],
                "ag_comm": ctx.ag_comm, "registered_buffers": not args.nccl_no_register,
            }

    # ---- NVLink bytes per step from the NVML counters (per direction, worst rank) -----------------
    nvlink_measured = None
    if world > 1:
        if nvl1 is not None:
            d = torch.tensor([(nvl1[k] - nvl0[k]) / K for k in ("data_tx", "data_rx", "raw_tx", "raw_rx")],
                             device=dev, dtype=torch.float64)
            dmax = d.clone()
            dist.all_reduce(dmax, op=dist.ReduceOp.MAX)
            nvlink_measured = {"data_tx_bytes_per_step": float(dmax[0]), "data_rx_bytes_per_step": float(dmax[1]),
                               "raw_tx_bytes_per_step": float(dmax[2]), "raw_rx_bytes_per_step": float(dmax[3]),
                               "algorithmic_bytes_per_dir_per_step": nvl_bytes,
                               "source": "NVML NVLINK_THROUGHPUT_DATA/RAW_TX/RX counters (sum over links) read "
                                         "before and after the timed region, max over ranks"}
            per_dir = max(float(dmax[0]), float(dmax[1]))
            if per_dir > 0:
                nvlink_measured["traffic_over_algorithmic"] = per_dir / nvl_bytes if nvl_bytes else None
                if roofline.get("bound") == "nvlink":
                    roofline["traffic"] = per_dir
                    roofline["traffic_source"] = "NVML NVLink data counters, per direction, worst rank, this run"
        else:
            nvlink_measured = {"error": nvl.err if nvl is not None else "unavailable"}

    # ---- parity: one extra outer step from saved inputs against the CPU oracle (outside the timed region)
    parity = None
    if not args.no_parity:
        try:
            from oracle import oracle as O               # the checker; never on the timed path
            from nanodiloco_b200.sync import shard_index_map, hybrid_index_map
            local.add_(torch.empty(n, device=dev).normal_(0, 1e-3, generator=gl).to(wdtype))   # workers drift apart
            S = int(args.parity_elems)
            if mode == "none":
                S = min(S, n)
                a0 = (n - S) // 2 // 4 * 4
                idx = torch.arange(a0, a0 + S, device=dev)
                mpos = bpos = idx
            elif mode == "p2p":
                lo, hi = pctx.shard_range()
                S = min(S, hi - lo)
                a0 = (hi - lo - S) // 2 // 4 * 4
                bpos = mpos = torch.arange(a0, a0 + S, device=dev)
                idx = mpos + lo
            else:
                full = (shard_index_map(n, world, rank, ctx.bucket) if mode == "nccl"
                        else hybrid_index_map(n, nctx.n_nvls, world, rank, nctx.bucket))
                S = min(S, full.numel())
                h = S // 2
                bpos = torch.cat([torch.arange(0, h), torch.arange(full.numel() - (S - h), full.numel())]).to(dev)
                idx = full.to(dev)[bpos]
                mpos = idx                                   # these modes keep a full master replica
            parity = parity_check(torch, dist, O, mode=mode, world=world, rank=rank, workers=workers,
                                  replicas=max(1, args.replicas), dev=dev, step_fn=step, local=local, master=master,
                                  mom=mom, idx=idx, master_pos=mpos, mom_pos=bpos, bf16=bf16, elems=S)
            del idx, mpos, bpos
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001  (reported, never hidden: ok stays false)
            parity = {"mode": mode, "ok": False, "error": repr(e)}

    # ---- e2e: last-sync copy in pinned HOST memory, H2D + D2H inside the timed region ------------
    e2e = None
    if world == 1:
        Ke = max(2, min(K, args.e2e_steps))
        chunk = max(4, int(args.offload_chunk_mb * (1 << 20)) // 4 // 4 * 4)
        octx = SyncContext(1, 0, bucket_elems=chunk, n_streams=3, device=dev)
        master_host = torch.empty(n, dtype=torch.float32, pin_memory=True)
        master_host.copy_(master)
        staging = torch.empty(3 * chunk, device=dev)
        for _ in range(2):
            octx.outer_step_offloaded(master_host, staging, chunk, mom, local, **hyper)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        if sampler:
            sampler.start()
        e0.record()
        for _ in range(Ke):
            octx.outer_step_offloaded(master_host, staging, chunk, mom, local, **hyper)
        e1.record()
        barrier()
        if sampler:
            sampler.stop()
        e2e = {"value": e0.elapsed_time(e1) / Ke, "unit": UNIT, "h2d_bytes_per_step": 4 * n,
               "d2h_bytes_per_step": 4 * n, "steps": Ke,
               "what": "nd_outer_step_offloaded: master copy in pinned host memory (the reference's layout), "
                       "streamed H2D -> fused kernel -> D2H in %.0f MiB chunks over 3 streams" % args.offload_chunk_mb}
        octx.close()
        del master_host, staging
        try:
            ceil = host_link_ceiling_ms(torch, dist, dev, 4 * n, 1)
            e2e["host_link_ceiling_ms"] = ceil
            e2e["frac_of_host_link_ceiling"] = ceil / e2e["value"]
            e2e["host_link_gbs_per_dir"] = 4 * n / (ceil * 1e-3) / 1e9
            e2e["ceiling_what"] = ("bare pinned cudaMemcpyAsync of the same bytes, H2D and D2H at once on two streams, "
                                   "no kernel: the PCIe/host-memory bound of this box for this step")
        except Exception as e:  # noqa: BLE001
            e2e["host_link_ceiling_error"] = repr(e)
    elif mode == "p2p" and not args.no_e2e:
        Ke = max(2, min(K, args.e2e_steps))
        shard = master.numel()
        # a 1/W shard is only a few hundred MB: smaller chunks so that it still pipelines over the 3 slots
        chunk_mb = min(args.offload_chunk_mb, 16.0)
        chunk = max(4, min(int(chunk_mb * (1 << 20)) // 4, shard) // 8 * 8)
        master_host = torch.empty(shard, dtype=torch.float32, pin_memory=True)
        master_host.copy_(master)
        staging = torch.empty(3 * chunk, device=dev)
        for _ in range(2):
            pctx.outer_sync_offloaded(master_host, staging, chunk, mom, **hyper)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        if sampler:
            sampler.start()
        e0.record()
        for _ in range(Ke):
            pctx.outer_sync_offloaded(master_host, staging, chunk, mom, **hyper)
        e1.record()
        barrier()
        if sampler:
            sampler.stop()
        e_ms = torch.tensor([e0.elapsed_time(e1) / Ke], device=dev, dtype=torch.float64)
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
        e2e = {"value": float(e_ms), "unit": UNIT, "h2d_bytes_per_step": 4 * shard * world,
               "d2h_bytes_per_step": 4 * shard * world, "steps": Ke,
               "what": "nd_outer_sync_p2p_offloaded: every rank's 1/W master shard in pinned host memory (the "
                       "reference keeps the last-sync copy on the host), streamed H2D -> fused peer kernel -> D2H in "
                       "%.0f MiB chunks over 3 streams; bytes are the whole job's (all ranks)" % (chunk * 4 / 2**20)}
        del master_host, staging
        try:
            ceil = host_link_ceiling_ms(torch, dist, dev, 4 * shard, world)
            e2e["host_link_ceiling_ms"] = ceil
            e2e["frac_of_host_link_ceiling"] = ceil / e2e["value"]
            e2e["host_link_gbs_per_dir_per_rank"] = 4 * shard / (ceil * 1e-3) / 1e9
            e2e["ceiling_what"] = ("bare pinned cudaMemcpyAsync of each rank's shard bytes, H2D and D2H at once on two "
                                   "streams, all ranks at once, no kernel: the PCIe/host-memory bound of this box")
        except Exception as e:  # noqa: BLE001
            e2e["host_link_ceiling_error"] = repr(e)
    if sampler:
        sampler.close()

    # ---- extras (N=1): the inner step, and the reference's own GPU path on this box --------------
    extras = None
    if world == 1 and not args.no_extras:
        extras = {}
        try:
            extras = _extras_n1(torch, _lib, dev, n, g, local, shapes, peak)
        except Exception as e:  # noqa: BLE001  (a failed side measurement must not lose the bench line)
            extras = {"error": repr(e)}
            torch.cuda.empty_cache()
    # ---- N>1: the reference's own multi-GPU outer step on this box (per-tensor NCCL all_reduce(AVG) +
    #      host-offloaded snapshot, diloco.py:46-54), the path to beat
    if world > 1 and not args.no_extras:
        try:
            extras = {"reference_gpu_outer_step_ms": _reference_gpu_outer_ms(torch, dist, dev, shapes, world)}
        except Exception as e:  # noqa: BLE001
            extras = {"error": repr(e)}

    # ---- configs[4]: the inner step of a worker that spans several GPUs (gradient exchange every step) ----
    replica_inner = None
    if mode == "p2p" and args.replicas > 1 and not bf16 and not args.no_replica_inner:
        try:
            replica_inner = _replica_inner_ms(torch, dist, _lib, dev, pctx, n, world, rank, args.replicas, peak)
        except Exception as e:  # noqa: BLE001
            replica_inner = {"error": repr(e)}
            torch.cuda.empty_cache()

    # ---- tokens/s: the drop-in training loop on synthetic tokens (BASELINE configs[1] model) -----
    tokens = None
    if not args.no_tokens:
        del master, local, mom
        if pctx is not None:
            pctx.close()
            pctx = None
        torch.cuda.empty_cache()
        try:
            tokens = _tokens_per_s(torch, dist, dev, world, args)
        except Exception as e:  # noqa: BLE001
            tokens = {"error": repr(e)}
    # ---- CPU baseline (rank 0, N=1 only) --------------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu_baseline = run_cpu_baseline(shapes, n_real, 1, 5, 1,
                                            max_elems=args.cpu_max_elems if args.cpu_max_elems > 0 else None)
        except Exception as e:  # noqa: BLE001
            cpu_baseline = {"error": repr(e)}

    if mode in ("nvls", "hybrid"):
        launches_per_step = 4                                          # K1, barrier, NVLS kernel, barrier
    elif mode == "p2p":
        folded = pctx_exchange == "pull" and not args.p2p_no_fold and os.environ.get("ND_P2P_FOLD", "1") != "0"
        # pull: ONE launch (rendezvous folded into the fused kernel) or barrier, kernel, barrier; push adds the copy kernel
        launches_per_step = 1 if folded else (4 if pctx_exchange == "push" else 3)
    elif world > 1:
        launches_per_step = 2 * ((n + ctx.bucket - 1) // ctx.bucket)   # K1 + K2 per bucket
    else:
        launches_per_step = 1
    if rank == 0:
        line = {
            "metric": METRIC, "value": ms, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if not bf16 else "f32 master/momentum, bf16 working copy", "data": "synthetic",
            "config": workload_config(args, n_real, len(shapes), world),
            "roofline": roofline, "step_roofline": step_roof, "cpu_baseline": cpu_baseline, "e2e": e2e,
            "parity_check": parity, "nvlink_measured": nvlink_measured, "nccl_timeline": nccl_timeline,
            "replica_inner_step": replica_inner,
            "p2p_exchange": pctx_exchange, "launches_per_step": launches_per_step,
            "clocks": sampler.summary() if sampler else None,
            "gpu_launches": int(launches_per_step * K),
            "params_per_worker": n_real, "worker_params_per_s": world * n_real / (ms * 1e-3),
            "nvls_fraction": args.nvls_fraction if mode == "hybrid" else None,
            "bucket_mb": args.bucket_mb if mode in ("nccl", "nvls", "hybrid") else None, "sync": mode, "extras": extras, "tokens_per_s": tokens,
        }
        print(json.dumps(line), flush=True)
    if ctx is not None:
        ctx.close()
    if pctx is not None:
        pctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="llama1b")
    ap.add_argument("--bucket-mb", type=float, default=64.0)
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--dtype", choices=["f32", "bf16"], default="f32",
                    help="working-copy dtype; bf16 = 'bf16 inner / fp32 outer' of BASELINE configs[1] (extension)")
    ap.add_argument("--replicas", type=int, default=1,
                    help="GPUs per DiLoCo worker (BASELINE configs[4]: 4 workers x 2 GPUs = --gpus 8 --replicas 2)")
    ap.add_argument("--nvls-fraction", type=float, default=0.75, help="--sync hybrid: share of the space reduced in-switch")
    ap.add_argument("--sync", choices=["p2p", "nccl", "nvls", "hybrid"], default="p2p",
                    help="N>1: fused peer-memory kernel (default) or bucketed NCCL reduce-scatter/all-gather")
    ap.add_argument("--offload-chunk-mb", type=float, default=64.0)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--cpu-max-elems", type=int, default=0,
                    help="cap the CPU reference at a prefix of the tensor list with this many params (0 = whole model "
                         "when host memory allows); a capped run says so in its line and is never scaled up")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="N>1 tuning runs: skip the host-resident-master measurement")
    ap.add_argument("--no-tokens", action="store_true", help="skip the tokens/s training-loop measurement")
    ap.add_argument("--tokens-inner-steps", type=int, default=10)
    ap.add_argument("--no-extras", action="store_true", help="skip the inner-step and reference-GPU-path timings")
    ap.add_argument("--nccl-single-comm", action="store_true", help="--sync nccl: all-gathers share the reduce-scatters' communicator")
    ap.add_argument("--nccl-no-register", action="store_true", help="--sync nccl: plain torch memory instead of ncclMemAlloc + ncclCommRegister")
    ap.add_argument("--no-replica-inner", action="store_true", help="--replicas: skip the inner-step measurement")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check of one extra step")
    ap.add_argument("--parity-elems", type=int, default=1 << 24, help="elements per rank the parity check compares")
    ap.add_argument("--p2p-no-fold", action="store_true",
                    help="--sync p2p: barrier launches around the fused kernel instead of rendezvous folded into it")
    ap.add_argument("--p2p-exchange", choices=["auto", "pull", "push"], default="auto",
                    help="--sync p2p: how the worker copies reach the shard owner (nd_peer_set_mode)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return native_arm(args)


if __name__ == "__main__":
    sys.exit(main())
