#!/usr/bin/env python
"""Outer-sync microbenchmark (BASELINE.json configs[2]): flat fp32 buffers of 10M .. 2B elements at
1/2/4/8 GPUs against the HBM / NVLink roofline.

    python tools/microbench.py                               # one GPU
    torchrun --nproc-per-node N tools/microbench.py          # N workers, peer-memory fused kernel

Per size: `master ~ N(0, 0.02^2)` seed 0, `local_w = master + N(0, 1e-3^2)` seed 100+w, momentum warmed by
one round; 5 warm-up + 20 timed iterations, CUDA events, max over ranks; L2 (126 MB) is flushed between
iterations by writing a 512 MB buffer when the working set is smaller than 1 GB.  Prints a markdown table
and one JSON line per row.
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

from nanodiloco_b200 import _lib
from nanodiloco_b200.sync import PeerContext

SIZES = [10_000_000, 30_000_000, 100_000_000, 300_000_000, 1_000_000_000, 2_000_000_000]
LR, MU = 0.7, 0.9


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lrank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for n_req in SIZES:
        q = 4 * world
        n = (n_req + q - 1) // q * q
        g = torch.Generator(device=dev).manual_seed(0)
        full_master = torch.empty(n, device=dev).normal_(0, 0.02, generator=g)
        gl = torch.Generator(device=dev).manual_seed(100 + rank)
        pctx = PeerContext(world, rank, n, device=dev)
        local = pctx.local()[:n]
        local.copy_(full_master)
        local.add_(torch.empty(n, device=dev).normal_(0, 1e-3, generator=gl))
        lo, hi = pctx.shard_range()
        master = full_master[lo:hi].clone()
        del full_master
        mom = torch.zeros(hi - lo, device=dev)
        need_flush = 12 * n < (1 << 30)

        def run(first):
            pctx.outer_sync(master, mom, lr=LR, momentum=MU, first_step=first)

        def timed(first, iters=20, warm=5):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
            for _ in range(warm):
                run(first)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            for a, b in ev:
                if need_flush:
                    flush.fill_(1)
                a.record()
                run(first)
                b.record()
            torch.cuda.synchronize()
            ms = sorted(a.elapsed_time(b) for a, b in ev)
            t = torch.tensor([ms[len(ms) // 2], ms[0]], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t[0]), float(t[1])

        first_med, first_best = timed(True)
        run(True)                       # leave a warmed momentum behind
        med, best = timed(False)
        hbm_b = (24.0 if world == 1 else 8.0 + 16.0 / world) * n
        nvl_b = 0.0 if world == 1 else 8.0 * (world - 1) / world * n
        t_hbm, t_nvl = hbm_b / (hbm_peak * 1e9) * 1e3, nvl_b / 900e9 * 1e3
        row = {"n": n_req, "world": world, "ms_median": med, "ms_best": best, "first_step_ms_median": first_med,
               "hbm_gbs": hbm_b / (med * 1e-3) / 1e9, "nvlink_gbs_per_dir": nvl_b / (med * 1e-3) / 1e9,
               "bound_ms": max(t_hbm, t_nvl), "frac_of_bound": max(t_hbm, t_nvl) / med, "l2_flushed": need_flush}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        del local, master, mom
        pctx.close()
        torch.cuda.empty_cache()
    if rank == 0:
        print("\n| elements | W | ms (median) | ms (best) | first-step ms | HBM GB/s | NVLink GB/s/dir | bound ms | frac |")
        print("|---|---|---|---|---|---|---|---|---|")
        for r in rows:
            print(f"| {r['n']:,} | {r['world']} | {r['ms_median']:.3f} | {r['ms_best']:.3f} | "
                  f"{r['first_step_ms_median']:.3f} | {r['hbm_gbs']:.0f} | {r['nvlink_gbs_per_dir']:.0f} | "
                  f"{r['bound_ms']:.3f} | {r['frac_of_bound']:.2f} |")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
