#!/usr/bin/env python
"""Small driver for ncu: the inner step the way the wrapper runs it by default (nd_sumsq_mt + nd_adamw_mt
over a device table of separately allocated gradient tensors, 150M-class Llama shapes), then the
pseudo-gradient / shard-update kernels of the NCCL pipeline on the same flat space."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nanodiloco_b200 import _lib
from nanodiloco_b200.flat import FlatLayout
from nanodiloco_b200.workloads import workload_shapes

shapes = workload_shapes("llama150m")
dev = torch.device("cuda", 0)
lay = FlatLayout.plan([torch.empty(s, device="meta") for s in shapes])
n = lay.total
g = torch.Generator(device=dev).manual_seed(0)
p = torch.empty(n, device=dev).normal_(0, 0.02, generator=g)
grads = [torch.empty(s, device=dev).normal_(0, 1e-3, generator=g) for s in shapes]
m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
ss, sc = torch.zeros(1, device=dev), _lib.scratch(dev)
tab, T, chunks = _lib.make_table(list(zip(lay.views(p), grads, lay.views(m), lay.views(v))), dev)
gtab, gT, gchunks = _lib.make_table([(x,) for x in grads], dev)
master, mom, pg = p.clone(), torch.zeros(n, device=dev), torch.empty(n, device=dev)
L, st = _lib.load(), _lib._stream()
for step in range(1, 5):
    _lib.check(L.nd_sumsq_mt(gtab.data_ptr(), gT, gchunks, _lib.ND_F32, ss.data_ptr(), sc.data_ptr(), st))
    _lib.check(L.nd_adamw_mt(tab.data_ptr(), T, chunks, _lib.ND_F32, 4e-4, 0.9, 0.999, 1e-8, 1e-2, step, ss.data_ptr(),
                             1.0, 0, st))
    _lib.pseudo_grad(master, p, pg)
    _lib.outer_nesterov(master, mom, pg, p, world=1.0, lr=0.7, momentum=0.9, first_step=(step == 1))
torch.cuda.synchronize()
print("ok", float(ss))
