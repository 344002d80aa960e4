#!/usr/bin/env python
"""Small driver for ncu: the inner-step kernels (sum of squares, fused clip+AdamW) and the
pseudo-gradient / shard-update kernels of the NCCL pipeline on a 150M-class flat space."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nanodiloco_b200 import _lib

n = 214_983_680
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(0)
p = torch.empty(n, device=dev).normal_(0, 0.02, generator=g)
grad = torch.empty(n, device=dev).normal_(0, 1e-3, generator=g)
m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
ss, sc = torch.zeros(1, device=dev), _lib.scratch(dev)
master, mom, pg = p.clone(), torch.zeros(n, device=dev), torch.empty(n, device=dev)
for step in range(1, 5):
    _lib.sumsq(grad, ss, sc)
    _lib.adamw(p, grad, m, v, lr=4e-4, step=step, sumsq_dev=ss, max_norm=1.0, zero_grad=False)
    _lib.pseudo_grad(master, p, pg)
    _lib.outer_nesterov(master, mom, pg, p, world=1.0, lr=0.7, momentum=0.9, first_step=(step == 1))
torch.cuda.synchronize()
print("ok", float(ss))
