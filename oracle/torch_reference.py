"""The reference's outer/inner step restated with torch CPU ops, statement by statement.

TEST INFRASTRUCTURE ONLY (see nd_oracle.c).  Used (a) by tests to pin the C oracle and the CUDA
path against what the reference computes, and (b) by bench.py as the timed CPU baseline
(`cpu_baseline`, `--impl reference`): the reference is pure Python over torch and cannot travel to
the GPU box (/root/reference does not exist there), so this port stands in for it.  In the build
container tests/test_torch_reference.py runs it side by side with the REAL class loaded from
/root/reference/nanodiloco/diloco/diloco.py and requires bit-identical results.

Each statement cites the reference line it restates (nanodiloco/diloco/diloco.py).
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch
import torch.distributed as dist


class ReferenceDilocoPort:
    def __init__(self, params: List[torch.nn.Parameter], inner_optimizer, outer_optimizer,
                 warmup_steps: int, total_steps: int, group=None):
        self.params = list(params)
        self.inner_optimizer = inner_optimizer
        self.outer_optimizer = outer_optimizer
        self.group = group
        self.distributed = dist.is_available() and dist.is_initialized()

        def lr_lambda(step):                                              # diloco.py:20 (transformers' cosine)
            if step < warmup_steps:
                return float(step) / float(max(1, warmup_steps))
            prog = float(step - warmup_steps) / float(max(1, total_steps - warmup_steps))
            return max(0.0, 0.5 * (1.0 + math.cos(math.pi * 0.5 * 2.0 * prog)))

        self.scheduler = torch.optim.lr_scheduler.LambdaLR(inner_optimizer, lr_lambda, -1)
        if self.distributed:
            for p in self.params:                                         # diloco.py:21-22
                dist.broadcast(p.data, src=0, group=group)
        self.offloaded_last_sync_parameters = self._get_offloaded_parameters()   # diloco.py:25

    def _get_offloaded_parameters(self):                                  # diloco.py:27-32
        return [p.data.detach().clone().to("cpu")
                for g in self.outer_optimizer.param_groups for p in g["params"]]

    def outer_step(self) -> None:                                         # diloco.py:34-54
        replica_params = [p for g in self.inner_optimizer.param_groups for p in g["params"]]   # :40-44
        for replica_param, last_sync_param in zip(replica_params, self.offloaded_last_sync_parameters):
            last_sync_param_on_device = last_sync_param.to(replica_param.device)                # :47
            replica_param.grad = last_sync_param_on_device - replica_param.data                 # :48
            if self.distributed:
                dist.all_reduce(tensor=replica_param.grad, op=dist.ReduceOp.AVG, group=self.group)   # :49
            replica_param.data = last_sync_param_on_device                                      # :50
        self.outer_optimizer.step()                                                             # :52
        self.outer_optimizer.zero_grad()                                                        # :53
        self.offloaded_last_sync_parameters = self._get_offloaded_parameters()                  # :54

    def inner_step(self) -> None:                                         # diloco.py:56-60
        torch.nn.utils.clip_grad_norm_(self.params, max_norm=1.0)
        self.inner_optimizer.step()
        self.scheduler.step()
        self.inner_optimizer.zero_grad()


def llama_shapes(hidden: int, inter: int, layers: int, vocab: int = 32000) -> List[tuple]:
    """Parameter shapes of HF LlamaForCausalLM(LlamaConfig(hidden_size, intermediate_size,
    num_hidden_layers, ...)) in `model.parameters()` order, untied embeddings (SURVEY.md section 8
    table: tiny 57 tensors / 9,766,528; 1B-class 201 tensors / 1,261,529,088)."""
    shapes = [(vocab, hidden)]
    for _ in range(layers):
        shapes += [(hidden, hidden)] * 4                        # q, k, v, o
        shapes += [(inter, hidden), (inter, hidden), (hidden, inter)]   # gate, up, down
        shapes += [(hidden,), (hidden,)]                        # input / post-attention norms
    shapes += [(hidden,), (vocab, hidden)]                      # final norm, lm_head
    return shapes


CONFIGS = {
    "tiny": dict(hidden=128, inter=512, layers=6),              # configs/llama_default.json
    "llama150m": dict(hidden=1024, inter=2688, layers=12),
    "llama1b": dict(hidden=2048, inter=5632, layers=22),
}


class ReferenceDilocoModelPort(ReferenceDilocoPort):
    """The port with the reference class's full surface (diloco.py:8-16,66-74): built from a model,
    callable, train()/eval() -- so it can stand where `Diloco` stands in the training loop."""

    def __init__(self, model, inner_optimizer, outer_optimizer, warmup_steps, total_steps,
                 inner_steps: int = 100, outer_steps: int = 10):
        self.model = model
        super().__init__(list(model.parameters()), inner_optimizer, outer_optimizer, warmup_steps, total_steps)
        self._sync_time, self._sync_calls = 0, 0                       # diloco.py:23-24 (never updated)

    @property
    def avg_sync_time(self):                                            # diloco.py:62-64
        return self._sync_time / self._sync_calls if self._sync_calls > 0 else 0

    def __call__(self, *args, **kwargs):                                # diloco.py:66-67
        return self.model(*args, **kwargs)

    def train(self):
        self.model.train()

    def eval(self):
        self.model.eval()
