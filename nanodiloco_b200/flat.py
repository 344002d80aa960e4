"""Flat parameter space: every tensor at a 16-byte-aligned offset of one buffer.

The reference keeps per-tensor lists and rebinds `param.data` to a fresh tensor
every outer step (nanodiloco/diloco/diloco.py:50), so parameter addresses move.
Here parameters are re-homed ONCE into a flat buffer and only ever updated in
place: `Parameter` identity and storage stay fixed, one kernel launch covers the
whole model, and optimizer state can be handed out as views.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence

import torch

ALIGN_ELEMS = 4  # 16 bytes of fp32; also keeps bf16 tensors 8-byte aligned


def _round_up(x: int, q: int) -> int:
    return (x + q - 1) // q * q


@dataclass
class FlatLayout:
    offsets: List[int]   # start of tensor i in elements
    numels: List[int]
    shapes: List[torch.Size]
    total: int           # elements including inter-tensor padding, multiple of `quantum`
    quantum: int

    @staticmethod
    def plan(tensors: Sequence[torch.Tensor], quantum: int = ALIGN_ELEMS, align: int = ALIGN_ELEMS) -> "FlatLayout":
        """Lays tensors out in order; `total` is padded to a multiple of `quantum`
        (4*world for the sharded sync, so every bucket splits into aligned slices).
        `align`: elements per 16 bytes of the narrowest buffer sharing this layout (4 for fp32,
        8 when the working copy is bf16)."""
        offsets, numels, shapes = [], [], []
        off = 0
        for t in tensors:
            off = _round_up(off, align)
            offsets.append(off)
            numels.append(t.numel())
            shapes.append(t.shape)
            off += t.numel()
        q = _round_up(max(quantum, align), align)
        return FlatLayout(offsets, numels, shapes, _round_up(max(off, 1), q), q)

    def views(self, flat: torch.Tensor) -> List[torch.Tensor]:
        return [flat[o:o + n].view(s) for o, n, s in zip(self.offsets, self.numels, self.shapes)]

    def span(self, first: int, last: int) -> tuple:
        """[start, end) element range covering tensors first..last inclusive."""
        return self.offsets[first], self.offsets[last] + self.numels[last]


def rehome_(params: Sequence[torch.nn.Parameter], flat: torch.Tensor, layout: FlatLayout) -> None:
    """Copies each parameter into its slot of `flat` and points `param.data` at the slot.

    Done once at construction; afterwards parameter storage never moves."""
    with torch.no_grad():
        for p, v in zip(params, layout.views(flat)):
            v.copy_(p.data)
            p.data = v


def trainable_runs(layout: FlatLayout, groups: Sequence[Sequence[torch.nn.Parameter]], first_index: int = 0):
    """For each parameter group (a list of parameters laid out consecutively from tensor index
    `first_index`), the maximal runs of consecutive TRAINABLE tensors as (group_index, start, end)
    element ranges.  torch's optimizers never touch a parameter without a gradient -- which is what a
    frozen (`requires_grad=False`) parameter has -- so the fused inner step must skip those tensors
    entirely (not even weight decay); they still take part in the outer sync."""
    out, i = [], first_index
    for gi, params in enumerate(groups):
        run_start = last = None
        for j, p in enumerate(params):
            k = i + j
            if p.requires_grad:
                if run_start is None:
                    run_start = k
                last = k
            if (not p.requires_grad or j == len(params) - 1) and run_start is not None:
                s, e = layout.span(run_start, last)
                out.append((gi, s, e))
                run_start = None
        i += len(params)
    return out
