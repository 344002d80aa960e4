"""GPU: nd_sumsq_static -- the entry point for callers that cannot provide scratch or allocate.

Some embedders of this library own no allocator (a thin ctypes/cgo caller forwarding somebody
else's tensors), or must not trigger one (all of HBM already belongs to a capped pool; a
latency-critical loop that cannot take the synchronisation hidden in cudaMalloc).  For them the
sum of squares keeps its partial sums in a fixed array inside the library instead of in
caller-provided scratch.  This file is collected last on purpose (pytest orders files by name),
and it must pass both on its own in a fresh process and after every other test has used the GPU.
"""
import numpy as np
import pytest
import torch

from nanodiloco_b200 import _lib
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _sumsq_static(g: torch.Tensor, out: torch.Tensor) -> None:
    # straight through the C ABI, no helper in between
    L = _lib.load()
    code = _lib.ND_BF16 if g.dtype == torch.bfloat16 else _lib.ND_F32
    rc = L.nd_sumsq_static(g.data_ptr(), code, g.numel(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.nd_last_error()


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    gen = torch.Generator(device="cuda").manual_seed(11)
    a = torch.randn(3_000_017, device="cuda", generator=gen) * 0.05
    b = torch.randn(70_001, device="cuda", generator=gen) * 3.0
    want = torch.zeros(2, device="cuda")
    sc = _lib.scratch("cuda")
    _lib.sumsq(a, want[0:1], sc)          # the existing entry point, caller-provided scratch
    _lib.sumsq(b, want[1:2], sc)
    got = torch.full((2,), float("nan"), device="cuda")
    _sumsq_static(a, got[0:1])            # two calls back to back share the library's array:
    _sumsq_static(b, got[1:2])            # the second must not see leftovers of the first
    torch.cuda.synchronize()
    # same tolerance as the other sum-of-squares tests (tests/test_gpu_inner.py)
    assert got[0].item() == pytest.approx(want[0].item(), rel=2e-6)
    assert got[1].item() == pytest.approx(want[1].item(), rel=2e-6)
    # and against the CPU oracle, so that this file does not only compare the kernel with itself
    assert got[0].item() == pytest.approx(O.sumsq(a.cpu().numpy()), rel=2e-6)
    assert got[1].item() == pytest.approx(O.sumsq(b.cpu().numpy()), rel=2e-6)
    # and again, now that the array holds the second call's partial sums
    again = torch.zeros(1, device="cuda")
    _sumsq_static(a, again)
    assert again.item() == pytest.approx(want[0].item(), rel=2e-6)
    # bf16 input and the empty input behave like the existing entry point too
    h = (torch.randn(12_345, device="cuda", generator=gen)).to(torch.bfloat16)
    w16, g16 = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    _lib.sumsq(h, w16, sc)
    _sumsq_static(h, g16)
    assert g16.item() == pytest.approx(w16.item(), rel=2e-6)
    e = torch.full((1,), 5.0, device="cuda")
    _sumsq_static(torch.zeros(0, device="cuda"), e)
    assert e.item() == 0.0


def test_static_scratch_rejects_sizes_beyond_its_limit_and_allocates_nothing():
    L = _lib.load()
    out = torch.zeros(1, device="cuda")
    g = torch.ones(1024, device="cuda")
    _sumsq_static(g, out)                 # set-up happens here, before the measurement below
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(50):
        _sumsq_static(g, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0                 # device memory in use did not move
    assert out.item() == 1024.0
    # one element past the documented maximum: refused before anything is launched
    rc = L.nd_sumsq_static(g.data_ptr(), _lib.ND_F32, (1 << 31) + 1, out.data_ptr(), 0)
    assert rc == 1 and b"ND_SUMSQ_STATIC_MAX_ELEMS" in L.nd_last_error()
