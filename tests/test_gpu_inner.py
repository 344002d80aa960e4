"""GPU parity of the inner-step kernels (sum of squares, fused clip + AdamW) through the C ABI."""
import ctypes
import os

import numpy as np
import pytest
import torch

from nanodiloco_b200 import _lib
from oracle import oracle as O
from gpu_util import dev, host, unaligned

pytestmark = pytest.mark.gpu


def _rand(n, seed, scale):
    return (np.random.default_rng(seed).standard_normal(n) * scale).astype(np.float32)


@pytest.mark.parametrize("n", [0, 1, 3, 5, 4096, 1_000_003, 30_000_001])
def test_sumsq_vs_oracle_and_deterministic(n):
    g_np = _rand(n, 5, 0.05)
    g = dev(g_np)
    sc = _lib.scratch("cuda")
    out = torch.full((2,), float("nan"), device="cuda")
    _lib.sumsq(g, out[0:1], sc)
    _lib.sumsq(g, out[1:2], sc)                     # scratch is reusable without re-zeroing
    a, b = out.tolist()
    assert a == b                                   # fixed reduction order: bit-identical reruns
    assert a == pytest.approx(O.sumsq(g_np), rel=2e-6, abs=0)
    if n:
        _lib.sumsq(unaligned(g), out[0:1], sc)      # scalar path
        assert float(out[0]) == pytest.approx(O.sumsq(g_np), rel=2e-6)


def test_sumsq_bf16_and_multi_tensor():
    g16 = torch.from_numpy(_rand(70_001, 6, 0.05)).to(torch.bfloat16).cuda()
    sc = _lib.scratch("cuda")
    out = torch.zeros(1, device="cuda")
    _lib.sumsq(g16, out, sc)
    assert float(out) == pytest.approx(O.sumsq(g16.float().cpu().numpy()), rel=2e-6)
    ts = [dev(_rand(s, 20 + i, 0.05)) for i, s in enumerate([128, 7, 16385, 1, 40_000])]
    tab, T, chunks = _lib.make_table([(t,) for t in ts], "cuda")
    _lib.check(_lib.load().nd_sumsq_mt(tab.data_ptr(), T, chunks, _lib.ND_F32, out.data_ptr(), sc.data_ptr(),
                                       _lib._stream()))
    want = sum(O.sumsq(host(t)) for t in ts)
    assert float(out) == pytest.approx(want, rel=2e-6)
    t16 = [t.to(torch.bfloat16) for t in ts]
    tab, T, chunks = _lib.make_table([(t,) for t in t16], "cuda")
    _lib.check(_lib.load().nd_sumsq_mt(tab.data_ptr(), T, chunks, _lib.ND_BF16, out.data_ptr(), sc.data_ptr(),
                                       _lib._stream()))
    want = sum(O.sumsq(t.float().cpu().numpy()) for t in t16)
    assert float(out) == pytest.approx(want, rel=2e-6)


@pytest.mark.parametrize("n", [1, 7, 4099, 1_000_003])
@pytest.mark.parametrize("scale", [0.05, 1e-4])          # clipping active / inactive
def test_adamw_bit_exact_vs_oracle_cuda_recipe(n, scale):
    p_np, m_np, v_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3), np.abs(_rand(n, 3, 1e-6))
    p, m, v = dev(p_np), dev(m_np), dev(v_np)
    sc = _lib.scratch("cuda")
    ss = torch.zeros(1, device="cuda")
    for step in range(1, 4):
        g_np = _rand(n, 10 + step, scale)
        g = dev(g_np)
        _lib.sumsq(g, ss, sc)
        _lib.adamw(p, g, m, v, lr=4e-4, step=step, sumsq_dev=ss, max_norm=1.0, zero_grad=True)
        coef = O.clip_coef(float(np.sqrt(np.float32(float(ss)))), 1.0)   # same fp32 sqrt as the kernel
        O.adamw_step(p_np, g_np, m_np, v_np, lr=4e-4, step=step, clip=coef, recipe="cuda")
        assert np.array_equal(host(m), m_np) and np.array_equal(host(v), v_np), step
        assert np.array_equal(host(p), p_np), step
        assert float(g.abs().max()) == 0.0                                # fused zero_grad (diloco.py:60)


def test_adamw_vs_torch_cuda_foreach():
    """What the reference executes on a GPU (diloco.py:57-58): clip_grad_norm_ + AdamW.step()."""
    sizes = [4096, 7, 100_000, 333]
    gen = torch.Generator(device="cuda").manual_seed(0)
    ref = [torch.nn.Parameter(torch.randn(s, device="cuda", generator=gen) * 0.02) for s in sizes]
    opt = torch.optim.AdamW(ref, lr=4e-4)
    n = sum(sizes)
    p = torch.cat([r.data.reshape(-1) for r in ref]).clone()
    m, v, ss = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda"), torch.zeros(1, device="cuda")
    sc = _lib.scratch("cuda")
    for step in range(1, 6):
        scale = 0.05 if step % 2 else 1e-4
        grads = [torch.randn(s, device="cuda", generator=gen) * scale for s in sizes]
        for r, g in zip(ref, grads):
            r.grad = g.clone()
        total = torch.nn.utils.clip_grad_norm_(ref, max_norm=1.0)
        opt.step()
        g = torch.cat(grads)
        _lib.sumsq(g, ss, sc)
        _lib.adamw(p, g, m, v, lr=4e-4, step=step, sumsq_dev=ss, max_norm=1.0)
        assert float(ss.sqrt()) == pytest.approx(float(total), rel=2e-6)
        want_p = torch.cat([r.data for r in ref])
        want_m = torch.cat([opt.state[r]["exp_avg"] for r in ref])
        want_v = torch.cat([opt.state[r]["exp_avg_sq"] for r in ref])
        # the clip coefficient differs by an ulp (different reduction order for the norm), so:
        # fp32 tolerance rtol 2e-6 with an atol floor of 1e-6 * max|x|
        for got, want in ((p, want_p), (m, want_m), (v, want_v)):
            torch.testing.assert_close(got, want, rtol=2e-6, atol=1e-6 * float(want.abs().max()))


def test_adamw_vs_reference_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "inner.npz"))
    p, n = dev(z["init"]), z["init"].size
    m, v, ss = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda"), torch.zeros(1, device="cuda")
    sc = _lib.scratch("cuda")
    for s in range(6):
        g = dev(z[f"grad_s{s}"])
        _lib.sumsq(g, ss, sc)
        assert float(ss.sqrt()) == pytest.approx(float(z[f"total_norm_s{s}"]), rel=2e-6)
        _lib.adamw(p, g, m, v, lr=float(z[f"lr_s{s}"]), step=s + 1, sumsq_dev=ss, max_norm=1.0)
        for got, key in ((p, "params"), (m, "exp_avg"), (v, "exp_avg_sq")):
            want = z[f"{key}_s{s}"]
            np.testing.assert_allclose(host(got), want, rtol=2e-6, atol=1e-6 * np.abs(want).max())


@pytest.mark.parametrize("bf16", [False, True])
@pytest.mark.parametrize("shift", [0, 1])
def test_adamw_multi_tensor_matches_flat(bf16, shift):
    """nd_adamw_mt (table-driven, TMA-staged) against nd_adamw tensor by tensor, bit for bit: sizes below,
    at and above one tile and one chunk, ragged tails, many tiles per CTA; `shift` makes every row
    unaligned so that the kernel's plain-access path is the one that runs."""
    sizes = [128, 7, 2048, 16384, 16385, 1, 40_000, 2048 * 148 * 2 * 3 + 12, 5]
    wd = torch.bfloat16 if bf16 else torch.float32
    al = (lambda t: unaligned(t, shift)) if shift else (lambda t: t)
    mk = lambda seed, sc, dt=torch.float32: [al(dev(_rand(s, seed + i, sc)).to(dt)) for i, s in enumerate(sizes)]
    ps, gs, ms, vs = mk(0, 0.02, wd), mk(10, 0.01, wd), mk(20, 1e-3), [t.abs() for t in mk(30, 1e-6)]
    ss = torch.full((1,), 9.0, device="cuda")          # norm 3 -> clip coefficient 1/3
    want = []
    for p, g, m, v in zip(ps, gs, ms, vs):
        p2, m2, v2 = p.clone(), m.clone(), v.clone()
        _lib.adamw(p2, g.clone(), m2, v2, lr=1e-3, step=3, sumsq_dev=ss, max_norm=1.0)
        want.append((p2, m2, v2))
    g0 = [g.clone() for g in gs]
    tab, T, chunks = _lib.make_table(list(zip(ps, gs, ms, vs)), "cuda")
    code = _lib.ND_BF16 if bf16 else _lib.ND_F32
    _lib.check(_lib.load().nd_adamw_mt(tab.data_ptr(), T, chunks, code, 1e-3, 0.9, 0.999, 1e-8, 1e-2, 3, ss.data_ptr(),
                                       1.0, 0, _lib._stream()))
    for (p2, m2, v2), p, m, v, g, gb in zip(want, ps, ms, vs, gs, g0):
        assert torch.equal(p, p2) and torch.equal(m, m2) and torch.equal(v, v2)
        assert torch.equal(g, gb)                       # zero_grad = 0: gradients untouched
    # zero_grad = 1 clears them (the flat-buffer mode of the wrapper relies on it)
    _lib.check(_lib.load().nd_adamw_mt(tab.data_ptr(), T, chunks, code, 1e-3, 0.9, 0.999, 1e-8, 1e-2, 4, None, 1.0, 1,
                                       _lib._stream()))
    assert all(int(torch.count_nonzero(g)) == 0 for g in gs)


# ---- bf16 working copy (extension: BASELINE config 2, no reference behaviour) -----------------

@pytest.mark.parametrize("n", [8, 13, 4099, 300_007])
def test_adamw_bf16_working_copy_vs_oracle(n):
    """bf16 parameters and gradients, fp32 moments: fp32 arithmetic on the exactly-widened values
    (the oracle's CUDA recipe), one rounding of the new parameter to bf16."""
    p16 = torch.from_numpy(_rand(n, 1, 0.02)).to(torch.bfloat16).cuda()
    m_np, v_np = _rand(n, 2, 1e-3), np.abs(_rand(n, 3, 1e-6))
    m, v = dev(m_np), dev(v_np)
    sc, ss = _lib.scratch("cuda"), torch.zeros(1, device="cuda")
    for step in range(1, 4):
        g16 = torch.from_numpy(_rand(n, 10 + step, 0.05 if step % 2 else 1e-4)).to(torch.bfloat16).cuda()
        p_np, g_np = p16.float().cpu().numpy().copy(), g16.float().cpu().numpy().copy()
        _lib.sumsq(g16, ss, sc)
        _lib.adamw(p16, g16, m, v, lr=4e-4, step=step, sumsq_dev=ss, max_norm=1.0, zero_grad=True)
        coef = O.clip_coef(float(np.sqrt(np.float32(float(ss)))), 1.0)
        O.adamw_step(p_np, g_np, m_np, v_np, lr=4e-4, step=step, clip=coef, recipe="cuda")
        assert np.array_equal(host(m), m_np) and np.array_equal(host(v), v_np), step
        assert torch.equal(p16.cpu(), torch.from_numpy(p_np).to(torch.bfloat16)), step
        assert float(g16.float().abs().max()) == 0.0


def test_adamw_nonfinite_gradients_match_torch():
    """An inf in the gradients: torch's clip makes the coefficient 0, inf*0 = NaN poisons exactly that
    element's moments and parameter, everything else sees a zero gradient.  Same here."""
    n = 10_000
    gen = torch.Generator(device="cuda").manual_seed(1)
    ref = torch.nn.Parameter(torch.randn(n, device="cuda", generator=gen) * 0.02)
    opt = torch.optim.AdamW([ref], lr=1e-3)
    p, m, v = ref.data.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    g = torch.randn(n, device="cuda", generator=gen) * 0.01
    g[123] = float("inf")
    ref.grad = g.clone()
    torch.nn.utils.clip_grad_norm_([ref], 1.0)
    opt.step()
    ss, sc = torch.zeros(1, device="cuda"), _lib.scratch("cuda")
    _lib.sumsq(g, ss, sc)
    _lib.adamw(p, g.clone(), m, v, lr=1e-3, step=1, sumsq_dev=ss, max_norm=1.0)
    assert torch.isinf(ss).all()
    assert torch.equal(torch.isnan(p), torch.isnan(ref.data)) and int(torch.isnan(p).sum()) == 1
    ok = ~torch.isnan(p)
    torch.testing.assert_close(p[ok], ref.data[ok], rtol=2e-6, atol=1e-9)
    assert torch.equal(torch.isnan(m), torch.isnan(opt.state[ref]["exp_avg"]))


# ---- a worker that spans several GPUs (extension: BASELINE config 5), emulated on one GPU ---------------

@pytest.mark.parametrize("R", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("bf16", [False, True])
def test_grad_avg_and_adamw_rw_match_allreduce_avg_then_adamw(R, bf16):
    """nd_grad_avg + nd_adamw_rw on each replica's 1/R of the space == average of the R gradients (what
    dist.all_reduce(AVG) leaves on every replica), one global norm, nd_adamw on the whole space."""
    n = 2048 * 148 * 2 + 8 * R * 5            # several tiles per CTA; divisible by 8*R
    n -= n % (8 * R)
    wd = torch.bfloat16 if bf16 else torch.float32
    code = _lib.ND_BF16 if bf16 else _lib.ND_F32
    L, st = _lib.load(), _lib._stream()
    p0 = dev(_rand(n, 1, 0.02)).to(wd)
    grads = [dev(_rand(n, 10 + k, 0.05)).to(wd) for k in range(R)]
    m0, v0 = dev(_rand(n, 2, 1e-3)), dev(_rand(n, 3, 1e-6)).abs()
    # reference: sum in replica order, divide, one rounding to the gradient dtype
    acc = grads[0].float()
    for g in grads[1:]:
        acc = acc + g.float()
    avg = (acc / float(R)).to(wd)
    want_ss = torch.zeros(1, device="cuda")
    sc = _lib.scratch("cuda")
    _lib.sumsq(avg, want_ss, sc)
    wp, wm, wv = p0.clone(), m0.clone(), v0.clone()
    _lib.adamw(wp, avg.clone(), wm, wv, lr=1e-3, step=2, sumsq_dev=want_ss, max_norm=1.0)
    # replicas: copies of the parameters, own gradient buffers, partial norms
    ps = [p0.clone() for _ in range(R)]
    gs = [g.clone() for g in grads]
    ms, vs = m0.clone(), v0.clone()
    parts = torch.zeros(R, device="cuda")
    part = n // R
    esz = 2 if bf16 else 4
    arr = lambda ptrs: (ctypes.c_void_p * len(ptrs))(*ptrs)
    for j in range(R):
        off = j * part
        _lib.check(L.nd_grad_avg(arr([g.data_ptr() + off * esz for g in gs]), R, gs[j].data_ptr() + off * esz, code, part,
                                 parts[j:j + 1].data_ptr(), sc.data_ptr(), st), "nd_grad_avg")
    for j in range(R):
        off = j * part
        _lib.check(L.nd_adamw_rw(ps[j].data_ptr() + off * esz, arr([p.data_ptr() + off * esz for p in ps]), R,
                                 gs[j].data_ptr() + off * esz, ms.data_ptr() + off * 4, vs.data_ptr() + off * 4, code, part,
                                 1e-3, 0.9, 0.999, 1e-8, 1e-2, 2, arr([parts[k:k + 1].data_ptr() for k in range(R)]), R,
                                 1.0, 1, st), "nd_adamw_rw")
    torch.cuda.synchronize()
    assert float(parts.sum()) == pytest.approx(float(want_ss), rel=2e-6)
    for j in range(R):
        # the clip coefficient comes from the sum of R partial sums instead of one reduction: allow 1 ulp on it
        if bf16:   # ... which can flip the rounding to bf16 of a handful of elements by one bf16 ulp
            torch.testing.assert_close(ps[j].float(), wp.float(), rtol=8e-3, atol=1e-6)
            assert float((ps[j] != wp).float().mean()) < 1e-3
        else:
            torch.testing.assert_close(ps[j], wp, rtol=2e-6, atol=1e-8)
        assert torch.equal(ps[j], ps[0])                                      # every replica identical, bit for bit
        off = j * part
        assert int(torch.count_nonzero(gs[j][off:off + part])) == 0            # own part zeroed
    torch.testing.assert_close(ms, wm, rtol=2e-6, atol=1e-10)
    torch.testing.assert_close(vs, wv, rtol=4e-6, atol=1e-14)
