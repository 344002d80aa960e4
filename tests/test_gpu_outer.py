"""GPU parity of the outer-sync kernels, called through the C ABI (nanodiloco_b200._lib is a
ctypes binding of include/nanodiloco_b200.h), against the oracle, the reference's golden
fixtures and torch's own CUDA SGD (what the reference executes on a GPU)."""
import os

import numpy as np
import pytest
import torch

from nanodiloco_b200 import _lib
from oracle import oracle as O
from gpu_util import dev, host, unaligned

pytestmark = pytest.mark.gpu
LR, MU = 0.7, 0.9


def _rand(n, seed, scale):
    return (np.random.default_rng(seed).standard_normal(n) * scale).astype(np.float32)


# ---- (1) pseudo-gradient --------------------------------------------------------------------

@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 127, 4096, 16384 * 3 + 7, 1_000_003])
@pytest.mark.parametrize("misalign", [False, True])
def test_pseudo_grad_bit_exact(n, misalign):
    master, local = _rand(n, 1, 0.02), _rand(n, 2, 0.02)
    m, l = dev(master), dev(local)
    out = torch.full((n,), float("nan"), device="cuda")
    if misalign and n:
        l = unaligned(l)
    _lib.pseudo_grad(m, l, out)
    torch.cuda.synchronize()
    assert np.array_equal(host(out), O.pseudo_grad(master, local))


def test_pseudo_grad_bf16_local():
    n = 100_003
    master = _rand(n, 1, 0.02)
    local = torch.from_numpy(_rand(n, 2, 0.02)).to(torch.bfloat16)
    out = torch.empty(n, device="cuda")
    _lib.pseudo_grad(dev(master), local.cuda(), out)
    want = O.pseudo_grad(master, local.float().numpy())      # extension: bf16 working copy widened exactly
    assert np.array_equal(host(out), want)


# ---- (1)+(2) fused, against the reference's golden vectors -------------------------------------

@pytest.mark.parametrize("W", [1, 2, 4])
def test_outer_fused_vs_reference_golden(golden_dir, W):
    z = np.load(os.path.join(golden_dir, f"outer_w{W}.npz"))
    master = dev(z["init"])
    mom = torch.zeros_like(master)
    om, ob = z["init"].copy(), np.zeros_like(z["init"])
    for r in range(3):
        locs_np = [z[f"local_r{r}_w{w}"].copy() for w in range(W)]
        locs = [dev(a) for a in locs_np]
        _lib.outer_fused(master, mom, locs, lr=LR, momentum=MU, first_step=(r == 0))
        O.outer_step(om, locs_np, ob, lr=LR, momentum=MU, first_step=(r == 0))
        torch.cuda.synchronize()
        # bit-exact against the oracle for every W (same summation order by construction)
        assert np.array_equal(host(master), om) and np.array_equal(host(mom), ob)
        for l in locs:
            assert torch.equal(l, master)                    # diloco.py:50,54: all workers rewound
        if W <= 2:   # and bit-exact against the reference itself where the average is order-free
            assert np.array_equal(host(master), z[f"params_r{r}"])
            assert np.array_equal(host(mom), z[f"mom_r{r}"])
        else:        # north-star tolerance: rtol 1e-6 (+ atol floor 1e-6*max|x|)
            for got, want in ((host(master), z[f"params_r{r}"]), (host(mom), z[f"mom_r{r}"])):
                np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6 * np.abs(want).max())


@pytest.mark.parametrize("W", [1, 2, 3, 4, 5, 8, 9, 16])
@pytest.mark.parametrize("n", [5, 4099, 300_001])
def test_outer_fused_any_w_vs_oracle(W, n):
    master_np, mom_np = _rand(n, 10, 0.02), _rand(n, 11, 1e-3)
    locs_np = [master_np + _rand(n, 100 + w, 1e-3) for w in range(W)]
    master, mom, locs = dev(master_np), dev(mom_np), [dev(a) for a in locs_np]
    sumsq = torch.zeros(1, device="cuda")
    _lib.outer_fused(master, mom, locs, lr=LR, momentum=MU, first_step=False, sumsq_out=sumsq,
                     scratch_buf=_lib.scratch("cuda"))
    g = O.average([O.pseudo_grad(master_np, l) for l in locs_np])
    O.outer_step(master_np, locs_np, mom_np, lr=LR, momentum=MU, first_step=False)
    assert np.array_equal(host(master), master_np) and np.array_equal(host(mom), mom_np)
    assert all(torch.equal(l, master) for l in locs)
    assert float(sumsq) == pytest.approx(O.sumsq(g), rel=1e-5)       # pseudo-gradient norm


@pytest.mark.parametrize("nesterov,dampening,wd,first", [(True, 0, 0, True), (False, 0, 0, False),
                                                         (False, 0.1, 0, False), (True, 0, 1e-2, False),
                                                         (True, 0, 1e-2, True)])
def test_outer_fused_hyper_variants(nesterov, dampening, wd, first):
    n = 70_001
    master_np, mom_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3)
    locs_np = [master_np + _rand(n, 3, 1e-3), master_np + _rand(n, 4, 1e-3)]
    master, mom, locs = dev(master_np), dev(mom_np), [dev(a) for a in locs_np]
    kw = dict(lr=0.5, momentum=0.8, dampening=dampening, weight_decay=wd, nesterov=nesterov, first_step=first)
    _lib.outer_fused(master, mom, locs, **kw)
    O.outer_step(master_np, locs_np, mom_np, **kw)
    assert np.array_equal(host(master), master_np) and np.array_equal(host(mom), mom_np)


def test_outer_fused_unaligned_and_bf16():
    n = 33_333
    master_np, mom_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3)
    loc_np = master_np + _rand(n, 3, 1e-3)
    master, mom, loc = unaligned(dev(master_np)), dev(mom_np), unaligned(dev(loc_np), 3)
    _lib.outer_fused(master, mom, [loc], lr=LR, momentum=MU)
    O.outer_step(master_np, [loc_np], mom_np, lr=LR, momentum=MU)
    assert np.array_equal(host(master), master_np) and np.array_equal(host(loc), master_np)
    # bf16 working copy (extension, no reference behaviour): fp32 master, bf16 local
    master_np, mom_np = _rand(n, 1, 0.02), np.zeros(n, np.float32)
    loc16 = torch.from_numpy(master_np + _rand(n, 3, 1e-3)).to(torch.bfloat16)
    master, mom, loc = dev(master_np), dev(mom_np), loc16.cuda()
    _lib.outer_fused(master, mom, [loc], lr=LR, momentum=MU, first_step=True)
    O.outer_step(master_np, [loc16.float().numpy()], mom_np, lr=LR, momentum=MU, first_step=True)
    assert np.array_equal(host(master), master_np)
    assert torch.equal(loc.cpu(), torch.from_numpy(master_np).to(torch.bfloat16))


def test_outer_fused_separate_read_and_write_sets():
    """Workers that span several GPUs (BASELINE config 5, an extension): average over one copy per worker,
    rewind every replica.  Replicas 1 and 3 hold stale garbage on purpose: they must not be read."""
    n = 50_003
    master_np, mom_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3)
    w0, w1 = master_np + _rand(n, 3, 1e-3), master_np + _rand(n, 4, 1e-3)
    copies = [dev(w0), dev(_rand(n, 9, 5.0)), dev(w1), dev(_rand(n, 10, 5.0))]
    master, mom = dev(master_np), dev(mom_np)
    _lib.outer_fused_rw(master, mom, [copies[0], copies[2]], copies, lr=LR, momentum=MU)
    O.outer_step(master_np, [w0, w1], mom_np, lr=LR, momentum=MU)
    assert np.array_equal(host(master), master_np) and np.array_equal(host(mom), mom_np)
    assert all(torch.equal(c, master) for c in copies)
    # no write set at all: master and momentum move, the copies stay
    before = copies[0].clone()
    copies[0].add_(1e-3)
    keep = copies[0].clone()
    _lib.outer_fused_rw(master, mom, [copies[0]], [], lr=LR, momentum=MU)
    assert torch.equal(copies[0], keep) and not torch.equal(master, before)


# ---- (2) shard update after an external SUM ---------------------------------------------------

@pytest.mark.parametrize("W", [2, 8])
@pytest.mark.parametrize("with_copy", [True, False])
def test_outer_nesterov_shard(W, with_copy):
    n = 123_457
    master_np, mom_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3)
    locs_np = [master_np + _rand(n, 10 + w, 1e-3) for w in range(W)]
    pg_sum = O.pseudo_grad(master_np, locs_np[0])
    for l in locs_np[1:]:
        pg_sum = (pg_sum + O.pseudo_grad(master_np, l)).astype(np.float32)
    master, mom = dev(master_np), dev(mom_np)
    out = torch.zeros(n, device="cuda") if with_copy else None
    sumsq = torch.zeros(1, device="cuda")
    _lib.outer_nesterov(master, mom, dev(pg_sum), out, world=float(W), lr=LR, momentum=MU,
                        sumsq_out=sumsq, scratch_buf=_lib.scratch("cuda"))
    O.outer_step(master_np, locs_np, mom_np, lr=LR, momentum=MU)
    assert np.array_equal(host(master), master_np) and np.array_equal(host(mom), mom_np)
    if with_copy:
        assert torch.equal(out, master)


@pytest.mark.parametrize("first", [True, False])
def test_outer_nesterov_shard_bf16_copy_back_tma_path(first):
    """bf16 copy-back of the shard update on the TMA-staged path (every buffer 16-byte aligned), several
    tiles per CTA, repeated: the bf16 image of the new parameters has a stage buffer of its own -- when
    it shared the pseudo-gradient's, one warp's store could overwrite elements another warp had yet to
    read, and only some runs would show it."""
    n = 148 * 2 * 2048 * 3 + 8 * 37            # > 2 tiles for every CTA of a full grid, plus a partial tile
    W = 4
    master_np, mom_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3)
    locs_np = [master_np + _rand(n, 10 + w, 1e-3) for w in range(W)]
    pg_sum = O.pseudo_grad(master_np, locs_np[0])
    for l in locs_np[1:]:
        pg_sum = (pg_sum + O.pseudo_grad(master_np, l)).astype(np.float32)
    want_m, want_b = master_np.copy(), mom_np.copy()
    O.outer_step(want_m, locs_np, want_b, lr=LR, momentum=MU, first_step=first)
    want16 = torch.from_numpy(want_m).to(torch.bfloat16)
    pg = dev(pg_sum)
    for _ in range(20):
        master, mom = dev(master_np), dev(mom_np)
        out = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
        assert out.data_ptr() % 16 == 0 and master.data_ptr() % 16 == 0
        _lib.outer_nesterov(master, mom, pg, out, world=float(W), lr=LR, momentum=MU, first_step=first)
        assert np.array_equal(host(master), want_m) and np.array_equal(host(mom), want_b)
        assert torch.equal(out.cpu(), want16)
    # no copy-back requested: master and momentum only
    master, mom = dev(master_np), dev(mom_np)
    L = _lib.load()
    rc = L.nd_outer_nesterov(master.data_ptr(), mom.data_ptr(), pg.data_ptr(), None, _lib.ND_BF16, n, float(W), LR, MU,
                             0.0, 0.0, 1, int(first), None, None, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.nd_last_error()
    assert np.array_equal(host(master), want_m) and np.array_equal(host(mom), want_b)


# ---- multi-tensor forms --------------------------------------------------------------------------

def test_multi_tensor_forms_match_flat():
    sizes = [128, 7, 165, 1, 16384, 16385, 50_000, 3]
    L = _lib.load()
    ms = [dev(_rand(s, 10 + i, 0.02)) for i, s in enumerate(sizes)]
    ls = [m + dev(_rand(s, 50 + i, 1e-3)) for i, (m, s) in enumerate(zip(ms, sizes))]
    bs = [dev(_rand(s, 90 + i, 1e-3)) for i, s in enumerate(sizes)]
    pgs = [torch.empty_like(m) for m in ms]
    tab, T, chunks = _lib.make_table(list(zip(ms, ls, pgs)), "cuda")
    _lib.check(L.nd_pseudo_grad_mt(tab.data_ptr(), T, chunks, 0, _lib._stream()))
    for m, l, pg in zip(ms, ls, pgs):
        assert torch.equal(pg, m - l)
    want = []
    for m, l, b in zip(ms, ls, bs):
        m2, l2, b2 = m.clone(), l.clone(), b.clone()
        _lib.outer_fused(m2, b2, [l2], lr=LR, momentum=MU)
        want.append((m2, b2))
    tab, T, chunks = _lib.make_table(list(zip(ms, bs, ls)), "cuda")
    _lib.check(L.nd_outer_fused_mt(tab.data_ptr(), T, chunks, 0, LR, MU, 0.0, 0.0, 1, 0, _lib._stream()))
    for (m2, b2), m, l, b in zip(want, ms, ls, bs):
        assert torch.equal(m, m2) and torch.equal(b, b2) and torch.equal(l, m2)


# ---- against what the reference executes on a GPU: torch's CUDA foreach SGD -----------------

@pytest.mark.parametrize("n", [1_000_003, 50_000_000])
def test_outer_fused_bit_exact_vs_torch_cuda_sgd(n):
    """diloco.py:46-54 restated with torch CUDA ops (sub, SGD.step) on the same inputs."""
    g = torch.Generator(device="cuda").manual_seed(0)
    master = torch.randn(n, device="cuda", generator=g) * 0.02
    ref_p = torch.nn.Parameter(master.clone())
    opt = torch.optim.SGD([ref_p], lr=LR, momentum=MU, nesterov=True)
    mom = torch.zeros(n, device="cuda")
    for r in range(3):
        local = master + torch.randn(n, device="cuda", generator=g) * 1e-3
        ref_p.grad = ref_p.data - local                    # diloco.py:48 (W=1: AVG is the identity)
        opt.step()                                         # diloco.py:52
        _lib.outer_fused(master, mom, [local], lr=LR, momentum=MU, first_step=(r == 0))
        assert torch.equal(master, ref_p.data), r
        assert torch.equal(mom, opt.state[ref_p]["momentum_buffer"]), r
        assert torch.equal(local, master)


# ---- full size (BASELINE config 4: 1B-class Llama, N = 1,261,529,088): properties -------------

def test_full_size_properties():
    n = 1_261_529_088
    g = torch.Generator(device="cuda").manual_seed(1)
    master = torch.empty(n, device="cuda").normal_(0, 0.02, generator=g)
    local = master.clone()
    mom = torch.zeros(n, device="cuda")
    # idempotence: local == master -> zero pseudo-gradient -> nothing moves, momentum stays zero
    _lib.outer_fused(master, mom, [local], lr=LR, momentum=MU, first_step=True)
    assert torch.equal(master, local) and float(mom.abs().max()) == 0.0
    before = master.clone()
    # a real step: every worker copy ends equal to the master, and the update has the closed form
    local.add_(torch.empty(n, device="cuda").normal_(0, 1e-3, generator=g))
    delta = before - local
    _lib.outer_fused(master, mom, [local], lr=LR, momentum=MU, first_step=True)
    assert torch.equal(master, local)
    assert torch.equal(mom, delta)                                    # first step: b = g
    # p = p - lr*(g + mu*g): the same sequence done with torch's own CUDA add(alpha=)
    d = torch.add(delta, mom, alpha=MU)                               # fma(mu, b, g), sgd.py:368
    want = torch.add(before, d, alpha=-LR)                            # sgd.py:380
    assert torch.equal(master, want)
    del before, delta, d, want
    # linearity checksum: sum of master == sum of local (bitwise equal arrays)
    assert float(master.double().sum()) == float(local.double().sum())


def test_more_than_2_31_elements():
    """Maximum sizes: BASELINE's microbench goes to 2e9 elements; one step past 2^31 checks that no index
    is 32-bit anywhere (fused TMA kernel, pseudo-gradient, sum of squares)."""
    n = (1 << 31) + 4100
    master = torch.full((n,), 0.5, device="cuda")
    local = torch.full((n,), 0.25, device="cuda")
    local[-3:] = 0.0                                    # make the far end distinguishable
    mom = torch.zeros(n, device="cuda")
    _lib.outer_fused(master, mom, [local], lr=LR, momentum=MU, first_step=True)
    # g = 0.25 (0.5 at the end); b = g; p = 0.5 - 0.7*(g + 0.9 g)
    want_body, want_end = 0.5 - 0.7 * (0.25 + 0.9 * 0.25), 0.5 - 0.7 * (0.5 + 0.9 * 0.5)
    assert torch.equal(master, local)
    assert float(master[0]) == pytest.approx(want_body, rel=1e-6) and float(master[-1]) == pytest.approx(want_end, rel=1e-6)
    assert float(master[(1 << 31) + 5]) == pytest.approx(want_body, rel=1e-6)
    assert float(mom[-1]) == 0.5 and float(mom[1 << 31]) == 0.25
    assert int((master == master[0]).sum()) == n - 3    # every element was visited exactly as expected
    pg = torch.empty(n, device="cuda")
    local.fill_(0.125)
    _lib.pseudo_grad(master, local, pg)
    assert torch.equal(pg, master - 0.125)
    del pg, mom
    ss, sc = torch.zeros(1, device="cuda"), _lib.scratch("cuda")
    _lib.sumsq(local, ss, sc)
    assert float(ss) == pytest.approx(n * 0.125 ** 2, rel=1e-6)


# ---- peer-memory entry point on one GPU (world size 1: plumbing of the shared buffer) ------------

def test_peer_context_world1_matches_oracle():
    from nanodiloco_b200.sync import PeerContext
    n = 40_004
    ctx = PeerContext(1, 0, n, device=torch.device("cuda", 0))
    local = ctx.local()[:n]
    assert float(local.abs().max()) == 0.0 and local.data_ptr() % 256 == 0       # zero-filled, aligned
    master_np, mom_np = _rand(n, 1, 0.02), _rand(n, 2, 1e-3)
    loc_np = master_np + _rand(n, 3, 1e-3)
    local.copy_(torch.from_numpy(loc_np))
    assert torch.equal(ctx.peer_view(0)[:n], local)                              # same memory
    master, mom = dev(master_np), dev(mom_np)
    ctx.outer_sync(master, mom, lr=LR, momentum=MU)
    O.outer_step(master_np, [loc_np], mom_np, lr=LR, momentum=MU)
    torch.cuda.synchronize()
    assert np.array_equal(host(master), master_np) and np.array_equal(host(local), master_np)
    assert ctx.last_ms() > 0
    ctx.barrier()                                                                # single-rank barrier returns
    torch.cuda.synchronize()
    ctx.close()


def test_multi_tensor_forms_with_bf16_working_copy():
    sizes = [128, 8, 16385, 40_000, 24]
    L = _lib.load()
    ms = [dev(_rand(s, 10 + i, 0.02)) for i, s in enumerate(sizes)]
    ls = [(m + dev(_rand(s, 50 + i, 1e-3))).to(torch.bfloat16) for i, (m, s) in enumerate(zip(ms, sizes))]
    bs = [dev(_rand(s, 90 + i, 1e-3)) for i, s in enumerate(sizes)]
    pgs = [torch.empty_like(m) for m in ms]
    tab, T, chunks = _lib.make_table(list(zip(ms, ls, pgs)), "cuda")
    _lib.check(L.nd_pseudo_grad_mt(tab.data_ptr(), T, chunks, _lib.ND_BF16, _lib._stream()))
    for m, l, pg in zip(ms, ls, pgs):
        assert torch.equal(pg, m - l.float())
    want = []
    for m, l, b in zip(ms, ls, bs):
        m2, l2, b2 = m.clone(), l.clone(), b.clone()
        _lib.outer_fused(m2, b2, [l2], lr=LR, momentum=MU)
        want.append((m2, b2, l2))
    tab, T, chunks = _lib.make_table(list(zip(ms, bs, ls)), "cuda")
    _lib.check(L.nd_outer_fused_mt(tab.data_ptr(), T, chunks, _lib.ND_BF16, LR, MU, 0.0, 0.0, 1, 0, _lib._stream()))
    for (m2, b2, l2), m, l, b in zip(want, ms, ls, bs):
        assert torch.equal(m, m2) and torch.equal(b, b2) and torch.equal(l, l2)
