"""GPU: the drop-in `Diloco` wrapper end to end against the reference's golden fixtures."""
import os

import numpy as np
import pytest
import torch

from nanodiloco_b200 import Diloco
from gpu_util import host

pytestmark = pytest.mark.gpu
SHAPES = [(128,), (64, 32), (1000,), (7,), (3, 5, 11), (1,), (2048,)]   # tests/golden/make_golden.py


class Bag(torch.nn.Module):
    def __init__(self, init):
        super().__init__()
        ps, off = [], 0
        for s in SHAPES:
            n = int(np.prod(s))
            ps.append(torch.nn.Parameter(torch.from_numpy(init[off:off + n].copy()).reshape(s)))
            off += n
        self.ps = torch.nn.ParameterList(ps)


def _flat(ts):
    return torch.cat([t.detach().reshape(-1) for t in ts])


def _make(z, **kw):
    model = Bag(z["init"]).cuda()
    inner = torch.optim.AdamW(model.parameters(), lr=4e-4)
    outer = torch.optim.SGD(model.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    return model, inner, outer, Diloco(model, inner, outer, **kw)


def test_outer_step_drop_in_w1(golden_dir):
    z = np.load(os.path.join(golden_dir, "outer_w1.npz"))
    model, inner, outer, d = _make(z, warmup_steps=2, total_steps=100)
    params = list(model.parameters())
    ptrs, ids = [p.data_ptr() for p in params], [id(p) for p in params]
    assert len(outer.state) == 0                                     # like torch: state appears at the first step
    for r in range(3):
        with torch.no_grad():                                        # stand-in for H inner steps
            off = 0
            loc = z[f"local_r{r}_w0"]
            for p in params:
                p.copy_(torch.from_numpy(loc[off:off + p.numel()]).reshape(p.shape))
                off += p.numel()
        d.outer_step()
        assert np.array_equal(host(_flat(params)), z[f"params_r{r}"])                # bit-exact vs reference
        mom = _flat([outer.state[p]["momentum_buffer"] for p in params])
        assert np.array_equal(host(mom), z[f"mom_r{r}"])
        assert np.array_equal(host(_flat(d.offloaded_last_sync_parameters)), z[f"params_r{r}"])
    assert ptrs == [p.data_ptr() for p in params] and ids == [id(p) for p in params]   # storage never moves
    assert d.avg_sync_time > 0                                        # diloco.py:62-64, now real
    c = d.sync_counters()                                             # counters of the last sync
    assert c["sync"] == "none" and c["ms"] > 0 and c["hbm_gbs"] > 0 and d.avg_sync_wait_time is None
    assert float(d.last_outer_grad_norm) > 0
    sd = outer.state_dict()
    assert len(sd["state"]) == len(params)


@pytest.mark.parametrize("grad_buffer", ["autograd", "flat"])
def test_inner_step_drop_in(golden_dir, grad_buffer):
    z = np.load(os.path.join(golden_dir, "inner.npz"))
    model, inner, outer, d = _make(z, warmup_steps=2, total_steps=10, grad_buffer=grad_buffer)
    params = list(model.parameters())
    flat = grad_buffer == "flat"
    assert d.grad_buffer == grad_buffer
    for s in range(6):
        assert inner.param_groups[0]["lr"] == pytest.approx(float(z[f"lr_s{s}"]), rel=1e-12)
        off = 0
        g = z[f"grad_s{s}"]
        for i, p in enumerate(params):
            gi = torch.from_numpy(g[off:off + p.numel()]).reshape(p.shape).cuda()
            if flat and s % 2 == 0:
                p.grad.copy_(gi)            # autograd accumulating into the flat views
            else:
                assert flat or p.grad is None   # default mode: no gradient outlives a step (diloco.py:60)
                p.grad = gi                 # what autograd does when .grad is None; or a caller that replaced it
            off += p.numel()
        d.inner_step()
        assert float(d.last_grad_norm) == pytest.approx(float(z[f"total_norm_s{s}"]), rel=2e-6)
        for got, key in ((_flat(params), "params"),
                         (_flat([inner.state[p]["exp_avg"] for p in params]), "exp_avg"),
                         (_flat([inner.state[p]["exp_avg_sq"] for p in params]), "exp_avg_sq")):
            want = z[f"{key}_s{s}"]
            np.testing.assert_allclose(host(got), want, rtol=2e-6, atol=1e-6 * np.abs(want).max())
        if flat:
            assert all(float(p.grad.abs().max()) == 0.0 for p in params)   # zero_grad, fused into the AdamW pass
        else:
            assert all(p.grad is None for p in params)                     # zero_grad(set_to_none), diloco.py:60
        assert float(inner.state[params[0]]["step"]) == s + 1


@pytest.mark.parametrize("grad_buffer", ["autograd", "flat"])
def test_inner_step_skips_parameters_without_gradient(grad_buffer):
    """torch.optim.AdamW leaves a parameter whose .grad is None alone -- no weight decay, no moment decay
    (torch:optim/adam.py:150-160) -- and clip_grad_norm_ ignores it.  Same here, checked against torch's
    own CUDA AdamW on a model with an unused branch."""
    torch.manual_seed(5)
    mk = lambda: torch.nn.ModuleDict({"a": torch.nn.Linear(24, 40), "unused": torch.nn.Linear(40, 40),
                                      "b": torch.nn.Linear(40, 3)}).cuda()
    ours, ref = mk(), mk()
    ref.load_state_dict(ours.state_dict())
    fwd = lambda m, x: m["b"](torch.tanh(m["a"](x)))
    inner = torch.optim.AdamW(ours.parameters(), lr=1e-2)
    outer = torch.optim.SGD(ours.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(ours, inner, outer, warmup_steps=0, total_steps=1000, grad_buffer=grad_buffer)
    if grad_buffer == "flat":           # what a partial backward after zero_grad(set_to_none=True) leaves behind
        for p in ours["unused"].parameters():
            p.grad = None
    ref_opt = torch.optim.AdamW(ref.parameters(), lr=1e-2, foreach=True)
    ref_sched = torch.optim.lr_scheduler.LambdaLR(ref_opt, lambda k: d.scheduler.lr_lambdas[0](k))
    before = [p.detach().clone() for p in ours["unused"].parameters()]
    for step in range(3):
        x = torch.randn(16, 24, device="cuda")
        fwd(ours, x).square().mean().backward()
        fwd(ref, x).square().mean().backward()
        torch.nn.utils.clip_grad_norm_(ref.parameters(), max_norm=1.0)
        ref_opt.step(); ref_sched.step(); ref_opt.zero_grad()
        d.inner_step()
        if grad_buffer == "flat":
            for p in ours["unused"].parameters():
                p.grad = None
    for p, q in zip(ours["unused"].parameters(), before):
        assert torch.equal(p, q)                                           # untouched, weight decay included
    for p, q in zip(ours.parameters(), ref.parameters()):
        torch.testing.assert_close(p, q, rtol=2e-6, atol=1e-7)


def test_training_loop_with_autograd():
    """A real model stepping through the wrapper: loss goes down, nothing moves in memory."""
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.Tanh(), torch.nn.Linear(64, 1)).cuda()
    inner = torch.optim.AdamW(model.parameters(), lr=1e-2)
    outer = torch.optim.SGD(model.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(model, inner, outer, warmup_steps=2, total_steps=60)
    x = torch.randn(256, 32, device="cuda")
    y = x[:, :1] * 2 - x[:, 1:2]
    losses = []
    d.train()
    for step in range(1, 41):
        loss = torch.nn.functional.mse_loss(d(x), y)
        loss.backward()
        d.inner_step()
        if step % 10 == 0:
            d.outer_step()
        losses.append(float(loss))
    assert losses[-1] < 0.5 * losses[0]
    for p, q in zip(model.parameters(), d.offloaded_last_sync_parameters):
        assert torch.equal(p.data, q)                                 # just synced


def test_outer_step_offloaded_master_matches_resident(golden_dir):
    """Reference semantics: the last-sync copy lives on the HOST (diloco.py:25-32,47,54).  Same bits
    as the HBM-resident path and as the reference's golden vectors, with chunks smaller than,
    equal to and larger than the parameter space."""
    z = np.load(os.path.join(golden_dir, "outer_w1.npz"))
    for chunk_mb in (0.001, 0.01, 64.0):
        model, inner, outer, d = _make(z, warmup_steps=2, total_steps=100, offload_master=True,
                                       offload_chunk_mb=chunk_mb)
        params = list(model.parameters())
        for r in range(3):
            loc, off = z[f"local_r{r}_w0"], 0
            with torch.no_grad():
                for p in params:
                    p.copy_(torch.from_numpy(loc[off:off + p.numel()]).reshape(p.shape))
                    off += p.numel()
            d.outer_step()
            torch.cuda.synchronize()
            assert np.array_equal(host(_flat(params)), z[f"params_r{r}"])
            snap = d.offloaded_last_sync_parameters
            assert all(not t.is_cuda for t in snap)                    # on the host, like the reference's
            assert np.array_equal(_flat(snap).numpy(), z[f"params_r{r}"])
            mom = _flat([outer.state[p]["momentum_buffer"] for p in params])
            assert np.array_equal(host(mom), z[f"mom_r{r}"])


def test_bf16_working_copy_wrapper():
    """bf16 inner / fp32 outer (BASELINE config 2; an extension -- the reference is fp32 end to end).
    Oracle: the same C restatement on the exactly-widened bf16 values; master and momentum stay fp32."""
    from oracle import oracle as O
    torch.manual_seed(3)
    model = torch.nn.Sequential(torch.nn.Linear(40, 33), torch.nn.Tanh(), torch.nn.Linear(33, 5)).cuda().bfloat16()
    inner = torch.optim.AdamW(model.parameters(), lr=1e-2)
    outer = torch.optim.SGD(model.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(model, inner, outer, warmup_steps=1, total_steps=40)
    params = list(model.parameters())
    assert all(p.dtype == torch.bfloat16 and p.data_ptr() % 16 == 0 for p in params)
    flat32 = lambda ts: torch.cat([t.detach().reshape(-1).float() for t in ts]).cpu().numpy()
    master = flat32(params).copy()
    mom = np.zeros_like(master)
    x = torch.randn(64, 40, device="cuda", dtype=torch.bfloat16)
    for rnd in range(2):
        for _ in range(3):
            d(x).float().square().mean().backward()
            assert all(p.grad.dtype == torch.bfloat16 for p in params)
            d.inner_step()
        local = flat32(params)
        assert not np.array_equal(local, master)                       # the inner steps moved the weights
        O.outer_step(master, [local.copy()], mom, lr=0.7, momentum=0.9, first_step=(rnd == 0))
        d.outer_step()
        got_master = flat32(d.offloaded_last_sync_parameters)          # fp32 master, bit-exact
        assert np.array_equal(got_master, master)
        want16 = torch.from_numpy(master).to(torch.bfloat16).float().numpy()
        assert np.array_equal(flat32(params), want16)                  # working copy = bf16(master)
        assert np.array_equal(flat32([outer.state[p]["momentum_buffer"] for p in params]), mom)


def test_state_dict_round_trip(golden_dir):
    """Save after round 1, resume in a fresh wrapper, run round 2: same bits as the uninterrupted run."""
    z = np.load(os.path.join(golden_dir, "outer_w1.npz"))

    def drift(params, r):
        loc, off = z[f"local_r{r}_w0"], 0
        with torch.no_grad():
            for p in params:
                p.copy_(torch.from_numpy(loc[off:off + p.numel()]).reshape(p.shape))
                off += p.numel()

    model, inner, outer, d = _make(z, warmup_steps=2, total_steps=100)
    drift(list(model.parameters()), 0)
    d.outer_step()
    sd = d.state_dict()
    assert sd["master"].device.type == "cpu" and not sd["first_outer"]
    model2, inner2, outer2, d2 = _make(z, warmup_steps=2, total_steps=100)
    d2.load_state_dict(sd)
    for r in (1, 2):
        drift(list(model2.parameters()), r)
        d2.outer_step()
        assert np.array_equal(host(_flat(model2.parameters())), z[f"params_r{r}"])
        mom = _flat([outer2.state[p]["momentum_buffer"] for p in model2.parameters()])
        assert np.array_equal(host(mom), z[f"mom_r{r}"])


def test_stock_optimizer_state_dicts_round_trip(golden_dir):
    """A checkpoint layer written against the reference saves `inner_optimizer.state_dict()` and
    `outer_optimizer.state_dict()` (plus the model) and loads them back.  torch's load_state_dict replaces the
    state tensors -- which here are views of the flat buffers -- by fresh ones; the wrapper's hooks copy them
    back, so a resumed run continues with the same bits as the uninterrupted one."""
    z = np.load(os.path.join(golden_dir, "inner.npz"))
    zo = np.load(os.path.join(golden_dir, "outer_w1.npz"))

    def grads(params, s):
        off = 0
        for p in params:
            p.grad = torch.from_numpy(z[f"grad_s{s}"][off:off + p.numel()].copy()).reshape(p.shape).cuda()
            off += p.numel()

    def drift(params, r):
        off = 0
        with torch.no_grad():
            for p in params:
                p.add_(torch.from_numpy(zo[f"local_r{r}_w0"][off:off + p.numel()] * 0.01).reshape(p.shape).cuda())
                off += p.numel()

    def run(d, model, first_step, steps):
        for s in range(first_step, first_step + steps):
            grads(list(model.parameters()), s)
            d.inner_step()
        drift(list(model.parameters()), 1)
        d.outer_step()

    model, inner, outer, d = _make(z, warmup_steps=2, total_steps=10)
    run(d, model, 0, 3)
    import copy
    # state_dict() hands out references (as in stock torch); a checkpoint is a copy of them
    ckpt = copy.deepcopy({"model": model.state_dict(), "inner": inner.state_dict(), "outer": outer.state_dict(),
                          "diloco": d.state_dict()})
    assert len(ckpt["outer"]["state"]) == len(list(model.parameters()))
    run(d, model, 3, 3)                                                   # the uninterrupted run goes on
    want = _flat(model.parameters()).clone()
    want_m = _flat([inner.state[p]["exp_avg"] for p in model.parameters()]).clone()
    want_b = _flat([outer.state[p]["momentum_buffer"] for p in model.parameters()]).clone()

    model2, inner2, outer2, d2 = _make(z, warmup_steps=2, total_steps=10)
    ptrs = [p.data_ptr() for p in model2.parameters()]
    model2.load_state_dict(ckpt["model"])
    inner2.load_state_dict(ckpt["inner"])
    outer2.load_state_dict(ckpt["outer"])
    d2.load_state_dict(ckpt["diloco"])                                    # theta_sync, scheduler, step counters
    assert ptrs == [p.data_ptr() for p in model2.parameters()]            # still living in the flat buffer
    assert float(inner2.state[next(iter(model2.parameters()))]["step"]) == 3
    run(d2, model2, 3, 3)
    assert torch.equal(_flat(model2.parameters()), want)
    assert torch.equal(_flat([inner2.state[p]["exp_avg"] for p in model2.parameters()]), want_m)
    assert torch.equal(_flat([outer2.state[p]["momentum_buffer"] for p in model2.parameters()]), want_b)


def test_frozen_parameters_are_left_alone_by_the_inner_step():
    """torch's AdamW skips parameters without a gradient: a frozen tensor must not even be weight-decayed."""
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(16, 16), torch.nn.Linear(16, 16), torch.nn.Linear(16, 4)).cuda()
    for p in model[1].parameters():
        p.requires_grad_(False)
    ref = [p.detach().clone() for p in model.parameters()]
    inner = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.1)
    outer = torch.optim.SGD(model.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(model, inner, outer, warmup_steps=0, total_steps=20)
    x = torch.randn(8, 16, device="cuda")
    for _ in range(3):
        d(x).square().mean().backward()
        d.inner_step()
    ps = list(model.parameters())
    assert torch.equal(ps[2], ref[2]) and torch.equal(ps[3], ref[3])          # frozen: bit-identical
    assert not torch.equal(ps[0], ref[0]) and not torch.equal(ps[4], ref[4])  # trainable: moved
    d.outer_step()
    assert torch.equal(ps[2], ref[2])                                         # zero pseudo-gradient
