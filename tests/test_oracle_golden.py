"""CPU: the oracle (oracle/nd_oracle.c) against the fixtures generated from the REFERENCE's own
`Diloco` class (tests/golden/make_golden.py) and against torch's CPU optimizers, which is where the
reference's arithmetic lives (nanodiloco/main.py:100-101, nanodiloco/diloco/diloco.py:52,57-58)."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

OUTER_LR, OUTER_MU = 0.7, 0.9


def _ulp_diff(a, b):
    ai = a.view(np.int32).astype(np.int64)
    bi = b.view(np.int32).astype(np.int64)
    return np.abs(ai - bi)


@pytest.mark.parametrize("W", [1, 2])
def test_outer_step_bit_exact_vs_reference(golden_dir, W):
    z = np.load(os.path.join(golden_dir, f"outer_w{W}.npz"))
    assert int(z["ranks_equal"]) == 1
    master = z["init"].copy()
    buf = np.zeros_like(master)
    for r in range(3):
        locs = [z[f"local_r{r}_w{w}"].copy() for w in range(W)]
        O.outer_step(master, locs, buf, lr=OUTER_LR, momentum=OUTER_MU, first_step=(r == 0))
        assert np.array_equal(master, z[f"params_r{r}"]), f"params differ in round {r}"
        assert np.array_equal(buf, z[f"mom_r{r}"]), f"momentum differs in round {r}"
        for l in locs:  # diloco.py:50: every worker is rewound to theta_sync
            assert np.array_equal(l, master)


@pytest.mark.parametrize("W", [4, 8])
def test_outer_step_w4_w8_within_tolerance(golden_dir, W):
    """W>=4: gloo's ring sums in a different order than worker order; the north-star tolerance
    (rtol 1e-6 on params and momentum) needs an atol floor because the pseudo-gradient is a
    difference of nearly equal numbers.  atol = 1e-6 * max|value|."""
    z = np.load(os.path.join(golden_dir, f"outer_w{W}.npz"))
    master = z["init"].copy()
    buf = np.zeros_like(master)
    for r in range(3):
        locs = [z[f"local_r{r}_w{w}"].copy() for w in range(W)]
        O.outer_step(master, locs, buf, lr=OUTER_LR, momentum=OUTER_MU, first_step=(r == 0))
        for got, want in ((master, z[f"params_r{r}"]), (buf, z[f"mom_r{r}"])):
            np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6 * np.abs(want).max())


def test_pieces_compose_to_outer_step(golden_dir):
    z = np.load(os.path.join(golden_dir, "outer_w2.npz"))
    master = z["init"].copy()
    buf = np.zeros_like(master)
    locs = [z[f"local_r0_w{w}"] for w in range(2)]
    g = O.average([O.pseudo_grad(master, l) for l in locs])
    O.sgd_step(master, buf, g, lr=OUTER_LR, momentum=OUTER_MU, first_step=True)
    assert np.array_equal(master, z["params_r0"]) and np.array_equal(buf, z["mom_r0"])


@pytest.mark.parametrize("nesterov,dampening,wd", [(True, 0.0, 0.0), (False, 0.0, 0.0), (False, 0.1, 0.0),
                                                   (True, 0.0, 1e-2), (False, 0.3, 5e-3)])
def test_sgd_step_bit_exact_vs_torch_cpu(nesterov, dampening, wd):
    g = torch.Generator().manual_seed(3)
    p = torch.randn(10007, generator=g) * 0.02
    tp = torch.nn.Parameter(p.clone())
    opt = torch.optim.SGD([tp], lr=0.7, momentum=0.9, nesterov=nesterov, dampening=dampening, weight_decay=wd)
    pn, bn = p.numpy().copy(), np.zeros(10007, np.float32)
    for step in range(4):
        grad = torch.randn(10007, generator=g) * 1e-3
        tp.grad = grad.clone()
        opt.step()
        O.sgd_step(pn, bn, grad.numpy().copy(), lr=0.7, momentum=0.9, dampening=dampening, weight_decay=wd,
                   nesterov=nesterov, first_step=(step == 0))
        assert np.array_equal(pn, tp.detach().numpy()), step
        assert np.array_equal(bn, opt.state[tp]["momentum_buffer"].numpy()), step


def test_inner_step_vs_reference(golden_dir):
    """clip + AdamW + cosine LR against the reference's inner_step (diloco.py:56-60) on CPU.

    exp_avg and exp_avg_sq are bit-exact.  params are bit-exact on > 99.8 % of elements and within
    1 ulp elsewhere: torch's CPU sqrt (MKL VML) is not correctly rounded on ~0.6 % of inputs, the
    oracle's sqrtf (like CUDA's sqrt.rn) is."""
    z = np.load(os.path.join(golden_dir, "inner.npz"))
    p = z["init"].copy()
    m = np.zeros_like(p)
    v = np.zeros_like(p)
    for s in range(6):
        g = z[f"grad_s{s}"]
        lr = float(z[f"lr_s{s}"])
        assert lr == pytest.approx(O.cosine_lr(4e-4, s, warmup=2, total=10), rel=1e-15, abs=0)
        tn = float(z[f"total_norm_s{s}"])
        assert np.sqrt(O.sumsq(g)) == pytest.approx(tn, rel=1e-6)
        O.adamw_step(p, g, m, v, lr=lr, step=s + 1, clip=O.clip_coef(tn), recipe="cpu")
        assert np.array_equal(m, z[f"exp_avg_s{s}"]), s
        assert np.array_equal(v, z[f"exp_avg_sq_s{s}"]), s
        want = z[f"params_s{s}"]
        d = _ulp_diff(p, want)
        assert d.max() <= 1 and (d > 0).mean() < 2e-3, (s, d.max(), (d > 0).mean())
        p = want.copy()  # re-pin so that a 1-ulp difference does not compound across steps


def test_cuda_recipe_close_to_cpu_recipe(golden_dir):
    """The two AdamW rounding recipes (torch CPU vs torch CUDA) differ by a few ulp at most."""
    z = np.load(os.path.join(golden_dir, "inner.npz"))
    outs = {}
    for recipe in ("cpu", "cuda"):
        p, m, v = z["init"].copy(), np.zeros_like(z["init"]), np.zeros_like(z["init"])
        for s in range(6):
            O.adamw_step(p, z[f"grad_s{s}"], m, v, lr=float(z[f"lr_s{s}"]), step=s + 1,
                         clip=O.clip_coef(float(z[f"total_norm_s{s}"])), recipe=recipe)
        outs[recipe] = (p, m, v)
    for a, b in zip(outs["cpu"], outs["cuda"]):
        np.testing.assert_allclose(a, b, rtol=2e-6, atol=1e-9)


def test_edge_sizes():
    for n in (0, 1, 3, 5):
        master = np.arange(n, dtype=np.float32)
        locs = [master - 1.0, master + 0.5]
        buf = np.zeros(n, np.float32)
        O.outer_step(master, locs, buf, lr=0.7, first_step=True)
        assert all(np.array_equal(l, master) for l in locs)
    assert O.sumsq(np.zeros(0, np.float32)) == 0.0
