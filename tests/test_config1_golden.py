"""BASELINE.json configs[0] -- tiny Llama (57 tensors, 9,766,528 parameters), 2 gloo workers, H = 10 -- pinned by the
reference itself: tests/golden/config1_tiny_llama_w2.npz holds what the reference's `Diloco` class computed
around the real `LlamaForCausalLM` on synthetic tokens (tests/golden/make_config1.py), as SHA-256 of the full
arrays plus a ~49 K-element sample of every array.  The path is elementwise, so an implementation run on the
sampled inputs must give the sampled outputs.  Checked here: the C oracle and the torch-CPU port (CPU); the
CUDA kernels and the drop-in wrapper (`-m gpu`); and, where the reference is mounted, that the fixture
regenerates bit for bit from it."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "config1_tiny_llama_w2.npz")
REF = "/root/reference/nanodiloco/diloco/diloco.py"
W, ROUNDS, RECORDED = 2, 3, 4


@pytest.fixture(scope="module")
def z():
    return np.load(FIX)


def test_fixture_is_what_the_generator_describes(z):
    assert int(z["numels"].sum()) == 9_766_528 and z["numels"].size == 57      # SURVEY section 8, "tiny"
    assert int(z["oracle_full_match"]) == 1          # oracle == reference on the FULL arrays at generation time
    n = z["sample_idx"].size
    assert n > 40_000 and np.all(np.diff(z["sample_idx"]) > 0) and z["sample_idx"][-1] < 9_766_528
    for r in range(ROUNDS):
        assert z[f"params_r{r}"].shape == (n,) and len(str(z[f"sha_params_r{r}"])) == 64
        assert not np.array_equal(z[f"local_r{r}_w0"], z[f"local_r{r}_w1"])    # the two workers really drifted apart


def test_oracle_outer_step_bit_exact_on_config1(z):
    om, ob = z["init"].copy(), np.zeros_like(z["init"])
    for r in range(ROUNDS):
        O.outer_step(om, [z[f"local_r{r}_w{w}"].copy() for w in range(W)], ob, lr=0.7, momentum=0.9,
                     first_step=(r == 0))
        assert np.array_equal(om, z[f"params_r{r}"]), r
        assert np.array_equal(ob, z[f"mom_r{r}"]), r


def test_oracle_inner_steps_on_real_llama_gradients(z):
    """clip_grad_norm_ + AdamW + cosine warm-up on the gradients the tiny Llama actually produced.  Moments
    bit-exact; parameters within 2 ulp of the largest of old value, new value and update (torch's CPU sqrt is not
    correctly rounded, see test_oracle_golden.py; where the update nearly cancels the parameter that one ulp of
    the operands is several ulps of the small result)."""
    p, m, v = z["init"].copy(), np.zeros_like(z["init"]), np.zeros_like(z["init"])
    for s in range(RECORDED):
        lr = float(z[f"lr_s{s}"])
        assert lr == pytest.approx(O.cosine_lr(4e-4, s, warmup=5, total=30), rel=1e-15, abs=0)
        O.adamw_step(p, z[f"grad_s{s}"], m, v, lr=lr, step=s + 1, clip=O.clip_coef(float(z[f"total_norm_s{s}"])),
                     recipe="cpu")
        assert np.array_equal(m, z[f"exp_avg_s{s}"]) and np.array_equal(v, z[f"exp_avg_sq_s{s}"]), s
        want = z[f"inner_params_s{s}"]
        prev = z["init"] if s == 0 else z[f"inner_params_s{s - 1}"]
        ulp = np.spacing(np.maximum(np.maximum(np.abs(prev), np.abs(want)), np.abs(want - prev)))
        assert np.all(np.abs(p - want) <= 2 * ulp), s      # sqrt off by one ulp, then one more rounding in the division
        assert (p != want).mean() < 2e-3, s
        p = want.copy()


def _port_worker(rank, world, store, fix, result):
    import torch.distributed as dist
    from oracle.torch_reference import ReferenceDilocoPort
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=rank, world_size=world)
    z = np.load(fix)
    p = torch.nn.Parameter(torch.from_numpy(z["init"].copy()))
    inner = torch.optim.AdamW([p], lr=4e-4)
    outer = torch.optim.SGD([p], lr=0.7, momentum=0.9, nesterov=True)
    d = ReferenceDilocoPort([p], inner, outer, 5, 30)
    ok = True
    for r in range(ROUNDS):
        with torch.no_grad():
            p.copy_(torch.from_numpy(z[f"local_r{r}_w{rank}"]))
        d.outer_step()
        ok &= np.array_equal(p.detach().numpy(), z[f"params_r{r}"])
        ok &= np.array_equal(outer.state[p]["momentum_buffer"].numpy(), z[f"mom_r{r}"])
    open(f"{result}.{rank}", "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


def test_torch_port_under_gloo_bit_exact_on_config1(tmp_path):
    """The statement-by-statement port (what bench.py times as the CPU baseline), 2 gloo workers."""
    import torch.multiprocessing as mp
    mp.spawn(_port_worker, args=(W, str(tmp_path / "store"), FIX, str(tmp_path / "res")), nprocs=W, join=True)
    assert [open(str(tmp_path / f"res.{r}")).read() for r in range(W)] == ["ok"] * W


@pytest.mark.skipif(not os.path.exists(REF), reason="the reference is only mounted in the build container")
def test_config1_fixture_regenerates_bit_for_bit_from_the_reference(tmp_path):
    """Runs the REAL reference class around the real tiny Llama again (2 gloo workers, ~30 s) and compares every
    array and every SHA-256 with the committed fixture."""
    sys.path.insert(0, os.path.join(HERE, "golden"))
    import make_config1 as mc
    fresh = np.load(mc.generate(str(tmp_path / "c1.npz")))
    kept = np.load(FIX)
    assert sorted(fresh.files) == sorted(kept.files)
    for k in kept.files:
        assert np.array_equal(fresh[k], kept[k]), k


# ---- the CUDA path on the same vectors ---------------------------------------------------------------------
@pytest.mark.gpu
def test_cuda_outer_kernel_bit_exact_on_config1(z):
    from nanodiloco_b200 import _lib
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    master, mom = dev(z["init"]), torch.zeros(z["init"].size, device="cuda")
    for r in range(ROUNDS):
        locs = [dev(z[f"local_r{r}_w{w}"]) for w in range(W)]
        _lib.outer_fused(master, mom, locs, lr=0.7, momentum=0.9, first_step=(r == 0))
        torch.cuda.synchronize()
        assert np.array_equal(master.cpu().numpy(), z[f"params_r{r}"]), r          # bit-exact vs the reference
        assert np.array_equal(mom.cpu().numpy(), z[f"mom_r{r}"]), r
        assert all(torch.equal(l, master) for l in locs)                           # diloco.py:50: both workers rewound
    # the two-kernel form the NCCL pipeline uses: K1 per worker, sum, K2 on the "shard"
    master, mom = dev(z["init"]), torch.zeros(z["init"].size, device="cuda")
    for r in range(ROUNDS):
        pgs = []
        for w in range(W):
            pg = torch.empty_like(master)
            _lib.pseudo_grad(master, dev(z[f"local_r{r}_w{w}"]), pg)
            pgs.append(pg)
        out = torch.empty_like(master)
        _lib.outer_nesterov(master, mom, pgs[0] + pgs[1], out, world=2.0, lr=0.7, momentum=0.9, first_step=(r == 0))
        assert np.array_equal(master.cpu().numpy(), z[f"params_r{r}"]) and torch.equal(out, master), r
        assert np.array_equal(mom.cpu().numpy(), z[f"mom_r{r}"]), r


@pytest.mark.gpu
@pytest.mark.parametrize("grad_buffer", ["autograd", "flat"])
def test_cuda_wrapper_inner_steps_on_real_llama_gradients(z, grad_buffer):
    """The drop-in wrapper stepping on the recorded Llama gradients.  The clip coefficient needs the norm of the
    WHOLE model's gradient, of which the fixture holds a sample: one extra one-element parameter carries the
    rest of the squared norm, so that the kernel sees the recorded total."""
    from nanodiloco_b200 import Diloco
    n = z["init"].size
    body = torch.nn.Parameter(torch.from_numpy(z["init"].copy()).cuda())
    pad = torch.nn.Parameter(torch.zeros(1, device="cuda"))
    holder = torch.nn.ParameterList([body, pad])
    inner = torch.optim.AdamW(holder.parameters(), lr=4e-4)
    outer = torch.optim.SGD(holder.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    d = Diloco(holder, inner, outer, 5, 30, grad_buffer=grad_buffer)
    for s in range(RECORDED):
        assert inner.param_groups[0]["lr"] == pytest.approx(float(z[f"lr_s{s}"]), rel=1e-12, abs=0)
        g = z[f"grad_s{s}"]
        rest = max(0.0, float(z[f"total_norm_s{s}"]) ** 2 - float(np.sum(g.astype(np.float64) ** 2)))
        gb, gp = torch.from_numpy(g.copy()).cuda(), torch.tensor([rest ** 0.5], device="cuda")
        if grad_buffer == "flat":
            body.grad.copy_(gb); pad.grad.copy_(gp)
        else:
            body.grad, pad.grad = gb, gp
        d.inner_step()
        assert float(d.last_grad_norm) == pytest.approx(float(z[f"total_norm_s{s}"]), rel=2e-6)
        for got, key in ((body, "inner_params"), (inner.state[body]["exp_avg"], "exp_avg"),
                         (inner.state[body]["exp_avg_sq"], "exp_avg_sq")):
            want = z[f"{key}_s{s}"]
            np.testing.assert_allclose(got.detach().cpu().numpy(), want, rtol=2e-6, atol=1e-6 * np.abs(want).max())
