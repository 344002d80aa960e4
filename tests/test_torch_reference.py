"""CPU: pins oracle/torch_reference.py (the port bench.py times as the CPU baseline) against
(a) the REAL reference class when /root/reference is mounted (build container only) and
(b) the committed golden fixtures (always)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle.torch_reference import CONFIGS, ReferenceDilocoPort, llama_shapes

REF = "/root/reference/nanodiloco/diloco/diloco.py"
SHAPES = [(128,), (64, 32), (1000,), (7,), (3, 5, 11), (1,), (2048,)]


def _params(init):
    ps, off = [], 0
    for s in SHAPES:
        n = int(np.prod(s))
        ps.append(torch.nn.Parameter(torch.from_numpy(init[off:off + n].copy()).reshape(s)))
        off += n
    return ps


def _flat(ts):
    return torch.cat([t.detach().reshape(-1) for t in ts]).numpy()


def _set(ps, flat):
    off = 0
    with torch.no_grad():
        for p in ps:
            p.copy_(torch.from_numpy(flat[off:off + p.numel()]).reshape(p.shape))
            off += p.numel()


def test_port_matches_golden_w1(golden_dir):
    z = np.load(os.path.join(golden_dir, "outer_w1.npz"))
    ps = _params(z["init"])
    inner = torch.optim.AdamW(ps, lr=4e-4)
    outer = torch.optim.SGD(ps, lr=0.7, momentum=0.9, nesterov=True)
    d = ReferenceDilocoPort(ps, inner, outer, 2, 100)
    for r in range(3):
        _set(ps, z[f"local_r{r}_w0"])
        d.outer_step()
        assert np.array_equal(_flat(ps), z[f"params_r{r}"])
        assert np.array_equal(_flat([outer.state[p]["momentum_buffer"] for p in ps]), z[f"mom_r{r}"])
        assert all(p.grad is None for p in ps)


def test_port_inner_matches_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "inner.npz"))
    ps = _params(z["init"])
    inner = torch.optim.AdamW(ps, lr=4e-4)
    outer = torch.optim.SGD(ps, lr=0.7, momentum=0.9, nesterov=True)
    d = ReferenceDilocoPort(ps, inner, outer, 2, 10)
    for s in range(6):
        off = 0
        for p in ps:
            p.grad = torch.from_numpy(z[f"grad_s{s}"][off:off + p.numel()].copy()).reshape(p.shape)
            off += p.numel()
        d.inner_step()
        assert np.array_equal(_flat(ps), z[f"params_s{s}"])
        assert np.array_equal(_flat([inner.state[p]["exp_avg_sq"] for p in ps]), z[f"exp_avg_sq_s{s}"])


@pytest.mark.skipif(not os.path.exists(REF), reason="the reference is only mounted in the build container")
def test_port_is_bit_identical_to_the_real_reference():
    import torch.distributed as dist
    import tempfile
    spec = importlib.util.spec_from_file_location("ref_diloco", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    with tempfile.TemporaryDirectory() as td:
        dist.init_process_group("gloo", init_method=f"file://{td}/s", rank=0, world_size=1)
        try:
            g = torch.Generator().manual_seed(1337)
            init = (torch.randn(sum(int(np.prod(s)) for s in SHAPES), generator=g) * 0.02).numpy()

            class Bag(torch.nn.Module):
                def __init__(self):
                    super().__init__()
                    self.ps = torch.nn.ParameterList(_params(init))

            a = Bag()
            ai = torch.optim.AdamW(a.parameters(), lr=4e-4)
            ao = torch.optim.SGD(a.parameters(), lr=0.7, momentum=0.9, nesterov=True)
            ref = mod.Diloco(a, ai, ao, 2, 20)
            bp = _params(init)
            bi = torch.optim.AdamW(bp, lr=4e-4)
            bo = torch.optim.SGD(bp, lr=0.7, momentum=0.9, nesterov=True)
            port = ReferenceDilocoPort(bp, bi, bo, 2, 20)
            for step in range(1, 9):
                for i, (p, q) in enumerate(zip(a.parameters(), bp)):
                    gg = torch.Generator().manual_seed(step * 100 + i)
                    grad = torch.randn(p.shape, generator=gg) * (0.05 if step % 2 else 0.002)
                    p.grad, q.grad = grad.clone(), grad.clone()
                ref.inner_step(); port.inner_step()
                if step % 4 == 0:
                    ref.outer_step(); port.outer_step()
                assert np.array_equal(_flat(a.parameters()), _flat(bp)), step
        finally:
            dist.destroy_process_group()


def test_llama_shape_tables():
    """tensor counts and element counts of the BASELINE model sizes."""
    for name, (tensors, elems) in {"tiny": (57, 9_766_528), "llama150m": (111, 214_983_680),
                                   "llama1b": (201, 1_261_529_088)}.items():
        shapes = llama_shapes(**CONFIGS[name])
        assert len(shapes) == tensors
        assert sum(int(np.prod(s)) for s in shapes) == elems
        from nanodiloco_b200.workloads import workload_shapes       # the product's own copy of the table
        assert workload_shapes(name) == shapes


def test_shape_table_matches_the_real_model():
    transformers = pytest.importorskip("transformers")
    cfg = transformers.LlamaConfig(hidden_size=128, intermediate_size=512, num_attention_heads=4,
                                   num_hidden_layers=6, rms_norm_eps=1e-5, use_cache=False)   # configs/llama_default.json
    with torch.device("meta"):
        model = transformers.LlamaForCausalLM(cfg)
    got = [tuple(p.shape) for p in model.parameters()]
    assert sorted(got) == sorted(llama_shapes(**CONFIGS["tiny"]))
    assert sum(int(np.prod(s)) for s in got) == 9_766_528


@pytest.mark.skipif(not os.path.exists(REF), reason="the reference is only mounted in the build container")
@pytest.mark.parametrize("W", [1, 2])
def test_committed_fixtures_regenerate_bit_for_bit_from_the_reference(tmp_path, W):
    """tests/golden/*.npz are outputs of the reference itself: re-run the generator (which drives the
    reference's own `Diloco` class under gloo) and compare with the committed files."""
    import sys
    import torch.multiprocessing as mp
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "golden"))    # spawned workers import the generator by name
    import make_golden as mg
    out = str(tmp_path / f"outer_w{W}.npz")
    mp.spawn(mg.outer_worker, args=(W, str(tmp_path / "store"), out), nprocs=W, join=True)
    fresh, kept = np.load(out), np.load(os.path.join(here, "golden", f"outer_w{W}.npz"))
    assert sorted(fresh.files) == sorted(kept.files)
    for k in kept.files:
        assert np.array_equal(fresh[k], kept[k]), k
