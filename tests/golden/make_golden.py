"""Generate the golden fixtures in tests/golden/ from the REFERENCE ITSELF.

Runs only in the build container, where /root/reference exists.  It loads the
reference's `Diloco` class by file path (`import nanodiloco` needs cyclopts,
which is not installed) and drives it, unmodified, under a gloo process group on
CPU -- the reference's own CPU-runnable configuration (BASELINE.json configs[0]).

    python tests/golden/make_golden.py            # writes tests/golden/*.npz

Fixtures (all float32, flat concatenation of the tensors in `shapes` order):
  outer_w{W}.npz  W in {1,2,4,8}; R=3 outer rounds
      init                 params every worker starts from (seed 1337, main.py:43)
      local_r{r}_w{w}      worker w's params just before outer round r (inputs)
      params_r{r}          every worker's params after round r  (outputs)
      mom_r{r}             outer SGD momentum_buffer after round r (outputs)
      ranks_equal          1 if all ranks ended bit-identical in every round
  inner.npz       W=1; 6 inner steps of clip_grad_norm_(1.0) + AdamW + cosine LR
      init, grad_s{s}, total_norm_s{s}, params_s{s}, exp_avg_s{s}, exp_avg_sq_s{s}, lr_s{s}
"""
from __future__ import annotations

import importlib.util
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REF = "/root/reference/nanodiloco/diloco/diloco.py"
HERE = os.path.dirname(os.path.abspath(__file__))

# ragged on purpose: sizes that are not multiples of 4 exercise the kernels' tails
SHAPES = [(128,), (64, 32), (1000,), (7,), (3, 5, 11), (1,), (2048,)]
ROUNDS = 3
OUTER_LR, OUTER_MU = 0.7, 0.9          # main.py:51, main.py:101
INNER_LR = 4e-4                        # main.py:50
INNER_STEPS_GOLDEN = 6


def load_reference_diloco():
    spec = importlib.util.spec_from_file_location("ref_diloco", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.Diloco


class Bag(torch.nn.Module):
    """A parameter container standing in for the model: the outer step only
    ever sees `model.parameters()` (diloco.py:21, main.py:100-101)."""

    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(1337)
        self.ps = torch.nn.ParameterList(
            [torch.nn.Parameter(torch.randn(s, generator=g) * 0.02) for s in SHAPES])


def flat(tensors) -> np.ndarray:
    return torch.cat([t.detach().reshape(-1).float() for t in tensors]).numpy().copy()



def outer_worker(rank: int, world: int, store_path: str, out_path: str):
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store_path}", rank=rank, world_size=world)
    Diloco = load_reference_diloco()
    model = Bag()
    inner = torch.optim.AdamW(model.parameters(), lr=INNER_LR)
    outer = torch.optim.SGD(model.parameters(), lr=OUTER_LR, momentum=OUTER_MU, nesterov=True)
    d = Diloco(model, inner, outer, warmup_steps=2, total_steps=100)
    out = {"init": flat(model.parameters()), "shapes": np.array([int(np.prod(s)) for s in SHAPES])}
    ranks_equal = 1
    for r in range(ROUNDS):
        with torch.no_grad():
            for i, p in enumerate(model.parameters()):
                g = torch.Generator().manual_seed(1000 * r + 100 + rank + 7919 * i)
                p.add_(torch.randn(p.shape, generator=g) * 1e-3)
        mine = torch.from_numpy(flat(model.parameters()))
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        for w in range(world):
            out[f"local_r{r}_w{w}"] = gathered[w].numpy().copy()
        d.outer_step()
        assert all(p.grad is None for p in model.parameters())
        after = torch.from_numpy(flat(model.parameters()))
        gathered = [torch.empty_like(after) for _ in range(world)]
        dist.all_gather(gathered, after)
        if not all(torch.equal(gathered[0], x) for x in gathered):
            ranks_equal = 0
        out[f"params_r{r}"] = after.numpy().copy()
        out[f"mom_r{r}"] = flat([outer.state[p]["momentum_buffer"] for p in model.parameters()])
        # host snapshot the reference keeps (diloco.py:54) equals the params
        assert np.array_equal(flat(d.offloaded_last_sync_parameters), out[f"params_r{r}"])
    out["ranks_equal"] = np.array(ranks_equal)
    if rank == 0:
        np.savez_compressed(out_path, **out)
    dist.barrier()
    dist.destroy_process_group()


def inner_worker(rank: int, world: int, store_path: str, out_path: str):
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store_path}", rank=rank, world_size=world)
    Diloco = load_reference_diloco()
    model = Bag()
    inner = torch.optim.AdamW(model.parameters(), lr=INNER_LR)
    outer = torch.optim.SGD(model.parameters(), lr=OUTER_LR, momentum=OUTER_MU, nesterov=True)
    d = Diloco(model, inner, outer, warmup_steps=2, total_steps=10)
    out = {"init": flat(model.parameters()), "shapes": np.array([int(np.prod(s)) for s in SHAPES])}
    for s in range(INNER_STEPS_GOLDEN):
        # alternate between gradients whose norm is above and below max_norm=1.0
        scale = 0.05 if s % 2 == 0 else 0.002
        for i, p in enumerate(model.parameters()):
            g = torch.Generator().manual_seed(5000 + 31 * s + i)
            p.grad = torch.randn(p.shape, generator=g) * scale
        out[f"grad_s{s}"] = flat([p.grad for p in model.parameters()])
        out[f"lr_s{s}"] = np.array(inner.param_groups[0]["lr"], dtype=np.float64)
        # what clip_grad_norm_ (diloco.py:57) is about to compute; inner_step() discards it
        out[f"total_norm_s{s}"] = torch.nn.utils.get_total_norm(
            [p.grad for p in model.parameters()], 2.0).numpy().copy()
        d.inner_step()
        out[f"params_s{s}"] = flat(model.parameters())
        out[f"exp_avg_s{s}"] = flat([inner.state[p]["exp_avg"] for p in model.parameters()])
        out[f"exp_avg_sq_s{s}"] = flat([inner.state[p]["exp_avg_sq"] for p in model.parameters()])
    np.savez_compressed(out_path, **out)
    dist.destroy_process_group()


def run(fn, world: int, out_path: str):
    with tempfile.TemporaryDirectory() as td:
        mp.spawn(fn, args=(world, os.path.join(td, "store"), out_path), nprocs=world, join=True)


if __name__ == "__main__":
    if not os.path.exists(REF):
        sys.exit("the reference is not mounted here; fixtures can only be generated in the build container")
    only_inner = "--inner-only" in sys.argv
    worlds = (1, 2, 4, 8)
    if "--only-w" in sys.argv:
        worlds = (int(sys.argv[sys.argv.index("--only-w") + 1]),)
    for W in () if only_inner else worlds:
        run(outer_worker, W, os.path.join(HERE, f"outer_w{W}.npz"))
        print("wrote", f"outer_w{W}.npz")
    if "--only-w" not in sys.argv:
        run(inner_worker, 1, os.path.join(HERE, "inner.npz"))
        print("wrote inner.npz")
