"""CPU: host-side logic of the drop-in wrapper (no kernels run)."""
import math
import os
import sys
import re

import pytest
import torch

from nanodiloco_b200 import _lib
from nanodiloco_b200.diloco import Diloco, _cosine_schedule_with_warmup
from nanodiloco_b200.flat import FlatLayout, rehome_
from nanodiloco_b200.sync import shard_index_map
from nanodiloco_b200.training_utils import create_run_name, set_seed_all, synthetic_token_batches


def test_flat_layout_alignment_and_views():
    ts = [torch.randn(s) for s in [(128,), (7,), (3, 5, 11), (1,), (64, 32)]]
    lay = FlatLayout.plan(ts, quantum=8)
    assert all(o % 4 == 0 for o in lay.offsets)
    assert lay.total % 8 == 0 and lay.total >= sum(t.numel() for t in ts)
    flat = torch.zeros(lay.total)
    ps = [torch.nn.Parameter(t.clone()) for t in ts]
    ids = [id(p) for p in ps]
    rehome_(ps, flat, lay)
    for p, t, v in zip(ps, ts, lay.views(flat)):
        assert torch.equal(p.data, t) and p.data_ptr() == v.data_ptr() and p.shape == t.shape
    assert ids == [id(p) for p in ps]          # Parameter identity survives
    flat.mul_(2.0)                              # in-place updates of the flat buffer are seen by the params
    assert torch.equal(ps[0].data, ts[0] * 2)
    # padding between tensors stays zero
    used = torch.zeros(lay.total, dtype=torch.bool)
    for o, n in zip(lay.offsets, lay.numels):
        used[o:o + n] = True
    assert float(flat[~used].abs().sum()) == 0.0


@pytest.mark.parametrize("world,bucket,n", [(2, 8, 40), (4, 32, 80), (8, 64, 64 * 5 + 32), (2, 1000, 16)])
def test_shard_index_map_partitions_the_space(world, bucket, n):
    seen = torch.zeros(n, dtype=torch.int64)
    sizes = set()
    for r in range(world):
        idx = shard_index_map(n, world, r, bucket)
        sizes.add(idx.numel())
        seen[idx] += 1
        assert torch.all(idx[1:] > idx[:-1])    # shard order == ascending global order
    assert torch.all(seen == 1)                 # every element owned exactly once
    assert sizes == {n // world}


@pytest.mark.parametrize("world,bucket,n,n_nvls", [(2, 8, 40, 24), (4, 32, 160, 96), (8, 64, 64 * 7, 64 * 3),
                                                    (8, 64, 640, 0), (4, 16, 64, 64)])
def test_hybrid_index_map_partitions_the_space(world, bucket, n, n_nvls):
    from nanodiloco_b200.sync import hybrid_index_map
    seen = torch.zeros(n, dtype=torch.int64)
    for r in range(world):
        idx = hybrid_index_map(n, n_nvls, world, r, bucket)
        assert idx.numel() == n // world
        seen[idx] += 1
    assert torch.all(seen == 1)


def test_trainable_runs_skip_frozen_tensors():
    from nanodiloco_b200.flat import trainable_runs
    ps = [torch.nn.Parameter(torch.zeros(n)) for n in (8, 5, 16, 4, 12, 7)]
    ps[1].requires_grad_(False)
    ps[5].requires_grad_(False)
    lay = FlatLayout.plan(ps)
    runs = trainable_runs(lay, [ps[:4], ps[4:]])
    # group 0: tensor 0 alone, then tensors 2..3; group 1: tensor 4 alone
    assert runs == [(0, lay.offsets[0], lay.offsets[0] + 8), (0, lay.offsets[2], lay.offsets[3] + 4),
                    (1, lay.offsets[4], lay.offsets[4] + 12)]
    for p in ps:
        p.requires_grad_(True)
    assert trainable_runs(lay, [ps]) == [(0, 0, lay.offsets[5] + 7)]


def test_cosine_schedule_matches_transformers():
    transformers = pytest.importorskip("transformers")
    a = torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))], lr=4e-4)
    b = torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))], lr=4e-4)
    sa = _cosine_schedule_with_warmup(a, 3, 20)
    sb = transformers.get_cosine_schedule_with_warmup(b, num_warmup_steps=3, num_training_steps=20)
    for _ in range(25):
        assert a.param_groups[0]["lr"] == b.param_groups[0]["lr"]
        a.step(); b.step(); sa.step(); sb.step()


def test_run_name_format():
    name = create_run_name("nanodiloco", {"nodes": 2, "location": "modal"}, is_debug=True)
    assert re.fullmatch(r"debug_nanodiloco_n2_modal_\d{4}_\d{4}_[0-9a-f]{8}", name)
    assert re.fullmatch(r"x_\d{4}_\d{4}_[0-9a-f]{8}", create_run_name("x", {}))


def test_synthetic_batches_are_seeded_and_shaped():
    a = list(synthetic_token_batches(100, 2, 8, seed=7, device="cpu", steps=3))
    b = list(synthetic_token_batches(100, 2, 8, seed=7, device="cpu", steps=3))
    assert len(a) == 3 and a[0]["input_ids"].shape == (2, 8)
    assert all(torch.equal(x["input_ids"], y["input_ids"]) for x, y in zip(a, b))
    assert torch.equal(a[0]["labels"], a[0]["input_ids"])       # main.py:87
    set_seed_all(1)


def test_wrapper_refuses_cpu_parameters():
    """No CPU fallback: the product path fails loudly without CUDA tensors."""
    m = torch.nn.Linear(4, 4)
    inner = torch.optim.AdamW(m.parameters(), lr=1e-3)
    outer = torch.optim.SGD(m.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    with pytest.raises(_lib.NDError, match="no CPU"):
        Diloco(m, inner, outer, 2, 10)


def test_wrapper_checks_optimizer_param_order():
    m = torch.nn.Linear(4, 4)
    inner = torch.optim.AdamW(list(m.parameters())[::-1], lr=1e-3)
    outer = torch.optim.SGD(m.parameters(), lr=0.7, momentum=0.9, nesterov=True)
    with pytest.raises(ValueError, match="same order"):
        Diloco(m, inner, outer, 2, 10)


def test_tensor_wrappers_refuse_cpu_tensors():
    x = torch.zeros(8)
    with pytest.raises(_lib.NDError, match="CUDA"):
        _lib.pseudo_grad(x, x, x)


def test_product_never_imports_the_oracle():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "nanodiloco_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "nd_oracle" not in src and "libnd_oracle" not in src, f


def test_reference_import_names_resolve_to_this_package():
    """`from nanodiloco import Diloco, main`, `nanodiloco.diloco`, `nanodiloco.training_utils`, `python -m nanodiloco`
    and `-m nanodiloco.main` (reference nanodiloco/__init__.py:1-4, __main__.py:1-4, scripts/train_modal.py:61)."""
    import subprocess
    import nanodiloco
    import nanodiloco_b200
    from nanodiloco import Diloco, main
    from nanodiloco.diloco import Diloco as D2
    from nanodiloco.diloco.diloco import Diloco as D3
    from nanodiloco.training_utils import (create_run_name, ddp_setup, get_tokenized_dataset, get_tokenizer,  # noqa: F401
                                           set_seed_all)
    assert Diloco is nanodiloco_b200.Diloco is D2 is D3
    assert callable(main) and callable(nanodiloco.main)            # the function, as in the reference, not the module
    assert sorted(nanodiloco.__all__) == ["Diloco", "main"]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for mod in ("nanodiloco", "nanodiloco.main"):
        out = subprocess.run([sys.executable, "-m", mod, "--help"], capture_output=True, text=True, cwd=root, timeout=300)
        assert out.returncode == 0 and "--inner-steps" in out.stdout and "--llama-config-file" in out.stdout, out.stderr[-500:]
