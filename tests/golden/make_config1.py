"""Golden fixture for BASELINE.json configs[0] -- the reference's own CPU-runnable case -- made BY THE REFERENCE.

    python tests/golden/make_config1.py            # writes tests/golden/config1_tiny_llama_w2.npz

Runs only in the build container (needs /root/reference).  What runs is the reference's `Diloco`
(nanodiloco/diloco/diloco.py, loaded by file path, unmodified) around `LlamaForCausalLM(LlamaConfig(
**configs/llama_default.json))` exactly as nanodiloco/main.py:97-102 builds it: model seed 1337 (main.py:43),
AdamW lr 4e-4 (main.py:50,100), SGD lr 0.7 momentum 0.9 nesterov (main.py:51,101), 2 workers under gloo,
H = 10 inner steps per outer step, 3 outer rounds, synthetic tokens `randint(0, 32000, (2, 64))` from a
generator seeded 7 + rank, labels = input ids (main.py:87), one micro-batch per step, loss.backward() as in
main.py:110-111.  57 tensors / 9,766,528 parameters.

The whole run is ~470 MB of float32, too much to commit.  Everything on this path is ELEMENTWISE given its
inputs (the only global quantities are the gradient norm, a scalar, and the cross-worker mean, taken
element by element), so the fixture keeps
  * SHA-256 of every full array (inputs and outputs of every outer round), and
  * the same SAMPLE of elements from every array: whole small tensors plus 2,048-element blocks spread
    over every tensor class, ~45 K elements, block starts chosen to straddle tile and chunk boundaries;
an implementation run on the sampled inputs must reproduce the sampled outputs bit for bit.  At
generation time the C oracle is run on the FULL inputs and must equal the reference's full outputs bit for
bit (`oracle_full_match`).

Arrays (float32 unless noted; `[S]` = restricted to the sample):
  sample_idx (int64)             flat element indices of the sample
  numels (int64)                 the 57 tensor sizes, model.parameters() order
  init[S]                        parameters every worker starts from
  local_r{r}_w{w}[S]             worker w's parameters just before outer round r        (inputs)
  params_r{r}[S], mom_r{r}[S]    parameters / SGD momentum_buffer after outer round r   (outputs)
  sha_*                          hex SHA-256 of the corresponding FULL arrays (numpy bytes)
  grad_s{s}[S], total_norm_s{s}, lr_s{s}, inner_params_s{s}[S], exp_avg_s{s}[S], exp_avg_sq_s{s}[S]
                                 worker 0, round 0, inner steps s = 0..3: what clip_grad_norm_ + AdamW saw and did
  loss_r{r}_w{w}                 last loss of the round (float64), for the record
"""
from __future__ import annotations

import hashlib
import importlib.util
import json
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REF = "/root/reference/nanodiloco/diloco/diloco.py"
REF_CFG = "/root/reference/configs/llama_default.json"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "config1_tiny_llama_w2.npz")
WORLD, H, ROUNDS, B, S = 2, 10, 3, 2, 64
BLOCK = 2048
INNER_RECORDED = 4          # inner steps of worker 0, round 0, whose gradients and results are kept


def load_reference_diloco():
    spec = importlib.util.spec_from_file_location("ref_diloco", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.Diloco


def flat(tensors) -> np.ndarray:
    return torch.cat([t.detach().reshape(-1).float() for t in tensors]).numpy().copy()


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sample_indices(numels) -> np.ndarray:
    """Whole tensors up to BLOCK elements; for larger ones BLOCK-element blocks at the start, at the end and at
    an odd interior offset (so that blocks cross the 1,024/2,048/4,096-element tile and 16,384-element chunk
    boundaries of the kernels), for every 8th tensor of a size class and the two embedding matrices."""
    idx, off, seen = [], 0, {}
    for n in numels:
        k = seen.get(n, 0)
        seen[n] = k + 1
        if n <= BLOCK:
            if k % 8 == 0:
                idx.append(np.arange(off, off + n))
        elif k % 8 == 0 or n > 1_000_000:
            starts = [0, n - BLOCK, 16384 - 1024 + 12 if n > 40000 else (n // 3) // 4 * 4 + 4]
            for s0 in sorted(set(max(0, min(s0, n - BLOCK)) for s0 in starts)):
                idx.append(np.arange(off + s0, off + s0 + BLOCK))
        off += n
    return np.unique(np.concatenate(idx)).astype(np.int64)


def worker(rank: int, world: int, store_path: str, out_path: str):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.set_num_threads(1)                      # one thread: the same reduction order on every machine
    dist.init_process_group("gloo", init_method=f"file://{store_path}", rank=rank, world_size=world)
    Diloco = load_reference_diloco()
    torch.manual_seed(1337)                       # set_seed_all(seed) of main.py:60 (training_utils/utils.py)
    np.random.seed(1337)
    cfg = json.load(open(REF_CFG))
    model = LlamaForCausalLM(LlamaConfig(**cfg))                                           # main.py:97-98
    inner = torch.optim.AdamW(model.parameters(), lr=4e-4)                                  # main.py:100
    outer = torch.optim.SGD(model.parameters(), lr=0.7, momentum=0.9, nesterov=True)        # main.py:101
    d = Diloco(model, inner, outer, warmup_steps=5, total_steps=H * ROUNDS, inner_steps=H,
               outer_steps=ROUNDS)                                                          # main.py:102
    params = list(model.parameters())
    numels = np.array([p.numel() for p in params], dtype=np.int64)
    sidx = sample_indices(numels.tolist())
    out = {"numels": numels, "sample_idx": sidx, "meta": np.array(json.dumps(
        {"world": world, "H": H, "rounds": ROUNDS, "batch": B, "seq": S, "token_seed": "7+rank",
         "model_seed": 1337, "torch": torch.__version__, "threads_per_worker": 1, "warmup_steps": 5,
         "total_steps": H * ROUNDS}))}
    full = {}
    init = flat(params)
    out["init"], out["sha_init"] = init[sidx], np.array(sha(init))
    full["init"] = init
    gen = torch.Generator().manual_seed(7 + rank)
    d.train()
    for r in range(ROUNDS):
        for s in range(H):
            ids = torch.randint(0, 32000, (B, S), generator=gen)
            output = d(input_ids=ids, labels=ids.clone())                                    # main.py:87,108
            output.loss.backward()                                                          # main.py:111
            if r == 0 and rank == 0 and s < INNER_RECORDED:
                grads = [p.grad for p in params]
                out[f"grad_s{s}"] = flat(grads)[sidx]
                out[f"total_norm_s{s}"] = torch.nn.utils.get_total_norm(grads, 2.0).numpy().copy()
                out[f"lr_s{s}"] = np.array(inner.param_groups[0]["lr"], dtype=np.float64)
            d.inner_step()                                                                  # main.py:113
            if r == 0 and rank == 0 and s < INNER_RECORDED:
                out[f"inner_params_s{s}"] = flat(params)[sidx]
                out[f"exp_avg_s{s}"] = flat([inner.state[p]["exp_avg"] for p in params])[sidx]
                out[f"exp_avg_sq_s{s}"] = flat([inner.state[p]["exp_avg_sq"] for p in params])[sidx]
        mine = torch.from_numpy(flat(params))
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        losses = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(losses, output.loss.detach().double().reshape(1))
        for w in range(world):
            a = gathered[w].numpy()
            full[f"local_r{r}_w{w}"] = a.copy()
            out[f"local_r{r}_w{w}"], out[f"sha_local_r{r}_w{w}"] = a[sidx], np.array(sha(a))
            out[f"loss_r{r}_w{w}"] = losses[w].numpy().copy()
        d.outer_step()                                                                      # main.py:116
        after = flat(params)
        mom = flat([outer.state[p]["momentum_buffer"] for p in params])
        theirs = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(theirs, torch.from_numpy(after))
        assert all(torch.equal(theirs[0], t) for t in theirs), "workers diverged after the outer step"
        full[f"params_r{r}"], full[f"mom_r{r}"] = after, mom
        out[f"params_r{r}"], out[f"sha_params_r{r}"] = after[sidx], np.array(sha(after))
        out[f"mom_r{r}"], out[f"sha_mom_r{r}"] = mom[sidx], np.array(sha(mom))
    if rank == 0:
        # the C oracle on the FULL inputs must reproduce the reference's FULL outputs bit for bit
        sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
        from oracle import oracle as O
        om, ob = full["init"].copy(), np.zeros_like(full["init"])
        match = 1
        for r in range(ROUNDS):
            O.outer_step(om, [full[f"local_r{r}_w{w}"].copy() for w in range(world)], ob, lr=0.7, momentum=0.9,
                         first_step=(r == 0))
            match &= int(np.array_equal(om, full[f"params_r{r}"]) and np.array_equal(ob, full[f"mom_r{r}"]))
        out["oracle_full_match"] = np.array(match)
        np.savez_compressed(out_path, **out)
        print("sample elements:", sidx.size, " oracle == reference on the full arrays:", bool(match))
    dist.barrier()
    dist.destroy_process_group()


def generate(out_path: str = OUT):
    with tempfile.TemporaryDirectory() as td:
        mp.spawn(worker, args=(WORLD, os.path.join(td, "store"), out_path), nprocs=WORLD, join=True)
    return out_path


if __name__ == "__main__":
    if not os.path.exists(REF):
        sys.exit("the reference is not mounted here; fixtures can only be generated in the build container")
    print("wrote", generate(sys.argv[1] if len(sys.argv) > 1 else OUT))
