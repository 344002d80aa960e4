"""The training loop that calls the hot path -- drop-in for the reference's nanodiloco/main.py.

Same command-line flags and defaults as the reference's cyclopts command (main.py:42-56) and the
same loop body (main.py:106-127), including its quirks (kept on purpose, they are upstream of the
path this package accelerates):
  * `loss = output.loss / gas` is computed for logging but `output.loss.backward()` is what runs
    (main.py:110-111), so accumulated gradients are a sum, not a mean;
  * `gradient_accumulation_steps` is a float (main.py:66);
  * the loop runs until the data is exhausted; `total_steps` only shapes the LR schedule.
What is new: `--synthetic-tokens` (default on: every BASELINE config uses synthetic tokens and the
reference's dataset/tokenizer need a network volume), `--max-steps`, `--backend`, `--sync`, and a
tokens/s + sync-time report.  wandb is optional and off unless `--wandb` is given.

    torchrun --nproc-per-node 8 -m nanodiloco_b200.main --llama-config-file cfg.json --inner-steps 500
"""
from __future__ import annotations

import argparse
import json
import os
import time
from typing import Callable, Optional

import torch
import torch.distributed as dist

from .diloco import Diloco
from .training_utils import create_run_name, ddp_setup, set_seed_all, synthetic_token_batches


def get_default_llama_config():
    """main.py:16-27"""
    return {"architectures": ["LlamaForCausalLM"], "hidden_size": 128, "intermediate_size": 512,
            "num_attention_heads": 4, "num_hidden_layers": 6, "rms_norm_eps": 1e-05, "use_cache": False}


def get_default_wandb_config():
    """main.py:29-35"""
    return {"nodes": 1, "location": "local", "backend": "nccl", "measure_comms": True}


def load_config_from_file(config_path: str) -> dict:
    with open(config_path, "r") as f:
        return json.load(f)


def train_model(
    seed: int = 1337,
    batch_size: int = 256,
    per_device_batch_size: int = 8,
    seq_length: int = 1024,
    warmup_steps: int = 100,
    total_steps: int = 10_000,
    inner_steps: int = 100,
    lr: float = 4e-4,
    outer_lr: float = 0.7,
    project: str = "nano-diloco",
    dataset_path: str = "/mnt/hf-c4-tiny/datasets/PrimeIntellect/c4-tiny/en/save_to_disk",
    llama_config_file: Optional[str] = None,
    wandb_config_file: Optional[str] = None,
    *,
    synthetic_tokens: bool = True,
    max_steps: Optional[int] = None,
    use_wandb: bool = False,
    sync: str = "auto",
    llama_config: Optional[dict] = None,
    diloco_factory: Optional[Callable] = None,
    log_every: int = 0,
) -> dict:
    """main.py:41-130.  Returns a dict of run statistics (tokens/s, mean outer-sync time, losses)."""
    from transformers import LlamaConfig, LlamaForCausalLM

    llama_config = llama_config or (load_config_from_file(llama_config_file) if llama_config_file
                                    else get_default_llama_config())
    wandb_config = load_config_from_file(wandb_config_file) if wandb_config_file else get_default_wandb_config()

    set_seed_all(seed)                                                   # main.py:60
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))                 # main.py:61-63
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    global_rank = int(os.environ.get("RANK", "0"))

    assert batch_size % per_device_batch_size == 0                       # main.py:65
    gradient_accumulation_steps = batch_size / per_device_batch_size     # a float, as in main.py:66
    outer_steps = total_steps // inner_steps
    assert total_steps % inner_steps == 0                                # main.py:69

    wandb = None
    if use_wandb and global_rank == 0:                                   # main.py:71-73
        import wandb as _wandb
        wandb = _wandb
        wandb.init(project=project, name=create_run_name("nanodiloco", wandb_config), config=wandb_config)
    device = torch.device(f"cuda:{local_rank}") if torch.cuda.is_available() else torch.device("cpu")

    config = LlamaConfig(**llama_config)                                 # main.py:97-99
    model = LlamaForCausalLM(config)
    model.to(device)
    if synthetic_tokens:
        micro_batches = None if max_steps is None else int(max_steps * gradient_accumulation_steps)
        if micro_batches is None:
            micro_batches = int(total_steps * gradient_accumulation_steps)
        batches = synthetic_token_batches(config.vocab_size, per_device_batch_size, seq_length,
                                          seed=7 + global_rank, device=device, steps=micro_batches)
    else:                                                                # main.py:75-96
        from datasets.distributed import split_dataset_by_node          # type: ignore
        from torch.utils.data import DataLoader
        from .training_utils import get_tokenized_dataset, get_tokenizer
        tokenizer = get_tokenizer()
        ds = split_dataset_by_node(get_tokenized_dataset(dataset_path, tokenizer), world_size=world_size,
                                   rank=global_rank)

        def collate_fn(batch):
            padded = tokenizer.pad(batch, padding="longest", max_length=seq_length, pad_to_multiple_of=8,
                                   return_tensors="pt")
            padded["labels"] = padded["input_ids"].clone()
            return padded

        batches = DataLoader(ds, batch_size=per_device_batch_size, collate_fn=collate_fn, drop_last=True,
                             shuffle=True)

    inner_optimizer = torch.optim.AdamW(model.parameters(), lr=lr)      # main.py:100
    outer_optimizer = torch.optim.SGD(model.parameters(), lr=outer_lr, momentum=0.9, nesterov=True)   # main.py:101
    if diloco_factory is None:
        diloco_model = Diloco(model, inner_optimizer, outer_optimizer, warmup_steps, total_steps,
                              inner_steps=inner_steps, outer_steps=outer_steps, sync=sync)        # main.py:102
    else:
        diloco_model = diloco_factory(model, inner_optimizer, outer_optimizer, warmup_steps, total_steps,
                                      inner_steps, outer_steps)
    diloco_model.train()

    losses, n_inner, n_outer, tokens = [], 0, 0, 0
    t_first_outer = None
    tokens_at_first_outer = 0
    sync_events = []
    t_start = time.perf_counter()
    for i, batch in enumerate(batches):                                  # main.py:106-127
        real_step = (i + 1) // gradient_accumulation_steps
        batch = {k: v.to(device) for k, v in batch.items()}
        output = diloco_model(**batch)
        loss = output.loss / gradient_accumulation_steps
        output.loss.backward()
        tokens += batch["input_ids"].numel()
        if (i + 1) % gradient_accumulation_steps == 0:
            diloco_model.inner_step()
            n_inner += 1
            if real_step % inner_steps == 0:
                if device.type == "cuda":
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                diloco_model.outer_step()
                if device.type == "cuda":
                    e1.record()
                    sync_events.append((e0, e1))
                n_outer += 1
                if t_first_outer is None:                                # exclude the first round (warm-up)
                    if device.type == "cuda":
                        torch.cuda.synchronize()
                    t_first_outer, tokens_at_first_outer = time.perf_counter(), tokens
        if log_every and (i + 1) % log_every == 0:
            losses.append(float(loss.detach()))
            if local_rank == 0:
                row = {"loss": losses[-1], "step": real_step,
                       "lr": [g["lr"] for g in diloco_model.inner_optimizer.param_groups][0],
                       "Perplexity": float(torch.exp(loss.detach())), "effective_step": real_step * world_size,
                       "total_samples": real_step * batch_size * world_size}
                if wandb is not None:
                    wandb.log(row)
                else:
                    print(json.dumps(row), flush=True)
        if max_steps is not None and n_inner >= max_steps:
            break
    if device.type == "cuda":
        torch.cuda.synchronize()
    t_end = time.perf_counter()
    if wandb is not None:
        wandb.finish()
    steady = t_first_outer is not None and tokens > tokens_at_first_outer
    wall = (t_end - t_first_outer) if steady else (t_end - t_start)
    toks = (tokens - tokens_at_first_outer) if steady else tokens
    sync_ms = [a.elapsed_time(b) for a, b in sync_events]
    return {
        "tokens_per_s_per_worker": toks / wall, "tokens_per_s": world_size * toks / wall,
        "inner_steps_run": n_inner, "outer_steps_run": n_outer, "wall_s": wall,
        "outer_sync_ms": sync_ms, "excluded_first_round": bool(steady),
        # of which: waiting for the slowest worker to arrive at the sync (skew, not cost of the sync)
        "avg_sync_wait_ms": (lambda w: None if w is None else w * 1e3)(getattr(diloco_model, "avg_sync_wait_time", None)),
        "avg_sync_ms": diloco_model.avg_sync_time * 1e3,
        "params": sum(p.numel() for p in model.parameters()), "losses": losses,
        "final_loss": float(loss.detach()),
    }


def _parse(argv=None):
    ap = argparse.ArgumentParser(prog="nanodiloco")
    d = train_model.__defaults__
    names = train_model.__code__.co_varnames[:len(d)]
    for name, default in zip(names, d):
        flag = "--" + name.replace("_", "-")
        ap.add_argument(flag, type=(str if default is None else type(default)), default=default)
    ap.add_argument("--synthetic-tokens", action=argparse.BooleanOptionalAction, default=True)
    ap.add_argument("--max-steps", type=int, default=None)
    ap.add_argument("--wandb", dest="use_wandb", action="store_true")
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--sync", default="auto", choices=["auto", "p2p", "nccl"])
    ap.add_argument("--log-every", type=int, default=0)
    return ap.parse_args(argv)


def main(argv=None):
    """main.py:133-137"""
    print("Training Diloco with NanoDiloco (nanodiloco_b200)...")
    args = vars(_parse(argv))
    backend = args.pop("backend")
    distributed = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    if distributed:
        ddp_setup(backend)
    stats = train_model(**args)
    if int(os.environ.get("RANK", "0")) == 0:
        stats.pop("losses", None)
        print(json.dumps(stats))
    print("Training completed!")
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
