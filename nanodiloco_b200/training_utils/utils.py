"""Helpers the training loop imports next to `Diloco`.

Same names and behaviour as the reference's nanodiloco/training_utils/utils.py:11-60, kept so
the loop stays drop-in; `synthetic_token_batches` is new (every BASELINE config uses synthetic
tokens -- the reference's dataset path needs a network volume).
"""
from __future__ import annotations

import os
import random
import uuid
from datetime import datetime

import numpy as np
import torch
import torch.distributed as dist


def set_seed_all(seed: int = 42) -> None:
    """utils.py:11-15"""
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    torch.cuda.manual_seed_all(seed)


def create_run_name(experiment_type, node_config, is_debug=False) -> str:
    """utils.py:18-39: "{type}_n{nodes}_{location}_{MMDD_HHMM}_{uuid8}", "debug_" first when asked."""
    fields = [experiment_type]
    nodes, location = node_config.get("nodes"), node_config.get("location")
    if nodes:
        fields.append(f"n{nodes}")
    if location:
        fields.append(location)
    fields.append(datetime.now().strftime("%m%d_%H%M"))
    if is_debug:
        fields = ["debug"] + fields
    return "_".join(fields) + "_" + str(uuid.uuid4())[:8]


def ddp_setup(backend: str = "nccl") -> None:
    """utils.py:41-43: one process per GPU, NCCL process group, device = LOCAL_RANK."""
    dist.init_process_group(backend=backend)
    if torch.cuda.is_available():
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))


def get_tokenized_dataset(dataset_path, tokenizer, seq_length=1024):
    """utils.py:45-56 (needs the `datasets` package and a dataset saved on disk)."""
    from datasets import load_from_disk  # type: ignore
    os.environ["HF_DATASETS_OFFLINE"] = "1"
    ds = load_from_disk(dataset_path)
    ds = ds.map(lambda rows: tokenizer(rows["text"], truncation=True, max_length=seq_length),
                batched=True, remove_columns=["text", "timestamp", "url"])
    return ds["train"]


def get_tokenizer():
    """utils.py:57-60 (needs the tokenizer files in the local HF cache)."""
    from transformers import AutoTokenizer
    tok = AutoTokenizer.from_pretrained("huggyllama/llama-7b")
    tok.pad_token = "</s>"
    return tok


def synthetic_token_batches(vocab_size: int, batch_size: int, seq_length: int, *, seed: int, device,
                            steps: int):
    """Yields `steps` batches shaped like the reference collate_fn's output (main.py:79-88):
    input_ids, attention_mask and labels = input_ids.clone(), drawn from a seeded generator."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    for _ in range(steps):
        ids = torch.randint(0, vocab_size, (batch_size, seq_length), generator=g)
        ids = ids.pin_memory().to(device, non_blocking=True) if torch.device(device).type == "cuda" else ids
        yield {"input_ids": ids, "attention_mask": torch.ones_like(ids), "labels": ids.clone()}
