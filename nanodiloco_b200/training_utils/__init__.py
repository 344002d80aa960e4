"""Helpers the training loop uses around the hot path (process-group set-up, seeding, run names, the
tokenizer/dataset loaders the reference ships, and the synthetic-token stream every BASELINE config uses).

The names the reference exports from `nanodiloco.training_utils` are all here with the reference's signatures;
`synthetic_token_batches` is the one addition.  Resolved lazily so that importing the package does not pull in
`datasets` / `transformers` before a loader is actually called.
"""
from . import utils as _utils

_EXPORTS = ("create_run_name", "ddp_setup", "get_tokenized_dataset", "get_tokenizer", "set_seed_all",
            "synthetic_token_batches")
__all__ = list(_EXPORTS)


def __getattr__(name):
    if name in _EXPORTS:
        return getattr(_utils, name)
    raise AttributeError(f"module 'nanodiloco_b200.training_utils' has no attribute {name!r}")


def __dir__():
    return sorted(list(globals()) + list(_EXPORTS))
