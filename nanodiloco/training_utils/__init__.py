"""reference nanodiloco/training_utils/__init__.py:1-3"""
from nanodiloco_b200.training_utils import (create_run_name, ddp_setup, get_tokenized_dataset, get_tokenizer,
                                            set_seed_all)

__all__ = ["ddp_setup", "create_run_name", "get_tokenized_dataset", "get_tokenizer", "set_seed_all"]
