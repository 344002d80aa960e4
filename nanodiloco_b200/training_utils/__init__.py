from .utils import (ddp_setup, create_run_name, get_tokenizer, get_tokenized_dataset, set_seed_all,
                    synthetic_token_batches)

__all__ = ["ddp_setup", "create_run_name", "get_tokenized_dataset", "get_tokenizer", "set_seed_all",
           "synthetic_token_batches"]
