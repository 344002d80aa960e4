"""`torchrun -m nanodiloco.main` (reference nanodiloco/main.py:133-141, launched by scripts/train_modal.py:61)."""
from nanodiloco_b200.main import (get_default_llama_config, get_default_wandb_config,  # noqa: F401
                                  load_config_from_file, main, train_model)

if __name__ == "__main__":
    main()
