"""`nanodiloco` -- the reference's import name, served by nanodiloco_b200.

A project written against SJCaldwell/NanoDiloco keeps every import and launch line it has:

    from nanodiloco import Diloco, main                      (nanodiloco/__init__.py:1-4)
    from nanodiloco.diloco import Diloco                     (nanodiloco/diloco/__init__.py:1)
    from nanodiloco.training_utils import ddp_setup, ...     (nanodiloco/training_utils/__init__.py:1-3)
    python -m nanodiloco ...                                 (nanodiloco/__main__.py:1-4)
    torchrun --nproc-per-node 8 -m nanodiloco.main ...       (scripts/train_modal.py:61)

and gets the sm_100a kernels underneath.  Put this repository's root on PYTHONPATH *instead of* the
reference package (both cannot own the name).
"""
from nanodiloco_b200.diloco import Diloco
from .main import main          # as in the reference: `nanodiloco.main` is the function once the package is imported

__all__ = ["Diloco", "main"]
