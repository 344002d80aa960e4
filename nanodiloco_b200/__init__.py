"""nanodiloco_b200 -- the DiLoCo outer-sync hot path of SJCaldwell/NanoDiloco on B200.

`from nanodiloco_b200 import Diloco` is the drop-in for `from nanodiloco import Diloco`
(reference nanodiloco/__init__.py:1-4).  Attributes resolve lazily so that
`python -m nanodiloco_b200.build` does not import torch.
"""
__all__ = ["Diloco", "main"]


def __getattr__(name):
    if name == "Diloco":
        from .diloco import Diloco
        return Diloco
    if name == "main":
        from .main import main as fn
        globals()["main"] = fn          # the submodule import bound the module under this name; rebind
        return fn
    raise AttributeError(f"module 'nanodiloco_b200' has no attribute {name!r}")
