"""nanodiloco_b200 -- the DiLoCo outer-sync hot path of SJCaldwell/NanoDiloco on B200.

`from nanodiloco_b200 import Diloco` is the drop-in for `from nanodiloco import Diloco`
(reference nanodiloco/__init__.py:1-4).
"""
from .diloco import Diloco

__all__ = ["Diloco", "main"]


def main():
    from .main import main as _main
    return _main()
