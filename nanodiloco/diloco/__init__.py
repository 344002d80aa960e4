"""`nanodiloco.diloco` under the reference's name (nanodiloco/diloco/__init__.py:1-3): the class is
nanodiloco_b200's drop-in wrapper around the sm_100a kernels."""
import nanodiloco_b200.diloco as _impl

Diloco = _impl.Diloco
__all__ = ("Diloco",)
