"""`nanodiloco.training_utils` under the reference's name: every helper the reference exports there
(nanodiloco/training_utils/__init__.py:1-3) resolves to nanodiloco_b200's implementation of it."""
import nanodiloco_b200.training_utils as _impl

__all__ = [n for n in _impl.__all__ if n != "synthetic_token_batches"]


def __getattr__(name):
    if name in _impl.__all__:
        return getattr(_impl, name)
    raise AttributeError(f"module 'nanodiloco.training_utils' has no attribute {name!r}")
