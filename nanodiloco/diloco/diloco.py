"""`from nanodiloco.diloco.diloco import Diloco` (reference nanodiloco/diloco/diloco.py:7)."""
from nanodiloco_b200.diloco import Diloco  # noqa: F401
