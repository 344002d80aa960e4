"""reference nanodiloco/diloco/__init__.py:1-3"""
from nanodiloco_b200.diloco import Diloco

__all__ = ["Diloco"]
