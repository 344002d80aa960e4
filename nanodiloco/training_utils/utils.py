"""reference nanodiloco/training_utils/utils.py"""
from nanodiloco_b200.training_utils.utils import *  # noqa: F401,F403
