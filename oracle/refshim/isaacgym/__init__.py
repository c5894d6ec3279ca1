"""Fake ``isaacgym`` package -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Lets the reference's unmodified ``legged_gym`` import and run on synthetic tensors.  The
simulator objects come from the repo's headless backend (no arithmetic there); the six
``torch_utils`` functions the hot path calls are restated in :mod:`.torch_utils`.
"""
from legged_gym_b200.sim import gymapi, gymtorch, gymutil, terrain_utils  # noqa: F401
from . import torch_utils  # noqa: F401
