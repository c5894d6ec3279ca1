"""Headless simulator backend (Isaac Gym tensor-API subset). See :mod:`.gymapi`."""
from . import gymapi, gymtorch, gymutil, terrain_utils  # noqa: F401
