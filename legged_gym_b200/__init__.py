"""legged_gym_b200 -- B200-native per-env step path of legged_gym.

Host side mirrors the reference's package layout (``envs``, ``utils``) so that a
task registered against the reference (``legged_gym/envs/__init__.py:52-59``) is
constructed the same way here; the arithmetic of ``LeggedRobot.step()`` /
``post_physics_step()`` runs in hand-written sm_100a CUDA kernels behind the C ABI
declared in ``include/lgk.h``.
"""
import os

LEGGED_GYM_ROOT_DIR = os.path.dirname(os.path.realpath(__file__))
LEGGED_GYM_ENVS_DIR = os.path.join(LEGGED_GYM_ROOT_DIR, "envs")

__version__ = "0.1.0"
