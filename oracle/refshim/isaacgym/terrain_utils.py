from legged_gym_b200.sim.terrain_utils import *  # noqa: F401,F403
