from legged_gym_b200.sim.gymapi import *  # noqa: F401,F403
from legged_gym_b200.sim.gymapi import acquire_gym, _Bag  # noqa: F401
