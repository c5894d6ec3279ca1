from legged_gym_b200.sim.gymutil import *  # noqa: F401,F403
