from legged_gym_b200.sim.gymtorch import *  # noqa: F401,F403
