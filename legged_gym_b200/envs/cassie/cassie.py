"""Cassie task.  Its extra reward ``no_fly`` (reference ``envs/cassie/cassie.py:43-46``: exactly one
foot with f_z > 0.1) is a built-in term of the fused kernel (LGK_REW_NO_FLY), so the class adds
nothing on the host side."""
from legged_gym_b200.envs.base.legged_robot import LeggedRobot


class Cassie(LeggedRobot):
    pass
