"""Aliengo task: identical to Go1 bar the config (reference ``envs/aliengo/aliengo.py``)."""
from legged_gym_b200.envs.go1.go1 import Go1


class Aliengo(Go1):
    pass
