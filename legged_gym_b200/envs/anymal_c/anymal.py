"""ANYmal task.  The reference overrides ``_compute_torques`` with an LSTM series-elastic actuator
net (``envs/anymal_c/anymal.py:46-81``) which this fork's ``step()`` never calls (position control,
``legged_robot.py:90-95``); the LSTM actuator is a SURVEY section-8f "next" row and is not built yet."""
from legged_gym_b200.envs.base.legged_robot import LeggedRobot


class Anymal(LeggedRobot):
    def _compute_torques(self, actions):
        if getattr(self.cfg.control, "use_actuator_network", False):
            raise NotImplementedError("ANYmal LSTM actuator net: not built yet (SURVEY 8f rank 1)")
        return super()._compute_torques(actions)
