"""A1 (alternate URDF) task config (values: reference ``envs/a1_src/a1_src_config.py``)."""
from legged_gym_b200.envs.base.legged_robot_config import LeggedRobotCfgPPO
from legged_gym_b200.envs.a1.a1_config import A1RoughCfg


class A1SrcRoughCfg(A1RoughCfg):
    class asset(A1RoughCfg.asset):
        file = '{LEGGED_GYM_ROOT_DIR}/resources/robots/a1_src/urdf/a1.urdf'
        name = "a1_src"

    class rewards(A1RoughCfg.rewards):
        class scales(A1RoughCfg.rewards.scales):
            torques = -0.00001


class A1SrcRoughCfgPPO(LeggedRobotCfgPPO):
    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01

    class runner(LeggedRobotCfgPPO.runner):
        run_name = ''
        experiment_name = 'rough_a1_src'
