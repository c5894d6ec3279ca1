"""Unitree Go1 task config (values: reference ``envs/go1/go1_config.py:34-97``).

As shipped the task is flat and blind (48 observations); rough terrain with the 187-point scan
is obtained by switching ``terrain.mesh_type`` / ``measure_heights`` (see utils/scenarios.py).
"""
from legged_gym_b200.envs.base.legged_robot_config import LeggedRobotCfg, LeggedRobotCfgPPO
from legged_gym_b200.envs._unitree import joint_angles, ppo_cfg, unitree_asset


class Go1RoughCfg(LeggedRobotCfg):
    class env(LeggedRobotCfg.env):
        num_observations = 48

    class terrain(LeggedRobotCfg.terrain):
        mesh_type = 'plane'
        measure_heights = False

    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 0.32]
        default_joint_angles = joint_angles(0.1, -0.1, 0.8, 1., -1.5)

    class control(LeggedRobotCfg.control):
        control_type = 'P'
        stiffness = {'hip_joint': 30, 'thigh_joint': 50., 'calf_joint': 50.}
        damping = {'hip_joint': 2., 'thigh_joint': 2., 'calf_joint': 2.}
        action_scale = 0.25
        decimation = 4
        use_actuator_network = True
        actuator_net_file = "{LEGGED_GYM_ROOT_DIR}/resources/actuator_nets/go1_net.pt"

    asset = unitree_asset(LeggedRobotCfg.asset, "go1")

    class domain_rand(LeggedRobotCfg.domain_rand):
        randomize_base_mass = True
        added_mass_range = [-1., 1.]
        randomize_limb_mass = True
        added_limb_percentage = [-0.2, 0.2]

    class rewards(LeggedRobotCfg.rewards):
        soft_dof_pos_limit = 0.9
        base_height_target = 0.25

        class scales(LeggedRobotCfg.rewards.scales):
            torques = -0.00025
            dof_pos_limits = -10.0


Go1RoughCfgPPO = ppo_cfg(LeggedRobotCfgPPO, 'rough_go1')
