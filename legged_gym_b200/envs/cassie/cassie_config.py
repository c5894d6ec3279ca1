"""Cassie biped task config (values: reference ``envs/cassie/cassie_config.py:33-112``)."""
from legged_gym_b200.envs.base.legged_robot_config import LeggedRobotCfg, LeggedRobotCfgPPO

_HALF_M = [round(0.1 * i, 1) for i in range(-5, 6)]     # 11 samples, -0.5 .. 0.5 m


def _mirrored(left: dict, flip=()):
    """joint table for both legs; joints named in ``flip`` change sign on the right side."""
    out = {f"{k}_left": v for k, v in left.items()}
    out.update({f"{k}_right": (-v if k in flip else v) for k, v in left.items()})
    return out


class CassieRoughCfg(LeggedRobotCfg):
    class env(LeggedRobotCfg.env):
        num_envs = 4096
        num_observations = 169
        num_actions = 12

    class terrain(LeggedRobotCfg.terrain):
        measured_points_x = _HALF_M     # 1 m x 1 m scan, 121 points
        measured_points_y = _HALF_M

    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 1.]
        default_joint_angles = _mirrored(
            {'hip_abduction': 0.1, 'hip_rotation': 0., 'hip_flexion': 1., 'thigh_joint': -1.8,
             'ankle_joint': 1.57, 'toe_joint': -1.57}, flip=('hip_abduction',))

    class control(LeggedRobotCfg.control):
        stiffness = {'hip_abduction': 100.0, 'hip_rotation': 100.0, 'hip_flexion': 200.,
                     'thigh_joint': 200., 'ankle_joint': 200., 'toe_joint': 40.}
        damping = {'hip_abduction': 3.0, 'hip_rotation': 3.0, 'hip_flexion': 6.,
                   'thigh_joint': 6., 'ankle_joint': 6., 'toe_joint': 1.}
        action_scale = 0.5
        decimation = 4

    class asset(LeggedRobotCfg.asset):
        file = '{LEGGED_GYM_ROOT_DIR}/resources/robots/cassie/urdf/cassie.urdf'
        name = "cassie"
        foot_name = 'toe'
        terminate_after_contacts_on = ['pelvis']
        flip_visual_attachments = False
        self_collisions = 1

    class rewards(LeggedRobotCfg.rewards):
        soft_dof_pos_limit = 0.95
        soft_dof_vel_limit = 0.9
        soft_torque_limit = 0.9
        max_contact_force = 300.
        only_positive_rewards = False

        class scales(LeggedRobotCfg.rewards.scales):
            termination = -200.
            tracking_ang_vel = 1.0
            torques = -5.e-6
            dof_acc = -2.e-7
            lin_vel_z = -0.5
            feet_air_time = 5.
            dof_pos_limits = -1.
            no_fly = 0.25
            dof_vel = -0.0
            ang_vel_xy = -0.0
            feet_contact_forces = -0.


class CassieRoughCfgPPO(LeggedRobotCfgPPO):
    class runner(LeggedRobotCfgPPO.runner):
        run_name = ''
        experiment_name = 'rough_cassie'

    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01
