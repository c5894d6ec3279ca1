"""Shared pieces of the Unitree-style quadruped configs (A1 / Go1 / Aliengo)."""

LEGS = ("FL", "RL", "FR", "RR")


def joint_angles(hip_left, hip_right, thigh_front, thigh_rear, calf):
    """default_joint_angles dict keyed by '<leg>_<joint>_joint' (target angle at zero action)."""
    out = {}
    for leg in LEGS:
        out[f"{leg}_hip_joint"] = hip_left if leg[1] == "L" else hip_right
    for leg in LEGS:
        out[f"{leg}_thigh_joint"] = thigh_front if leg[0] == "F" else thigh_rear
    for leg in LEGS:
        out[f"{leg}_calf_joint"] = calf
    return out


def unitree_asset(base, robot):
    """asset block shared by the Unitree quadrupeds: urdf under resources/robots/<robot>, feet named '*foot',
    contacts penalised on thighs and calves, episode ends on base contact."""
    return type("asset", (base,), dict(
        file='{LEGGED_GYM_ROOT_DIR}/resources/robots/%s/urdf/%s.urdf' % (robot, robot), name=robot, foot_name="foot",
        penalize_contacts_on=["thigh", "calf"], terminate_after_contacts_on=["base"], self_collisions=1))


def ppo_cfg(base, experiment_name, entropy_coef=0.01):
    """Train config of a Unitree task: the base PPO settings with the entropy bonus and the run's name."""
    return type(experiment_name + "_ppo", (base,), dict(
        algorithm=type("algorithm", (base.algorithm,), dict(entropy_coef=entropy_coef)),
        runner=type("runner", (base.runner,), dict(run_name='', experiment_name=experiment_name))))
