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
