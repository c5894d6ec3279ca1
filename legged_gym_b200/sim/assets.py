"""Robot asset descriptions served by the headless simulator backend.

The reference gets DoF / rigid-body names, counts and limits from Isaac Gym's asset
loader (``legged_robot.py:673-683``).  The headless backend answers the same queries
from (a) the built-in tables below -- numbers taken from the reference's URDF
``<limit>`` tags (go1.urdf:94,136,164; cassie.urdf:320-416) -- or (b) a URDF file on
disk, parsed with the same conventions Isaac Gym applies: fixed joints are collapsed
into the parent unless tagged ``dont_collapse``, and siblings are visited in
alphabetical order, depth first.
"""
from __future__ import annotations

import os
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np

DOF_PROP_DTYPE = np.dtype([
    ("hasLimits", np.bool_), ("lower", np.float32), ("upper", np.float32),
    ("driveMode", np.int32), ("velocity", np.float32), ("effort", np.float32),
    ("stiffness", np.float32), ("damping", np.float32), ("friction", np.float32),
    ("armature", np.float32),
])


@dataclass
class RobotAsset:
    name: str
    dof_names: List[str]
    body_names: List[str]
    lower: List[float]
    upper: List[float]
    velocity: List[float]
    effort: List[float]
    body_mass: List[float] = field(default_factory=list)

    @property
    def num_dof(self) -> int:
        return len(self.dof_names)

    @property
    def num_bodies(self) -> int:
        return len(self.body_names)

    def dof_properties(self) -> np.ndarray:
        props = np.zeros(self.num_dof, dtype=DOF_PROP_DTYPE)
        props["hasLimits"] = True
        props["lower"] = self.lower
        props["upper"] = self.upper
        props["velocity"] = self.velocity
        props["effort"] = self.effort
        props["driveMode"] = 3
        return props


def _quadruped(name, legs, lim_hip, lim_thigh, lim_calf, vel, eff, base="base",
               joints=("hip", "thigh", "calf"), suffix="_joint", foot="foot"):
    dofs, bodies = [], [base]
    lo, hi = [], []
    for leg in legs:
        for j, lim in zip(joints, (lim_hip, lim_thigh, lim_calf)):
            dofs.append(f"{leg}_{j}{suffix}")
            lo.append(lim[0])
            hi.append(lim[1])
        bodies += [f"{leg}_{j}" for j in joints] + [f"{leg}_{foot}"]
    n = len(dofs)
    return RobotAsset(name, dofs, bodies, lo, hi, [vel] * n, [eff] * n,
                      body_mass=[5.0] + [0.5] * (len(bodies) - 1))


def _cassie():
    side_joints = ["hip_abduction", "hip_rotation", "hip_flexion", "thigh_joint", "ankle_joint", "toe_joint"]
    side_bodies = ["pelvis_rotation", "hip", "thigh", "shin", "tarsus", "toe"]
    lim = {
        "left": [(-0.2618, 0.3927), (-0.3927, 0.3927), (-0.8727, 1.3963),
                 (-2.8623, -0.6458), (0.6458, 2.8623), (-2.4435, -0.5236)],
        "right": [(-0.3927, 0.2618), (-0.3927, 0.3927), (-0.8727, 1.3963),
                  (-2.8623, -0.6458), (0.6458, 2.8623), (-2.4435, -0.5236)],
    }
    vel = [20.1475, 20.1475, 20.5085, 20.5085, 20.5085, 20.5192]
    eff = [112.0, 112.0, 195.0, 195.0, 195.0, 45.0]
    dofs, bodies, lo, hi, v, e = [], ["pelvis"], [], [], [], []
    for side in ("left", "right"):
        for k, j in enumerate(side_joints):
            dofs.append(f"{j}_{side}")
            lo.append(lim[side][k][0])
            hi.append(lim[side][k][1])
            v.append(vel[k])
            e.append(eff[k])
        bodies += [f"{side}_{b}" for b in side_bodies]
    return RobotAsset("cassie", dofs, bodies, lo, hi, v, e, body_mass=[10.0] + [1.0] * 12)


_UNITREE_LEGS = ("FL", "FR", "RL", "RR")

BUILTIN_ASSETS: Dict[str, RobotAsset] = {
    # go1.urdf:94,136,164 (hip / thigh / calf limits), effort 23.7, velocity 30.1
    "go1": _quadruped("go1", _UNITREE_LEGS, (-1.0471975512, 1.0471975512),
                      (-0.663225115758, 2.96705972839), (-2.72271363311, -0.837758040957), 30.1, 23.7),
    "aliengo": _quadruped("aliengo", _UNITREE_LEGS, (-1.2217304764, 1.2217304764),
                          (-3.14159265359, 3.14159265359), (-2.77507351067, -0.645771823238), 40.0, 44.0),
    "a1": _quadruped("a1", _UNITREE_LEGS, (-0.802851455917, 0.802851455917),
                     (-1.0471975512, 4.18879020479), (-2.69653369433, -0.916297857297), 21.0, 33.5),
    "anymal_c": _quadruped("anymal_c", ("LF", "LH", "RF", "RH"), (-9.42, 9.42), (-9.42, 9.42), (-9.42, 9.42),
                           7.5, 80.0, joints=("HAA", "HFE", "KFE"), suffix="", foot="FOOT"),
    "cassie": _cassie(),
}
# ANYmal body names use HIP/THIGH/SHANK/FOOT per leg in the reference URDF
BUILTIN_ASSETS["anymal_c"].body_names = ["base"] + [
    f"{leg}_{b}" for leg in ("LF", "LH", "RF", "RH") for b in ("HIP", "THIGH", "SHANK", "FOOT")]
BUILTIN_ASSETS["anymal_b"] = BUILTIN_ASSETS["anymal_c"]


def parse_urdf(path: str, collapse_fixed_joints: bool = True) -> RobotAsset:
    """Minimal URDF reader (links, joints, limits) following Isaac Gym's conventions."""
    root = ET.parse(path).getroot()
    links = [l.attrib["name"] for l in root.findall("link")]
    children: Dict[str, List[Tuple[str, ET.Element]]] = {l: [] for l in links}
    has_parent = set()
    for j in root.findall("joint"):
        parent = j.find("parent").attrib["link"]
        child = j.find("child").attrib["link"]
        children[parent].append((child, j))
        has_parent.add(child)
    roots = [l for l in links if l not in has_parent]
    if not roots:
        raise ValueError(f"URDF {path}: no root link")
    dofs, bodies, lo, hi, vel, eff = [], [], [], [], [], []

    def visit(link, emit):
        if emit:
            bodies.append(link)
        for child, j in sorted(children[link], key=lambda cj: cj[0]):
            jt = j.attrib.get("type", "fixed")
            keep = j.attrib.get("dont_collapse", "false").lower() == "true"
            if jt == "fixed":
                visit(child, emit=(not collapse_fixed_joints) or keep)
            else:
                dofs.append(j.attrib["name"])
                lim = j.find("limit")
                a = lim.attrib if lim is not None else {}
                lo.append(float(a.get("lower", -np.pi)))
                hi.append(float(a.get("upper", np.pi)))
                vel.append(float(a.get("velocity", 100.0)))
                eff.append(float(a.get("effort", 100.0)))
                visit(child, emit=True)

    visit(roots[0], emit=True)
    name = root.attrib.get("name", os.path.splitext(os.path.basename(path))[0])
    return RobotAsset(name, dofs, bodies, lo, hi, vel, eff, body_mass=[1.0] * len(bodies))


def resolve_asset(asset_root: str, asset_file: str, collapse_fixed_joints: bool = True) -> RobotAsset:
    key = os.path.splitext(os.path.basename(asset_file))[0]
    if key in BUILTIN_ASSETS:
        return BUILTIN_ASSETS[key]
    path = os.path.join(asset_root, asset_file)
    if os.path.isfile(path):
        return parse_urdf(path, collapse_fixed_joints)
    raise FileNotFoundError(
        f"asset '{asset_file}' is neither a built-in robot ({sorted(BUILTIN_ASSETS)}) nor a URDF under {asset_root}")
