"""Headless stand-in for the subset of ``isaacgym.gymapi`` that LeggedRobot touches.

The reference drives the closed Isaac Gym simulator through its tensor API
(``legged_robot.py:95-99,114-116,411,432,441,507-524``).  This backend keeps the same
call surface and the same tensor shapes / dtypes but performs no physics: the four
state tensors (root ``[N,13]``, dof ``[N*D,2]``, net contact ``[N*B,3]``, dof force
``[N*D]``) are plain torch tensors that a test or the benchmark fills with
synthetic frames.  ``simulate`` pulls the next frame from an optional frame source.
Every ``set_*`` call is recorded so tests can check the compacted int32 id lists.
"""
from __future__ import annotations

from types import SimpleNamespace
from typing import Callable, Dict, List, Optional

import numpy as np
import torch

from .assets import RobotAsset, resolve_asset

SIM_PHYSX = 0
SIM_FLEX = 1
UP_AXIS_Y = 0
UP_AXIS_Z = 1
DOF_MODE_NONE = 0
DOF_MODE_POS = 1
DOF_MODE_VEL = 2
DOF_MODE_EFFORT = 3
KEY_ESCAPE = "ESCAPE"
KEY_V = "V"


class Vec3:
    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)

    def __iter__(self):
        return iter((self.x, self.y, self.z))


class Quat:
    def __init__(self, x=0.0, y=0.0, z=0.0, w=1.0):
        self.x, self.y, self.z, self.w = float(x), float(y), float(z), float(w)


class Transform:
    def __init__(self, p=None, r=None):
        self.p = p if p is not None else Vec3()
        self.r = r if r is not None else Quat()


class _Bag(SimpleNamespace):
    """Attribute bag used for the various *Params / *Options structs."""


def AssetOptions():
    return _Bag()


def PlaneParams():
    return _Bag(normal=Vec3(0, 0, 1))


def HeightFieldParams():
    return _Bag(transform=Transform())


def TriangleMeshParams():
    return _Bag(transform=Transform())


def CameraProperties():
    return _Bag()


class PhysXParams(_Bag):
    pass


class SimParams:
    def __init__(self):
        self.dt = 1.0 / 60.0
        self.substeps = 2
        self.up_axis = UP_AXIS_Z
        self.gravity = Vec3(0.0, 0.0, -9.81)
        self.use_gpu_pipeline = False
        self.physx = PhysXParams(use_gpu=False, num_subscenes=0, num_threads=0)
        self.flex = _Bag()


class _RigidBodyProps:
    __slots__ = ("mass",)

    def __init__(self, mass):
        self.mass = mass


class _ShapeProps:
    __slots__ = ("friction",)

    def __init__(self):
        self.friction = 1.0


class Sim:
    def __init__(self, device: str, sim_params):
        self.device = device
        self.sim_params = sim_params
        self.asset: Optional[RobotAsset] = None
        self.num_actors = 0
        self.envs: List[int] = []
        self.tensors: Dict[str, torch.Tensor] = {}
        self.calls: List[tuple] = []          # recorded set_* calls (name, args...)
        self.frame_source: Optional[Callable[["Sim"], None]] = None
        self.sim_steps = 0
        self.terrain = None

    def _tensor(self, name, shape):
        if name not in self.tensors:
            self.tensors[name] = torch.zeros(*shape, dtype=torch.float32, device=self.device)
        return self.tensors[name]


class SyntheticGym:
    """Object returned by :func:`acquire_gym`."""

    record_calls = True
    _last_asset: Optional[RobotAsset] = None

    # ---- sim / terrain / assets -------------------------------------------------
    def create_sim(self, compute_device, graphics_device, physics_engine, sim_params):
        use_gpu = bool(getattr(sim_params, "use_gpu_pipeline", False))
        device = f"cuda:{compute_device}" if use_gpu else "cpu"
        return Sim(device, sim_params)

    def add_ground(self, sim, params):
        sim.terrain = ("plane", params)

    def add_heightfield(self, sim, samples, params):
        sim.terrain = ("heightfield", params)

    def add_triangle_mesh(self, sim, vertices, triangles, params):
        sim.terrain = ("trimesh", params)

    def load_asset(self, sim, asset_root, asset_file, options=None):
        collapse = bool(getattr(options, "collapse_fixed_joints", True)) if options is not None else True
        sim.asset = resolve_asset(asset_root, asset_file, collapse)
        self._last_asset = sim.asset   # per-actor calls below do not receive ``sim``
        return sim.asset

    def get_asset_dof_count(self, asset):
        return asset.num_dof

    def get_asset_rigid_body_count(self, asset):
        return asset.num_bodies

    def get_asset_dof_properties(self, asset):
        return asset.dof_properties()

    def get_asset_rigid_shape_properties(self, asset):
        return [_ShapeProps() for _ in range(asset.num_bodies)]

    def get_asset_rigid_body_names(self, asset):
        return list(asset.body_names)

    def get_asset_dof_names(self, asset):
        return list(asset.dof_names)

    def set_asset_rigid_shape_properties(self, asset, props):
        pass

    # ---- envs / actors ----------------------------------------------------------
    def create_env(self, sim, lower, upper, num_per_row):
        sim.envs.append(len(sim.envs))
        return sim.envs[-1]

    def create_actor(self, env, asset, pose, name, group, collision_filter, segmentation_id=0):
        return 0

    def create_actors(self, sim, num):
        """Bulk form of create_env/create_actor used by the B200 host class."""
        sim.envs = list(range(num))
        sim.num_actors = num

    def set_actor_dof_properties(self, env, actor, props):
        pass

    def enable_actor_dof_force_sensors(self, env, actor):
        pass

    def get_actor_rigid_body_properties(self, env, actor):
        asset = self._last_asset
        return [_RigidBodyProps(m) for m in asset.body_mass]

    def set_actor_rigid_body_properties(self, env, actor, props, recomputeInertia=True):
        pass

    def find_actor_rigid_body_handle(self, env, actor, name):
        return self._last_asset.body_names.index(name)

    def prepare_sim(self, sim):
        sim.num_actors = max(sim.num_actors, len(sim.envs))

    # ---- tensor API -------------------------------------------------------------
    def _n(self, sim):
        return max(sim.num_actors, len(sim.envs))

    def acquire_actor_root_state_tensor(self, sim):
        fresh = "root" not in sim.tensors
        t = sim._tensor("root", (self._n(sim), 13))
        if fresh:
            t[:, 6] = 1.0  # identity quaternion (xyzw)
        return t

    def acquire_dof_state_tensor(self, sim):
        return sim._tensor("dof", (self._n(sim) * sim.asset.num_dof, 2))

    def acquire_net_contact_force_tensor(self, sim):
        return sim._tensor("contact", (self._n(sim) * sim.asset.num_bodies, 3))

    def acquire_dof_force_tensor(self, sim):
        return sim._tensor("dof_force", (self._n(sim) * sim.asset.num_dof,))

    def refresh_dof_state_tensor(self, sim):
        pass

    def refresh_actor_root_state_tensor(self, sim):
        pass

    def refresh_net_contact_force_tensor(self, sim):
        pass

    def refresh_dof_force_tensor(self, sim):
        pass

    def refresh_rigid_body_state_tensor(self, sim):
        pass

    def simulate(self, sim):
        sim.sim_steps += 1
        if sim.frame_source is not None:
            sim.frame_source(sim)

    def fetch_results(self, sim, wait):
        pass

    def _record(self, sim, *entry):
        if self.record_calls:
            sim.calls.append(entry)
            if len(sim.calls) > 64:
                del sim.calls[:-64]

    def set_dof_position_target_tensor(self, sim, targets):
        sim.tensors["dof_target"] = targets
        return True

    def set_dof_actuation_force_tensor(self, sim, forces):
        sim.tensors["dof_actuation"] = forces
        return True

    def set_dof_state_tensor_indexed(self, sim, state, ids, count):
        self._record(sim, "set_dof_state_tensor_indexed", ids, count)
        return True

    def set_actor_root_state_tensor_indexed(self, sim, state, ids, count):
        self._record(sim, "set_actor_root_state_tensor_indexed", ids, count)
        return True

    def set_actor_root_state_tensor(self, sim, state):
        self._record(sim, "set_actor_root_state_tensor")
        return True

    # ---- viewer (never available: no graphics stack) ------------------------------
    def create_viewer(self, sim, props):
        raise RuntimeError("the headless backend has no viewer; run with headless=True")

    def set_frame_source(self, sim, fn):
        sim.frame_source = fn


_GYM = None


def acquire_gym():
    global _GYM
    if _GYM is None:
        _GYM = SyntheticGym()
    return _GYM
