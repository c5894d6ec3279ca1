"""``isaacgym.gymutil`` stand-in: device-string / CLI / sim-config helpers (``utils/helpers.py:79-101,152-178``)."""
from __future__ import annotations

import argparse

from . import gymapi


def parse_device_str(device: str):
    device = device.lower()
    if device in ("cpu", "cuda", "gpu"):
        return ("cuda" if device != "cpu" else "cpu"), 0
    kind, _, idx = device.partition(":")
    if kind not in ("cpu", "cuda") or not idx.isdigit():
        raise ValueError(f"invalid device string: {device}")
    return kind, int(idx)


def parse_arguments(description="", headless=False, no_graphics=False, custom_parameters=(), argv=None):
    p = argparse.ArgumentParser(description=description)
    p.add_argument("--sim_device", type=str, default="cuda:0")
    p.add_argument("--pipeline", type=str, default="gpu")
    p.add_argument("--graphics_device_id", type=int, default=0)
    p.add_argument("--physx", action="store_true")
    p.add_argument("--flex", action="store_true")
    p.add_argument("--num_threads", type=int, default=0)
    p.add_argument("--subscenes", type=int, default=0)
    p.add_argument("--slices", type=int, default=None)
    for prm in custom_parameters:
        kw = {k: v for k, v in prm.items() if k != "name"}
        p.add_argument(prm["name"], **kw)
    args = p.parse_args(argv)
    args.sim_device_type, args.compute_device_id = parse_device_str(args.sim_device)
    args.use_gpu = args.sim_device_type == "cuda"
    args.use_gpu_pipeline = args.pipeline.lower() in ("gpu", "cuda") and args.use_gpu
    args.physics_engine = gymapi.SIM_FLEX if args.flex else gymapi.SIM_PHYSX
    if args.slices is None:
        args.slices = args.subscenes
    return args


def parse_sim_config(cfg: dict, sim_params):
    for key, val in cfg.items():
        if key == "physx" and isinstance(val, dict):
            for k, v in val.items():
                setattr(sim_params.physx, k, v)
        elif key == "gravity":
            sim_params.gravity = gymapi.Vec3(*val)
        else:
            setattr(sim_params, key, val)
    return sim_params
