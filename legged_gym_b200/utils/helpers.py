"""Config / CLI helpers with the reference's names (``utils/helpers.py:41-178``).

Only ``class_to_dict`` matters numerically: it enumerates attributes through ``dir()``, i.e.
alphabetically, and that order is the evaluation *and fp32 summation* order of the reward terms
(``legged_robot.py:588-593``).
"""
from __future__ import annotations

import os
import random

import numpy as np
import torch

from legged_gym_b200.sim.backend import gymapi, gymutil


def class_to_dict(obj) -> dict:
    if not hasattr(obj, "__dict__"):
        return obj
    out = {}
    for key in dir(obj):                      # dir() sorts -> alphabetical keys
        if key.startswith("_"):
            continue
        val = getattr(obj, key)
        out[key] = [class_to_dict(v) for v in val] if isinstance(val, list) else class_to_dict(val)
    return out


def update_class_from_dict(obj, dict):
    for key, val in dict.items():
        attr = getattr(obj, key, None)
        if isinstance(attr, type):
            update_class_from_dict(attr, val)
        else:
            setattr(obj, key, val)


def set_seed(seed):
    if seed == -1:
        seed = np.random.randint(0, 10000)
    print("Setting seed: {}".format(seed))
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    os.environ['PYTHONHASHSEED'] = str(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def parse_sim_params(args, cfg):
    sim_params = gymapi.SimParams()
    if args.physics_engine == gymapi.SIM_PHYSX:
        sim_params.physx.use_gpu = args.use_gpu
        sim_params.physx.num_subscenes = args.subscenes
    sim_params.use_gpu_pipeline = args.use_gpu_pipeline
    if "sim" in cfg:
        gymutil.parse_sim_config(cfg["sim"], sim_params)
    if args.physics_engine == gymapi.SIM_PHYSX and args.num_threads > 0:
        sim_params.physx.num_threads = args.num_threads
    return sim_params


def get_load_path(root, load_run=-1, checkpoint=-1):
    try:
        runs = sorted(os.listdir(root))
        if 'exported' in runs:
            runs.remove('exported')
        last_run = os.path.join(root, runs[-1])
    except Exception:
        raise ValueError("No runs in this directory: " + root)
    load_run = last_run if load_run == -1 else os.path.join(root, load_run)
    if checkpoint == -1:
        models = sorted((f for f in os.listdir(load_run) if 'model' in f), key=lambda m: '{0:0>15}'.format(m))
        model = models[-1]
    else:
        model = "model_{}.pt".format(checkpoint)
    return os.path.join(load_run, model)


def update_cfg_from_args(env_cfg, cfg_train, args):
    if env_cfg is not None and getattr(args, "num_envs", None) is not None:
        env_cfg.env.num_envs = args.num_envs
    if cfg_train is not None:
        if getattr(args, "seed", None) is not None:
            cfg_train.seed = args.seed
        for name in ("max_iterations", "experiment_name", "run_name", "load_run", "checkpoint"):
            val = getattr(args, name, None)
            if val is not None:
                setattr(cfg_train.runner, name, val)
        if getattr(args, "resume", False):
            cfg_train.runner.resume = args.resume
    return env_cfg, cfg_train


def get_args(argv=None):
    custom_parameters = [
        {"name": "--task", "type": str, "default": "anymal_c_flat", "help": "Registered task name."},
        {"name": "--resume", "action": "store_true", "default": False, "help": "Resume training from a checkpoint"},
        {"name": "--experiment_name", "type": str, "help": "Name of the experiment to run or load."},
        {"name": "--run_name", "type": str, "help": "Name of the run."},
        {"name": "--load_run", "type": str, "help": "Run to load when resume=True (-1: last)."},
        {"name": "--checkpoint", "type": int, "help": "Saved model checkpoint number (-1: last)."},
        {"name": "--headless", "action": "store_true", "default": False, "help": "Force display off at all times"},
        {"name": "--horovod", "action": "store_true", "default": False, "help": "Declared by the reference, never read"},
        {"name": "--rl_device", "type": str, "default": "cuda:0", "help": "Device used by the RL algorithm"},
        {"name": "--num_envs", "type": int, "help": "Number of environments to create."},
        {"name": "--seed", "type": int, "help": "Random seed."},
        {"name": "--max_iterations", "type": int, "help": "Maximum number of training iterations."},
    ]
    args = gymutil.parse_arguments(description="RL Policy", custom_parameters=custom_parameters, argv=argv)
    args.sim_device_id = args.compute_device_id
    args.sim_device = args.sim_device_type
    if args.sim_device == 'cuda':
        args.sim_device += f":{args.sim_device_id}"
    return args


def export_policy_as_jit(actor_critic, path):
    os.makedirs(path, exist_ok=True)
    model = torch.jit.script(__import__("copy").deepcopy(actor_critic.actor).to('cpu'))
    model.save(os.path.join(path, 'policy_1.pt'))
