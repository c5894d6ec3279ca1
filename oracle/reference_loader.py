"""Import the reference's unmodified ``legged_gym`` over the fake-Isaac-Gym shim -- TEST INFRASTRUCTURE.

Only works where /root/reference exists (the build container).  Also provides the RNG "tap" harness:
the reference draws random numbers from torch's global generator in six places per step (SURVEY
A-4); for parity both sides must consume the *same* uniforms, so the draws are redirected to
caller-provided tensors with one row per env -- row ``i`` of a tap belongs to env ``i``, whatever
compacted position it has in the reference's ``env_ids``.
"""
from __future__ import annotations

import contextlib
import copy
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SHIM = os.path.join(_HERE, "refshim")
STAGED_ROOT = os.path.join(_HERE, "_ref")      # oracle/make_ref.sh: unmodified copy that travels to the GPU box
_loaded = None


def _has_reference(root) -> bool:
    return os.path.isfile(os.path.join(root, "legged_gym", "envs", "base", "legged_robot.py"))


def _find_root() -> str:
    """LEGGED_GYM_REFERENCE if set, else /root/reference (build container), else the staged copy oracle/_ref."""
    env = os.environ.get("LEGGED_GYM_REFERENCE")
    if env:
        return env
    for root in ("/root/reference", STAGED_ROOT):
        if _has_reference(root):
            return root
    return "/root/reference"


REFERENCE_ROOT = _find_root()


def reference_available() -> bool:
    return _has_reference(REFERENCE_ROOT)


def reference_kind() -> str:
    """where the imported reference lives: 'source tree' (/root/reference) or 'staged copy' (oracle/_ref)"""
    return "staged copy oracle/_ref" if os.path.abspath(REFERENCE_ROOT) == os.path.abspath(STAGED_ROOT) else REFERENCE_ROOT


def load_reference():
    """Returns the imported reference package ``legged_gym`` (with ``legged_gym.envs`` registered)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    for p in (_SHIM, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    # go1_net.pt was saved from CUDA; torch.jit.load(path) without map_location fails on a GPU-less host
    if not torch.cuda.is_available() and not getattr(torch.jit.load, "_lgk_patched", False):
        orig = torch.jit.load

        def load_cpu(f, map_location=None, *a, **k):
            return orig(f, map_location or "cpu", *a, **k)
        load_cpu._lgk_patched = True
        torch.jit.load = load_cpu
    import legged_gym            # noqa: F401  (reference)
    import legged_gym.envs       # noqa: F401  registers the 8 tasks
    _loaded = legged_gym
    return legged_gym


def make_reference_env(task: str, mutate=None, num_envs: int = 64, device: str = "cpu", seed: int = 0):
    """Instantiate the reference's task class on the fake simulator."""
    lg = load_reference()
    from legged_gym.utils.task_registry import task_registry
    from legged_gym.utils.helpers import class_to_dict
    from isaacgym import gymapi, gymutil
    task_class = task_registry.get_task_class(task)
    cfg = copy.deepcopy(task_registry.env_cfgs[task])
    cfg.env.num_envs = num_envs
    cfg.seed = seed
    if mutate is not None:
        mutate(cfg)
    sim_params = gymapi.SimParams()
    gymutil.parse_sim_config(class_to_dict(cfg.sim), sim_params)
    sim_params.use_gpu_pipeline = device.startswith("cuda")
    import numpy as np
    np.random.seed(seed)
    torch.manual_seed(seed)
    with contextlib.redirect_stdout(open(os.devnull, "w")):
        env = task_class(cfg=cfg, sim_params=sim_params, physics_engine=gymapi.SIM_PHYSX, sim_device=device,
                         headless=True)
    return env


class TapHarness:
    """Redirects the reference's random draws to per-env tap tensors for one or more steps.

    taps: dict with uniform [0,1) float tensors ``cmd_callback [N,3]``, ``cmd_reset [N,3]``,
    ``push [N,2]``, ``dof [N,D]``, ``root_xy [N,2]``, ``root_vel [N,6]``, ``noise [N,O]`` and the int64
    tensor ``level [N]`` (values in [0, max_terrain_level)).
    """

    def __init__(self, env):
        self.env = env
        self.taps = None
        self._site = None
        self._ids = None
        self._col = 0
        self._in_reset = False
        self._wrap_methods()

    # -- wrap the bound methods on the instance so each draw knows its site and its env ids
    def _wrap_methods(self):
        env = self.env
        h = self

        def wrap(name, site_fn):
            orig = getattr(env, name)

            def wrapped(*args, **kw):
                prev = (h._site, h._ids, h._col)
                h._site, h._ids, h._col = site_fn(), (args[0] if args else None), 0
                try:
                    return orig(*args, **kw)
                finally:
                    h._site, h._ids, h._col = prev
            setattr(env, name, wrapped)

        orig_reset = env.reset_idx

        def reset_idx(env_ids):
            h._in_reset = True
            try:
                return orig_reset(env_ids)
            finally:
                h._in_reset = False
        env.reset_idx = reset_idx
        wrap("_resample_commands", lambda: "cmd_reset" if h._in_reset else "cmd_callback")
        wrap("_reset_dofs", lambda: "dof")
        wrap("_reset_root_states", lambda: "root")
        wrap("_update_terrain_curriculum", lambda: "level")
        wrap("_push_robots", lambda: "push")

    def _rand_float(self, lower, upper, shape, device):
        t = self.taps
        site, ids = self._site, self._ids
        if site in ("cmd_callback", "cmd_reset"):
            u = t[site][ids, self._col].unsqueeze(1)
            self._col += 1
        elif site == "dof":
            u = t["dof"][ids]
        elif site == "root":
            u = t["root_xy"][ids] if shape[1] == 2 else t["root_vel"][ids]
        elif site == "push":
            u = t["push"]
        else:
            raise RuntimeError(f"unexpected torch_rand_float call (site={site}, shape={shape})")
        assert tuple(u.shape) == tuple(shape), (site, u.shape, shape)
        return (upper - lower) * u.to(device) + lower

    @contextlib.contextmanager
    def active(self, taps):
        import legged_gym.envs.base.legged_robot as ref_lr
        self.taps = taps
        saved = (ref_lr.torch_rand_float, torch.rand_like, torch.randint_like)
        h = self

        def rand_like(x, *a, **k):
            assert tuple(x.shape) == tuple(taps["noise"].shape)
            return taps["noise"].to(x.device).clone()

        def randint_like(x, high, *a, **k):
            assert h._site == "level"
            return taps["level"][h._ids].to(x.device).clone()
        ref_lr.torch_rand_float = self._rand_float
        torch.rand_like = rand_like
        torch.randint_like = randint_like
        try:
            yield self
        finally:
            ref_lr.torch_rand_float, torch.rand_like, torch.randint_like = saved
            self.taps = None
