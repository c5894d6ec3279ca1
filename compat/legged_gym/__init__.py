"""`legged_gym` import path for user code written against the reference
(`from legged_gym.envs.base.legged_robot import LeggedRobot`, `from legged_gym.utils.task_registry import task_registry`,
reference legged_gym/envs/__init__.py:33-59): every `legged_gym.<x>` resolves to the SAME module object as
`legged_gym_b200.<x>`.  Opt-in: put this directory's parent on the path (PYTHONPATH=<repo>/compat:<repo>); it is not on
the path by default so that it can never shadow a real `legged_gym` (the reference itself is imported under that name by
this repo's test oracle)."""
import importlib
import importlib.abc
import importlib.machinery
import sys

import legged_gym_b200 as _impl
from legged_gym_b200 import *  # noqa: F401,F403
from legged_gym_b200 import LEGGED_GYM_ROOT_DIR  # noqa: F401

_PREFIX, _REAL = __name__ + ".", _impl.__name__ + "."


class _Alias(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(_PREFIX):
            return None
        try:
            importlib.import_module(_REAL + fullname[len(_PREFIX):])
        except ModuleNotFoundError:
            return None
        return importlib.machinery.ModuleSpec(fullname, self, is_package=True)

    def create_module(self, spec):
        return sys.modules[_REAL + spec.name[len(_PREFIX):]]     # the very same module object

    def exec_module(self, module):
        pass


sys.meta_path.insert(0, _Alias())
__path__ = []          # submodules come from the finder above, never from this directory
LEGGED_GYM_ENVS_DIR = getattr(_impl, "LEGGED_GYM_ENVS_DIR", None)
