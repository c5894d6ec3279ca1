"""Task registrations, same names as the reference (``envs/__init__.py:52-59``)."""
from legged_gym_b200 import LEGGED_GYM_ROOT_DIR, LEGGED_GYM_ENVS_DIR  # noqa: F401
from .base.legged_robot import LeggedRobot
from .base.legged_robot_config import LeggedRobotCfg, LeggedRobotCfgPPO  # noqa: F401
from .anymal_c.anymal import Anymal
from .anymal_c.mixed_terrains.anymal_c_rough_config import AnymalCRoughCfg, AnymalCRoughCfgPPO
from .anymal_c.flat.anymal_c_flat_config import AnymalCFlatCfg, AnymalCFlatCfgPPO
from .anymal_b.anymal_b_config import AnymalBRoughCfg, AnymalBRoughCfgPPO
from .cassie.cassie import Cassie
from .cassie.cassie_config import CassieRoughCfg, CassieRoughCfgPPO
from .a1.a1_config import A1RoughCfg, A1RoughCfgPPO
from .a1_src.a1_src_config import A1SrcRoughCfg, A1SrcRoughCfgPPO
from .go1.go1_config import Go1RoughCfg, Go1RoughCfgPPO
from .go1.go1 import Go1
from .aliengo.aliengo_config import AliengoRoughCfg, AliengoRoughCfgPPO
from .aliengo.aliengo import Aliengo

from legged_gym_b200.utils.task_registry import task_registry

task_registry.register("anymal_c_rough", Anymal, AnymalCRoughCfg(), AnymalCRoughCfgPPO())
task_registry.register("anymal_c_flat", Anymal, AnymalCFlatCfg(), AnymalCFlatCfgPPO())
task_registry.register("anymal_b", Anymal, AnymalBRoughCfg(), AnymalBRoughCfgPPO())
task_registry.register("a1", LeggedRobot, A1RoughCfg(), A1RoughCfgPPO())
task_registry.register("cassie", Cassie, CassieRoughCfg(), CassieRoughCfgPPO())
task_registry.register("a1_src", LeggedRobot, A1SrcRoughCfg(), A1SrcRoughCfgPPO())
task_registry.register("go1", Go1, Go1RoughCfg(), Go1RoughCfgPPO())
task_registry.register("aliengo", Aliengo, AliengoRoughCfg(), AliengoRoughCfgPPO())
