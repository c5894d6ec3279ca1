"""Config base class: nested config *classes* become nested *instances* on construction.

Same contract as the reference's ``BaseConfig`` (``envs/base/base_config.py:33-55``): user task
configs subclass nested classes (``class env(LeggedRobotCfg.env)``) and read plain attributes.
"""
import inspect


def _instantiate_nested(node):
    for name in dir(node):
        if name == "__class__":
            continue
        member = getattr(node, name)
        if inspect.isclass(member):
            inst = member()
            setattr(node, name, inst)
            _instantiate_nested(inst)


class BaseConfig:
    def __init__(self) -> None:
        self.init_member_classes(self)

    @staticmethod
    def init_member_classes(obj):
        _instantiate_nested(obj)
