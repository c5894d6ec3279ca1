"""The one place that decides which simulator binding the package talks to: the real ``isaacgym`` when it is
importable (the reference's own dependency, `legged_robot.py:37-38`), else the tensor-only stand-in of this
directory.  Every module imports gymapi / gymutil / gymtorch from here, so that one process never mixes the two
(a stand-in SimParams handed to the real gym.create_sim)."""
try:
    from isaacgym import gymapi, gymtorch, gymutil  # type: ignore  # noqa: F401
    # (the test oracle's shim re-exports the stand-in under the name isaacgym: that is not the real one)
    REAL_ISAACGYM = not hasattr(gymapi, "SyntheticGym")
except Exception:   # noqa: BLE001  (not installed here: closed source)
    from legged_gym_b200.sim import gymapi, gymtorch, gymutil  # noqa: F401
    REAL_ISAACGYM = False
