"""``isaacgym.gymtorch`` stand-in: the headless backend's tensors already are torch tensors."""


def wrap_tensor(t):
    return t


def unwrap_tensor(t):
    return t
