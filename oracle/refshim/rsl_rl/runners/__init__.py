class OnPolicyRunner:
    def __init__(self, *a, **k):
        raise RuntimeError("rsl_rl is not available in this container")
