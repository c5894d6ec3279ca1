def subplots(*a, **k):
    raise RuntimeError("matplotlib is not available")
