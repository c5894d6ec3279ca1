class VecEnv:
    pass
