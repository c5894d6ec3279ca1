"""Empty stand-in for rsl_rl (caller of the env; imported by utils/task_registry.py:37-38). TEST INFRASTRUCTURE."""
