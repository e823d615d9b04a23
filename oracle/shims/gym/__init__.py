"""Minimal stand-in for `gym` (reference setup.py:15), oracle shim only.

The reference needs `gym.Env` as a base class (base_task.py:14) and `gym.Wrapper` with
attribute forwarding (history_wrapper.py:6-16).
"""
from . import spaces  # noqa: F401


class Env:
    pass


class Wrapper(Env):
    def __init__(self, env):
        self.env = env

    def __getattr__(self, name):
        if name.startswith("_") or name == "env":
            raise AttributeError(name)
        return getattr(self.env, name)

    def reset(self, **kwargs):
        return self.env.reset(**kwargs)

    def step(self, action):
        return self.env.step(action)


def make(*a, **k):
    raise NotImplementedError
