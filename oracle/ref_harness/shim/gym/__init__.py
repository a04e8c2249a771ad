"""Minimal stand-in for the absent `gym` package (test infrastructure only).

The reference DCA env only subclasses ``gym.Env`` and imports ``gym.spaces``
(/root/reference/MISSION/dca/collective_assault_parallel_run.py:1-2,
/root/reference/MISSION/dca/collective_assault_env.py:1). Nothing else is used.
"""
from . import spaces  # noqa: F401


class Env(object):
    pass
