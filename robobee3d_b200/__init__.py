"""robobee3d_b200 -- B200-native batched UprightMPC2 / WLQP / rigid-body closed loop.

Public surface (mirrors the reference's pybind modules, template/uprightmpc2/py/uprightmpc2py.cpp
and template/wlcontroller/pybind/wlqppy.cpp, plus the batched API):
  robobee3d_b200.uprightmpc2py.UprightMPC2C / WLCon      drop-in single-controller classes
  robobee3d_b200.wlqppy.WLController                       drop-in true-WLQP class
  robobee3d_b200.batch.BatchUMPC                           B environments per launch, closed-loop rollout
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
