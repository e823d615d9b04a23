"""Fake `isaacgym` package -- TEST INFRASTRUCTURE ONLY (oracle side).

Lets the *unmodified* reference at /root/reference import and run on CPU in a
container without Isaac Gym / PhysX.  The simulator is replaced by a replay
source that writes synthetic state into tensors with Isaac Gym's
acquire_*_tensor layouts.  Nothing under oracle/ is
imported by the product package.
"""
from . import gymapi, gymtorch, gymutil, torch_utils, terrain_utils  # noqa: F401
