"""Fake `isaacgym.gymtorch` (oracle shim): sim tensors already are torch tensors."""


def wrap_tensor(t):
    return t


def unwrap_tensor(t):
    return t
