"""fem_b200: B200-native explicit RK-DG time stepping for the dg1d/python path of cpraveen/fem.

Host orchestration in Python/PyTorch; the hot path is hand-written sm_100a CUDA behind a C ABI
(include/dg1d.h, fem_b200/libdg1d.so).  Nothing here falls back to the CPU.
"""
from . import _lib  # noqa: F401

__all__ = ['DG', 'Euler', 'Dg1dDevice']


def __getattr__(name):
    if name == 'DG':
        from .dgsolver import DG
        return DG
    if name == 'Euler':
        from .euler import Euler
        return Euler
    if name == 'Dg1dDevice':
        from .device import Dg1dDevice
        return Dg1dDevice
    raise AttributeError(name)
