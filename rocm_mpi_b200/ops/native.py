"""ctypes bindings to the in-tree native library ``rocm_mpi_b200/_lib/libigg_b200.so``.

Filled in below; ``available()`` is the single switch the Python layer consults.
"""
from __future__ import annotations


def available() -> bool:
    return False
