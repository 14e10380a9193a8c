"""Scalar / element-wise functions shared by `taichi` and `taichi.math` of the shim."""
import ctypes
import os
import sys

import numpy as np

from .matrix import Matrix

_ORACLE = None


def _oracle_exp():
    """float gso_expf(float) of oracle/libgs_oracle.so (built on demand)."""
    global _ORACLE
    if _ORACLE is None:
        here = os.path.abspath(__file__)
        root = here
        for _ in range(5):  # tests/golden/ti_shim/taichi/_scalar.py -> repo root
            root = os.path.dirname(root)
        if root not in sys.path:
            sys.path.insert(0, root)
        from oracle import gs_oracle
        gs_oracle.build()
        lib = ctypes.CDLL(os.path.join(root, "oracle", "libgs_oracle.so"))
        lib.gso_expf.restype = ctypes.c_float
        lib.gso_expf.argtypes = [ctypes.c_float]
        _ORACLE = lib.gso_expf
    return _ORACLE


def _map(fn, x):
    if isinstance(x, Matrix):
        return x.map(fn)
    return fn(x)


def exp(x):
    return _map(lambda v: np.float32(_oracle_exp()(float(np.float32(v)))), x)


def sqrt(x):
    return _map(lambda v: np.sqrt(np.float32(v)), x)


def log(x):
    return _map(lambda v: np.float32(np.log(np.float32(v))), x)


def sin(x):
    return _map(lambda v: np.float32(np.sin(np.float32(v))), x)


def cos(x):
    return _map(lambda v: np.float32(np.cos(np.float32(v))), x)


def floor(x):
    return _map(lambda v: np.floor(np.float32(v)), x)
