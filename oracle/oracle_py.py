"""TEST INFRASTRUCTURE — ctypes loader for the CPU oracle (oracle/liboracle.so).

PARITY UNPINNED (see oracle/minimath.h).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np
from ba_b200.abi import Binding

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    srcs.append(os.path.join(_HERE, "..", "include", "ba_b200.h"))
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
    return _LIB


class Oracle(Binding):
    def __init__(self, lm_dim, pose_dim, options=None):
        super().__init__(lib(), "baoracle_", lm_dim, pose_dim, options)


def _arr(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def math_call(name, ins, out_shapes):
    f = getattr(lib(), "baoracle_math_" + name)
    f.restype = None
    ins = [_arr(x) for x in ins]
    outs = [np.zeros(s) for s in out_shapes]
    f(*[x.ctypes.data_as(C.c_void_p) for x in ins + outs])
    return outs[0] if len(outs) == 1 else outs
