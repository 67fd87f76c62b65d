"""Loader for the CUDA product library ba_b200/libba_b200.so (C ABI: include/ba_b200.h).

There is no CPU fallback: if the library has not been built the import of the solver
fails loudly, and on a box without a CUDA device every compute entry point returns
BAGPU_ERR_NO_DEVICE (raised here as RuntimeError).
"""
import ctypes as C
import os
import subprocess
from .abi import Binding

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libba_b200.so")
_LIB = None

EXPORTS = [
    "create", "destroy", "last_error", "set_options", "default_options", "set_cameras", "set_poses",
    "set_landmarks", "set_proj_residuals", "set_unary_residuals", "set_binary_residuals",
    "set_imu_residuals", "solve", "build", "get_poses", "get_landmarks", "get_proj_residuals", "get_other_residuals",
    "get_trace", "get_deltas", "get_pattern", "get_system", "dump", "nccl_unique_id", "comm_init",
    "get_phase_ms", "sync_state", "host_alloc", "host_free", "symbolic_tile_fill", "pack_warp_tasks", "build_work_items", "select_kth_static", "chol_task_plan",
]


def build(force=False):
    """Compile the sm_100a kernels in-tree (nvcc cross-compiles without a GPU)."""
    src = os.path.join(_HERE, "csrc")
    if force:
        subprocess.check_call(["make", "-C", src, "-s", "clean"])
    subprocess.check_call(["make", "-C", src, "-s", "-j8"])
    return SO_PATH


def build_cpp_driver():
    """tests/cpp/host_api_driver: a C++ program that drives the drop-in class include/ba/BundleAdjuster.h the way the
    reference's applications drive ba::BundleAdjuster (used by the tests and by bench.py's e2e_cpp leg)."""
    root = os.path.dirname(_HERE)
    src = os.path.join(root, "tests", "cpp", "host_api_driver.cpp")
    exe = os.path.join(root, "tests", "cpp", "build", "host_api_driver")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    deps = [src, os.path.join(root, "include", "ba", "BundleAdjuster.h"), os.path.join(root, "include", "ba", "MiniTypes.h"),
            os.path.join(root, "include", "ba_b200.h"), SO_PATH]
    if os.path.exists(exe) and all(os.path.getmtime(d) <= os.path.getmtime(exe) for d in deps if os.path.exists(d)):
        return exe
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(root, "include"), src, "-o", exe,
                           "-L" + _HERE, "-lba_b200", "-Wl,-rpath," + _HERE])
    return exe


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO_PATH):
            raise ImportError("ba_b200/libba_b200.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(the CUDA extension is the only implementation; there is no fallback)")
        _LIB = C.CDLL(SO_PATH)
    return _LIB


class BundleAdjusterGPU(Binding):
    """ba::BundleAdjuster<double, LmSize, PoseSize, 0, false> on the GPU, through the C ABI."""

    def __init__(self, lm_dim, pose_dim, options=None):
        super().__init__(lib(), "bagpu_", lm_dim, pose_dim, options)

    def comm_init(self, rank, world, unique_id):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        self._check(self._f("comm_init")(self.h, C.c_int(rank), C.c_int(world), buf))

    @staticmethod
    def nccl_unique_id():
        buf = (C.c_char * 128)()
        rc = lib().bagpu_nccl_unique_id(buf)
        if rc != 0:
            f = lib().bagpu_last_error
            f.restype = C.c_char_p
            raise RuntimeError("bagpu_nccl_unique_id: %s" % f().decode())
        return bytes(buf)


class _PinnedBlock:
    """One page-locked host allocation (bagpu_host_alloc); freed when the last numpy view goes away."""

    def __init__(self, nbytes):
        self.p = C.c_void_p()
        rc = lib().bagpu_host_alloc(C.c_size_t(max(1, nbytes)), C.byref(self.p))
        if rc != 0 or not self.p:
            raise MemoryError("bagpu_host_alloc(%d) failed" % nbytes)
        self.nbytes = nbytes

    def __del__(self):
        try:
            lib().bagpu_host_free(self.p)
        except Exception:
            pass


def pinned_like(a):
    """Copy of the numpy array `a` in page-locked host memory (the host buffers of an end-to-end Solve()).
    The block lives until release_pinned()."""
    import numpy as np
    a = np.ascontiguousarray(a)
    blk = _PinnedBlock(a.nbytes)
    buf = (C.c_char * max(1, a.nbytes)).from_address(blk.p.value)
    out = np.frombuffer(buf, dtype=a.dtype, count=a.size).reshape(a.shape)
    out[...] = a
    _PINNED_KEEP.append(blk)
    return out


def release_pinned():
    """Free every block handed out by pinned_like (the arrays must not be used afterwards)."""
    del _PINNED_KEEP[:]


_PINNED_KEEP = []
