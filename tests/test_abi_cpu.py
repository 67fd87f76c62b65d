"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/ba_b200.h declares, mirrors the reference's bookkeeping rules, and refuses to compute
without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re
import numpy as np
import pytest

from ba_b200 import synth
from ba_b200.problem import Problem, UINT_MAX
from ba_b200 import capi, abi

ROOT = os.path.join(os.path.dirname(__file__), "..")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "ba_b200.h")).read()
    declared = sorted(set(re.findall(r"^(?:int|void|const char\*)\s+(bagpu_[a-z0-9_]+)\s*\(", hdr, re.M)))
    assert len(declared) >= 25
    lib = capi.lib()
    for name in declared:
        assert hasattr(lib, name), name
    for name in capi.EXPORTS:
        assert "bagpu_" + name in declared


def test_struct_sizes_match_header():
    # the ctypes mirrors must agree with the compiled structs: default options round-trip
    lib = capi.lib()
    o = abi.Options()
    lib.bagpu_default_options.restype = None
    lib.bagpu_default_options(C.byref(o))
    assert o.trust_region_size == -1 and o.dogleg_max_inner_iterations == 100
    assert o.error_change_threshold == 0.01 and o.param_change_threshold == 1e-3
    assert o.use_dogleg == 1 and o.use_robust_norm_for_proj_residuals == 1 and o.use_robust_norm_for_inertial_residuals == 0
    assert abs(o.gyro_sigma - 5.3088444e-5) < 1e-20 and abs(o.accel_bias_sigma - 1.2589254e-2) < 1e-18
    assert o.pcg_max_iterations == 1000


def test_create_rejects_unbuilt_instantiations():
    lib = capi.lib()
    h = C.c_void_p()
    for dims in (abi.Dims(2, 1, 6, 0, 0), abi.Dims(8, 1, 6, 5, 0), abi.Dims(8, 2, 6, 0, 0), abi.Dims(8, 1, 7, 0, 0)):
        assert lib.bagpu_create(C.byref(dims), C.byref(h)) == 5  # BAGPU_ERR_UNSUPPORTED
    # REAL_TYPE=float callers (scalar_bytes 4) are accepted: their host class converts at the boundary
    assert lib.bagpu_create(C.byref(abi.Dims(4, 1, 6, 0, 0)), C.byref(h)) == 0
    lib.bagpu_destroy(h)


def test_add_projection_residual_rules():
    # reference BundleAdjuster.h:484-501: a landmark's own reference view is not a residual (LmSize 1)
    p = Problem(1, 6)
    p.AddCamera(synth.CAM)
    a = p.AddPose([0, 0, 0, 1, 0, 0, 0]); b = p.AddPose([0, 0, 0, 1, 1, 0, 0])
    lm = p.AddLandmark([0, 0, 5, 1], a, 0)
    assert p.AddProjectionResidual([1, 2], a, lm, 0) == UINT_MAX
    assert p.AddProjectionResidual([1, 2], b, lm, 0) == 0
    p3 = Problem(3, 6)
    p3.AddCamera(synth.CAM)
    a = p3.AddPose([0, 0, 0, 1, 0, 0, 0])
    lm = p3.AddLandmark([0, 0, 5, 1], a, 0)
    assert p3.AddProjectionResidual([1, 2], a, lm, 0) == 0   # LmSize 3 keeps every measurement


def test_set_residuals_validates_ids():
    g = capi.BundleAdjusterGPU(1, 6)
    p = synth.make_c1(n_poses=6, n_landmarks=20, obs_per_lm=3)
    a = p.arrays()
    bad = Problem(1, 6)
    bad.AddCamera(synth.CAM)
    bad.set_bulk(**{k: a[k] for k in a if k.startswith(("pose_", "lm_"))})
    bad.set_bulk(proj_z=a["proj_z"][:1], proj_meas_pose=np.array([99]), proj_lm=a["proj_lm"][:1], proj_cam=a["proj_cam"][:1], proj_weight=a["proj_weight"][:1])
    with pytest.raises(RuntimeError):
        g.load(bad)


@pytest.mark.skipif(_has_gpu(), reason="checks the behaviour of a box without a CUDA device")
def test_no_cpu_fallback():
    g = capi.BundleAdjusterGPU(1, 6)
    g.load(synth.make_c1(n_poses=6, n_landmarks=20, obs_per_lm=3))
    with pytest.raises(RuntimeError) as e:
        g.solve(1)
    assert "no CUDA device" in str(e.value)
    for call in (g.build, lambda: g.dump("."), g.get_system):           # every computing entry point, not just solve
        with pytest.raises(RuntimeError) as e:
            call()
        assert "no CUDA device" in str(e.value)


def test_static_select_validates_its_arguments_before_touching_the_device():
    # bagpu_select_kth_static: sizes and pointers are checked on the host first (no device needed for the refusals)
    L = capi.lib()
    f = L.bagpu_select_kth_static
    f.restype = C.c_int
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
    L.bagpu_last_error.restype = C.c_char_p
    fake = C.c_void_p(256)                                              # never dereferenced by a refused call
    hdr = open(os.path.join(ROOT, "include", "ba_b200.h")).read()
    assert re.search(r"#define\s+BAGPU_SELECT_STATIC_MAX_N\s+\(1ull\s*<<\s*31\)", hdr)
    for n, k, word in (((1 << 31) + 1, 0, b"BAGPU_SELECT_STATIC_MAX_N"), (0, 0, b"n = 0"), (5, 5, b"below n")):
        assert f(fake, n, k, fake, None) == 1                           # BAGPU_ERR_INVALID
        assert word in L.bagpu_last_error()
    assert f(None, 5, 0, fake, None) == 1 and f(fake, 5, 0, None, None) == 1
    assert b"null device pointer" in L.bagpu_last_error()
    if not _has_gpu():
        assert f(fake, 5, 0, fake, None) != 0                           # a valid request without a device fails loudly
        assert b"bagpu_select_kth_static" in L.bagpu_last_error()


def test_patterns_from_ids_need_no_device():
    # block patterns are symbolic (host): compare against a numpy derivation from the id arrays
    p = synth.make_c1(n_poses=8, n_landmarks=40, obs_per_lm=4, n_inactive_poses=2)
    g = capi.BundleAdjusterGPU(1, 6).load(p)
    a = p.arrays()
    opt = np.cumsum(a["pose_active"]) - 1
    act = a["pose_active"].astype(bool)
    cols = {}
    for m, l in zip(a["proj_meas_pose"], a["proj_lm"]):
        r = a["lm_ref_pose"][l]
        for x, y in ((m, m), (r, r), (m, r)):
            if act[x] and act[y]:
                i, j = sorted((opt[x], opt[y]))
                cols.setdefault(j, set()).add(i)
    cp, ri = g.get_pattern(abi.PATTERN_U)
    n = int(act.sum())
    assert len(cp) == n + 1
    for j in range(n):
        assert list(ri[cp[j]:cp[j + 1]]) == sorted(cols.get(j, set()))
