"""ctypes mirror of include/ba_b200.h (the C ABI of the hot path).

Struct layouts follow include/ba_b200.h field by field; the enum values follow the
reference's ba::OptimizationResult (reference include/ba/BundleAdjuster.h:38-46).
The same class binds any library exporting these entry points under a prefix
(``bagpu_`` for the CUDA product library).
"""
import ctypes as C
import os
import numpy as np

PH_NAMES = ["residual", "median", "accumulate", "other_residuals", "schur", "allreduce", "solve", "step", "total"]
RESULT_NAMES = ["Success", "ErrorIncreased", "ErrorChangeBelowThreshold",
                "ParamChangeBelowThreshold", "FactorizationError", "SolverError"]
PATTERN_U, PATTERN_W, PATTERN_S = 0, 1, 2


class Dims(C.Structure):
    _fields_ = [("scalar_bytes", C.c_int32), ("lm_dim", C.c_int32), ("pose_dim", C.c_int32),
                ("calib_dim", C.c_int32), ("do_tvs", C.c_int32)]


class Options(C.Structure):
    _fields_ = [("trust_region_size", C.c_double),
                ("gyro_sigma", C.c_double), ("accel_sigma", C.c_double),
                ("gyro_bias_sigma", C.c_double), ("accel_bias_sigma", C.c_double),
                ("projection_outlier_threshold", C.c_double),
                ("error_change_threshold", C.c_double),
                ("param_change_threshold", C.c_double),
                ("dogleg_max_inner_iterations", C.c_uint32),
                ("apply_results", C.c_int32), ("use_dogleg", C.c_int32),
                ("use_triangular_matrices", C.c_int32), ("use_sparse_solver", C.c_int32),
                ("write_reduced_camera_matrix", C.c_int32),
                ("calculate_calibration_marginals", C.c_int32),
                ("use_per_pose_cam_params", C.c_int32),
                ("regularize_biases_in_batch", C.c_int32),
                ("enable_auto_regularization", C.c_int32),
                ("use_robust_norm_for_proj_residuals", C.c_int32),
                ("use_robust_norm_for_inertial_residuals", C.c_int32),
                ("calculate_inertial_covariance_once", C.c_int32),
                ("gravity", C.c_double * 3),
                ("root_pose_id", C.c_uint32),
                ("reduced_solver", C.c_int32), ("record_deltas", C.c_int32),
                ("pcg_tolerance", C.c_double), ("pcg_max_iterations", C.c_uint32),
                ("keep_on_device", C.c_int32)]


class Summary(C.Structure):
    _fields_ = [("num_proj_residuals", C.c_uint32), ("num_inertial_residuals", C.c_uint32),
                ("num_cond_proj_residuals", C.c_uint32), ("num_cond_inertial_residuals", C.c_uint32),
                ("cond_proj_error", C.c_double), ("cond_inertial_error", C.c_double),
                ("proj_error", C.c_double), ("inertial_error", C.c_double),
                ("unary_error", C.c_double), ("binary_error", C.c_double),
                ("delta_norm", C.c_double), ("pre_solve_norm", C.c_double),
                ("post_solve_norm", C.c_double), ("trust_region_size", C.c_double),
                ("result", C.c_int32), ("iterations", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class IterTrace(C.Structure):
    _fields_ = [("proj_error_build", C.c_double), ("binary_error_build", C.c_double),
                ("unary_error_build", C.c_double), ("inertial_error_build", C.c_double),
                ("median_maha", C.c_double),
                ("sd_numerator", C.c_double), ("sd_denominator", C.c_double), ("sd_norm", C.c_double),
                ("gn_norm", C.c_double), ("trust_region_after", C.c_double),
                ("pre_solve_norm", C.c_double), ("post_solve_norm", C.c_double),
                ("delta_norm", C.c_double), ("delta_p_norm", C.c_double), ("delta_l_norm", C.c_double),
                ("inner_iterations", C.c_uint32), ("result", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def _p(a, dtype):
    """numpy array (or None) -> typed pointer, keeping C-contiguity and dtype exact."""
    if a is None:
        return None
    assert a.dtype == dtype and a.flags["C_CONTIGUOUS"], (a.dtype, dtype)
    return a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


def _u32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint32)


def _u8(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint8)


class Binding:
    """One solver handle behind the C ABI of include/ba_b200.h."""

    def __init__(self, lib, prefix, lm_dim, pose_dim, options=None):
        self.lib, self.prefix = lib, prefix
        self.lm_dim, self.pose_dim = int(lm_dim), int(pose_dim)
        self.dims = Dims(8, lm_dim, pose_dim, 0, 0)
        self.h = C.c_void_p()
        self._check(self._f("create")(C.byref(self.dims), C.byref(self.h)))
        self.n_poses = self.n_landmarks = self.n_proj = 0
        self.n_unary = self.n_binary = self.n_imu = 0
        self.n_active_poses = self.n_active_landmarks = 0
        self.options = options if options is not None else self.default_options()
        self.set_options(self.options)

    # -- plumbing -------------------------------------------------------------------
    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def _check(self, status):
        if status != 0:
            f = self._f("last_error")
            f.restype = C.c_char_p
            msg = f()
            raise RuntimeError("%s status %d: %s" % (self.prefix, status, msg.decode() if msg else ""))

    def close(self):
        if self.h:
            f = self._f("destroy")
            f.restype = None
            f(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def default_options(self):
        o = Options()
        f = self._f("default_options")
        f.restype = None
        f(C.byref(o))
        return o

    def set_options(self, o):
        self.options = o
        self._check(self._f("set_options")(self.h, C.byref(o)))

    # -- upload ---------------------------------------------------------------------
    def load(self, prob):
        """Upload a ba_b200.problem.Problem (SoA arrays) through the per-category setters."""
        assert prob.lm_dim == self.lm_dim and prob.pose_dim == self.pose_dim
        a = prob.arrays()
        self._keep = a  # keep the arrays alive during the calls
        n = len(a["cam_params"])
        self._check(self._f("set_cameras")(self.h, C.c_uint32(n), _p(a["cam_params"], np.float64), _p(a["cam_t_vs"], np.float64)))
        n = len(a["pose_t_wp"])
        self.n_poses = n
        self.n_active_poses = int(a["pose_active"].sum())
        self._check(self._f("set_poses")(self.h, C.c_uint32(n), _p(a["pose_t_wp"], np.float64), _p(a["pose_v_w"], np.float64),
                                         _p(a["pose_b"], np.float64), _p(a["pose_active"], np.uint8),
                                         _p(a["pose_mask_used"], np.uint8), _p(a["pose_mask"], np.uint8)))
        n = len(a["lm_x_w"])
        self.n_landmarks = n
        self.n_active_landmarks = int(a["lm_active"].sum()) if n else 0
        self._check(self._f("set_landmarks")(self.h, C.c_uint32(n), _p(a["lm_x_w"], np.float64), _p(a["lm_ref_pose"], np.uint32),
                                             _p(a["lm_ref_cam"], np.uint32), _p(a["lm_active"], np.uint8)))
        n = len(a["proj_z"])
        self.n_proj = n
        self._check(self._f("set_proj_residuals")(self.h, C.c_uint32(n), _p(a["proj_z"], np.float64), _p(a["proj_meas_pose"], np.uint32),
                                                  _p(a["proj_lm"], np.uint32), _p(a["proj_cam"], np.uint32), _p(a["proj_weight"], np.float64)))
        n = len(a["un_pose"])
        self.n_unary = n
        self._check(self._f("set_unary_residuals")(self.h, C.c_uint32(n), _p(a["un_pose"], np.uint32), _p(a["un_t_wp"], np.float64),
                                                   _p(a["un_cov_inv"], np.float64), _p(a["un_use_rot"], np.uint8)))
        n = len(a["bin_p1"])
        self.n_binary = n
        self._check(self._f("set_binary_residuals")(self.h, C.c_uint32(n), _p(a["bin_p1"], np.uint32), _p(a["bin_p2"], np.uint32),
                                                    _p(a["bin_t_12"], np.float64), _p(a["bin_cov_inv"], np.float64),
                                                    _p(a["bin_weight"], np.float64), _p(a["bin_use_rot"], np.uint8)))
        n = len(a["imu_p1"])
        self.n_imu = n
        self._check(self._f("set_imu_residuals")(self.h, C.c_uint32(n), _p(a["imu_p1"], np.uint32), _p(a["imu_p2"], np.uint32),
                                                 _p(a["imu_meas_start"], np.uint32), _p(a["imu_meas"], np.float64),
                                                 _p(a["imu_weight"], np.float64)))
        return self

    # -- the hot path -----------------------------------------------------------------
    def solve(self, max_iter, gn_damping=1.0, error_increase_allowed=False):
        s = Summary()
        self._check(self._f("solve")(self.h, C.c_uint32(max_iter), C.c_double(gn_damping),
                                     C.c_int(1 if error_increase_allowed else 0), C.byref(s)))
        return s

    def build(self):
        ms = C.c_float(0)
        self._check(self._f("build")(self.h, C.byref(ms)))
        return ms.value

    # -- results ------------------------------------------------------------------------
    def get_poses(self):
        n, d = self.n_poses, self.pose_dim
        t = np.zeros((n, 7)); v = np.zeros((n, 3)); b = np.zeros((n, 6))
        mu = np.zeros(n, np.uint8); m = np.zeros((n, d), np.uint8)
        self._check(self._f("get_poses")(self.h, _p(t, np.float64), _p(v, np.float64), _p(b, np.float64), _p(mu, np.uint8), _p(m, np.uint8)))
        return dict(t_wp=t, v_w=v, b=b, mask_used=mu, mask=m)

    def get_landmarks(self, out=None):
        """Refined landmarks.  `out` (optional): dict with preallocated C-contiguous arrays x_w, x_s (n x 4 float64),
        is_reliable (uint8), num_outlier_residuals (uint32) — e.g. page-locked ones — that are filled in place."""
        n = self.n_landmarks
        if out is not None:
            xw, xs, rel, nout = out["x_w"], out["x_s"], out["is_reliable"], out["num_outlier_residuals"]
            assert xw.shape == (n, 4) and xs.shape == (n, 4) and rel.shape == (n,) and nout.shape == (n,)
            assert all(a.flags.c_contiguous for a in (xw, xs, rel, nout))
            assert xw.dtype == np.float64 and xs.dtype == np.float64 and rel.dtype == np.uint8 and nout.dtype == np.uint32
        else:
            xw = np.empty((n, 4)); xs = np.empty((n, 4)); rel = np.empty(n, np.uint8); nout = np.empty(n, np.uint32)
        self._check(self._f("get_landmarks")(self.h, _p(xw, np.float64), _p(xs, np.float64), _p(rel, np.uint8), _p(nout, np.uint32)))
        return dict(x_w=xw, x_s=xs, is_reliable=rel, num_outlier_residuals=nout)

    def get_proj_residuals(self):
        n = self.n_proj
        r = np.zeros((n, 2)); w = np.zeros(n); m = np.zeros(n)
        self._check(self._f("get_proj_residuals")(self.h, _p(r, np.float64), _p(w, np.float64), _p(m, np.float64)))
        return dict(residual=r, weight=w, maha=m)

    def get_other_residuals(self):
        """residual / mahalanobis_distance / cov_inv of the unary, binary and inertial residuals as Solve() left them."""
        nu, nb, ni, D = self.n_unary, self.n_binary, self.n_imu, self.pose_dim
        ur = np.zeros((nu, 6)); um = np.zeros(nu); uc = np.zeros((nu, 6, 6)); br = np.zeros((nb, 6)); bm = np.zeros(nb)
        ir = np.zeros((ni, D)); im = np.zeros(ni); ic = np.zeros((ni, D, D))
        self._check(self._f("get_other_residuals")(self.h, *[_p(x, np.float64) for x in (ur, um, uc, br, bm, ir, im, ic)]))
        return dict(un_residual=ur, un_maha=um, un_cov_inv=uc, bin_residual=br, bin_maha=bm, imu_residual=ir, imu_maha=im, imu_cov_inv=ic)

    def get_trace(self):
        n = C.c_uint32(0)
        self._check(self._f("get_trace")(self.h, C.c_uint32(0), None, C.byref(n)))
        arr = (IterTrace * max(1, n.value))()
        self._check(self._f("get_trace")(self.h, C.c_uint32(n.value), arr, C.byref(n)))
        return [arr[i].as_dict() for i in range(n.value)]

    def get_deltas(self, it):
        dp = np.zeros(self.n_active_poses * self.pose_dim)
        dl = np.zeros(self.n_active_landmarks * max(1, self.lm_dim))
        self._check(self._f("get_deltas")(self.h, C.c_uint32(it), _p(dp, np.float64), _p(dl, np.float64)))
        return dp, dl

    def get_pattern(self, which):
        nc, nz = C.c_uint32(0), C.c_uint32(0)
        self._check(self._f("get_pattern")(self.h, C.c_int(which), C.byref(nc), C.byref(nz), None, None))
        cp = np.zeros(nc.value + 1, np.uint32); ri = np.zeros(max(1, nz.value), np.uint32)
        self._check(self._f("get_pattern")(self.h, C.c_int(which), C.byref(nc), C.byref(nz), _p(cp, np.uint32), _p(ri, np.uint32)))
        return cp, ri[:nz.value]

    def get_system(self, dense_s=True):
        n = self.n_active_poses * self.pose_dim
        L = max(1, self.lm_dim)
        s = np.zeros((n, n)) if dense_s else None     # column-major storage: s[j, i] = S(i, j)
        rsc = np.zeros(n); rp = np.zeros(n)
        rl = np.zeros(self.n_active_landmarks * L); vi = np.zeros((self.n_active_landmarks, L, L))
        self._check(self._f("get_system")(self.h, _p(s, np.float64), _p(rsc, np.float64), _p(rp, np.float64),
                                          _p(rl, np.float64) if self.lm_dim else None, _p(vi, np.float64) if self.lm_dim else None))
        return dict(S=(s.T.copy() if dense_s else None), rhs_sc=rsc, rhs_p=rp, rhs_l=rl, v_inv=vi)

    def dump(self, directory):
        """s.txt / rhs.txt (the oracle adds j_pr.txt, r_pr.txt, j_l.txt) of the last build, in the reference's CSV
        format (reference src/BundleAdjuster.cpp:600-620); read them back with ba_b200.dumpio.read_dump."""
        self._check(self._f("dump")(self.h, C.c_char_p(os.fsencode(directory))))

    def sync_state(self):
        self._check(self._f("sync_state")(self.h))

    def get_phase_ms(self):
        ms = (C.c_float * len(PH_NAMES))()
        launches = C.c_uint64(0)
        self._check(self._f("get_phase_ms")(self.h, ms, C.byref(launches)))
        return {k: ms[i] for i, k in enumerate(PH_NAMES)}, launches.value
