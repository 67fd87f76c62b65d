"""Seeded synthetic problems of the BASELINE.json shapes.

Everything is generated once on the host with numpy (seed = 1000 + config number) and
the identical arrays are fed to the CUDA path and, in the tests, to the CPU oracle.
The pinhole camera constants are the ones the reference's own simulator uses
(reference matlab/simulate_vins.py:81-87); the visual-inertial trajectory follows the
conventions of the reference integrator (reference include/ba/Types.h:376-416:
w_world = R (w_meas + b_g), a_world = R (a_meas + b_a) - g).
"""
import numpy as np
from .problem import Problem

CAM = (198.969, 198.1284, 329.9368, 240.1017)


# ---- small SE3 toolbox (quaternions x y z w, Hamilton) ---------------------------------
def quat_to_R(q):
    q = np.asarray(q, float)
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1 - 2 * (y * y + z * z); R[..., 0, 1] = 2 * (x * y - w * z); R[..., 0, 2] = 2 * (x * z + w * y)
    R[..., 1, 0] = 2 * (x * y + w * z); R[..., 1, 1] = 1 - 2 * (x * x + z * z); R[..., 1, 2] = 2 * (y * z - w * x)
    R[..., 2, 0] = 2 * (x * z - w * y); R[..., 2, 1] = 2 * (y * z + w * x); R[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def R_to_quat(R):
    R = np.asarray(R, float)
    single = R.ndim == 2
    R = R.reshape(-1, 3, 3)
    q = np.empty((len(R), 4))
    for i, m in enumerate(R):
        t = np.trace(m)
        if t > 0:
            s = np.sqrt(t + 1.0) * 2
            q[i] = [(m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s, 0.25 * s]
        elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
            s = np.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
            q[i] = [0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s, (m[2, 1] - m[1, 2]) / s]
        elif m[1, 1] > m[2, 2]:
            s = np.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
            q[i] = [(m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s, (m[0, 2] - m[2, 0]) / s]
        else:
            s = np.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
            q[i] = [(m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s, (m[1, 0] - m[0, 1]) / s]
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q[0] if single else q


def so3_exp(w):
    w = np.asarray(w, float)
    th = np.linalg.norm(w, axis=-1, keepdims=True)
    k = np.where(th < 1e-12, 0.5, np.sin(th / 2) / np.where(th < 1e-12, 1.0, th))
    return np.concatenate([k * w, np.cos(th / 2)], axis=-1)


def quat_mul(a, b):
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx,
                     aw * bw - ax * bx - ay * by - az * bz], axis=-1)


def se3(q, t):
    return np.concatenate([np.asarray(q, float), np.asarray(t, float)], axis=-1)


def se3_inv(T):
    q = T[..., :4] * np.array([-1, -1, -1, 1.0])
    t = -np.einsum("...ij,...j->...i", quat_to_R(q), T[..., 4:])
    return se3(q, t)


def se3_mul(A, B):
    q = quat_mul(A[..., :4], B[..., :4])
    q = q / np.linalg.norm(q, axis=-1, keepdims=True)
    t = A[..., 4:] + np.einsum("...ij,...j->...i", quat_to_R(A[..., :4]), B[..., 4:])
    return se3(q, t)


def look_at_R(cam_pos, target, up=(0, 0, 1.0)):
    """Rotation world<-camera with camera z towards target, x right, y down."""
    z = target - cam_pos
    z = z / np.linalg.norm(z)
    x = np.cross(z, np.asarray(up, float)); x /= np.linalg.norm(x)
    y = np.cross(z, x)
    return np.stack([x, y, z], axis=1)


def project(T_wp, X, cam=CAM, t_vs=None):
    """Pixel of world points X (n,3) in poses T_wp (n,7); returns (z (n,2), depth (n,))."""
    T_ws = T_wp if t_vs is None else se3_mul(T_wp, np.broadcast_to(t_vs, T_wp.shape))
    R = quat_to_R(T_ws[..., :4])
    Xc = np.einsum("...ji,...j->...i", R, X - T_ws[..., 4:])
    fx, fy, u0, v0 = cam
    z = np.stack([fx * Xc[..., 0] / Xc[..., 2] + u0, fy * Xc[..., 1] / Xc[..., 2] + v0], axis=-1)
    return z, Xc[..., 2]


def _perturb_poses(rng, T, sig_t, sig_r, fixed=()):
    dq = so3_exp(rng.normal(0, sig_r, (len(T), 3)))
    dt = rng.normal(0, sig_t, (len(T), 3))
    for f in fixed:
        dq[f] = [0, 0, 0, 1]; dt[f] = 0
    q = quat_mul(T[:, :4], dq)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return se3(q, T[:, 4:] + dt)


def _ragged_window_picks(rng, n_items, k, lo, hi):
    """For each item i pick k[i] distinct increasing integers in [lo[i], hi[i]] (hi-lo+1 >= k)."""
    total = int(k.sum())
    owner = np.repeat(np.arange(n_items), k)
    start = np.concatenate([[0], np.cumsum(k)[:-1]])
    rank = np.arange(total) - start[owner]
    gaps = rng.exponential(1.0, total) + 1e-9
    csum = np.cumsum(gaps)
    seg0 = csum[start] - gaps[start]
    cum = csum - seg0[owner]                       # inclusive cumulative gap within the item
    tot = cum[start + k - 1] + rng.exponential(1.0, n_items)
    free = (hi - lo + 1 - k).astype(float)         # slack to distribute
    off = np.floor(cum / tot[owner] * (free[owner] + 1) - 1e-12).astype(np.int64)
    off = np.clip(off, 0, free[owner].astype(np.int64))
    return owner, lo[owner] + off + rank


# ---- C1: monocular inverse-depth BA  <double,1,6,0> ----------------------------------------
def make_c1(n_poses=50, n_landmarks=5000, obs_per_lm=20, seed=1001, t_vs=None, noise_px=0.5,
            n_inactive_poses=2, outlier_frac=0.0):
    rng = np.random.default_rng(seed)
    ang = np.linspace(0, 2 * np.pi, n_poses, endpoint=False)
    C = np.stack([7.5 * np.cos(ang), 7.5 * np.sin(ang), np.zeros(n_poses)], axis=1)
    Rw = np.stack([look_at_R(c, np.zeros(3)) for c in C])
    T_ws_true = se3(R_to_quat(Rw), C)                        # sensor poses
    tvs = np.array([0, 0, 0, 1, 0, 0, 0.0]) if t_vs is None else np.asarray(t_vs, float)
    T_true = se3_mul(T_ws_true, np.broadcast_to(se3_inv(tvs), T_ws_true.shape))   # vehicle poses
    X = rng.uniform(-5, 5, (n_landmarks, 3))
    # visibility: in front of the camera with margin
    lm_id, pose_id = [], []
    ref_pose = np.zeros(n_landmarks, np.int64)
    Xs = []
    for l in range(n_landmarks):
        _, depth = project(T_ws_true, np.broadcast_to(X[l], (n_poses, 3)))
        vis = np.nonzero(depth > 1.5)[0]
        if len(vis) < 3:
            continue
        ref_pose[l] = vis[0]
        others = vis[1:]
        k = min(obs_per_lm, len(others))
        pick = np.sort(rng.choice(others, k, replace=False))
        lm_id.append(np.full(k, len(Xs))); pose_id.append(pick)
        Xs.append(l)
    X = X[Xs]; ref_pose = ref_pose[Xs]
    lm_id = np.concatenate(lm_id); pose_id = np.concatenate(pose_id)
    z, _ = project(T_true[pose_id], X[lm_id], t_vs=tvs)
    z = z + rng.normal(0, noise_px, z.shape)
    if outlier_frac > 0:
        bad = rng.random(len(z)) < outlier_frac
        z[bad] += rng.normal(0, 30.0, (int(bad.sum()), 2))
    fixed = list(range(n_inactive_poses))
    T_init = _perturb_poses(rng, T_true, 0.02, 0.01, fixed=fixed)
    # initial landmark: true ray from the (perturbed) reference sensor, inverse depth x U(0.9,1.1)
    T_ws_init = se3_mul(T_init, np.broadcast_to(tvs, T_init.shape))
    Cr = T_ws_init[ref_pose, 4:]
    Ctrue = se3_mul(T_true, np.broadcast_to(tvs, T_true.shape))[ref_pose, 4:]
    u = rng.uniform(0.9, 1.1, (len(X), 1))
    Xw = Cr + (X - Ctrue) / u
    p = Problem(1, 6)
    p.AddCamera(CAM, tvs)
    active = np.ones(n_poses, np.uint8); active[:n_inactive_poses] = 0
    p.set_bulk(pose_t_wp=T_init, pose_v_w=np.zeros((n_poses, 3)), pose_b=np.zeros((n_poses, 6)), pose_active=active,
               pose_mask_used=np.zeros(n_poses, np.uint8), pose_mask=np.ones((n_poses, 6), np.uint8),
               lm_x_w=np.concatenate([Xw, np.ones((len(X), 1))], axis=1), lm_ref_pose=ref_pose,
               lm_ref_cam=np.zeros(len(X)), lm_active=np.ones(len(X), np.uint8),
               proj_z=z, proj_meas_pose=pose_id, proj_lm=lm_id, proj_cam=np.zeros(len(z)), proj_weight=np.ones(len(z)))
    p.truth = dict(T=T_true, X=X)
    return p


# ---- windowed generator shared by C2 (L=3) and C4 (L=1/3) ----------------------------------
def _windowed(rng, n_cams, n_points, window, track_len, lm_dim, noise_px=0.5, step=0.15, n_inactive=2, anchor_every=0):
    # smooth 3-D random walk along world +x with the camera looking sideways (+y): every
    # window of consecutive cameras has a real baseline, so depth is well constrained
    R0 = np.array([[1.0, 0, 0], [0, 0, 1.0], [0, -1.0, 0]])      # columns: cam x=+x, cam y=-z (down), cam z=+y
    yaw = np.cumsum(rng.normal(0, 0.004, n_cams))
    pitch = np.cumsum(rng.normal(0, 0.001, n_cams))
    heading = np.stack([np.cos(yaw) * np.cos(pitch), np.sin(yaw) * np.cos(pitch), np.sin(pitch)], axis=1)
    C = np.cumsum(heading * step + rng.normal(0, 0.01, (n_cams, 3)), axis=0)
    cy, sy = np.cos(yaw), np.sin(yaw)
    Rz = np.zeros((n_cams, 3, 3)); Rz[:, 0, 0] = cy; Rz[:, 0, 1] = -sy; Rz[:, 1, 0] = sy; Rz[:, 1, 1] = cy; Rz[:, 2, 2] = 1
    Rw = Rz @ R0
    T_true = se3(R_to_quat(Rw), C)
    center = np.sort(rng.integers(0, n_cams, n_points))
    half = window // 2
    lo = np.clip(center - half, 0, n_cams - window - 1)
    hi = lo + window
    center = lo + half                              # keep the point in front of its whole window
    k = np.asarray(track_len(n_points), np.int64)
    owner, cam = _ragged_window_picks(rng, n_points, k, lo, hi)
    # point in front of its centre camera
    local = np.stack([rng.uniform(-3, 3, n_points), rng.uniform(-2, 2, n_points), rng.uniform(8, 25, n_points)], axis=1)
    X = C[center] + np.einsum("nij,nj->ni", Rw[center], local)
    fixed = list(range(n_inactive))
    if anchor_every:        # surveyed keyframes along the chain: inactive poses at their true value
        fixed = sorted(set(fixed) | set(range(0, n_cams, anchor_every)))
    T_init = _perturb_poses(rng, T_true, 0.02, 0.002, fixed=fixed)
    active = np.ones(n_cams, np.uint8); active[fixed] = 0
    start = np.concatenate([[0], np.cumsum(k)[:-1]])
    ref_pose = cam[start]
    if lm_dim == 1:
        keep = np.ones(len(cam), bool); keep[start] = False   # the reference frame's own view is not a residual
        owner_r, cam_r = owner[keep], cam[keep]
        u = rng.uniform(0.9, 1.1, (n_points, 1))
        Xw = T_init[ref_pose, 4:] + (X - T_true[ref_pose, 4:]) / u
    else:
        owner_r, cam_r = owner, cam
        Xw = X + rng.normal(0, 0.05, X.shape)
        # The reference zeroes dz/dx_meas whenever the measuring pose is the landmark's reference pose, for every
        # LmSize (parallel_algos.h:86-99).  A 3-D world point has no privileged frame, so the caller names a pose
        # that is not optimised (the fixed first camera): no optimised Jacobian is dropped.
        ref_pose = np.zeros(n_points, np.int64)
    z, depth = project(T_true[cam_r], X[owner_r])
    assert depth.min() > 0.5, depth.min()
    z = z + rng.normal(0, noise_px, z.shape)
    p = Problem(lm_dim, 6)
    p.AddCamera(CAM)
    p.set_bulk(pose_t_wp=T_init, pose_v_w=np.zeros((n_cams, 3)), pose_b=np.zeros((n_cams, 6)), pose_active=active,
               pose_mask_used=np.zeros(n_cams, np.uint8), pose_mask=np.ones((n_cams, 6), np.uint8),
               lm_x_w=np.concatenate([Xw, np.ones((n_points, 1))], axis=1), lm_ref_pose=ref_pose,
               lm_ref_cam=np.zeros(n_points), lm_active=np.ones(n_points, np.uint8),
               proj_z=z, proj_meas_pose=cam_r, proj_lm=owner_r, proj_cam=np.zeros(len(z)), proj_weight=np.ones(len(z)))
    p.truth = dict(T=T_true, X=X)
    return p


def make_c2(n_cams=1778, n_points=993000, seed=1002, window=64):
    """BAL-shaped 3-D landmark BA <double,3,6,0>: track length 2 + Geometric, mean 5.04."""
    rng = np.random.default_rng(seed)
    window = min(window, n_cams - 2)

    def track(n):
        return np.minimum(2 + rng.geometric(1.0 / (1.0 + 3.04), n) - 1, min(40, window))
    return _windowed(rng, n_cams, n_points, window, track, 3)


def make_c4(n_cams=10000, n_points=2000000, seed=1004, lm_dim=1, obs_per_point=5, window=8, anchor_every=0):
    """Large BA <double,1,6,0> (or L=3): obs_per_point residuals per point from `window` consecutive cameras.
    `anchor_every` > 0 holds every anchor_every-th camera fixed (a 10k-camera monocular chain with an 8-camera
    window and two fixed cameras drifts along its scale gauge, cond(S) grows with the chain length)."""
    rng = np.random.default_rng(seed)
    k = obs_per_point + (1 if lm_dim == 1 else 0)
    return _windowed(rng, n_cams, n_points, window, lambda n: np.full(n, k), lm_dim, anchor_every=anchor_every)


# ---- C3: visual-inertial <double,1,15,0> -----------------------------------------------------
def make_c3(n_keyframes=500, n_landmarks=50000, obs_per_lm=40, imu_per_kf=10, seed=1003, pose_dim=15, lm_dim=1,
            kf_dt=0.1, noise_px=0.5):
    rng = np.random.default_rng(seed)
    g = np.array([0, 0, -9.806])
    R0 = np.array([[0, 0, 1.0], [-1, 0, 0], [0, -1, 0]])
    rad = 20.0
    om = 0.0126 / kf_dt        # 0.72 degrees of arc (0.25 m) per keyframe, camera looking radially outward

    def traj(t):
        t = np.asarray(t, float)
        th = om * t
        p = np.stack([rad * np.cos(th), rad * np.sin(th), 0.5 * np.sin(3 * th)], axis=-1)
        v = np.stack([-rad * om * np.sin(th), rad * om * np.cos(th), 1.5 * om * np.cos(3 * th)], axis=-1)
        a = np.stack([-rad * om * om * np.cos(th), -rad * om * om * np.sin(th), -4.5 * om * om * np.sin(3 * th)], axis=-1)
        yaw = th
        cy, sy = np.cos(yaw), np.sin(yaw)
        Rz = np.zeros(t.shape + (3, 3)); Rz[..., 0, 0] = cy; Rz[..., 0, 1] = -sy; Rz[..., 1, 0] = sy; Rz[..., 1, 1] = cy; Rz[..., 2, 2] = 1
        R = Rz @ R0
        w_world = np.zeros(t.shape + (3,)); w_world[..., 2] = om
        w_body = np.einsum("...ji,...j->...i", R, w_world)
        a_body = np.einsum("...ji,...j->...i", R, a + g)
        return p, v, R, w_body, a_body

    t_kf = np.arange(n_keyframes) * kf_dt
    p, v, R, _, _ = traj(t_kf)
    T_true = se3(R_to_quat(R), p)
    prob = Problem(lm_dim, pose_dim)
    prob.AddCamera(CAM)
    T_init = _perturb_poses(rng, T_true, 0.02, 0.002)
    v_init = v + rng.normal(0, 0.01, v.shape)
    # landmarks seen from a window of keyframes around a centre keyframe
    window = min(max(obs_per_lm + 20, 16), n_keyframes - 2)
    if lm_dim > 0 and n_landmarks > 0:
        center = np.sort(rng.integers(0, n_keyframes, n_landmarks))
        lo = np.clip(center - window // 2, 0, n_keyframes - window - 1)
        center = lo + window // 2
        k = np.full(n_landmarks, obs_per_lm + 1)
        owner, cam = _ragged_window_picks(rng, n_landmarks, k, lo, lo + window)
        local = np.stack([rng.uniform(-6, 6, n_landmarks), rng.uniform(-3, 3, n_landmarks), rng.uniform(15, 40, n_landmarks)], axis=1)
        X = p[center] + np.einsum("nij,nj->ni", R[center], local)
        start = np.arange(n_landmarks) * (obs_per_lm + 1)
        ref_pose = cam[start]
        keep = np.ones(len(cam), bool); keep[start] = False
        owner_r, cam_r = owner[keep], cam[keep]
        z, depth = project(T_true[cam_r], X[owner_r])
        ok = depth > 1.0
        assert ok.all(), float(depth.min())
        z = z + rng.normal(0, noise_px, z.shape)
        u = rng.uniform(0.9, 1.1, (n_landmarks, 1))
        Xw = T_init[ref_pose, 4:] + (X - T_true[ref_pose, 4:]) / u
        prob.set_bulk(lm_x_w=np.concatenate([Xw, np.ones((n_landmarks, 1))], axis=1), lm_ref_pose=ref_pose,
                      lm_ref_cam=np.zeros(n_landmarks), lm_active=np.ones(n_landmarks, np.uint8),
                      proj_z=z, proj_meas_pose=cam_r, proj_lm=owner_r, proj_cam=np.zeros(len(z)), proj_weight=np.ones(len(z)))
    prob.set_bulk(pose_t_wp=T_init, pose_v_w=v_init, pose_b=np.zeros((n_keyframes, 6)), pose_active=np.ones(n_keyframes, np.uint8),
                  pose_mask_used=np.zeros(n_keyframes, np.uint8), pose_mask=np.ones((n_keyframes, pose_dim), np.uint8))
    # IMU residuals between consecutive keyframes: imu_per_kf + 1 samples including both ends
    n_res = n_keyframes - 1
    ts = (t_kf[:-1, None] + np.linspace(0, kf_dt, imu_per_kf + 1)[None, :])          # (n_res, m)
    _, _, _, wb, ab = traj(ts)
    wb = wb + rng.normal(0, 1e-4, wb.shape); ab = ab + rng.normal(0, 2e-3, ab.shape)
    meas = np.concatenate([wb, ab, ts[..., None]], axis=-1).reshape(-1, 7)
    prob.set_bulk(imu_p1=np.arange(n_res), imu_p2=np.arange(1, n_res + 1),
                  imu_meas_start=np.arange(n_res + 1) * (imu_per_kf + 1), imu_meas=meas, imu_weight=np.ones(n_res))
    prob.gravity = g
    prob.truth = dict(T=T_true, v=v)
    return prob


# ---- C5: pose graph <double,0,6,0> -----------------------------------------------------------
def make_c5(n_poses=20000, seed=1005, closure_every=20, closure_span=100, prior_every=100, with_unary=True):
    rng = np.random.default_rng(seed)
    yaw = np.cumsum(rng.normal(0, 0.05, n_poses))
    C = np.cumsum(np.stack([np.cos(yaw), np.sin(yaw), np.zeros(n_poses)], axis=1) * 0.5, axis=0)
    q = so3_exp(np.stack([np.zeros(n_poses), np.zeros(n_poses), yaw], axis=1))
    T_true = se3(q, C)
    i1 = np.arange(n_poses - 1); i2 = i1 + 1
    c2 = np.arange(closure_every, n_poses, closure_every)
    c1 = np.maximum(0, c2 - rng.integers(2, closure_span + 1, len(c2)))
    p1 = np.concatenate([i1, c1]); p2 = np.concatenate([i2, c2])
    T12 = se3_mul(se3_inv(T_true[p1]), T_true[p2])
    T12 = _perturb_poses(rng, T12, 0.01, 0.002)
    prob = Problem(0, 6)
    T_init = _perturb_poses(rng, T_true, 0.05, 0.01)
    prob.set_bulk(pose_t_wp=T_init, pose_v_w=np.zeros((n_poses, 3)), pose_b=np.zeros((n_poses, 6)), pose_active=np.ones(n_poses, np.uint8),
                  pose_mask_used=np.zeros(n_poses, np.uint8), pose_mask=np.ones((n_poses, 6), np.uint8),
                  bin_p1=p1, bin_p2=p2, bin_t_12=T12, bin_cov_inv=np.tile(np.eye(6).reshape(1, 36), (len(p1), 1)),
                  bin_weight=np.ones(len(p1)), bin_use_rot=np.ones(len(p1), np.uint8))
    if with_unary:
        up = np.arange(0, n_poses, prior_every)
        prob.set_bulk(un_pose=up, un_t_wp=T_true[up], un_cov_inv=np.tile((np.eye(6) * 100.0).reshape(1, 36), (len(up), 1)),
                      un_use_rot=np.ones(len(up), np.uint8))
    prob.truth = dict(T=T_true)
    return prob


def make(name, scale=1.0, **kw):
    """Named configs of BASELINE.json; `scale` shrinks counts for tests / bounded CPU samples."""
    s = float(scale)
    if name == "c1":
        return make_c1(n_poses=max(6, int(50 * min(1.0, s * 4))), n_landmarks=max(20, int(5000 * s)), **kw)
    if name == "c2":
        return make_c2(n_cams=max(80, int(1778 * s)), n_points=max(100, int(993000 * s)), **kw)
    if name == "c3":
        return make_c3(n_keyframes=max(40, int(500 * s)), n_landmarks=max(50, int(50000 * s)), **kw)
    if name == "c4":
        kw.setdefault("anchor_every", 64)   # every 64th camera surveyed (fixed): see make_c4
        return make_c4(n_cams=max(20, int(10000 * s)), n_points=max(100, int(2000000 * s)), **kw)
    if name == "c5":
        return make_c5(n_poses=max(40, int(20000 * s)), **kw)
    raise ValueError(name)
