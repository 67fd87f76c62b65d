"""Host-side problem container: the bookkeeping half of ba::BundleAdjuster.

Mirrors the builder calls of the reference class (reference
include/ba/BundleAdjuster.h:177-546: Init / AddCamera / AddPose / AddLandmark /
AddUnaryConstraint / AddBinaryConstraint / AddProjectionResidual / AddImuResidual)
with the same names, argument meaning, returned ids and skip rules, but stores
structure-of-arrays that go straight through the C ABI (include/ba_b200.h) instead of
the reference's AoS records.  SE3 arguments are 7-vectors qx qy qz qw tx ty tz.

This is the Python twin of the C++ template in include/ba/BundleAdjuster.h (this repo);
the tests use it the way the reference's applications use the C++ class.
"""
import numpy as np

UINT_MAX = 0xFFFFFFFF


class Problem:
    def __init__(self, lm_dim=1, pose_dim=6):
        assert lm_dim in (0, 1, 3) and pose_dim in (6, 9, 15)
        self.lm_dim, self.pose_dim = lm_dim, pose_dim
        self.Init()

    # reference BundleAdjuster.h:177-237
    def Init(self, num_poses=0, num_measurements=0, num_landmarks=0):
        self.cam_params, self.cam_t_vs = [], []
        self.pose_t_wp, self.pose_v_w, self.pose_b, self.pose_active = [], [], [], []
        self.pose_mask_used, self.pose_mask = [], []
        self.lm_x_w, self.lm_ref_pose, self.lm_ref_cam, self.lm_active = [], [], [], []
        self.proj_z, self.proj_meas_pose, self.proj_lm, self.proj_cam, self.proj_weight = [], [], [], [], []
        self.un_pose, self.un_t_wp, self.un_cov_inv, self.un_use_rot = [], [], [], []
        self.bin_p1, self.bin_p2, self.bin_t_12, self.bin_cov_inv, self.bin_weight, self.bin_use_rot = [], [], [], [], [], []
        self.imu_p1, self.imu_p2, self.imu_meas_start, self.imu_meas, self.imu_weight = [], [], [0], [], []
        self._bulk = {}

    # reference BundleAdjuster.h:259-263 (calibu linear camera: fx fy u0 v0, Pose() = T_vs)
    def AddCamera(self, params4, t_vs7=(0, 0, 0, 1, 0, 0, 0)):
        self.cam_params.append(np.asarray(params4, float))
        self.cam_t_vs.append(np.asarray(t_vs7, float))
        return len(self.cam_params)

    # reference BundleAdjuster.h:267-323
    def AddPose(self, t_wp7, is_active=True, v_w=(0, 0, 0), b=(0, 0, 0, 0, 0, 0)):
        self.pose_t_wp.append(np.asarray(t_wp7, float))
        self.pose_v_w.append(np.asarray(v_w, float))
        self.pose_b.append(np.asarray(b, float))
        self.pose_active.append(1 if is_active else 0)
        self.pose_mask_used.append(0)
        self.pose_mask.append(np.ones(self.pose_dim, np.uint8))
        return len(self.pose_t_wp) - 1

    # reference BundleAdjuster.h:608-631 (including its rotation quirk: indices 2,4,5)
    def RegularizePose(self, pose_id, translation, gravity, bias, rotation, gravity_dim=None):
        m = np.ones(self.pose_dim, np.uint8)
        if translation:
            m[0:3] = 0
        if rotation:
            m[2] = m[4] = m[5] = 0
        if gravity:
            m[gravity_dim] = 0
        if bias and self.pose_dim >= 15:
            m[9:15] = 0
        self.pose_mask_used[pose_id] = 1
        self.pose_mask[pose_id] = m

    # reference BundleAdjuster.h:326-367
    def AddLandmark(self, x_w4, ref_pose_id, ref_cam_id, is_active=True):
        assert ref_pose_id < len(self.pose_t_wp)
        self.lm_x_w.append(np.asarray(x_w4, float))
        self.lm_ref_pose.append(ref_pose_id)
        self.lm_ref_cam.append(ref_cam_id)
        self.lm_active.append(1 if is_active else 0)
        return len(self.lm_x_w) - 1

    # reference BundleAdjuster.h:377-407
    def AddUnaryConstraint(self, pose_id, t_wv7, covariance, use_rotation=True):
        assert pose_id < len(self.pose_t_wp)
        cov = np.array(covariance, float).reshape(6, 6)
        if not use_rotation:
            cov[3, 3] = cov[4, 4] = cov[5, 5] = 1.0
        self.un_pose.append(pose_id)
        self.un_t_wp.append(np.asarray(t_wv7, float))
        self.un_cov_inv.append(np.linalg.inv(cov))
        self.un_use_rot.append(1 if use_rotation else 0)
        return len(self.un_pose) - 1

    # reference BundleAdjuster.h:410-456
    def AddBinaryConstraint(self, pose1_id, pose2_id, t_12_7, covariance=None, weight=1.0, use_rotation=True):
        assert pose1_id < len(self.pose_t_wp) and pose2_id < len(self.pose_t_wp)
        cov = np.eye(6) if covariance is None else np.array(covariance, float).reshape(6, 6)
        self.bin_p1.append(pose1_id); self.bin_p2.append(pose2_id)
        self.bin_t_12.append(np.asarray(t_12_7, float))
        self.bin_cov_inv.append(np.linalg.inv(cov))
        self.bin_weight.append(weight)
        self.bin_use_rot.append(1 if use_rotation else 0)
        return len(self.bin_p1) - 1

    # reference BundleAdjuster.h:459-513
    def AddProjectionResidual(self, z2, meas_pose_id, landmark_id, cam_id, weight=1.0):
        assert landmark_id < len(self.lm_x_w) and meas_pose_id < len(self.pose_t_wp)
        ref = self.lm_ref_pose[landmark_id]
        diff_poses = meas_pose_id != ref
        if not (diff_poses or cam_id != self.lm_ref_cam[landmark_id] or self.lm_dim != 1):
            return UINT_MAX  # the reference returns (uint32_t)-1 and stores nothing
        self.proj_z.append(np.asarray(z2, float))
        self.proj_meas_pose.append(meas_pose_id)
        self.proj_lm.append(landmark_id)
        self.proj_cam.append(cam_id)
        self.proj_weight.append(weight)
        return len(self.proj_z) - 1

    # reference BundleAdjuster.h:516-546; imu_meas rows = wx wy wz ax ay az time
    def AddImuResidual(self, pose1_id, pose2_id, imu_meas, weight=1.0):
        assert pose1_id < len(self.pose_t_wp) and pose2_id < len(self.pose_t_wp)
        m = np.asarray(imu_meas, float).reshape(-1, 7)
        self.imu_p1.append(pose1_id); self.imu_p2.append(pose2_id)
        self.imu_meas.extend(list(m))
        self.imu_meas_start.append(len(self.imu_meas))
        self.imu_weight.append(weight)
        return len(self.imu_p1) - 1

    # -- bulk paths for the large synthetic configurations (no per-item Python loop) ---------
    def set_bulk(self, **arrays):
        """Replace whole categories by ready numpy arrays (keys as in arrays())."""
        self._bulk.update(arrays)

    def arrays(self):
        d = self.pose_dim

        def f64(lst, shape):
            return np.ascontiguousarray(np.array(lst, dtype=np.float64).reshape(shape))

        def u32(lst):
            return np.ascontiguousarray(np.array(lst, dtype=np.uint32).reshape(-1))

        def u8(lst, shape=(-1,)):
            return np.ascontiguousarray(np.array(lst, dtype=np.uint8).reshape(shape))

        a = dict(
            cam_params=f64(self.cam_params, (-1, 4)), cam_t_vs=f64(self.cam_t_vs, (-1, 7)),
            pose_t_wp=f64(self.pose_t_wp, (-1, 7)), pose_v_w=f64(self.pose_v_w, (-1, 3)), pose_b=f64(self.pose_b, (-1, 6)),
            pose_active=u8(self.pose_active), pose_mask_used=u8(self.pose_mask_used), pose_mask=u8(self.pose_mask, (-1, d)),
            lm_x_w=f64(self.lm_x_w, (-1, 4)), lm_ref_pose=u32(self.lm_ref_pose), lm_ref_cam=u32(self.lm_ref_cam), lm_active=u8(self.lm_active),
            proj_z=f64(self.proj_z, (-1, 2)), proj_meas_pose=u32(self.proj_meas_pose), proj_lm=u32(self.proj_lm),
            proj_cam=u32(self.proj_cam), proj_weight=f64(self.proj_weight, (-1,)),
            un_pose=u32(self.un_pose), un_t_wp=f64(self.un_t_wp, (-1, 7)), un_cov_inv=f64(self.un_cov_inv, (-1, 36)), un_use_rot=u8(self.un_use_rot),
            bin_p1=u32(self.bin_p1), bin_p2=u32(self.bin_p2), bin_t_12=f64(self.bin_t_12, (-1, 7)),
            bin_cov_inv=f64(self.bin_cov_inv, (-1, 36)), bin_weight=f64(self.bin_weight, (-1,)), bin_use_rot=u8(self.bin_use_rot),
            imu_p1=u32(self.imu_p1), imu_p2=u32(self.imu_p2), imu_meas_start=u32(self.imu_meas_start),
            imu_meas=f64(self.imu_meas, (-1, 7)), imu_weight=f64(self.imu_weight, (-1,)),
        )
        for k, v in self._bulk.items():
            ref = a[k]
            a[k] = np.ascontiguousarray(v, dtype=ref.dtype).reshape((-1,) + ref.shape[1:])
        return a

    @property
    def counts(self):
        a = self.arrays()
        return dict(poses=len(a["pose_t_wp"]), landmarks=len(a["lm_x_w"]), proj=len(a["proj_z"]),
                    unary=len(a["un_pose"]), binary=len(a["bin_p1"]), imu=len(a["imu_p1"]))


def write_driver_problem(path, p):
    """Problem file of tests/cpp/host_api_driver.cpp (the C++ drop-in class driven like the reference's applications):
    int32 L, D, then per category a uint32 count followed by the raw arrays."""
    import struct
    a = p.arrays()
    g = getattr(p, "gravity", np.zeros(3))
    with open(path, "wb") as f:
        f.write(struct.pack("<ii", p.lm_dim, p.pose_dim))

        def blk(n, *arrs):
            f.write(struct.pack("<I", n))
            for x in arrs:
                f.write(np.ascontiguousarray(x).tobytes())
        blk(len(a["cam_params"]), a["cam_params"], a["cam_t_vs"])
        blk(len(a["pose_t_wp"]), a["pose_t_wp"], a["pose_v_w"], a["pose_b"], a["pose_active"])
        blk(len(a["lm_x_w"]), a["lm_x_w"], a["lm_ref_pose"], a["lm_ref_cam"], a["lm_active"])
        blk(len(a["proj_z"]), a["proj_z"], a["proj_meas_pose"], a["proj_lm"], a["proj_cam"], a["proj_weight"])
        # the C++ API takes covariances (it inverts them, reference BundleAdjuster.h:398,444)
        ucov = np.array([np.linalg.inv(c.reshape(6, 6)).reshape(36) for c in a["un_cov_inv"]]).reshape(-1, 36)
        blk(len(a["un_pose"]), a["un_pose"], a["un_t_wp"], ucov, a["un_use_rot"])
        bcov = np.array([np.linalg.inv(c.reshape(6, 6)).reshape(36) for c in a["bin_cov_inv"]]).reshape(-1, 36)
        blk(len(a["bin_p1"]), a["bin_p1"], a["bin_p2"], a["bin_t_12"], bcov, a["bin_weight"], a["bin_use_rot"])
        blk(len(a["imu_p1"]), a["imu_p1"], a["imu_p2"], a["imu_meas_start"], a["imu_meas"], a["imu_weight"])
        f.write(np.asarray(g, np.float64).tobytes())
