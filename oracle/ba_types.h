// TEST INFRASTRUCTURE — CPU oracle (see minimath.h header; PARITY UNPINNED).
// Restatement of reference include/ba/Types.h: state records (Types.h:41-321) and the
// RK4 IMU pre-integration with Jacobians and covariance (Types.h:324-738).
#pragma once
#include "ba_utils.h"
#include <vector>

namespace orc {

static const double IMU_GYRO_SIGMA = 5.3088444e-5;        // Types.h:33
static const double IMU_GYRO_BIAS_SIGMA = 1.4125375e-4;   // Types.h:34
static const double IMU_ACCEL_SIGMA = 0.001883649;        // Types.h:35
static const double IMU_ACCEL_BIAS_SIGMA = 1.2589254e-2;  // Types.h:36

// Types.h:41-71
struct Pose {
  SE3 t_wp;
  Vec3 v_w;
  Vec6 b;
  std::vector<bool> param_mask;
  bool is_param_mask_used = false;
  bool is_active = true;
  uint32_t id = 0, opt_id = 0;
  std::vector<int> proj_residuals, inertial_residuals, binary_residuals, unary_residuals;
  std::vector<SE3> t_sw;
  const SE3& GetTsw(uint32_t cam_id, const std::vector<LinearCamera>& rig) {
    while (t_sw.size() <= cam_id) t_sw.push_back((t_wp * rig[t_sw.size()].t_vs).inverse());
    return t_sw[cam_id];
  }
};

// Types.h:73-89
template <int LM>
struct Landmark {
  Vec2 z_ref;
  Vec4 x_s, x_w;
  std::vector<int> proj_residuals;
  uint32_t num_outlier_residuals = 0, id = 0, opt_id = 0, ref_pose_id = 0, ref_cam_id = 0;
  bool is_active = true, is_reliable = true;
  Mat<(LM > 0 ? LM : 1), (LM > 0 ? LM : 1)> jtj;
};

// Types.h:182-220
struct ImuPose {
  SE3 t_wp; Vec3 v_w, w_w; double time = 0;
  ImuPose() {}
  ImuPose(const SE3& T, const Vec3& v, const Vec3& w, double tm) : t_wp(T), v_w(v), w_w(w), time(tm) {}
};
// Types.h:222-244
struct ImuMeasurement { Vec3 w, a; double time = 0; };

// Types.h:246-253
struct ResidualBase { uint32_t residual_id = 0, residual_offset = 0; double mahalanobis_distance = 0, weight = 0, orig_weight = 0; };
// Types.h:255-266
struct UnaryResidual : ResidualBase {
  uint32_t pose_id = 0; SE3 t_wp; Mat<6, 6> dz_dx; Vec6 residual; Mat<6, 6> cov_inv, cov_inv_sqrt; bool use_rotation = true; };
// Types.h:268-280
struct BinaryResidual : ResidualBase {
  uint32_t x1_id = 0, x2_id = 0; SE3 t_12; Mat<6, 6> dz_dx1, dz_dx2; Vec6 residual; Mat<6, 6> cov_inv, cov_inv_sqrt; bool use_rotation = true; };
// Types.h:282-298
template <int LM>
struct ProjectionResidual : ResidualBase {
  Vec2 z; uint32_t x_meas_id = 0, x_ref_id = 0, landmark_id = 0, cam_id = 0;
  Mat<2, (LM > 0 ? LM : 1)> dz_dlm; Mat<2, 6> dz_dx_meas, dz_dx_ref; Vec2 residual; bool is_conditioning = false; };

typedef Mat<6, 6> Diag6;  // Eigen::DiagonalMatrix<Scalar,6> imu_.r stored dense

// Types.h:300-738
template <int D>
struct ImuResidual : ResidualBase {
  uint32_t pose1_id = 0, pose2_id = 0;
  std::vector<ImuMeasurement> measurements;
  std::vector<ImuPose> poses;
  Mat<D, D> dz_dx1, dz_dx2, cov, cov_inv, cov_inv_sqrt;
  Mat<D, 6> dz_db;
  Mat<D, 1> residual;
  Mat<10, 6> dintegration_db;
  Mat<10, 10> c_integration;
  bool covariance_computed = false;

  // Types.h:324-373
  static ImuPose IntegratePose(const ImuPose& pose, const Mat<9, 1>& k, const double dt,
                               Mat<10, 9>* pdy_dk = 0, Mat<4, 4>* pdy_dy = 0) {
    Vec3 kv, kw, ka;
    for (int i = 0; i < 3; ++i) { kv[i] = k[i]; kw[i] = k[3 + i]; ka[i] = k[6 + i]; }
    const SO3 r_v2_v1(SO3::exp(kw * dt));
    ImuPose y = pose;
    y.t_wp.translation() += kv * dt;
    const Quat q = r_v2_v1.unit_quaternion() * pose.t_wp.so3().unit_quaternion();
    y.t_wp.so3() = SO3::raw(q);  // memcpy, no normalisation (Types.h:338-339)
    y.v_w += ka * dt;
    if (pdy_dk != 0) {
      pdy_dk->setZero();
      pdy_dk->setBlock(0, 0, Mat3::Identity() * dt);
      Mat<4, 3> dq = dq1q2_dq1(pose.t_wp.so3().unit_quaternion()) * dq_exp_dw(kw * dt) * dt;
      pdy_dk->setBlock(3, 3, dq);
      pdy_dk->setBlock(7, 6, Mat3::Identity() * dt);
    }
    if (pdy_dy != 0) *pdy_dy = dq1q2_dq2(r_v2_v1.unit_quaternion());
    return y;
  }

  // Types.h:376-416
  static Mat<9, 1> GetPoseDerivative(const ImuPose& pose, const Vec3& g_w, const ImuMeasurement& z_start,
                                     const ImuMeasurement& z_end, const Vec3& bg, const Vec3& ba,
                                     const double dt, Mat<9, 6>* dk_db = 0, Mat<9, 10>* dk_dx = 0) {
    double alpha = (z_end.time - (z_start.time + dt)) / (z_end.time - z_start.time);
    Vec3 zg = z_start.w * alpha + z_end.w * (1.0 - alpha);
    Vec3 za = z_start.a * alpha + z_end.a * (1.0 - alpha);
    Mat<9, 1> deriv;
    Vec3 wv = pose.t_wp.so3().Adj() * (zg + bg);
    Vec3 av = pose.t_wp.so3() * (za + ba) - g_w;
    for (int i = 0; i < 3; ++i) { deriv[i] = pose.v_w[i]; deriv[3 + i] = wv[i]; deriv[6 + i] = av[i]; }
    if (dk_db != 0) {
      dk_db->setZero();
      dk_db->setBlock(3, 0, pose.t_wp.so3().Adj());
      dk_db->setBlock(6, 3, pose.t_wp.so3().matrix());
    }
    if (dk_dx != 0) {
      dk_dx->setZero();
      dk_dx->setBlock(0, 7, Mat3::Identity());
      dk_dx->setBlock(3, 3, dqx_dq(pose.t_wp.so3().unit_quaternion(), zg) + dqx_dq(pose.t_wp.so3().unit_quaternion(), bg));
      dk_dx->setBlock(6, 3, dqx_dq(pose.t_wp.so3().unit_quaternion(), za) + dqx_dq(pose.t_wp.so3().unit_quaternion(), ba));
    }
    return deriv;
  }

  // Types.h:419-643 (euler_covariance = true is the only mode the loop uses)
  static ImuPose IntegrateImu(const ImuPose& pose, const ImuMeasurement& z_start, const ImuMeasurement& z_end,
                              const Vec3& bg, const Vec3& ba, const Vec3& g, Mat<10, 6>* dy_db_ptr = 0,
                              Mat<10, 10>* dy_dpose_ptr = 0, Mat<10, 10>* c_prior = 0, const Diag6* r = 0) {
    double dt = z_end.time - z_start.time;
    if (dt == 0) return pose;
    ImuPose res = pose;
    Mat<9, 1> k;
    if (dy_db_ptr != 0 && dy_dpose_ptr != 0 && r != 0) {
      const Mat3 I3 = Mat3::Identity();
      Mat<10, 6>& dy_db = *dy_db_ptr;
      Mat<10, 10>& dy_dy0 = *dy_dpose_ptr;
      Mat<9, 6> dk_db; Mat<9, 10> dk_dy; Mat<10, 9> dy_dk; Mat<4, 4> dy_dy;
      dy_db.setZero();
      dy_dy0.setIdentity();
      const Mat<9, 1> k1 = GetPoseDerivative(pose, g, z_start, z_end, bg, ba, 0, &dk_db, &dk_dy);
      const Mat<9, 6> dk1_db = dk_db;
      const Mat<9, 10> dk1_dy = dk_dy;
      const ImuPose y1 = IntegratePose(pose, k1, dt * 0.5, &dy_dk, &dy_dy);
      dy_db = dy_dk * dk1_db;
      dy_dy0 = dy_dk * dk1_dy;
      dy_dy0.addBlock(0, 0, I3); dy_dy0.addBlock(7, 7, I3); dy_dy0.addBlock(3, 3, dy_dy);

      const Mat<9, 1> k2 = GetPoseDerivative(y1, g, z_start, z_end, bg, ba, dt / 2, &dk_db, &dk_dy);
      const Mat<9, 6> dk2_db = dk_db + dk_dy * dy_db;
      const Mat<9, 10> dk2_dy = dk_dy * dy_dy0;
      const ImuPose y2 = IntegratePose(pose, k2, dt * 0.5, &dy_dk, &dy_dy);
      dy_db = dy_dk * dk2_db;
      dy_dy0 = dy_dk * dk2_dy;
      dy_dy0.addBlock(0, 0, I3); dy_dy0.addBlock(7, 7, I3); dy_dy0.addBlock(3, 3, dy_dy);

      const Mat<9, 1> k3 = GetPoseDerivative(y2, g, z_start, z_end, bg, ba, dt / 2, &dk_db, &dk_dy);
      const Mat<9, 6> dk3_db = dk_db + dk_dy * dy_db;
      const Mat<9, 10> dk3_dy = dk_dy * dy_dy0;
      const ImuPose y3 = IntegratePose(pose, k3, dt, &dy_dk, &dy_dy);
      dy_db = dy_dk * dk3_db;
      dy_dy0 = dy_dk * dk3_dy;
      dy_dy0.addBlock(0, 0, I3); dy_dy0.addBlock(7, 7, I3); dy_dy0.addBlock(3, 3, dy_dy);

      const Mat<9, 1> k4 = GetPoseDerivative(y3, g, z_start, z_end, bg, ba, dt, &dk_db, &dk_dy);
      const Mat<9, 6> dk4_db = dk_db + dk_dy * dy_db;
      const Mat<9, 10> dk4_dy = dk_dy * dy_dy0;

      k = (k1 + 2 * k2 + 2 * k3 + k4);
      const Mat<9, 6> dk_total_db = dk1_db + 2 * dk2_db + 2 * dk3_db + dk4_db;
      const Mat<9, 10> dk_total_dy = dk1_dy + 2 * dk2_dy + 2 * dk3_dy + dk4_dy;

      res = IntegratePose(pose, k, dt / 6.0, &dy_dk, &dy_dy);
      dy_db = dy_dk * dk_total_db;
      dy_dy0 = dy_dk * dk_total_dy;
      dy_dy0.addBlock(0, 0, I3); dy_dy0.addBlock(7, 7, I3); dy_dy0.addBlock(3, 3, dy_dy);

      if (c_prior != 0) {  // Types.h:601-606
        const Mat<10, 10> c_prop = dy_dy0 * (*c_prior) * dy_dy0.transpose();
        *c_prior = c_prop + dy_db * (*r) * dy_db.transpose();
      }
    } else {
      const Mat<9, 1> k1 = GetPoseDerivative(pose, g, z_start, z_end, bg, ba, 0);
      const ImuPose y1 = IntegratePose(pose, k1, dt * 0.5);
      const Mat<9, 1> k2 = GetPoseDerivative(y1, g, z_start, z_end, bg, ba, dt / 2);
      const ImuPose y2 = IntegratePose(pose, k2, dt * 0.5);
      const Mat<9, 1> k3 = GetPoseDerivative(y2, g, z_start, z_end, bg, ba, dt / 2);
      const ImuPose y3 = IntegratePose(pose, k3, dt);
      const Mat<9, 1> k4 = GetPoseDerivative(y3, g, z_start, z_end, bg, ba, dt);
      k = (k1 + 2 * k2 + 2 * k3 + k4);
      res = IntegratePose(pose, k, dt / 6.0);
    }
    for (int i = 0; i < 3; ++i) res.w_w[i] = k[3 + i];
    res.time = z_end.time;
    return res;
  }

  // Types.h:646-738
  static ImuPose IntegrateResidual(ImuPose pose, const std::vector<ImuMeasurement>& measurements, const Vec3& bg,
                                   const Vec3& ba, const Vec3& g, std::vector<ImuPose>& poses,
                                   Mat<10, 6>* dpose_db = 0, Mat<10, 10>* dpose_dpose = 0,
                                   Mat<10, 10>* c_res = 0, const Diag6* r = 0) {
    const ImuMeasurement* prev_meas = 0;
    poses.clear();
    poses.reserve(measurements.size() + 1);
    poses.push_back(pose);
    if (dpose_db != 0) dpose_db->setZero();
    if (dpose_dpose != 0) dpose_dpose->setIdentity();
    Mat<10, 6> dy_db;
    Mat<10, 10> dy_dy;
    for (const ImuMeasurement& meas : measurements) {
      if (prev_meas != 0) {
        if ((dpose_db != 0 || dpose_dpose != 0) && r != 0) {
          pose = IntegrateImu(pose, *prev_meas, meas, bg, ba, g, &dy_db, &dy_dy, c_res, r);
          if (dpose_db != 0) (*dpose_db) = dy_db + dy_dy * (*dpose_db);
          if (dpose_dpose != 0) *dpose_dpose = dy_dy * (*dpose_dpose);
        } else {
          pose = IntegrateImu(pose, *prev_meas, meas, bg, ba, g);
        }
        poses.push_back(pose);
      }
      prev_meas = &meas;
    }
    return pose;
  }
};

}  // namespace orc
