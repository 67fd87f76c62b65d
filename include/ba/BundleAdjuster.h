/*
 * ba/BundleAdjuster.h — host side of the B200 drop-in for ba::BundleAdjuster.
 *
 * Same class template, method names, argument meaning, returned ids, skip rules and result
 * records as the reference's public header (reference include/ba/BundleAdjuster.h:111-753).
 * What differs is below the surface:
 *   * the builder calls (AddPose / AddLandmark / AddProjectionResidual / ...) append straight into
 *     structure-of-arrays buffers that go through the C ABI unchanged —
 *     the reference keeps AoS records with embedded Jacobians and per-pose std::vector<int>
 *     lists (reference include/ba/Types.h:41-321);
 *   * Solve() hands those buffers to include/ba_b200.h (bagpu_*), the whole optimisation loop
 *     (reference src/BundleAdjuster.cpp:278-1803) runs on the GPU, and the refined state is
 *     copied back into the same host containers so the getters and a second Solve() behave as in
 *     the reference.
 * There is no CPU implementation behind this header: without libba_b200.so / a CUDA device
 * Solve() reports SolverError and prints bagpu_last_error().
 *
 * Value types: Eigen / Sophus / Calibu do not exist in this image; ba/MiniTypes.h carries the
 * minimum (see there).  GetLandmarkObj / GetProjectionResidual / GetImuResidual return records by
 * value (assembled from the SoA buffers) instead of references into AoS vectors.
 */
#ifndef BA_B200_BUNDLEADJUSTER_H
#define BA_B200_BUNDLEADJUSTER_H

#include <cassert>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <memory>
#include <new>
#include <vector>

#include "../ba_b200.h"
#include "MiniTypes.h"

namespace ba {
namespace detail {
/* Allocator of the observation buffers: page-locked through the C ABI (bagpu_host_alloc), so Solve() hands the
 * GPU arrays it can DMA from directly.  Init(num_poses, num_measurements, num_landmarks) reserves them once, like
 * the reference's reserve calls (reference BundleAdjuster.h:226-233). */
template <typename T>
struct HostAllocator {
  typedef T value_type;
  HostAllocator() {}
  template <typename U> HostAllocator(const HostAllocator<U>&) {}
  T* allocate(size_t n) {
    void* p = nullptr;
    if (bagpu_host_alloc(n * sizeof(T), &p) != 0 || !p) throw std::bad_alloc();
    return static_cast<T*>(p);
  }
  void deallocate(T* p, size_t) { bagpu_host_free(p); }
  template <typename U> bool operator==(const HostAllocator<U>&) const { return true; }
  template <typename U> bool operator!=(const HostAllocator<U>&) const { return false; }
};
template <typename T> using HostVector = std::vector<T, HostAllocator<T>>;
}  // namespace detail
}  // namespace ba

/* reference include/ba/Types.h:33-36 */
#ifndef IMU_GYRO_SIGMA
#define IMU_GYRO_SIGMA 5.3088444e-5
#define IMU_GYRO_BIAS_SIGMA 1.4125375e-4
#define IMU_ACCEL_SIGMA 0.001883649
#define IMU_ACCEL_BIAS_SIGMA 1.2589254e-2
#endif

/* ---- stand-in for the slice of Calibu 0.1 the class touches (reference BundleAdjuster.h:225,
 * 259-263; parallel_algos.h:44-62): a camera = linear intrinsics fx fy u0 v0 + T_vs -------------- */
namespace calibu {
template <typename Scalar>
class CameraInterface {
 public:
  typedef ba::mini::SE3Group<Scalar> SE3t;
  virtual ~CameraInterface() {}
  virtual std::vector<Scalar> GetParams() const = 0;
  virtual void SetParams(const std::vector<Scalar>& p) = 0;
  const SE3t& Pose() const { return t_vs_; }
  void SetPose(const SE3t& t) { t_vs_ = t; }
 protected:
  SE3t t_vs_;
};
template <typename Scalar>
class LinearCamera : public CameraInterface<Scalar> {
 public:
  LinearCamera(Scalar fx, Scalar fy, Scalar u0, Scalar v0, const ba::mini::SE3Group<Scalar>& t_vs = ba::mini::SE3Group<Scalar>())
      : p_{fx, fy, u0, v0} { this->t_vs_ = t_vs; }
  std::vector<Scalar> GetParams() const override { return p_; }
  void SetParams(const std::vector<Scalar>& p) override { p_ = p; }
 private:
  std::vector<Scalar> p_;
};
template <typename Scalar>
struct Rig {
  std::vector<std::shared_ptr<CameraInterface<Scalar>>> cameras_;
  void AddCamera(std::shared_ptr<CameraInterface<Scalar>> c) { cameras_.push_back(c); }
  size_t NumCams() const { return cameras_.size(); }
};
}  // namespace calibu

namespace ba {
constexpr int kTrustRegionAuto = -1.0;   /* reference BundleAdjuster.h:29 (quirk Q5: an int) */

/* reference BundleAdjuster.h:37-45 */
enum OptimizationResult { Success, ErrorIncreased, ErrorChangeBelowThreshold, ParamChangeBelowThreshold, FactorizationError, SolverError };

/* reference BundleAdjuster.h:47-69 */
template <typename Scalar = double>
struct SolutionSummary {
  uint32_t num_proj_residuals = 0, num_inertial_residuals = 0, num_cond_proj_residuals = 0, num_cond_inertial_residuals = 0;
  Scalar cond_proj_error = 0, cond_inertial_error = 0, proj_error_ = 0, inertial_error = 0;
  Scalar delta_norm = 0, pre_solve_norm = 0, post_solve_norm = 0;
  std::vector<Scalar> calibration_marginals;   /* CalibSize > 0 only: not in this build */
  OptimizationResult result = Success;
  bool IsResultGood() { return (result != SolverError) && (result != FactorizationError); }
};

/* reference BundleAdjuster.h:71-106 (same names and defaults) */
template <typename Scalar = double>
struct Options {
  Scalar trust_region_size = kTrustRegionAuto;
  Scalar gyro_sigma = IMU_GYRO_SIGMA;
  Scalar accel_sigma = IMU_ACCEL_SIGMA;
  Scalar gyro_bias_sigma = IMU_GYRO_BIAS_SIGMA;
  Scalar accel_bias_sigma = IMU_ACCEL_BIAS_SIGMA;
  Scalar projection_outlier_threshold = 1.0;
  Scalar error_change_threshold = 0.01;
  Scalar param_change_threshold = 1e-3;
  uint32_t dogleg_max_inner_iterations = 100;
  bool apply_results = true;
  bool use_dogleg = true;
  bool use_triangular_matrices = true;
  bool use_sparse_solver = true;
  bool write_reduced_camera_matrix = false;
  bool calculate_calibration_marginals = false;
  bool use_per_pose_cam_params = false;
  bool regularize_biases_in_batch = true;
  bool enable_auto_regularization = true;
  bool use_robust_norm_for_proj_residuals = true;
  bool use_robust_norm_for_inertial_residuals = false;
  bool calculate_inertial_covariance_once = false;
};

/* state / result records (reference include/ba/Types.h:41-321), without the embedded Jacobians */
template <typename Scalar = double>
struct PoseT {
  mini::SE3Group<Scalar> t_wp;
  mini::Matrix<Scalar, 3, 1> v_w;
  mini::Matrix<Scalar, 6, 1> b;
  std::vector<Scalar> cam_params;
  std::vector<bool> param_mask;
  bool is_param_mask_used = false;
  bool is_active = true;
  int external_id = -1;
  uint32_t id = 0, opt_id = UINT_MAX;
  double time = -1;
};
template <typename Scalar = double, int LmSize = 1>
struct LandmarkT {
  mini::Matrix<Scalar, 2, 1> z_ref;
  mini::Matrix<Scalar, 4, 1> x_s, x_w;
  int external_id = -1;
  uint32_t num_outlier_residuals = 0, num_proj_residuals = 0;
  uint32_t id = 0, opt_id = UINT_MAX, ref_pose_id = 0, ref_cam_id = 0;
  bool is_active = true, is_reliable = true;
};
template <typename Scalar = double>
struct ImuMeasurementT {
  ImuMeasurementT() {}
  ImuMeasurementT(const mini::Matrix<Scalar, 3, 1>& w_, const mini::Matrix<Scalar, 3, 1>& a_, const double time_) : w(w_), a(a_), time(time_) {}
  mini::Matrix<Scalar, 3, 1> w, a;
  double time = 0;
  ImuMeasurementT operator*(const Scalar& rhs) const { return ImuMeasurementT(w * rhs, a * rhs, time); }
  ImuMeasurementT operator+(const ImuMeasurementT& rhs) const { return ImuMeasurementT(w + rhs.w, a + rhs.a, time); }
};
template <typename Scalar = double, int LmSize = 1>
struct ProjectionResidualT {
  static const uint32_t kResSize = 2;
  uint32_t residual_id = 0, residual_offset = 0;
  Scalar mahalanobis_distance = 0, weight = 0, orig_weight = 1;
  mini::Matrix<Scalar, 2, 1> z, residual;
  uint32_t x_meas_id = 0, x_ref_id = 0, landmark_id = 0, cam_id = 0;
  bool is_conditioning = false;   /* always false in the reference as well (quirk Q2) */
};
template <typename Scalar = double>
struct ImuCalibrationT {
  mini::SE3Group<Scalar> t_vs;
  mini::Matrix<Scalar, 3, 1> b_g, b_a, g_vec;
  mini::Matrix<Scalar, 6, 1> r, r_b;   /* diagonals */
};

template <typename Scalar = double, int LmSize = 1, int PoseSize = 6, int CalibSize = 0, bool DoTvs = false>
class BundleAdjuster {
 public:
  int debug_level_threshold;
  int debug_level;

  static constexpr uint32_t kPrPoseDim = 6;
  static constexpr uint32_t kLmDim = LmSize;
  static constexpr uint32_t kPoseDim = PoseSize;
  static constexpr uint32_t kCalibDim = CalibSize + (DoTvs ? 6 : 0);
  static constexpr bool kVelInState = (kPoseDim >= 9);
  static constexpr bool kBiasInState = (kPoseDim >= 15);

  typedef PoseT<Scalar> Pose;
  typedef LandmarkT<Scalar, LmSize> Landmark;
  typedef ProjectionResidualT<Scalar, LmSize> ProjectionResidual;
  typedef ImuMeasurementT<Scalar> ImuMeasurement;
  typedef ImuCalibrationT<Scalar> ImuCalibration;
  typedef mini::Matrix<Scalar, 2, 1> Vector2t;
  typedef mini::Matrix<Scalar, 3, 1> Vector3t;
  typedef mini::Matrix<Scalar, 4, 1> Vector4t;
  typedef mini::Matrix<Scalar, 6, 1> Vector6t;
  typedef mini::Matrix<Scalar, 6, 6> Matrix6t;
  typedef std::vector<Scalar> VectorXt;
  typedef mini::SE3Group<Scalar> SE3t;

  BundleAdjuster() : debug_level_threshold(0), debug_level(0) { Init(Options<Scalar>()); }
  ~BundleAdjuster() { if (h_) bagpu_destroy(h_); }
  BundleAdjuster(const BundleAdjuster&) = delete;
  BundleAdjuster& operator=(const BundleAdjuster&) = delete;

  /* reference BundleAdjuster.h:177-237 */
  void Init(const Options<Scalar>& options, uint32_t num_poses = 0, uint32_t num_measurements = 0, uint32_t num_landmarks = 0,
            const SE3t& t_vs = SE3t()) {
    num_poses = std::max(1u, num_poses); num_measurements = std::max(1u, num_measurements); num_landmarks = std::max(1u, num_landmarks);
    options_ = options;
    root_pose_id_ = 0; num_active_poses_ = 0; num_active_landmarks_ = 0;
    imu_.t_vs = t_vs;
    for (int k = 0; k < 3; ++k) {
      imu_.r[k] = options.gyro_sigma * options.gyro_sigma; imu_.r[3 + k] = options.accel_sigma * options.accel_sigma;
      imu_.r_b[k] = options.gyro_bias_sigma * options.gyro_bias_sigma; imu_.r_b[3 + k] = options.accel_bias_sigma * options.accel_bias_sigma;
    }
    rig_.reset(new calibu::Rig<Scalar>());
    poses_.clear(); poses_.reserve(num_poses);
    lm_x_w_.clear(); lm_x_s_.clear(); lm_ref_pose_.clear(); lm_ref_cam_.clear(); lm_active_.clear(); lm_reliable_.clear();
    lm_outliers_.clear(); lm_num_res_.clear(); lm_external_.clear(); lm_opt_.clear(); lm_z_ref_.clear();
    lm_x_w_.reserve(4 * (size_t)num_landmarks); lm_x_s_.reserve(4 * (size_t)num_landmarks); lm_ref_pose_.reserve(num_landmarks);
    lm_outliers_.reserve(num_landmarks); lm_reliable_.reserve(num_landmarks);
    pr_z_.clear(); pr_pose_.clear(); pr_lm_.clear(); pr_cam_.clear(); pr_w0_.clear(); pr_res_.clear(); pr_w_.clear(); pr_maha_.clear();
    pr_z_.reserve(2 * (size_t)num_measurements); pr_pose_.reserve(num_measurements); pr_lm_.reserve(num_measurements);
    pr_cam_.reserve(num_measurements); pr_w0_.reserve(num_measurements);
    un_pose_.clear(); un_t_.clear(); un_cov_inv_.clear(); un_rot_.clear();
    bin_p1_.clear(); bin_p2_.clear(); bin_t_.clear(); bin_cov_inv_.clear(); bin_w_.clear(); bin_rot_.clear();
    imu_p1_.clear(); imu_p2_.clear(); imu_ms_.assign(1, 0u); imu_meas_.clear(); imu_w_.clear();
    conditioning_proj_residuals_.clear(); conditioning_inertial_residuals_.clear();
    summary_ = SolutionSummary<Scalar>();
    proj_error_ = binary_error_ = unary_error_ = inertial_error_ = 0;
    if (h_) { bagpu_destroy(h_); h_ = nullptr; }   /* trust_region_size_ restarts from the options (BundleAdjuster.h:193) */
    dirty_ = true;
  }

  void SetGravity(const Vector3t& g) { imu_.g_vec = g; }          /* reference :243-252 (kGravityInCalib == false) */
  Vector3t GetGravity() const { return imu_.g_vec; }

  /* reference :259-263 */
  uint32_t AddCamera(std::shared_ptr<calibu::CameraInterface<Scalar>> cam) { rig_->AddCamera(cam); dirty_ = true; return (uint32_t)rig_->NumCams(); }

  /* reference :267-274 */
  uint32_t AddPose(const SE3t& t_wp, const bool is_active = true, const double time = -1, const int external_id = -1) {
    return AddPose(t_wp, VectorXt(5, Scalar(0)), Vector3t::Zero(), Vector6t::Zero(), is_active, time, external_id);
  }
  /* reference :292-323 */
  uint32_t AddPose(const SE3t& t_wv, const VectorXt cam_params, const Vector3t& v_w, const Vector6t& b, const bool is_active = true,
                   const double time = -1, const int external_id = -1) {
    Pose pose;
    pose.external_id = external_id; pose.time = time; pose.t_wp = t_wv; pose.v_w = v_w; pose.b = b; pose.cam_params = cam_params;
    pose.is_active = is_active; pose.is_param_mask_used = false;
    pose.id = (uint32_t)poses_.size();
    pose.opt_id = is_active ? num_active_poses_++ : UINT_MAX;
    poses_.push_back(pose);
    dirty_ = true;
    return pose.id;
  }

  /* reference :326-367 */
  uint32_t AddLandmark(const Vector4t& x_w, const uint32_t ref_pose_id, const uint32_t ref_cam_id, const bool is_active, const int external_id = -1) {
    assert(ref_pose_id < poses_.size());
    const uint32_t id = (uint32_t)lm_ref_pose_.size();
    for (int k = 0; k < 4; ++k) { lm_x_w_.push_back((double)x_w[k]); lm_x_s_.push_back(0.0); }
    lm_ref_pose_.push_back(ref_pose_id); lm_ref_cam_.push_back(ref_cam_id); lm_active_.push_back(is_active ? 1 : 0);
    lm_reliable_.push_back(1); lm_outliers_.push_back(0); lm_num_res_.push_back(0); lm_external_.push_back(external_id);
    lm_opt_.push_back(is_active ? num_active_landmarks_++ : UINT_MAX);
    lm_z_ref_.push_back(0.0); lm_z_ref_.push_back(0.0);
    dirty_ = true;
    return id;
  }

  /* reference :377-407 */
  uint32_t AddUnaryConstraint(const uint32_t pose_id, const SE3t& t_wv, Matrix6t covariance, bool use_rotation = true) {
    assert(pose_id < poses_.size());
    if (!use_rotation) covariance(3, 3) = covariance(4, 4) = covariance(5, 5) = 1.0;
    const Matrix6t ci = covariance.inverse();
    un_pose_.push_back(pose_id);
    push_se3(un_t_, t_wv); push_rowmajor66(un_cov_inv_, ci); un_rot_.push_back(use_rotation ? 1 : 0);
    dirty_ = true;
    return (uint32_t)un_pose_.size() - 1;
  }

  /* reference :410-456 */
  uint32_t AddBinaryConstraint(const uint32_t pose1_id, const uint32_t pose2_id, const SE3t& t_12, Scalar weight = 1.0, bool use_rotation = true) {
    return AddBinaryConstraint(pose1_id, pose2_id, t_12, Matrix6t::Identity(), weight, use_rotation);
  }
  uint32_t AddBinaryConstraint(const uint32_t pose1_id, const uint32_t pose2_id, const SE3t& t_12, Matrix6t covariance, Scalar weight = 1.0,
                               bool use_rotation = true) {
    assert(pose1_id < poses_.size());
    assert(pose2_id < poses_.size());
    const Matrix6t ci = covariance.inverse();
    bin_p1_.push_back(pose1_id); bin_p2_.push_back(pose2_id);
    push_se3(bin_t_, t_12); push_rowmajor66(bin_cov_inv_, ci); bin_w_.push_back((double)weight); bin_rot_.push_back(use_rotation ? 1 : 0);
    dirty_ = true;
    return (uint32_t)bin_p1_.size() - 1;
  }

  /* reference :459-513 */
  uint32_t AddProjectionResidual(const Vector2t z, const uint32_t meas_pose_id, const uint32_t landmark_id, const uint32_t cam_id,
                                 const Scalar weight = 1.0) {
    assert(landmark_id < lm_ref_pose_.size());
    assert(meas_pose_id < poses_.size());
    const uint32_t x_ref_id = lm_ref_pose_[landmark_id];
    if (meas_pose_id == x_ref_id && cam_id == lm_ref_cam_[landmark_id]) { lm_z_ref_[2 * landmark_id] = (double)z[0]; lm_z_ref_[2 * landmark_id + 1] = (double)z[1]; }
    const bool diff_poses = meas_pose_id != x_ref_id;
    if (!(diff_poses || cam_id != lm_ref_cam_[landmark_id] || LmSize != 1)) return (uint32_t)-1;   /* not stored */
    const uint32_t res_id = (uint32_t)pr_pose_.size();
    lm_num_res_[landmark_id]++;
    pr_z_.push_back((double)z[0]); pr_z_.push_back((double)z[1]);
    pr_pose_.push_back(meas_pose_id); pr_lm_.push_back(landmark_id); pr_cam_.push_back(cam_id); pr_w0_.push_back((double)weight);
    if (poses_[x_ref_id].is_active == false && poses_[meas_pose_id].is_active == true) conditioning_proj_residuals_.push_back(res_id);
    dirty_ = true;
    return res_id;
  }

  /* reference :516-546 */
  uint32_t AddImuResidual(const uint32_t pose1_id, const uint32_t pose2_id, const std::vector<ImuMeasurement>& imu_meas, const Scalar weight = 1.0) {
    assert(pose1_id < poses_.size());
    assert(pose2_id < poses_.size());
    imu_p1_.push_back(pose1_id); imu_p2_.push_back(pose2_id); imu_w_.push_back((double)weight);
    for (const ImuMeasurement& m : imu_meas) {
      for (int k = 0; k < 3; ++k) imu_meas_.push_back((double)m.w[k]);
      for (int k = 0; k < 3; ++k) imu_meas_.push_back((double)m.a[k]);
      imu_meas_.push_back(m.time);
    }
    imu_ms_.push_back((uint32_t)(imu_meas_.size() / 7));
    if (poses_[pose1_id].is_active == false && poses_[pose2_id].is_active == true) conditioning_inertial_residuals_.push_back((uint32_t)imu_p1_.size() - 1);
    dirty_ = true;
    return (uint32_t)imu_p1_.size() - 1;
  }

  /* reference :549 / src/BundleAdjuster.cpp:278-705 — the whole loop runs behind the C ABI */
  void Solve(const uint32_t uMaxIter, const Scalar gn_damping = 1.0, const bool error_increase_allowed = false) {
    if (pr_pose_.empty() && bin_p1_.empty() && un_pose_.empty() && imu_p1_.empty()) return;   /* cpp:282-285 */
    if (!Upload()) { summary_.result = SolverError; return; }
    bagpu_summary s;
    if (bagpu_solve(h_, uMaxIter, (double)gn_damping, error_increase_allowed ? 1 : 0, &s) != BAGPU_OK) {
      std::cerr << "ba_b200: " << bagpu_last_error() << std::endl;
      summary_.result = SolverError;
      return;
    }
    Download();
    summary_.num_proj_residuals = s.num_proj_residuals; summary_.num_inertial_residuals = s.num_inertial_residuals;
    summary_.num_cond_proj_residuals = s.num_cond_proj_residuals; summary_.num_cond_inertial_residuals = s.num_cond_inertial_residuals;
    summary_.cond_proj_error = (Scalar)s.cond_proj_error; summary_.cond_inertial_error = (Scalar)s.cond_inertial_error;
    summary_.proj_error_ = (Scalar)s.proj_error; summary_.inertial_error = (Scalar)s.inertial_error;
    summary_.delta_norm = (Scalar)s.delta_norm; summary_.pre_solve_norm = (Scalar)s.pre_solve_norm; summary_.post_solve_norm = (Scalar)s.post_solve_norm;
    summary_.result = (OptimizationResult)s.result;
    proj_error_ = (Scalar)s.proj_error; unary_error_ = (Scalar)s.unary_error; binary_error_ = (Scalar)s.binary_error; inertial_error_ = (Scalar)s.inertial_error;
    iterations_ = s.iterations;
  }

  void SetRootPoseId(const uint32_t id) { root_pose_id_ = id; }
  uint32_t GetRootPoseId() { return root_pose_id_; }
  bool IsTranslationEnabled() { return true; }
  uint32_t GetNumPoses() const { return (uint32_t)poses_.size(); }
  uint32_t GetNumImuResiduals() const { return (uint32_t)imu_p1_.size(); }
  uint32_t GetNumProjResiduals() const { return (uint32_t)pr_pose_.size(); }
  uint32_t GetNumLandmarks() const { return (uint32_t)lm_ref_pose_.size(); }
  uint32_t GetNumUnaryResiduals() const { return (uint32_t)un_pose_.size(); }
  const ImuCalibration& GetImuCalibration() const { return imu_; }
  void SetImuCalibration(const ImuCalibration& calib) { imu_ = calib; }

  ProjectionResidual GetProjectionResidual(uint32_t id) const {
    ProjectionResidual r;
    r.residual_id = id; r.residual_offset = 2 * id; r.orig_weight = (Scalar)pr_w0_[id];
    r.z[0] = (Scalar)pr_z_[2 * id]; r.z[1] = (Scalar)pr_z_[2 * id + 1];
    r.x_meas_id = pr_pose_[id]; r.landmark_id = pr_lm_[id]; r.x_ref_id = lm_ref_pose_[pr_lm_[id]]; r.cam_id = pr_cam_[id];
    FetchProjResiduals();
    if (pr_w_.size() == pr_pose_.size()) {
      r.residual[0] = (Scalar)pr_res_[2 * id]; r.residual[1] = (Scalar)pr_res_[2 * id + 1];
      r.weight = (Scalar)pr_w_[id]; r.mahalanobis_distance = (Scalar)pr_maha_[id];
    }
    return r;
  }

  /* reference :571-582 (same error behaviour: message + throw 0) */
  const Pose& GetPose(const uint32_t id) const {
    if (id >= poses_.size()) {
      std::cerr << "Attempted to get pose with id " << id << " from BA.  when poses_.size() is only " << poses_.size() << "Aborting..." << std::endl;
      throw 0;
    }
    return poses_[id];
  }

  Landmark GetLandmarkObj(const uint32_t id) const {
    Landmark l;
    for (int k = 0; k < 4; ++k) { l.x_w[k] = (Scalar)lm_x_w_[4 * (size_t)id + k]; l.x_s[k] = (Scalar)lm_x_s_[4 * (size_t)id + k]; }
    l.z_ref[0] = (Scalar)lm_z_ref_[2 * (size_t)id]; l.z_ref[1] = (Scalar)lm_z_ref_[2 * (size_t)id + 1];
    l.external_id = lm_external_[id]; l.num_outlier_residuals = lm_outliers_[id]; l.num_proj_residuals = lm_num_res_[id];
    l.id = id; l.opt_id = lm_opt_[id]; l.ref_pose_id = lm_ref_pose_[id]; l.ref_cam_id = lm_ref_cam_[id];
    l.is_active = lm_active_[id] != 0; l.is_reliable = lm_reliable_[id] != 0;
    return l;
  }
  Vector4t GetLandmark(const uint32_t id) const { Vector4t x; for (int k = 0; k < 4; ++k) x[k] = (Scalar)lm_x_w_[4 * (size_t)id + k]; return x; }
  bool IsLandmarkReliable(const uint32_t id) const { return lm_reliable_[id] != 0; }
  /* reference src/BundleAdjuster.cpp:1806-1812 */
  double LandmarkOutlierRatio(const uint32_t id) const { return lm_num_res_[id] == 0 ? 0 : (double)lm_outliers_[id] / lm_num_res_[id]; }

  void GetErrors(Scalar& proj_error, Scalar& unary_error, Scalar& binary_error, Scalar& inertial_error) {
    proj_error = proj_error_; unary_error = unary_error_; binary_error = binary_error_; inertial_error = inertial_error_;
  }
  const SolutionSummary<Scalar>& GetSolutionSummary() const { return summary_; }
  Options<Scalar>& options() { return options_; }
  const std::shared_ptr<calibu::Rig<Scalar>> rig() const { return rig_; }
  uint32_t GetNumIterations() const { return iterations_; }   /* extension: outer iterations of the last Solve */
  bagpu_handle native_handle() { return h_; }                 /* extension: trace / phase timers / multi-GPU set-up */

  /* reference :608-631 (including its rotation indices 2,4,5) */
  void RegularizePose(uint32_t pose_id, bool translation, bool gravity, bool bias, bool rotation) {
    Pose& pose = poses_[pose_id];
    pose.is_param_mask_used = true;
    pose.param_mask.assign(kPoseDim, true);
    if (translation) pose.param_mask[0] = pose.param_mask[1] = pose.param_mask[2] = false;
    if (rotation) pose.param_mask[2] = pose.param_mask[4] = pose.param_mask[5] = false;
    if (gravity) pose.param_mask[GetGravityRegularizationDimension(pose_id)] = false;
    if (bias && kBiasInState) for (int k = 9; k < 15; ++k) pose.param_mask[k] = false;
    dirty_ = true;
  }

 private:
  /* reference :634-652 */
  uint32_t GetGravityRegularizationDimension(uint32_t pose_id) {
    const mini::Matrix<Scalar, 3, 3> rot = poses_[pose_id].t_wp.rotationMatrix();
    double max_dot = 0; uint32_t max_dim = 0;
    for (uint32_t ii = 0; ii < 3; ++ii) {
      const double dot = std::fabs((double)(rot(0, ii) * imu_.g_vec[0] + rot(1, ii) * imu_.g_vec[1] + rot(2, ii) * imu_.g_vec[2]));
      if (dot > max_dot) { max_dot = dot; max_dim = ii; }
    }
    return max_dim + 3;
  }

  static void push_se3(std::vector<double>& v, const SE3t& T) { double b[7]; detail::se3_to7(T, b); v.insert(v.end(), b, b + 7); }
  static void push_rowmajor66(std::vector<double>& v, const Matrix6t& m) { for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) v.push_back((double)m(r, c)); }

  bagpu_options MakeOptions() const {
    bagpu_options o;
    bagpu_default_options(&o);
    o.trust_region_size = (double)options_.trust_region_size;
    o.gyro_sigma = std::sqrt((double)imu_.r[0]); o.accel_sigma = std::sqrt((double)imu_.r[3]);
    o.gyro_bias_sigma = std::sqrt((double)imu_.r_b[0]); o.accel_bias_sigma = std::sqrt((double)imu_.r_b[3]);
    o.projection_outlier_threshold = (double)options_.projection_outlier_threshold;
    o.error_change_threshold = (double)options_.error_change_threshold; o.param_change_threshold = (double)options_.param_change_threshold;
    o.dogleg_max_inner_iterations = options_.dogleg_max_inner_iterations;
    o.apply_results = options_.apply_results; o.use_dogleg = options_.use_dogleg; o.use_triangular_matrices = options_.use_triangular_matrices;
    o.use_sparse_solver = options_.use_sparse_solver; o.write_reduced_camera_matrix = options_.write_reduced_camera_matrix;
    o.calculate_calibration_marginals = options_.calculate_calibration_marginals; o.use_per_pose_cam_params = options_.use_per_pose_cam_params;
    o.regularize_biases_in_batch = options_.regularize_biases_in_batch; o.enable_auto_regularization = options_.enable_auto_regularization;
    o.use_robust_norm_for_proj_residuals = options_.use_robust_norm_for_proj_residuals;
    o.use_robust_norm_for_inertial_residuals = options_.use_robust_norm_for_inertial_residuals;
    o.calculate_inertial_covariance_once = options_.calculate_inertial_covariance_once;
    for (int k = 0; k < 3; ++k) o.gravity[k] = (double)imu_.g_vec[k];
    o.root_pose_id = root_pose_id_;
    return o;
  }

  bool Check(int rc) { if (rc != BAGPU_OK) { std::cerr << "ba_b200: " << bagpu_last_error() << std::endl; return false; } return true; }

  bool Upload() {
    if (!h_) {
      bagpu_dims d;
      d.scalar_bytes = (int32_t)sizeof(Scalar);   /* 4 = the REAL_TYPE=float instantiation: it converts at this boundary, the device computes in FP64 */
      d.lm_dim = LmSize; d.pose_dim = PoseSize; d.calib_dim = CalibSize; d.do_tvs = DoTvs ? 1 : 0;
      if (!Check(bagpu_create(&d, &h_))) { h_ = nullptr; return false; }
      dirty_ = true;
    }
    const bagpu_options o = MakeOptions();
    if (!Check(bagpu_set_options(h_, &o))) return false;
    if (!dirty_) return true;
    const uint32_t nc = (uint32_t)rig_->NumCams();
    std::vector<double> cp(4 * (size_t)nc), ct(7 * (size_t)nc);
    for (uint32_t c = 0; c < nc; ++c) {
      const std::vector<Scalar> p = rig_->cameras_[c]->GetParams();
      for (int k = 0; k < 4; ++k) cp[4 * c + k] = k < (int)p.size() ? (double)p[k] : 0.0;
      detail::se3_to7(rig_->cameras_[c]->Pose(), &ct[7 * c]);
    }
    const uint32_t np = (uint32_t)poses_.size();
    std::vector<double> t(7 * (size_t)np), v(3 * (size_t)np), b(6 * (size_t)np);
    std::vector<uint8_t> act(np), mu(np), mk((size_t)np * PoseSize, 1);
    for (uint32_t i = 0; i < np; ++i) {
      const Pose& p = poses_[i];
      detail::se3_to7(p.t_wp, &t[7 * (size_t)i]);
      for (int k = 0; k < 3; ++k) v[3 * (size_t)i + k] = (double)p.v_w[k];
      for (int k = 0; k < 6; ++k) b[6 * (size_t)i + k] = (double)p.b[k];
      act[i] = p.is_active ? 1 : 0; mu[i] = p.is_param_mask_used ? 1 : 0;
      if (p.is_param_mask_used) for (int k = 0; k < PoseSize && k < (int)p.param_mask.size(); ++k) mk[(size_t)i * PoseSize + k] = p.param_mask[k] ? 1 : 0;
    }
    if (nc && !Check(bagpu_set_cameras(h_, nc, cp.data(), ct.data()))) return false;
    if (!Check(bagpu_set_poses(h_, np, t.data(), v.data(), b.data(), act.data(), mu.data(), mk.data()))) return false;
    if (!Check(bagpu_set_landmarks(h_, (uint32_t)lm_ref_pose_.size(), lm_x_w_.data(), lm_ref_pose_.data(), lm_ref_cam_.data(), lm_active_.data()))) return false;
    if (!Check(bagpu_set_proj_residuals(h_, (uint32_t)pr_pose_.size(), pr_z_.data(), pr_pose_.data(), pr_lm_.data(), pr_cam_.data(), pr_w0_.data()))) return false;
    if (!Check(bagpu_set_unary_residuals(h_, (uint32_t)un_pose_.size(), un_pose_.data(), un_t_.data(), un_cov_inv_.data(), un_rot_.data()))) return false;
    if (!Check(bagpu_set_binary_residuals(h_, (uint32_t)bin_p1_.size(), bin_p1_.data(), bin_p2_.data(), bin_t_.data(), bin_cov_inv_.data(), bin_w_.data(), bin_rot_.data()))) return false;
    if (!Check(bagpu_set_imu_residuals(h_, (uint32_t)imu_p1_.size(), imu_p1_.data(), imu_p2_.data(), imu_ms_.data(), imu_meas_.data(), imu_w_.data()))) return false;
    dirty_ = false;
    return true;
  }

  void FetchProjResiduals() const {
    const size_t nr = pr_pose_.size();
    if (pr_fetched_ || !h_ || !nr) return;
    pr_res_.resize(2 * nr); pr_w_.resize(nr); pr_maha_.resize(nr);
    if (bagpu_get_proj_residuals(h_, pr_res_.data(), pr_w_.data(), pr_maha_.data()) != BAGPU_OK) { pr_res_.clear(); pr_w_.clear(); pr_maha_.clear(); }
    pr_fetched_ = true;
  }

  void Download() {
    const uint32_t np = (uint32_t)poses_.size();
    std::vector<double> t(7 * (size_t)np), v(3 * (size_t)np), b(6 * (size_t)np);
    std::vector<uint8_t> mu(np), mk((size_t)np * PoseSize);
    bagpu_get_poses(h_, t.data(), v.data(), b.data(), mu.data(), mk.data());
    for (uint32_t i = 0; i < np; ++i) {
      Pose& p = poses_[i];
      p.t_wp = detail::se3_from7<SE3t, Scalar>(&t[7 * (size_t)i]);
      for (int k = 0; k < 3; ++k) p.v_w[k] = (Scalar)v[3 * (size_t)i + k];
      for (int k = 0; k < 6; ++k) p.b[k] = (Scalar)b[6 * (size_t)i + k];
      p.is_param_mask_used = mu[i] != 0;   /* BuildProblem writes masks into the poses (cpp:1250-1330) */
      if (p.is_param_mask_used) { p.param_mask.assign(PoseSize, true); for (int k = 0; k < PoseSize; ++k) p.param_mask[k] = mk[(size_t)i * PoseSize + k] != 0; }
    }
    if (!lm_ref_pose_.empty()) bagpu_get_landmarks(h_, lm_x_w_.data(), lm_x_s_.data(), lm_reliable_.data(), lm_outliers_.data());
    /* residual / weight / mahalanobis_distance of the projection residuals stay on the device until somebody asks for
       one (GetProjectionResidual): three arrays of N_o doubles and one more evaluation pass that most callers never need */
    pr_fetched_ = false;
    if (kBiasInState && np) {   /* cpp:666-669: the last pose's biases become the IMU calibration's */
      for (int k = 0; k < 3; ++k) { imu_.b_g[k] = poses_.back().b[k]; imu_.b_a[k] = poses_.back().b[3 + k]; }
    }
  }

  ImuCalibration imu_;
  Scalar proj_error_ = 0, binary_error_ = 0, unary_error_ = 0, inertial_error_ = 0;
  uint32_t root_pose_id_ = 0, num_active_poses_ = 0, num_active_landmarks_ = 0, iterations_ = 0;
  std::shared_ptr<calibu::Rig<Scalar>> rig_;
  std::vector<Pose> poses_;
  /* landmarks, SoA (what the reference keeps in std::vector<Landmark> landmarks_) */
  /* landmark state the GPU reads from and writes back into directly: page-locked like the observation buffers */
  detail::HostVector<double> lm_x_w_, lm_x_s_; std::vector<double> lm_z_ref_;
  std::vector<uint32_t> lm_ref_pose_, lm_ref_cam_, lm_num_res_, lm_opt_; detail::HostVector<uint32_t> lm_outliers_;
  std::vector<uint8_t> lm_active_; detail::HostVector<uint8_t> lm_reliable_;
  std::vector<int> lm_external_;
  /* projection residuals, SoA (proj_residuals_) */
  detail::HostVector<double> pr_z_, pr_w0_; mutable std::vector<double> pr_res_, pr_w_, pr_maha_; mutable bool pr_fetched_ = true;
  detail::HostVector<uint32_t> pr_pose_, pr_lm_, pr_cam_;
  std::vector<uint32_t> conditioning_proj_residuals_, conditioning_inertial_residuals_;
  /* unary / binary / inertial residuals, SoA */
  std::vector<uint32_t> un_pose_; std::vector<double> un_t_, un_cov_inv_; std::vector<uint8_t> un_rot_;
  std::vector<uint32_t> bin_p1_, bin_p2_; std::vector<double> bin_t_, bin_cov_inv_, bin_w_; std::vector<uint8_t> bin_rot_;
  std::vector<uint32_t> imu_p1_, imu_p2_, imu_ms_; std::vector<double> imu_meas_, imu_w_;
  SolutionSummary<Scalar> summary_;
  Options<Scalar> options_;
  bagpu_handle h_ = nullptr;
  bool dirty_ = true;
};

static const int NOT_USED = 0;
/* reference BundleAdjuster.h:757-763 */
template <typename Scalar> using VisualBundleAdjuster = BundleAdjuster<Scalar, 1, 6, 0>;
template <typename Scalar> using VisualInertialBundleAdjuster = BundleAdjuster<Scalar, 1, 15, 0>;

}  // namespace ba

#endif /* BA_B200_BUNDLEADJUSTER_H */
