/* LogReplay.h — the caller-side input preparation of the reference's integration application
 * (reference applications/unary_binary_imu_test/main.cpp): the ODO / UTM / IMU text-log reader (parse_file, :236-279)
 * and the replay that turns the records into AddPose / AddUnaryConstraint / AddImuResidual calls on a
 * ba::BundleAdjuster<double,0,9,0> (f_gps :137-209, add_imu :56-64, add_gyro_and_speed :67-85,
 * update_incremental_pose :88-133, setup :212-230, solve :282-287).  SURVEY.md §8f rank 4.
 *
 * Same arithmetic and the same order of calls as the application; its file-scope globals and function-local statics
 * are members here, so several logs can be replayed in one process.  The application exit(0)s when a GPS interval
 * holds no IMU measurement; OnUtm returns false instead and ParseFile stops.
 *
 * Log format (whitespace separated tokens, one record per line in practice):
 *   ODO <time> <right wheel speed> <left wheel speed>
 *   UTM <time> <easting> <northing> <altitude>
 *   IMU <time> <wx> <wy> <wz> <ax> <ay> <az>
 */
#ifndef BA_B200_LOGREPLAY_H
#define BA_B200_LOGREPLAY_H
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>
#include "BundleAdjuster.h"
#include "InterpolationBuffer.h"

namespace ba {

template <typename BA = BundleAdjuster<double, 0, 9, 0>>
class UnaryBinaryImuReplay {
 public:
  typedef typename BA::SE3t SE3t;
  typedef typename BA::Vector3t Vector3t;
  typedef typename BA::ImuMeasurement ImuMeasurement;

  explicit UnaryBinaryImuReplay(BA& slam) : slam_(slam) {}

  /* setup(), main.cpp:212-230 */
  void Setup() {
    Options<double> options;
    options.trust_region_size = 100000;
    slam_.Init(options);
    Vector3t gravity; gravity[0] = 0; gravity[1] = 0; gravity[2] = 9.8;
    slam_.SetGravity(gravity);
    incremental_gyro_pose_ = SE3t();
  }

  /* parse_file(), main.cpp:236-279.  Returns the number of records read, or -1 if the file cannot be opened. */
  long ParseFile(const char* filename, size_t max_nodes = 10000) {
    FILE* input = fopen(filename, "r");
    if (!input) return -1;
    long n = 0;
    for (;;) {
      char name[16];
      if (fscanf(input, "%15s", name) == EOF) break;
      if (strncmp(name, "ODO", 3) == 0) {
        double time, rr, rl;
        if (fscanf(input, "%lf %lf %lf", &time, &rr, &rl) == 3) { OnOdo(time, rr, rl); ++n; }
      } else if (strncmp(name, "UTM", 3) == 0) {
        double time, e, nn, alt;
        if (fscanf(input, "%lf %lf %lf %lf", &time, &e, &nn, &alt) == 4) {
          if (nodes.size() >= max_nodes) break;      /* main.cpp:254-258 */
          ++n;
          if (!OnUtm(time, e, nn, alt)) break;
        }
      } else if (strncmp(name, "IMU", 3) == 0) {
        double time, w[3], a[3];
        if (fscanf(input, "%lf %lf %lf %lf %lf %lf %lf", &time, w, w + 1, w + 2, a, a + 1, a + 2) == 7) { OnImu(time, w, a); ++n; }
      } else {
        fprintf(stderr, "Unknown symbol <%s>\n", name);
      }
    }
    fclose(input);
    return n;
  }

  /* update_incremental_pose(), main.cpp:88-133: differential-drive odometry between GPS fixes, wheel speed for the gyro path */
  void OnOdo(double timestamp, double rr, double rl) {
    if (odo_first_) { incremental_timestamp_ = timestamp; odo_first_ = false; return; }
    speed_ = 0.5 * (rr + rl);
    const double dt = timestamp - incremental_timestamp_;
    const double trackwidth = 1.5, TINY = 0.0001;
    if (std::fabs(rr) > TINY || std::fabs(rl) > TINY) {
      if (std::fabs(rr - rl) < TINY) {
        incremental_x_ += std::cos(incremental_yaw_) * rr * dt;
        incremental_y_ += std::sin(incremental_yaw_) * rr * dt;
      } else {
        const double w = (rr - rl) / trackwidth;
        const double R = trackwidth * 0.5 * (rr + rl) / (rr - rl);
        const double icc_x = incremental_x_ - R * std::sin(incremental_yaw_);
        const double icc_y = incremental_y_ + R * std::cos(incremental_yaw_);
        const double wdt = w * dt;
        const double new_x = std::cos(wdt) * (incremental_x_ - icc_x) - std::sin(wdt) * (incremental_y_ - icc_y) + icc_x;
        const double new_y = std::sin(wdt) * (incremental_x_ - icc_x) + std::cos(wdt) * (incremental_y_ - icc_y) + icc_y;
        incremental_yaw_ += wdt;
        incremental_x_ = new_x; incremental_y_ = new_y;
      }
    }
    Vector3t t; t[0] = incremental_x_; t[1] = incremental_y_; t[2] = 0;
    incremental_differential_update_ = SE3t(AxisAngle(2, incremental_yaw_), t);
    incremental_timestamp_ = timestamp;
  }

  /* add_gyro_and_speed() + add_imu(), main.cpp:56-85 */
  void OnImu(double timestamp, const double* w, const double* a) {
    if (gyro_last_timestamp_ != 0) {
      const double dt = timestamp - gyro_last_timestamp_;
      const double distance = speed_ * dt;
      const mini::SO3Group<double> q = AxisAngle(2, w[2] * dt) * AxisAngle(1, w[1] * dt) * AxisAngle(0, w[0] * dt);
      Vector3t t; t[0] = 0; t[1] = distance; t[2] = 0;
      incremental_gyro_update_ = incremental_gyro_update_ * SE3t(q, t);
    }
    gyro_last_timestamp_ = timestamp;
    Vector3t wv, av;
    for (int k = 0; k < 3; ++k) { wv[k] = w[k]; av[k] = a[k]; }
    imu_buffer.AddElement(ImuMeasurement(wv, av, timestamp));
  }

  /* f_gps(), main.cpp:137-209: one pose per fix (dead-reckoned from the previous one along the gyro path), a unary
     constraint at the fix, an inertial residual from the previous pose */
  bool OnUtm(double timestamp, double utm_e, double utm_n, double altitude) {
    if (gps_first_) { offset_e_ = utm_e; offset_n_ = utm_n; offset_u_ = altitude; gps_first_ = false; }
    incremental_gyro_pose_ = incremental_gyro_pose_ * incremental_gyro_update_;
    incremental_differential_pose_ = incremental_differential_pose_ * incremental_differential_update_;
    Vector3t fix; fix[0] = utm_e; fix[1] = utm_n; fix[2] = altitude;
    gps.push_back(fix);
    Vector3t pt; pt[0] = utm_e - offset_e_; pt[1] = utm_n - offset_n_; pt[2] = altitude - offset_u_;
    const SE3t utm_prior(mini::SO3Group<double>(), pt);
    if (nodes.empty()) {
      nodes.push_back(slam_.AddPose(SE3t(), true, timestamp));
    } else {
      const SE3t& recent_pose = slam_.GetPose(nodes.back()).t_wp;
      nodes.push_back(slam_.AddPose(recent_pose * incremental_gyro_update_, true, timestamp));
    }
    {
      typename BA::Matrix6t cov;
      const double d[6] = {1000, 1000, 30000, DBL_MAX, DBL_MAX, DBL_MAX};
      for (int k = 0; k < 6; ++k) cov(k, k) = d[k];
      slam_.AddUnaryConstraint(nodes.back(), utm_prior, cov);
    }
    bool ok = true;
    if (nodes.size() >= 2) {
      std::vector<ImuMeasurement> imu_meas = imu_buffer.GetRange(last_gps_timestamp_, timestamp);
      if (imu_meas.size() == 0) {
        std::cerr << "Could not find imu measurements between : " << last_gps_timestamp_ << " and " << timestamp << std::endl;
        ok = false;
      } else {
        slam_.AddImuResidual(nodes.back() - 1, nodes.back(), imu_meas);
        imu_counts.push_back((uint32_t)imu_meas.size());
      }
    }
    incremental_x_ = incremental_y_ = incremental_yaw_ = 0;
    incremental_differential_update_ = SE3t();
    incremental_gyro_update_ = SE3t();
    last_gps_timestamp_ = timestamp;
    return ok;
  }

  /* solve(), main.cpp:282-287 */
  void Solve() { slam_.Solve(25, 0.2); }

  std::vector<unsigned int> nodes;
  std::vector<Vector3t> gps;
  std::vector<uint32_t> imu_counts;       /* measurements handed to each AddImuResidual */
  InterpolationBufferT<ImuMeasurement, double> imu_buffer;

 private:
  /* Eigen::AngleAxisd(angle, unit axis) as a rotation */
  static mini::SO3Group<double> AxisAngle(int axis, double angle) {
    Vector3t w; w[0] = w[1] = w[2] = 0; w[axis] = angle;
    return mini::SO3Group<double>::exp(w);
  }
  BA& slam_;
  double offset_e_ = 0, offset_n_ = 0, offset_u_ = 0;
  double incremental_x_ = 0, incremental_y_ = 0, incremental_yaw_ = 0, incremental_timestamp_ = 0;
  SE3t incremental_differential_update_, incremental_differential_pose_, incremental_gyro_pose_, incremental_gyro_update_;
  double speed_ = 0, gyro_last_timestamp_ = 0, last_gps_timestamp_ = 0;
  bool odo_first_ = true, gps_first_ = true;
};

}  // namespace ba
#endif
