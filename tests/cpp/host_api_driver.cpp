// host_api_driver.cpp — drives the C++ drop-in class (include/ba/BundleAdjuster.h) the way the
// reference's applications drive ba::BundleAdjuster (reference
// applications/unary_binary_imu_test/main.cpp:243-296: Init -> SetGravity -> AddPose ->
// Add*Constraint / Add*Residual -> Solve -> GetPose).
//
//   host_api_driver <problem.bin> <out.bin> <iterations> [use_dogleg [gyro_sigma accel_sigma gyro_bias_sigma accel_bias_sigma]]
//
// problem.bin (little endian, written by tests/test_cpp_host_api.py): int32 L, D, then per
// category a uint32 count followed by the raw arrays in the order read below.  out.bin: the
// refined poses / landmarks / summary as doubles.  Exit code 3 = Solve() reported SolverError
// (e.g. no CUDA device: there is no CPU path behind this class).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ba/BundleAdjuster.h"

namespace {
struct Reader {
  FILE* f;
  template <typename T> std::vector<T> arr(size_t n) { std::vector<T> v(n); if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); } return v; }
  uint32_t u32() { return arr<uint32_t>(1)[0]; }
};

template <typename Scalar, int L, int D>
int run(Reader& rd, const char* out_path, int iters, bool dogleg, const double* sig) {
  typedef ba::BundleAdjuster<Scalar, L, D, 0> BA;
  typedef typename BA::SE3t SE3t;
  auto se3 = [](const double* p) { return ba::detail::se3_from7<SE3t, Scalar>(p); };
  BA ba;
  ba::Options<Scalar> opt;
  opt.error_change_threshold = 0; opt.param_change_threshold = 0; opt.use_dogleg = dogleg;
  if (sig) { opt.gyro_sigma = (Scalar)sig[0]; opt.accel_sigma = (Scalar)sig[1]; opt.gyro_bias_sigma = (Scalar)sig[2]; opt.accel_bias_sigma = (Scalar)sig[3]; }
  const uint32_t nc = rd.u32();
  auto cp = rd.arr<double>(4 * nc), ct = rd.arr<double>(7 * nc);
  const uint32_t np = rd.u32();
  auto pt = rd.arr<double>(7 * np), pv = rd.arr<double>(3 * np), pb = rd.arr<double>(6 * np); auto pa = rd.arr<uint8_t>(np);
  const uint32_t nl = rd.u32();
  auto lx = rd.arr<double>(4 * nl); auto lrp = rd.arr<uint32_t>(nl), lrc = rd.arr<uint32_t>(nl); auto la = rd.arr<uint8_t>(nl);
  const uint32_t nr = rd.u32();
  auto rz = rd.arr<double>(2 * nr); auto rp = rd.arr<uint32_t>(nr), rl = rd.arr<uint32_t>(nr), rc = rd.arr<uint32_t>(nr); auto rw = rd.arr<double>(nr);
  const uint32_t nu = rd.u32();
  auto up = rd.arr<uint32_t>(nu); auto ut = rd.arr<double>(7 * nu), ucov = rd.arr<double>(36 * nu); auto ur = rd.arr<uint8_t>(nu);
  const uint32_t nb = rd.u32();
  auto b1 = rd.arr<uint32_t>(nb), b2 = rd.arr<uint32_t>(nb); auto bt = rd.arr<double>(7 * nb), bcov = rd.arr<double>(36 * nb), bw = rd.arr<double>(nb); auto br = rd.arr<uint8_t>(nb);
  const uint32_t ni = rd.u32();
  auto i1 = rd.arr<uint32_t>(ni), i2 = rd.arr<uint32_t>(ni), ims = rd.arr<uint32_t>(ni + 1);
  auto im = rd.arr<double>(7 * (size_t)(ni ? ims[ni] : 0)); auto iw = rd.arr<double>(ni);
  auto g = rd.arr<double>(3);

  const auto t_start = std::chrono::steady_clock::now();
  auto since = [](std::chrono::steady_clock::time_point t0) { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
  ba.Init(opt, np, nr, nl);
  typename BA::Vector3t gv; gv[0] = (Scalar)g[0]; gv[1] = (Scalar)g[1]; gv[2] = (Scalar)g[2];
  ba.SetGravity(gv);
  for (uint32_t c = 0; c < nc; ++c)
    ba.AddCamera(std::make_shared<calibu::LinearCamera<Scalar>>((Scalar)cp[4 * c], (Scalar)cp[4 * c + 1], (Scalar)cp[4 * c + 2], (Scalar)cp[4 * c + 3], se3(&ct[7 * c])));
  for (uint32_t i = 0; i < np; ++i) {
    typename BA::Vector3t v; typename BA::Vector6t b;
    for (int k = 0; k < 3; ++k) v[k] = (Scalar)pv[3 * i + k];
    for (int k = 0; k < 6; ++k) b[k] = (Scalar)pb[6 * i + k];
    const uint32_t id = ba.AddPose(se3(&pt[7 * i]), typename BA::VectorXt(5, 0), v, b, pa[i] != 0, (double)i);
    if (id != i) return 2;
  }
  for (uint32_t i = 0; i < nl; ++i) {
    typename BA::Vector4t x; for (int k = 0; k < 4; ++k) x[k] = (Scalar)lx[4 * i + k];
    if (ba.AddLandmark(x, lrp[i], lrc[i], la[i] != 0) != i) return 2;
  }
  for (uint32_t i = 0; i < nr; ++i) {
    typename BA::Vector2t z; z[0] = (Scalar)rz[2 * i]; z[1] = (Scalar)rz[2 * i + 1];
    if (ba.AddProjectionResidual(z, rp[i], rl[i], rc[i], (Scalar)rw[i]) != i) return 2;
  }
  if (L == 1 && nl) {   // the skip rule: a measurement in the landmark's own reference frame is not stored
    typename BA::Vector2t z;
    if (ba.AddProjectionResidual(z, lrp[0], 0, lrc[0]) != (uint32_t)-1) return 2;
    if (ba.GetNumProjResiduals() != nr) return 2;
  }
  for (uint32_t i = 0; i < nu; ++i) {
    typename BA::Matrix6t cov; for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) cov(r, c) = (Scalar)ucov[36 * i + 6 * r + c];
    ba.AddUnaryConstraint(up[i], se3(&ut[7 * i]), cov, ur[i] != 0);
  }
  for (uint32_t i = 0; i < nb; ++i) {
    typename BA::Matrix6t cov; for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) cov(r, c) = (Scalar)bcov[36 * i + 6 * r + c];
    ba.AddBinaryConstraint(b1[i], b2[i], se3(&bt[7 * i]), cov, (Scalar)bw[i], br[i] != 0);
  }
  for (uint32_t i = 0; i < ni; ++i) {
    std::vector<typename BA::ImuMeasurement> ms;
    for (uint32_t m = ims[i]; m < ims[i + 1]; ++m) {
      typename BA::Vector3t w, a;
      for (int k = 0; k < 3; ++k) { w[k] = (Scalar)im[7 * m + k]; a[k] = (Scalar)im[7 * m + 3 + k]; }
      ms.push_back(typename BA::ImuMeasurement(w, a, im[7 * m + 6]));
    }
    ba.AddImuResidual(i1[i], i2[i], ms, (Scalar)iw[i]);
  }
  try { ba.GetPose(np + 5); return 2; } catch (int) {}   // reference BundleAdjuster.h:573-580

  const double t_add = since(t_start);
  const auto t_solve0 = std::chrono::steady_clock::now();
  ba.Solve((uint32_t)iters);
  const double t_solve = since(t_solve0);
  const ba::SolutionSummary<Scalar>& s = ba.GetSolutionSummary();
  if (s.result == ba::SolverError || s.result == ba::FactorizationError) { fprintf(stderr, "Solve: result %d\n", (int)s.result); return 3; }

  const auto t_get0 = std::chrono::steady_clock::now();
  // BA_DRIVER_NO_RESIDUALS: a caller that reads poses and landmarks only (the per-residual records stay on the device)
  const bool want_residuals = getenv("BA_DRIVER_NO_RESIDUALS") == nullptr;
  FILE* o = fopen(out_path, "wb");
  if (!o) return 2;
  std::vector<double> out;
  out.reserve(10 + 16 * (size_t)np + 6 * (size_t)nl + 3 * (size_t)nr);
  Scalar e0, e1, e2, e3;
  ba.GetErrors(e0, e1, e2, e3);
  const double head[] = {(double)s.result, (double)ba.GetNumIterations(), (double)s.pre_solve_norm, (double)s.post_solve_norm, (double)s.delta_norm,
                         (double)e0, (double)e1, (double)e2, (double)e3, (double)s.num_proj_residuals};
  out.insert(out.end(), head, head + 10);
  for (uint32_t i = 0; i < np; ++i) {
    double b[7]; ba::detail::se3_to7(ba.GetPose(i).t_wp, b);
    out.insert(out.end(), b, b + 7);
    for (int k = 0; k < 3; ++k) out.push_back((double)ba.GetPose(i).v_w[k]);
    for (int k = 0; k < 6; ++k) out.push_back((double)ba.GetPose(i).b[k]);
  }
  for (uint32_t i = 0; i < nl; ++i) {
    const typename BA::Vector4t x = ba.GetLandmark(i);
    for (int k = 0; k < 4; ++k) out.push_back((double)x[k]);
    out.push_back(ba.IsLandmarkReliable(i) ? 1.0 : 0.0);
    out.push_back(ba.LandmarkOutlierRatio(i));
  }
  const double t_get_state = since(t_get0);
  for (uint32_t i = 0; want_residuals && i < nr; ++i) {
    const typename BA::ProjectionResidual r = ba.GetProjectionResidual(i);
    out.push_back((double)r.residual[0]); out.push_back((double)r.residual[1]); out.push_back((double)r.weight);
  }
  const double t_get = since(t_get0);
  fwrite(out.data(), sizeof(double), out.size(), o);
  fclose(o);
  // one JSON line for bench.py's e2e_cpp leg: the Add* calls (host-side SoA marshalling), Solve() (upload + set-up + iterations +
  // read-back of poses / landmarks), the getters
  fprintf(stderr, "{\"driver_timing\": {\"add_s\": %.6f, \"solve_s\": %.6f, \"get_state_s\": %.6f, \"get_all_s\": %.6f, \"iterations\": %u}}\n",
          t_add, t_solve, t_get_state, t_get, ba.GetNumIterations());
  printf("host_api_driver: result %d, %u iterations, cost %.12g -> %.12g\n", (int)s.result, ba.GetNumIterations(), (double)s.pre_solve_norm, (double)s.post_solve_norm);
  return 0;
}
}  // namespace

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: %s problem.bin out.bin iterations [use_dogleg]\n", argv[0]); return 2; }
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 2; }
  Reader rd{f};
  const int iters = atoi(argv[3]);
  const bool dogleg = argc > 4 ? atoi(argv[4]) != 0 : true;
  double sigv[4]; const double* sig = nullptr;
  if (argc > 8) { for (int k = 0; k < 4; ++k) sigv[k] = atof(argv[5 + k]); sig = sigv; }
  int hdr[2];
  if (fread(hdr, sizeof(int), 2, f) != 2) return 2;
  const int L = hdr[0], D = hdr[1];
  // BA_DRIVER_SCALAR=float: the REAL_TYPE=float build of the reference (CMakeLists.txt:16-20), instantiations <float, ...>
  const char* sc = getenv("BA_DRIVER_SCALAR");
  if (sc && sc[0] == 'f') {
    if (L == 1 && D == 6) return run<float, 1, 6>(rd, argv[2], iters, dogleg, sig);
    if (L == 3 && D == 6) return run<float, 3, 6>(rd, argv[2], iters, dogleg, sig);
    if (L == 0 && D == 6) return run<float, 0, 6>(rd, argv[2], iters, dogleg, sig);
    if (L == 1 && D == 15) return run<float, 1, 15>(rd, argv[2], iters, dogleg, sig);
    fprintf(stderr, "instantiation <float,%d,%d,0> not compiled into this driver\n", L, D);
    return 2;
  }
  if (L == 1 && D == 6) return run<double, 1, 6>(rd, argv[2], iters, dogleg, sig);
  if (L == 3 && D == 6) return run<double, 3, 6>(rd, argv[2], iters, dogleg, sig);
  if (L == 0 && D == 6) return run<double, 0, 6>(rd, argv[2], iters, dogleg, sig);
  if (L == 1 && D == 15) return run<double, 1, 15>(rd, argv[2], iters, dogleg, sig);
  if (L == 0 && D == 9) return run<double, 0, 9>(rd, argv[2], iters, dogleg, sig);
  if (L == 0 && D == 15) return run<double, 0, 15>(rd, argv[2], iters, dogleg, sig);
  fprintf(stderr, "instantiation <double,%d,%d,0> not compiled into this driver\n", L, D);
  return 2;
}
