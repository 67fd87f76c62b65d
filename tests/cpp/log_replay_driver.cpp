// log_replay_driver.cpp — replays an ODO / UTM / IMU text log (reference applications/unary_binary_imu_test/main.cpp)
// through include/ba/LogReplay.h into ba::BundleAdjuster<double,0,9,0>.
//
//   log_replay_driver <log.dat> <out.bin> <solve: 0|1>
//
// out.bin (doubles): n_nodes, n_imu_residuals, records read, then per node its time and the initial pose estimate (7), then
// the number of IMU measurements of every inertial residual; with solve = 1 additionally result, iterations, pre / post
// solve norm, unary and inertial error and the refined poses (7 each).  Exit code 3 = Solve() reported a solver error
// (e.g. no CUDA device).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ba/LogReplay.h"

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: %s log.dat out.bin solve\n", argv[0]); return 2; }
  typedef ba::BundleAdjuster<double, 0, 9, 0> BA;
  BA slam;
  ba::UnaryBinaryImuReplay<BA> replay(slam);
  replay.Setup();
  const long n = replay.ParseFile(argv[1]);
  if (n < 0) { perror(argv[1]); return 2; }
  std::vector<double> out;
  out.push_back((double)replay.nodes.size()); out.push_back((double)replay.imu_counts.size()); out.push_back((double)n);
  for (unsigned id : replay.nodes) {
    double b[7]; ba::detail::se3_to7(slam.GetPose(id).t_wp, b);
    out.push_back(slam.GetPose(id).time);
    out.insert(out.end(), b, b + 7);
  }
  for (uint32_t c : replay.imu_counts) out.push_back((double)c);
  int rc = 0;
  if (atoi(argv[3])) {
    replay.Solve();
    const ba::SolutionSummary<double>& s = slam.GetSolutionSummary();
    if (s.result == ba::SolverError || s.result == ba::FactorizationError) { fprintf(stderr, "Solve: result %d\n", (int)s.result); rc = 3; }
    double e0, e1, e2, e3;
    slam.GetErrors(e0, e1, e2, e3);
    const double head[] = {(double)s.result, (double)slam.GetNumIterations(), s.pre_solve_norm, s.post_solve_norm, e1, e3};
    out.insert(out.end(), head, head + 6);
    for (unsigned id : replay.nodes) { double b[7]; ba::detail::se3_to7(slam.GetPose(id).t_wp, b); out.insert(out.end(), b, b + 7); }
  }
  FILE* o = fopen(argv[2], "wb");
  if (!o) return 2;
  fwrite(out.data(), sizeof(double), out.size(), o);
  fclose(o);
  return rc;
}
