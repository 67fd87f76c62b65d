// Drives include/ba/InterpolationBuffer.h the way the reference's applications do (reference
// applications/unary_binary_imu_test/main.cpp:243-280: AddElement per IMU sample, GetRange per pose pair).
// in:  uint32 n, n rows of 7 doubles (w a time); uint32 q, q rows of (op, t0, t1) doubles
//      op 0 = GetElement(t0), 1 = GetRange(t0, t1), 2 = AppendRangeFlat(t0, t1)
// out: per query: op 0 -> 7 doubles + index; op 1 / 2 -> count then count rows of 7 doubles
#include <cstdio>
#include <cstdint>
#include <vector>
#include "ba/BundleAdjuster.h"
#include "ba/InterpolationBuffer.h"

typedef ba::ImuMeasurementT<double> Meas;

static void put(std::vector<double>& o, const Meas& m) {
  for (int k = 0; k < 3; ++k) o.push_back(m.w[k]);
  for (int k = 0; k < 3; ++k) o.push_back(m.a[k]);
  o.push_back(m.time);
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  uint32_t n = 0, q = 0;
  if (fread(&n, 4, 1, f) != 1) return 2;
  ba::InterpolationBufferT<Meas, double> buf;
  for (uint32_t i = 0; i < n; ++i) {
    double r[7];
    if (fread(r, 8, 7, f) != 7) return 2;
    Meas m; for (int k = 0; k < 3; ++k) { m.w[k] = r[k]; m.a[k] = r[3 + k]; } m.time = r[6];
    buf.AddElement(m);
  }
  std::vector<double> out;
  out.push_back(buf.start_time); out.push_back(buf.end_time); out.push_back(buf.average_dt);
  if (fread(&q, 4, 1, f) != 1) return 2;
  for (uint32_t i = 0; i < q; ++i) {
    double c[3];
    if (fread(c, 8, 3, f) != 3) return 2;
    if (c[0] == 0) {
      size_t idx = 0;
      put(out, buf.GetElement(c[1], &idx));
      out.push_back((double)idx);
    } else if (c[0] == 1) {
      const std::vector<Meas> r = buf.GetRange(c[1], c[2]);
      out.push_back((double)r.size());
      for (const Meas& m : r) put(out, m);
    } else {
      std::vector<double> flat;
      const size_t rows = buf.AppendRangeFlat(c[1], c[2], flat);
      out.push_back((double)rows);
      out.insert(out.end(), flat.begin(), flat.end());
    }
  }
  fclose(f);
  buf.Clear();
  out.push_back(buf.start_time); out.push_back((double)buf.elements.size());
  f = fopen(argv[2], "wb");
  if (!f) return 2;
  fwrite(out.data(), 8, out.size(), f);
  fclose(f);
  return 0;
}
