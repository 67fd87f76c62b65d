// select.cu — exact k-th order statistic of non-negative doubles on the device.
//
// Replaces std::nth_element(errors_.begin(), errors_.begin() + floor(n * 0.5), errors_.end())
// (reference src/BundleAdjuster.cpp:1355-1358, 1457-1460, 1499-1502).  The Huber weights of
// every residual depend on this value, so it has to be the exact element, not an estimate:
// MSD radix select on the IEEE-754 bit pattern (monotone for x >= 0), 12 bits per pass,
// shared-memory histograms merged into one global histogram.  With several GPUs the
// histogram is all-reduced between `hist` and `scan`, which makes the selection global
// (SURVEY.md §8e).
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"

namespace bagpu {

constexpr int kSelBits = 12;
constexpr int kSelBins = 1 << kSelBits;

__global__ void __launch_bounds__(512) k_select_hist(const double* x, size_t n, SelectState* st, int pass) {
  __shared__ unsigned int h[kSelBins];
  for (int i = threadIdx.x; i < kSelBins; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const int shift = max(0, 64 - kSelBits * (pass + 1));
  const int bits = min(kSelBits, 64 - kSelBits * pass);
  const unsigned long long prefix = st->prefix;
  const int hi_shift = shift + bits;  // bits above the current digit must equal the prefix
  const unsigned long long* k = reinterpret_cast<const unsigned long long*>(x);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned long long key = k[i];
    const bool match = (hi_shift >= 64) ? true : ((key >> hi_shift) == (prefix >> hi_shift));
    if (match) atomicAdd(&h[(unsigned)((key >> shift) & ((1u << bits) - 1u))], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kSelBins; i += blockDim.x)
    if (h[i]) atomicAdd(&st->hist[i], h[i]);
}

__global__ void __launch_bounds__(1024) k_select_scan(SelectState* st, int pass, double* out) {
  __shared__ unsigned long long cum[1024];
  const int shift = max(0, 64 - kSelBits * (pass + 1));
  const int bits = min(kSelBits, 64 - kSelBits * pass);
  const int nb = 1 << bits;
  // each thread owns 4 consecutive bins
  const int b0 = threadIdx.x * 4;
  unsigned long long loc[4] = {0, 0, 0, 0};
  unsigned long long s = 0;
  for (int j = 0; j < 4; ++j) { if (b0 + j < nb) loc[j] = st->hist[b0 + j]; s += loc[j]; }
  cum[threadIdx.x] = s;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    unsigned long long v = (threadIdx.x >= off) ? cum[threadIdx.x - off] : 0;
    __syncthreads();
    cum[threadIdx.x] += v;
    __syncthreads();
  }
  const unsigned long long before = cum[threadIdx.x] - s;
  const unsigned long long krem = st->k_rem;
  if (krem >= before && krem < before + s) {
    unsigned long long c = before;
    for (int j = 0; j < 4; ++j) {
      if (krem < c + loc[j]) {
        st->prefix_next = st->prefix | ((unsigned long long)(b0 + j) << shift);
        st->k_rem_next = krem - c;
        break;
      }
      c += loc[j];
    }
  }
  __syncthreads();
  for (int j = 0; j < 4; ++j) st->hist[b0 + j] = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    st->prefix = st->prefix_next;
    st->k_rem = st->k_rem_next;
    if (shift == 0 && out) *out = __longlong_as_double((long long)st->prefix);
  }
}

__global__ void k_select_init(SelectState* st, unsigned long long k) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kSelBins) st->hist[i] = 0;
  if (i == 0) { st->prefix = 0; st->k_rem = k; st->prefix_next = 0; st->k_rem_next = 0; }
}

int select_num_passes() { return (64 + kSelBits - 1) / kSelBits; }

void launch_select_init(SelectState* st, unsigned long long k, cudaStream_t s) {
  k_select_init<<<(kSelBins + 255) / 256, 256, 0, s>>>(st, k);
}
void launch_select_hist(const double* x, size_t n, SelectState* st, int pass, int sms, cudaStream_t s) {
  if (n == 0) return;
  size_t blocks = (n + 512 * 8 - 1) / (512 * 8);
  if (blocks > (size_t)sms * 4) blocks = (size_t)sms * 4;
  if (blocks < 1) blocks = 1;
  k_select_hist<<<(int)blocks, 512, 0, s>>>(x, n, st, pass);
}
void launch_select_scan(SelectState* st, int pass, double* out, cudaStream_t s) {
  k_select_scan<<<1, 1024, 0, s>>>(st, pass, out);
}

}  // namespace bagpu
