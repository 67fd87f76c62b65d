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
        st->bucket_count = loc[j];
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
  if (i == 0) { st->prefix = 0; st->k_rem = k; st->prefix_next = 0; st->k_rem_next = 0; st->bucket_count = 0; st->cand_n = 0; st->all_n = 0; }
}

// After two 12-bit passes the bucket of the k-th value holds a few thousand of 10^7 values (its width is 2^-12 of a binade):
// instead of four more passes over all the data (and four more histogram exchanges between GPUs) the bucket's members are
// collected, gathered over the ranks and the selection is finished by one CTA.
__global__ void __launch_bounds__(256) k_select_compact(const double* x, size_t n, SelectState* st, int hi_shift, unsigned long long* cand) {
  const unsigned long long want = st->prefix >> hi_shift;
  const unsigned long long* k = reinterpret_cast<const unsigned long long*>(x);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned long long key = k[i];
    if ((key >> hi_shift) == want) {
      const unsigned int slot = atomicAdd(&st->cand_n, 1u);
      if (slot < (unsigned int)kSelCandCap) cand[slot] = key;
    }
  }
}

// k-th smallest (k = st->k_rem, 0-based, within the bucket) of at most kSelCandCap keys: LSD-free MSD radix select in shared
// memory, 8 bits per round over the bits below the known prefix.  The order in which the candidates arrived does not matter.
__global__ void __launch_bounds__(1024) k_select_final(const unsigned long long* all, const unsigned int* all_n, SelectState* st, int low_bits, double* out) {
  extern __shared__ unsigned long long keys[];
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long s_prefix, s_k;
  const unsigned int n = min(*all_n, (unsigned int)kSelCandCap);
  for (unsigned int i = threadIdx.x; i < n; i += blockDim.x) keys[i] = all[i];
  if (threadIdx.x == 0) { s_prefix = st->prefix; s_k = st->k_rem; }
  __syncthreads();
  for (int shift = low_bits - 8; shift > -8; shift -= 8) {
    const int sh = shift < 0 ? 0 : shift;
    const int bits = shift < 0 ? 8 + shift : 8;
    if (threadIdx.x < 256) hist[threadIdx.x] = 0;
    __syncthreads();
    const unsigned long long prefix = s_prefix;
    const int hi = sh + bits;
    for (unsigned int i = threadIdx.x; i < n; i += blockDim.x) {
      const unsigned long long key = keys[i];
      if ((key >> hi) == (prefix >> hi)) atomicAdd(&hist[(unsigned)((key >> sh) & ((1u << bits) - 1u))], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 32) {   // warp 0 finds the bin of the k-th key: 8 bins per lane, warp scan of the lane totals
      const int lane = threadIdx.x;
      unsigned int loc[8], tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { loc[j] = (lane * 8 + j < (1 << bits)) ? hist[lane * 8 + j] : 0u; tot += loc[j]; }
      unsigned int inc = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const unsigned int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
      const unsigned long long k = s_k;
      __syncwarp();
      const unsigned long long before = inc - tot;
      if (k >= before && k < before + tot) {
        unsigned long long c = before;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (k < c + loc[j]) { s_prefix = prefix | ((unsigned long long)(lane * 8 + j) << sh); s_k = k - c; break; }
          c += loc[j];
        }
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { st->prefix = s_prefix; st->k_rem = s_k; if (out) *out = __longlong_as_double((long long)s_prefix); }
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
void launch_select_compact(const double* x, size_t n, SelectState* st, int passes_done, unsigned long long* cand, int sms, cudaStream_t s) {
  if (n == 0) return;
  size_t blocks = (n + 256 * 8 - 1) / (256 * 8);
  if (blocks > (size_t)sms * 8) blocks = (size_t)sms * 8;
  if (blocks < 1) blocks = 1;
  k_select_compact<<<(int)blocks, 256, 0, s>>>(x, n, st, 64 - kSelBits * passes_done, cand);
}
void launch_select_final(const unsigned long long* all, const unsigned int* all_n, SelectState* st, double* out, cudaStream_t s) {
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(k_select_final, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(unsigned long long) * (kSelCandCap + 1))); attr = true; }
  k_select_final<<<1, 1024, sizeof(unsigned long long) * (kSelCandCap + 1), s>>>(all, all_n, st, 64 - 2 * kSelBits, out);
}

}  // namespace bagpu
