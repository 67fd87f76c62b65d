// symbolic.cu — pair lists of the pose-pair-major Schur reduction, built on the device once per
// Solve() (they depend only on the ids, like the reference's block patterns).
//
// For every landmark the reference's (W V^-1) W^T product (SparseBlockMatrixOps.h:182-254 called
// at src/BundleAdjuster.cpp:464-470) touches every pair of poses observing it.  Here each such
// pair becomes one list entry keyed by its S block (row pose, column pose); a stable radix sort by
// key groups the entries per block, a run-length encode yields the block table.  The sort / scan /
// run-length primitives are CUB's (set-up code, outside the per-iteration hot path).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_run_length_encode.cuh>
#include <cub/device/device_scan.cuh>
#include <string>
#include <vector>
#include "ba_gpu.h"
#include "dev_alloc.h"
#include "launch.h"

namespace bagpu {

namespace {

// an observation takes part in the reduced system when its measurement pose is optimised and is not the landmark's
// reference pose (parallel_algos.h:86-99: dz/dx_meas is zero there)
__device__ __forceinline__ int part_opt(const DeviceProblem& P, int i, uint32_t ref_pose) {
  const uint32_t m = P.obs_pose[i];
  return (m == ref_pose) ? -1 : P.pose_opt[m];
}

__global__ void k_pair_count(DeviceProblem P, unsigned long long* count) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
  const uint32_t rp = P.lm_ref_pose[lm];
  unsigned long long np = 0;
  for (int i = s; i < e; ++i) np += (part_opt(P, i, rp) >= 0);
  unsigned long long n = 0;
  if (P.lm_opt[lm] >= 0) n += np * (np - (np ? 1 : 0)) / 2;
  if (P.L == 1 && P.pose_opt[P.lm_ref_pose[lm]] >= 0) n += np;
  count[lm] = n;
}

__global__ void k_pair_emit(DeviceProblem P, const unsigned long long* __restrict__ offs, unsigned long long* __restrict__ keys,
                            unsigned long long* __restrict__ vals) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
  unsigned long long o = offs[lm];
  const unsigned long long np = (unsigned long long)P.n_active_poses;
  const bool lm_on = P.lm_opt[lm] >= 0;
  const uint32_t rp = P.lm_ref_pose[lm];
  const int ref_opt = (P.L == 1) ? P.pose_opt[rp] : -1;
  for (int i = s; i < e; ++i) {
    const int oa = part_opt(P, i, rp);
    if (oa < 0) continue;
    const uint32_t sa = (uint32_t)i;
    const unsigned long long pa = (unsigned long long)oa;
    if (ref_opt >= 0) {   // (measurement pose, reference pose) block of this observation
      const unsigned long long pr = (unsigned long long)ref_opt;
      const bool meas_row = pa < pr;
      keys[o] = ((meas_row ? pa * np + pr : pr * np + pa) << 1) | 1ull;
      vals[o] = ((unsigned long long)(meas_row ? kPairTagMeasRow : kPairTagRefRow) << 32) | sa;
      ++o;
    }
    if (!lm_on) continue;
    for (int j = i + 1; j < e; ++j) {
      const int ob = part_opt(P, j, rp);
      if (ob < 0) continue;
      const uint32_t sb = (uint32_t)j;
      const unsigned long long pb = (unsigned long long)ob;
      if (pa <= pb) { keys[o] = (pa * np + pb) << 1; vals[o] = ((unsigned long long)sb << 32) | sa; }
      else { keys[o] = (pb * np + pa) << 1; vals[o] = ((unsigned long long)sa << 32) | sb; }
      ++o;
    }
  }
}

// sorted (first, second) entries -> three arrays: observation of the row pose, observation of the column pose (or the
// tag of a measurement/reference entry), landmark
__global__ void k_pair_split(const unsigned long long* __restrict__ v, size_t n, const uint32_t* __restrict__ obs_lm, uint32_t* __restrict__ a,
                             uint32_t* __restrict__ b, uint32_t* __restrict__ lm) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long x = v[i];
  const uint32_t lo = (uint32_t)x;
  a[i] = lo; b[i] = (uint32_t)(x >> 32); lm[i] = obs_lm[lo];
}

#define SYM_CU(x)                                                                    \
  do {                                                                               \
    cudaError_t e_ = (x);                                                            \
    if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); goto done; } \
  } while (0)

}  // namespace

int build_pair_lists(const DeviceProblem& P, cudaStream_t st, uint32_t** pairs_out, size_t* n_pairs_out,
                     std::vector<unsigned long long>& block_key, std::vector<uint32_t>& block_count, std::string& err) {
  *pairs_out = nullptr; *n_pairs_out = 0;
  uint32_t* d_split = nullptr;
  block_key.clear(); block_count.clear();
  if (P.n_lms == 0 || P.n_obs == 0 || P.n_active_poses == 0) return 0;
  typedef unsigned long long u64;
  u64 *d_count = nullptr, *d_offs = nullptr, *d_keys = nullptr, *d_vals = nullptr, *d_keys2 = nullptr, *d_vals2 = nullptr, *d_unique = nullptr;
  uint32_t* d_runs = nullptr; int* d_nruns = nullptr; void* d_tmp = nullptr;
  size_t tmp_bytes = 0, tb = 0;
  u64 last_off = 0, last_cnt = 0, total = 0;
  int n_runs = 0, bits = 1, rc = 1;
  const int n = P.n_lms;
  SYM_CU(dev_alloc((void**)&d_count, sizeof(u64) * n));
  SYM_CU(dev_alloc((void**)&d_offs, sizeof(u64) * n));
  k_pair_count<<<(n + 255) / 256, 256, 0, st>>>(P, d_count);
  SYM_CU(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_count, d_offs, n, st));
  SYM_CU(dev_alloc((void**)&d_tmp, tmp_bytes ? tmp_bytes : 1));
  SYM_CU(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_count, d_offs, n, st));
  SYM_CU(cudaMemcpyAsync(&last_off, d_offs + (n - 1), sizeof(u64), cudaMemcpyDeviceToHost, st));
  SYM_CU(cudaMemcpyAsync(&last_cnt, d_count + (n - 1), sizeof(u64), cudaMemcpyDeviceToHost, st));
  SYM_CU(cudaStreamSynchronize(st));
  total = last_off + last_cnt;
  if (total == 0) { rc = 0; goto done; }
  if (total >= 0xfffffff0ull) { err = "pair list exceeds 2^32 entries"; goto done; }
  SYM_CU(dev_alloc((void**)&d_keys, sizeof(u64) * total));
  SYM_CU(dev_alloc((void**)&d_vals, sizeof(u64) * total));
  SYM_CU(dev_alloc((void**)&d_keys2, sizeof(u64) * total));
  SYM_CU(dev_alloc((void**)&d_vals2, sizeof(u64) * total));
  k_pair_emit<<<(n + 255) / 256, 256, 0, st>>>(P, d_offs, d_keys, d_vals);
  while ((1ull << bits) < 2ull * (u64)P.n_active_poses * (u64)P.n_active_poses && bits < 64) ++bits;
  SYM_CU(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_keys, d_keys2, d_vals, d_vals2, (u64)total, 0, bits, st));
  if (tb > tmp_bytes) { dev_free(d_tmp); d_tmp = nullptr; SYM_CU(dev_alloc((void**)&d_tmp, tb)); tmp_bytes = tb; }
  SYM_CU(cub::DeviceRadixSort::SortPairs(d_tmp, tb, d_keys, d_keys2, d_vals, d_vals2, (u64)total, 0, bits, st));
  // block table: unique keys and their run lengths (reuse d_keys / d_vals as outputs)
  d_unique = d_keys; d_runs = reinterpret_cast<uint32_t*>(d_vals);
  SYM_CU(dev_alloc((void**)&d_nruns, sizeof(int)));
  SYM_CU(cub::DeviceRunLengthEncode::Encode(nullptr, tb, d_keys2, d_unique, d_runs, d_nruns, (u64)total, st));
  if (tb > tmp_bytes) { dev_free(d_tmp); d_tmp = nullptr; SYM_CU(dev_alloc((void**)&d_tmp, tb)); tmp_bytes = tb; }
  SYM_CU(cub::DeviceRunLengthEncode::Encode(d_tmp, tb, d_keys2, d_unique, d_runs, d_nruns, (u64)total, st));
  SYM_CU(cudaMemcpyAsync(&n_runs, d_nruns, sizeof(int), cudaMemcpyDeviceToHost, st));
  SYM_CU(cudaStreamSynchronize(st));
  block_key.resize(n_runs); block_count.resize(n_runs);
  SYM_CU(cudaMemcpyAsync(block_key.data(), d_unique, sizeof(u64) * n_runs, cudaMemcpyDeviceToHost, st));
  SYM_CU(cudaMemcpyAsync(block_count.data(), d_runs, sizeof(uint32_t) * n_runs, cudaMemcpyDeviceToHost, st));
  SYM_CU(cudaStreamSynchronize(st));
  SYM_CU(dev_alloc((void**)&d_split, sizeof(uint32_t) * 3 * total));
  k_pair_split<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_vals2, (size_t)total, P.obs_lm, d_split, d_split + total, d_split + 2 * total);
  SYM_CU(cudaStreamSynchronize(st));
  *pairs_out = d_split; d_split = nullptr;   // ownership moves to the caller: [a (total) | b (total) | lm (total)]
  *n_pairs_out = (size_t)total;
  rc = 0;
done:
  dev_free(d_count); dev_free(d_offs); dev_free(d_keys); dev_free(d_vals); dev_free(d_keys2); dev_free(d_vals2);
  dev_free(d_nruns); dev_free(d_tmp); dev_free(d_split);
  return rc;
}

}  // namespace bagpu
