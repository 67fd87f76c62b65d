// setup_device.cu — symbolic set-up of one Solve() on the device.
//
// The reference keeps per-pose / per-landmark residual lists that its Add* calls fill one residual at a
// time (include/ba/BundleAdjuster.h:326-546) and walks them in BuildProblem (src/BundleAdjuster.cpp:1552-1802).
// Here the observations arrive as flat arrays in residual-id order, already in HBM; everything the hot path
// needs from the ids is derived there with a few sorts and scans instead of host loops over 10^7 residuals:
//   * landmark-major order (stable radix sort by landmark id  ->  ties keep ascending residual ids, the order
//     of `landmarks_[ii].proj_residuals`), CSR starts, gathered z / pose / cam / weight arrays
//   * pose-major observation lists (stable sort of the landmark-major stream by the measurement pose's opt id)
//   * LmSize 1: the landmarks referenced to every optimised pose
//   * which poses own a residual (BundleAdjuster.h:490-497), the conditioning-residual count (cpp:698)
//   * the tile pattern of S (which T x T tiles the pose pairs of any landmark touch)
// The radix sort is CUB's (set-up code, outside the per-iteration hot path).
#include <cub/device/device_radix_sort.cuh>
#include <string>
#include "ba_gpu.h"
#include "dev_alloc.h"
#include "launch.h"

namespace bagpu {

namespace {

__global__ void k_iota(uint32_t* v, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = i;
}

// keys ascending in [0, n_keys] (n_keys itself = "no key"): starts[l] = first position whose key is >= l, l = 0 .. n_keys
__global__ void k_key_starts(const uint32_t* __restrict__ keys, uint32_t n, uint32_t n_keys, uint32_t* __restrict__ starts) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  const uint32_t lo = (i == 0) ? 0u : keys[i - 1] + 1u;
  uint32_t hi = (i == n) ? n_keys : keys[i];
  if (hi > n_keys) hi = n_keys;
  for (uint32_t l = lo; l <= hi && l <= n_keys; ++l) starts[l] = i;
}

__global__ void k_validate_obs(const uint32_t* __restrict__ pose, const uint32_t* __restrict__ lm, const uint32_t* __restrict__ cam,
                               uint32_t n, uint32_t n_poses, uint32_t n_lms, int L, const uint32_t* __restrict__ lm_ref_pose,
                               const uint32_t* __restrict__ lm_ref_cam, int* __restrict__ flag) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t m = pose[i], l = lm[i];
  if (m >= n_poses || l >= n_lms) { *flag = 1; return; }
  const uint32_t c = cam ? cam[i] : 0u;
  if (L == 1 && m == lm_ref_pose[l] && c == lm_ref_cam[l]) atomicMax(flag + 1, 1);
}

// residual membership of the poses and the number of conditioning residuals (reference pose inactive, measurement
// pose active: BundleAdjuster.h:503-510)
__global__ void k_obs_flags(const uint32_t* __restrict__ pose, const uint32_t* __restrict__ lm, uint32_t n, int L,
                            const uint32_t* __restrict__ lm_ref_pose, const int* __restrict__ pose_opt,
                            uint8_t* __restrict__ has_res, unsigned int* __restrict__ n_cond) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool cond = false;
  if (i < n) {
    const uint32_t m = pose[i], r = lm_ref_pose[lm[i]];
    if (m != r || L != 1) {
      if (!has_res[m]) has_res[m] = 1;
      if (L == 1 && !has_res[r]) has_res[r] = 1;
    }
    cond = pose_opt[r] < 0 && pose_opt[m] >= 0;
  }
  const unsigned b = __ballot_sync(0xffffffffu, cond);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(n_cond, (unsigned int)__popc(b));
}

__global__ void k_gather_obs(const uint32_t* __restrict__ perm, uint32_t n, const double2* __restrict__ z, const uint32_t* __restrict__ pose,
                             const uint32_t* __restrict__ cam, const double* __restrict__ w0, double2* __restrict__ oz,
                             uint32_t* __restrict__ opose, uint32_t* __restrict__ ocam, double* __restrict__ ow0, uint32_t* __restrict__ operm) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = perm[i];
  oz[i] = z[r]; opose[i] = pose[r]; ow0[i] = w0 ? w0[r] : 1.0;
  if (operm) operm[i] = r;
  if (ocam) ocam[i] = cam ? cam[r] : 0u;
}

__global__ void k_local_starts(const uint32_t* __restrict__ gstarts, uint32_t n_lms, uint32_t s_lo, uint32_t s_hi, uint32_t* __restrict__ out) {
  const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l > n_lms) return;
  uint32_t v = gstarts[l];
  v = v < s_lo ? s_lo : (v > s_hi ? s_hi : v);
  out[l] = v - s_lo;
}

// key of the pose-major record slot of a (local, landmark-major) observation; lm_any[l] = the landmark has an
// observation from a pose other than its reference pose
__global__ void k_slot_keys(const uint32_t* __restrict__ obs_pose, const uint32_t* __restrict__ obs_lm, uint32_t n,
                            const uint32_t* __restrict__ lm_ref_pose, const int* __restrict__ pose_opt, uint32_t n_ap,
                            uint32_t* __restrict__ keys, uint8_t* __restrict__ lm_any) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t m = obs_pose[i], l = obs_lm[i];
  uint32_t k = n_ap;
  if (m != lm_ref_pose[l]) {
    if (lm_any && !lm_any[l]) lm_any[l] = 1;
    const int o = pose_opt[m];
    if (o >= 0) k = (uint32_t)o;
  }
  keys[i] = k;
}

// the pose-major observation list: position j of the pose-sorted order -> observation index and its landmark
__global__ void k_pose_lists(const uint32_t* __restrict__ vals_sorted, uint32_t n, const uint32_t* __restrict__ obs_lm,
                             uint32_t* __restrict__ pose_obs, uint32_t* __restrict__ pose_obs_lm) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t i = vals_sorted[j];
  pose_obs[j] = i; pose_obs_lm[j] = obs_lm[i];
}


__global__ void k_rslot_keys(const uint8_t* __restrict__ lm_any, const uint32_t* __restrict__ lm_ref_pose, const int* __restrict__ pose_opt,
                             uint32_t n_lms, uint32_t n_ap, uint32_t* __restrict__ keys) {
  const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n_lms) return;
  const int o = pose_opt[lm_ref_pose[l]];
  keys[l] = (lm_any[l] && o >= 0) ? (uint32_t)o : n_ap;
}

// tiles of S touched by the pose pairs of each landmark (all landmarks, all ranks: the pattern is global)
__global__ void k_tile_pattern(const uint32_t* __restrict__ starts, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ pose,
                               const uint32_t* __restrict__ lm_ref_pose, const int* __restrict__ pose_opt, uint32_t n_lms, int L, int nT,
                               uint8_t* __restrict__ present) {
  const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n_lms) return;
  const uint32_t s = starts[l], e = starts[l + 1], rp = lm_ref_pose[l];
  const int tr = (L == 1 && pose_opt[rp] >= 0) ? pose_opt[rp] / kTilePoses : -1;
  bool any = false;
  for (uint32_t i = s; i < e; ++i) {
    const uint32_t mi = pose[perm[i]];
    if (mi == rp || pose_opt[mi] < 0) continue;
    const int ti = pose_opt[mi] / kTilePoses;
    any = true;
    if (!present[(size_t)ti * nT + ti]) present[(size_t)ti * nT + ti] = 1;
    if (tr >= 0) {
      const size_t x = (size_t)min(ti, tr) * nT + max(ti, tr);
      if (!present[x]) present[x] = 1;
    }
    for (uint32_t j = i + 1; j < e; ++j) {
      const uint32_t mj = pose[perm[j]];
      if (mj == rp || pose_opt[mj] < 0) continue;
      const int tj = pose_opt[mj] / kTilePoses;
      const size_t x = (size_t)min(ti, tj) * nT + max(ti, tj);
      if (!present[x]) present[x] = 1;
    }
  }
  if (any && tr >= 0 && !present[(size_t)tr * nT + tr]) present[(size_t)tr * nT + tr] = 1;
}

// sum over the local conditioning residuals of maha / weight (cpp:699-702); maha is indexed by residual id
__global__ void k_cond_error(const uint32_t* __restrict__ obs_pose, const uint32_t* __restrict__ obs_lm, const uint32_t* __restrict__ obs_perm,
                             uint32_t n, const uint32_t* __restrict__ lm_ref_pose, const int* __restrict__ pose_opt,
                             const double* __restrict__ maha_by_id, const double* __restrict__ w, double* __restrict__ out) {
  double acc = 0.0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (pose_opt[lm_ref_pose[obs_lm[i]]] < 0 && pose_opt[obs_pose[i]] >= 0) acc += maha_by_id[obs_perm[i]] / w[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc != 0.0) atomicAdd(out, acc);
}

__global__ void k_scatter_by_perm(const uint32_t* __restrict__ perm, const double* __restrict__ in, uint32_t n, double* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[perm[i]] = in[i];
}

inline unsigned blocks_for(size_t n, int b = 256) { return (unsigned)((n + b - 1) / b); }

}  // namespace

// stable sort of (keys[i], i): keys_out ascending, vals_out = original positions.  0 on success.
int sort_positions_by_key(const uint32_t* keys_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t n, uint32_t max_key, cudaStream_t st,
                          std::string& err) {
  if (n == 0) return 0;
  uint32_t* iota = nullptr; void* tmp = nullptr; size_t tb = 0;
  int bits = 1, rc = 1;
  cudaError_t e;
  while (bits < 32 && (1ull << bits) <= (unsigned long long)max_key) ++bits;
  if ((e = dev_alloc((void**)&iota, sizeof(uint32_t) * (size_t)n)) != cudaSuccess) goto fail;
  k_iota<<<blocks_for(n), 256, 0, st>>>(iota, n);
  if ((e = cub::DeviceRadixSort::SortPairs(nullptr, tb, keys_in, keys_out, iota, vals_out, (unsigned long long)n, 0, bits, st)) != cudaSuccess) goto fail;
  if ((e = dev_alloc((void**)&tmp, tb ? tb : 1)) != cudaSuccess) goto fail;
  if ((e = cub::DeviceRadixSort::SortPairs(tmp, tb, keys_in, keys_out, iota, vals_out, (unsigned long long)n, 0, bits, st)) != cudaSuccess) goto fail;
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) goto fail;
  rc = 0;
fail:
  if (rc) err = std::string("sort_positions_by_key: ") + cudaGetErrorString(e);
  dev_free(iota); dev_free(tmp);
  return rc;
}

void launch_key_starts(const uint32_t* keys_sorted, uint32_t n, uint32_t n_keys, uint32_t* starts, cudaStream_t st) {
  k_key_starts<<<blocks_for((size_t)n + 1), 256, 0, st>>>(keys_sorted, n, n_keys, starts);
}
void launch_validate_obs(const uint32_t* pose, const uint32_t* lm, const uint32_t* cam, uint32_t n, uint32_t n_poses, uint32_t n_lms, int L,
                         const uint32_t* lm_ref_pose, const uint32_t* lm_ref_cam, int* flag2, cudaStream_t st) {
  if (n) k_validate_obs<<<blocks_for(n), 256, 0, st>>>(pose, lm, cam, n, n_poses, n_lms, L, lm_ref_pose, lm_ref_cam, flag2);
}
void launch_obs_flags(const uint32_t* pose, const uint32_t* lm, uint32_t n, int L, const uint32_t* lm_ref_pose, const int* pose_opt,
                      uint8_t* has_res, unsigned int* n_cond, cudaStream_t st) {
  if (n) k_obs_flags<<<blocks_for(n), 256, 0, st>>>(pose, lm, n, L, lm_ref_pose, pose_opt, has_res, n_cond);
}
void launch_gather_obs(const uint32_t* perm, uint32_t n, const double2* z, const uint32_t* pose, const uint32_t* cam, const double* w0,
                       double2* oz, uint32_t* opose, uint32_t* ocam, double* ow0, uint32_t* operm, cudaStream_t st) {
  if (n) k_gather_obs<<<blocks_for(n), 256, 0, st>>>(perm, n, z, pose, cam, w0, oz, opose, ocam, ow0, operm);
}
void launch_local_starts(const uint32_t* gstarts, uint32_t n_lms, uint32_t s_lo, uint32_t s_hi, uint32_t* out, cudaStream_t st) {
  k_local_starts<<<blocks_for((size_t)n_lms + 1), 256, 0, st>>>(gstarts, n_lms, s_lo, s_hi, out);
}
void launch_slot_keys(const uint32_t* obs_pose, const uint32_t* obs_lm, uint32_t n, const uint32_t* lm_ref_pose, const int* pose_opt,
                      uint32_t n_ap, uint32_t* keys, uint8_t* lm_any, cudaStream_t st) {
  if (n) k_slot_keys<<<blocks_for(n), 256, 0, st>>>(obs_pose, obs_lm, n, lm_ref_pose, pose_opt, n_ap, keys, lm_any);
}
void launch_pose_lists(const uint32_t* vals_sorted, uint32_t n, const uint32_t* obs_lm, uint32_t* pose_obs, uint32_t* pose_obs_lm, cudaStream_t st) {
  if (n) k_pose_lists<<<blocks_for(n), 256, 0, st>>>(vals_sorted, n, obs_lm, pose_obs, pose_obs_lm);
}
void launch_rslot_keys(const uint8_t* lm_any, const uint32_t* lm_ref_pose, const int* pose_opt, uint32_t n_lms, uint32_t n_ap, uint32_t* keys,
                       cudaStream_t st) {
  if (n_lms) k_rslot_keys<<<blocks_for(n_lms), 256, 0, st>>>(lm_any, lm_ref_pose, pose_opt, n_lms, n_ap, keys);
}
namespace {
__global__ void k_fill_tile_map(const int* __restrict__ row_ptr, const int* __restrict__ row_col, int nT, int* __restrict__ tile_map) {
  const int k = blockIdx.x;
  if (k >= nT) return;
  const int b = row_ptr[k], e = row_ptr[k + 1], base = k + b;
  if (threadIdx.x == 0) tile_map[(size_t)k * nT + k] = base;
  for (int q = b + threadIdx.x; q < e; q += blockDim.x) tile_map[(size_t)k * nT + row_col[q]] = base + 1 + (q - b);
}
}  // namespace
void launch_fill_tile_map(const int* row_ptr, const int* row_col, int nT, int* tile_map, cudaStream_t st) {
  if (nT <= 0) return;
  cudaMemsetAsync(tile_map, 0xFF, sizeof(int) * (size_t)nT * nT, st);
  k_fill_tile_map<<<nT, 64, 0, st>>>(row_ptr, row_col, nT, tile_map);
}
void launch_tile_pattern(const uint32_t* starts, const uint32_t* perm, const uint32_t* pose, const uint32_t* lm_ref_pose, const int* pose_opt,
                         uint32_t n_lms, int L, int nT, uint8_t* present, cudaStream_t st) {
  if (n_lms) k_tile_pattern<<<blocks_for(n_lms, 128), 128, 0, st>>>(starts, perm, pose, lm_ref_pose, pose_opt, n_lms, L, nT, present);
}
void launch_cond_error(const uint32_t* obs_pose, const uint32_t* obs_lm, const uint32_t* obs_perm, uint32_t n, const uint32_t* lm_ref_pose,
                       const int* pose_opt, const double* maha_by_id, const double* w, double* out, cudaStream_t st) {
  if (n) k_cond_error<<<std::min(blocks_for(n), 148u * 8u), 256, 0, st>>>(obs_pose, obs_lm, obs_perm, n, lm_ref_pose, pose_opt, maha_by_id, w, out);
}
void launch_scatter_by_perm(const uint32_t* perm, const double* in, uint32_t n, double* out, cudaStream_t st) {
  if (n) k_scatter_by_perm<<<blocks_for(n), 256, 0, st>>>(perm, in, n, out);
}
// is_reliable over ranks: phase 0 writes 1 for every landmark this rank marked unreliable, phase 1 (after the sum over
// ranks) clears the flag of every landmark any rank marked
__global__ void k_reliable_merge(uint8_t* rel, unsigned int* unrel, uint32_t n, int phase) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (phase == 0) unrel[i] = rel[i] ? 0u : 1u;
  else if (unrel[i]) rel[i] = 0;
}
void launch_iota_i32(int* v, uint32_t n, cudaStream_t st) {
  if (n) k_iota<<<blocks_for(n), 256, 0, st>>>(reinterpret_cast<uint32_t*>(v), n);
}
void launch_reliable_merge(uint8_t* rel, unsigned int* unrel, uint32_t n, int phase, cudaStream_t st) {
  if (n) k_reliable_merge<<<blocks_for(n), 256, 0, st>>>(rel, unrel, n, phase);
}

}  // namespace bagpu
