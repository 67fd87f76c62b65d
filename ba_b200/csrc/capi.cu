// capi.cu — C ABI of the B200 BundleAdjuster hot path (include/ba_b200.h) and the host
// orchestration of one Solve() call.
//
// Host side of (reference file:line):
//   Solve            src/BundleAdjuster.cpp:278-705     -> bagpu_solve
//   BuildProblem     src/BundleAdjuster.cpp:1166-1803   -> Handle::build (mask logic literal, kernels for the rest)
//   SolveInternal    src/BundleAdjuster.cpp:838-1161    -> Handle::solve_internal (scalars on host, vectors on device)
// Everything numerical runs in the kernels of proj_kernels.cu / other_kernels.cu / select.cu /
// chol.cu; there is no CPU fallback: without a CUDA device every entry point that would
// compute returns BAGPU_ERR_NO_DEVICE.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_set>
#include <iterator>
#include <thread>
#include <vector>
#include <dlfcn.h>
#include "ba_gpu.h"
#include "device_math.cuh"
#include "dev_alloc.h"
#include "peer_reduce.h"
#include "launch.h"

using namespace bagpu;

namespace {

thread_local std::string g_err;
std::mutex g_pinned_mu;
std::unordered_set<void*> g_pinned;   // buffers handed out by bagpu_host_alloc that are page-locked
// BAGPU_TIMING=1: wall-clock of the set-up phases on stderr (development aid)
struct PhaseClock {
  bool on; std::chrono::steady_clock::time_point t;
  PhaseClock() : on(getenv("BAGPU_TIMING") != nullptr), t(std::chrono::steady_clock::now()) {}
  void lap(const char* what) {
    if (!on) return;
    const auto n = std::chrono::steady_clock::now();
    fprintf(stderr, "[bagpu timing] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
    t = n;
  }
};
struct ApiClock { const char* what; PhaseClock pc; explicit ApiClock(const char* w) : what(w) {} ~ApiClock() { pc.lap(what); } };
int fail(int code, const std::string& msg) { g_err = msg; return code; }

#define CU(x)                                                                                        \
  do {                                                                                               \
    cudaError_t e_ = (x);                                                                            \
    if (e_ != cudaSuccess) throw std::string(#x) + ": " + cudaGetErrorString(e_);                    \
  } while (0)

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  bool view = false;   // p points into memory owned by somebody else (the peer-mapped system window)
  ~DevBuf() { release(); }
  void release() { if (p && !view) dev_free(p); p = nullptr; n = 0; view = false; }
  void adopt(T* ptr, size_t count) { release(); p = ptr; n = count; view = true; }
  void alloc(size_t count) {
    if (view) release();
    if (count <= n && p) return;
    release();
    if (count == 0) count = 1;
    CU(dev_alloc((void**)&p, count * sizeof(T)));
    n = count;
  }
  void upload(const std::vector<T>& h, cudaStream_t st) {
    alloc(h.size());
    if (!h.empty()) CU(cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, st));
  }
  void upload(const T* h, size_t count, cudaStream_t st) {
    alloc(count);
    if (count) CU(cudaMemcpyAsync(p, h, count * sizeof(T), cudaMemcpyHostToDevice, st));
  }
  void download(T* h, size_t count, cudaStream_t st) const {
    if (count) CU(cudaMemcpyAsync(h, p, count * sizeof(T), cudaMemcpyDeviceToHost, st));
  }
  void zero(cudaStream_t st, size_t count = (size_t)-1) { if (p) CU(cudaMemsetAsync(p, 0, std::min(count, n) * sizeof(T), st)); }
};

// ---- NCCL through dlopen: the library must load (and export its symbols) on a box without NCCL
struct UniqueId { char b[128]; };
struct NcclApi {
  void* lib = nullptr;
  int (*getUniqueId)(UniqueId*) = nullptr;
  int (*commInitRank)(void**, int, UniqueId, int) = nullptr;
  int (*allReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*commDestroy)(void*) = nullptr;
  int (*groupStart)() = nullptr;
  int (*groupEnd)() = nullptr;
  const char* (*getErrorString)(int) = nullptr;
  bool load() {
    if (lib) return true;
    const char* env = getenv("BAGPU_NCCL_LIB");
    const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      if (!nm) continue;
      lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) return false;
    getUniqueId = (int (*)(UniqueId*))dlsym(lib, "ncclGetUniqueId");
    commInitRank = (int (*)(void**, int, UniqueId, int))dlsym(lib, "ncclCommInitRank");
    allReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(lib, "ncclAllReduce");
    commDestroy = (int (*)(void*))dlsym(lib, "ncclCommDestroy");
    getErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
    groupStart = (int (*)())dlsym(lib, "ncclGroupStart");
    groupEnd = (int (*)())dlsym(lib, "ncclGroupEnd");
    return getUniqueId && commInitRank && allReduce && commDestroy;
  }
};
NcclApi g_nccl;
constexpr int kNcclFloat64 = 8, kNcclUint32 = 3, kNcclSum = 0;  // ncclDataType_t / ncclRedOp_t values (nccl.h)

}  // namespace

// Work items of the pose-pair-major reduction (k_reduce_blocks): every S block — the diagonal block of a pose over its
// record slots, an off-diagonal block over its run(s) of the pair list — is cut into chunks of at most kWorkChunk
// entries, one warp each; the chunks of one block share a group of partial-sum slots (mfirst, mcount, mself, mgroup).
// block_key (sorted) = (row * n_ap + col) * 2 + (1 for LmSize-1 measurement/reference entries), block_count = run
// lengths in the pair list.  Blocks with row == col from the pair list (one pose observing a landmark twice) go to `sym`
// (a second launch: they add into diagonal blocks that the first launch owns).
constexpr uint32_t kWorkChunk = 512;
static void build_work_items(int n_ap, int L, const uint32_t* pose_rec_start, const uint32_t* pose_lmrec_start,
                             const std::vector<unsigned long long>& bkey, const std::vector<uint32_t>& bcnt, size_t n_pairs,
                             std::vector<ReduceItem>& items, std::vector<ReduceItem>& sym, uint32_t& n_partial, uint32_t& n_groups) {
  constexpr uint32_t kChunk = kWorkChunk;
  items.clear(); sym.clear(); n_partial = 0; n_groups = 0;
  auto push_block = [&](std::vector<ReduceItem>& out, std::vector<ReduceItem>& chunks) {
    if (chunks.empty()) return;
    const uint32_t mc = (uint32_t)chunks.size();
    for (uint32_t k = 0; k < mc; ++k) {
      ReduceItem& it = chunks[k];
      it.mcount = mc; it.mfirst = 0; it.mself = 0; it.mgroup = 0;
      if (mc > 1) { it.mfirst = n_partial; it.mself = n_partial + k; it.mgroup = n_groups; }
      out.push_back(it);
    }
    if (mc > 1) { n_partial += mc; ++n_groups; }
    chunks.clear();
  };
  auto cut = [&](std::vector<ReduceItem>& chunks, uint32_t lo, uint32_t hi, uint32_t row, uint32_t col, uint32_t kind) {
    for (uint32_t b = lo; b < hi; b += kChunk) {
      ReduceItem it; memset(&it, 0, sizeof(it));
      it.begin = b; it.end = std::min(hi, b + kChunk); it.row = row; it.col = col; it.kind = kind;
      chunks.push_back(it);
    }
  };
  std::vector<ReduceItem> chunks;
  items.reserve(bkey.size() + n_pairs / kChunk + 4 * (size_t)n_ap + (size_t)(pose_rec_start[n_ap] + (L == 1 ? pose_lmrec_start[n_ap] : 0)) / kChunk + 16);
  size_t bi = 0; uint32_t pair_off = 0;
  for (int p = 0; p < n_ap; ++p) {
    cut(chunks, pose_rec_start[p], pose_rec_start[p + 1], p, p, kItemDiagObs);
    if (L == 1) cut(chunks, pose_lmrec_start[p], pose_lmrec_start[p + 1], p, p, kItemDiagLm);
    push_block(items, chunks);
    const unsigned long long row_lo = (unsigned long long)p * (unsigned long long)n_ap, row_hi = row_lo + (unsigned long long)n_ap;
    while (bi < bkey.size() && (bkey[bi] >> 1) < row_hi) {   // keys are sorted by (row, column, kind): no division per block
      // all runs of one block (entries between measurement poses, then measurement/reference entries) form one chunk chain
      const unsigned long long blk = bkey[bi] >> 1;
      const uint32_t col = (uint32_t)(blk - row_lo);
      const bool is_sym = (col == (uint32_t)p);
      if (!is_sym && bcnt[bi] <= kChunk && (bi + 1 == bkey.size() || (bkey[bi + 1] >> 1) != blk)) {
        // the common case, one run that fits one chunk: written in place
        const ReduceItem it = {pair_off, pair_off + bcnt[bi], (uint32_t)p, col, (bkey[bi] & 1ull) ? (uint32_t)kItemOffRef : (uint32_t)kItemOff, 0u, 1u, 0u, 0u, 0u};
        items.push_back(it);
        pair_off += bcnt[bi]; ++bi;
        continue;
      }
      while (bi < bkey.size() && (bkey[bi] >> 1) == blk) {
        const bool is_ref = (bkey[bi] & 1ull) != 0;
        cut(chunks, pair_off, pair_off + bcnt[bi], p, col, is_ref ? kItemOffRef : (is_sym ? kItemSym : kItemOff));
        pair_off += bcnt[bi]; ++bi;
      }
      push_block(is_sym ? sym : items, chunks);
    }
  }
}

// Warp tasks of the landmark-major kernels: greedy runs of whole landmarks [lo, hi) with at most 32 observations in
// total, as (first observation, count) relative to s_lo; landmarks with more than 32 observations are listed apart.
// gstarts = CSR bounds of the landmark-major observation stream, clamped to the shard [s_lo, s_hi).
static void pack_warp_tasks(const uint32_t* gstarts, int lo, int hi, uint32_t s_lo, uint32_t s_hi, std::vector<uint2>& tasks,
                            std::vector<uint32_t>& longs) {
  auto lstart = [&](int l) { const uint32_t v = gstarts[l]; return (v < s_lo ? s_lo : (v > s_hi ? s_hi : v)) - s_lo; };
  tasks.clear(); longs.clear();
  tasks.reserve((size_t)(s_hi - s_lo) / 24 + 16);
  int l = lo;
  uint32_t b = lstart(l);
  while (l < hi) {
    uint32_t e1 = lstart(l + 1);
    if (e1 - b > 32) { longs.push_back((uint32_t)l); ++l; b = e1; continue; }
    const uint32_t first = b;
    uint32_t end = b;
    while (l < hi) {
      e1 = lstart(l + 1);
      if (e1 - end > 32 || e1 - first > 32) break;
      end = e1; ++l;
    }
    if (end > first) tasks.push_back(make_uint2(first, end - first));
    b = end;
  }
}

struct bagpu_handle_s {
  bagpu_dims dims;
  bagpu_options opt;
  int L = 0, D = 6;
  bool has_device = false;
  int sms = 148;
  cudaStream_t st = nullptr;

  // ---- host copies of the problem (the reference's poses_/landmarks_/residual vectors) ----
  std::vector<double> cam_params, cam_tvs;
  std::vector<double> pose_t7, pose_v, pose_b;
  std::vector<uint8_t> pose_active, pose_mask_used, pose_mask;   // mask: D per pose, 1 = free
  std::vector<double> lm_xw, lm_xs;
  bool lm_host_fresh = true;   // false: the current landmarks live on the device only (sent there by the setter, or
                               // refined by a Solve()); the host vectors are filled on demand
  bool lm_rel_on_device = false;
  bool lm_ids_on_device = false;   // d_lm_ref_pose / d_lm_ref_cam hold the current arrays (sent by bagpu_set_proj_residuals)
  size_t n_lm_set = 0;         // landmarks handed to bagpu_set_landmarks
  std::vector<uint32_t> lm_ref_pose, lm_ref_cam;
  std::vector<uint8_t> lm_active, lm_reliable;
  std::vector<uint32_t> lm_outliers;
  std::vector<double> proj_z, proj_w0;
  std::vector<uint32_t> proj_pose, proj_lm, proj_cam;
  std::vector<uint32_t> un_pose; std::vector<double> un_t, un_cov_inv; std::vector<uint8_t> un_rot;
  std::vector<uint32_t> bin_p1, bin_p2; std::vector<double> bin_t, bin_cov_inv, bin_w; std::vector<uint8_t> bin_rot;
  std::vector<uint32_t> imu_p1, imu_p2, imu_ms; std::vector<double> imu_meas, imu_w;
  bool dirty = true;
  bool solved_once = false;

  // ---- derived host data ----
  int n_cams = 0, n_poses = 0, n_lms = 0, n_obs = 0, n_obs_total = 0, n_ap = 0, n_al = 0;
  int lm_owner_lo = 0, lm_owner_hi = 0;
  std::vector<int> pose_opt, lm_opt;
  bool lm_opt_valid = false;
  void host_lm_opt() {
    if (lm_opt_valid) return;
    lm_opt.assign(n_lms, -1);
    int k = 0;
    for (int i = 0; i < n_lms; ++i) if (lm_active[i]) lm_opt[i] = k++;
    lm_opt_valid = true;
  }
  std::vector<uint8_t> pose_has_res;     // any residual in the pose's lists
  std::vector<uint8_t> pose_has_imu;
  std::vector<uint32_t> lm_obs_start, obs_perm;   // host landmark-major order (pattern queries only, host_order())
  bool host_order_valid = false;
  bool maha0_valid = false;         // obs_maha0 holds |r|^2 w0 of the current state (written by the last accepted evaluation)
  size_t n_raw = 0;                 // projection residuals handed to set_proj_residuals
  bool raw_on_device = false, raw_has_cam = false, raw_has_w0 = false, ids_on_host = true;
  uint32_t s_lo = 0, s_hi = 0;      // this rank's slice of the landmark-major stream
  unsigned int n_cond_proj = 0;
  DevBuf<uint2> d_tasks; DevBuf<uint32_t> d_long_lms;   // packed warp tasks of the landmark-major stream (schur_kernels.cu)
  int n_tasks = 0, n_long = 0;
  std::thread packer;                       // greedy task packing, runs beside the device set-up
  std::vector<uint32_t> pk_gstarts, pk_longs; std::vector<uint2> pk_tasks;
  void join_packer() { if (packer.joinable()) packer.join(); }
  ~bagpu_handle_s() { join_packer(); }
  void finish_task_packing() {
    if (!packer.joinable()) return;
    packer.join();
    n_tasks = (int)pk_tasks.size(); n_long = (int)pk_longs.size();
    d_tasks.upload(pk_tasks, st); d_long_lms.upload(pk_longs, st);
    sync();
    std::vector<uint32_t>().swap(pk_gstarts);
  }
  std::vector<uint32_t> pose_rec_start, pose_lmrec_start;   // pose-major record slots (ba_gpu.h)
  uint32_t n_rec = 0, n_lmrec = 0;
  int n_items = 0, n_items_sym = 0;
  size_t n_pairs = 0;
  uint32_t* d_pairs = nullptr;         // from build_pair_lists: [a | b | landmark], n_pairs entries each
  int G = 8;
  int nT = 0, T = 48, n_tiles = 0, max_row = 0;
  std::vector<int> row_ptr, row_col, tile_perm;
  double trust_region = -1;
  bool is_param_mask_used = false;

  // ---- device buffers ----
  DevBuf<double> d_cam_params, d_cam_Rvs, d_cam_tvs;
  DevBuf<double> d_pose_q, d_pose_t, d_pose_v, d_pose_b, d_Tsw, d_Tws, d_Rwp, d_kd_pack;
  DevBuf<double> d_bk_q, d_bk_t, d_bk_v, d_bk_b, d_bk_lm_x; DevBuf<uint8_t> d_bk_rel; DevBuf<uint32_t> d_bk_outliers;
  DevBuf<int> d_pose_opt, d_lm_opt; DevBuf<uint32_t> d_pose_free; DevBuf<uint8_t> d_pose_mask_used;
  DevBuf<double> d_lm_x, d_lm_xw; DevBuf<uint32_t> d_lm_ref_pose, d_lm_ref_cam, d_lm_outliers, d_lm_obs_start;
  DevBuf<uint8_t> d_lm_reliable;
  DevBuf<double2> d_raw_z; DevBuf<uint32_t> d_raw_pose, d_raw_lm, d_raw_cam; DevBuf<double> d_raw_w0;   // residual-id order, as set
  DevBuf<uint32_t> d_sorted_lm, d_perm_global, d_gstarts;   // landmark-major: landmark id, residual id, CSR starts (all ranks' observations)
  DevBuf<double2> d_obs_z; DevBuf<uint32_t> d_obs_pose, d_obs_cam; DevBuf<double> d_obs_w0, d_obs_maha0, d_obs_w;
  DevBuf<uint32_t> d_pose_obs, d_pose_obs_lm, d_pose_lms; DevBuf<double> d_obs_G, d_obs_h, d_obs_gv; DevBuf<int> d_opt_to_pose;
  DevBuf<double> d_lm_rec, d_lm_recC, d_partial, d_red_partial; DevBuf<unsigned int> d_counters, d_red_counter;
  DevBuf<ReduceItem> d_items, d_items_sym;
  DevBuf<double> d_lm_bl;
  DevBuf<double> d_S, d_S_copy, d_uinv, d_scratch, d_bp, d_rhs_sc, d_rhs_work;
  DevBuf<int> d_tile_map, d_row_ptr, d_row_col, d_chol_sync, d_tile_perm;
  DevBuf<CholTask> d_chol_tasks; DevBuf<CholTerm> d_chol_terms; int n_chol_tasks = 0;
  DevBuf<Scalars> d_sc; DevBuf<SelectState> d_sel; DevBuf<unsigned long long> d_sel_cand, d_sel_all;
  DevBuf<double> d_bl_compact, d_gn_p, d_gn_l, d_dl_p, d_dl_l, d_tmp_p, d_tmp_l;
  DevBuf<double> d_res_out, d_maha_out;
  // other residuals
  DevBuf<uint32_t> d_un_pose, d_bin_p1, d_bin_p2, d_imu_p1, d_imu_p2, d_imu_ms;
  DevBuf<double> d_un_t, d_un_cov_inv, d_un_maha, d_bin_t, d_bin_cov_inv, d_bin_w, d_imu_meas, d_imu_w;
  DevBuf<double> d_imu_cov_inv, d_imu_J1, d_imu_J2, d_imu_res, d_un_out, d_bin_out, d_imu_out, d_imu_once; DevBuf<uint8_t> d_imu_once_done;
  DevBuf<uint8_t> d_un_rot, d_bin_rot;

  DeviceProblem P;
  OtherResiduals R;

  // ---- results ----
  bagpu_summary summary;
  std::vector<bagpu_iter_trace> trace;
  std::vector<std::vector<double>> rec_dp, rec_dl;
  double proj_error = 0, binary_error = 0, unary_error = 0, inertial_error = 0;
  float phase_ms[BAGPU_PH_COUNT] = {0};
  float dev_sd_ms = 0;   // development aid (BAGPU_TIMING): device time of the steepest-descent section, not part of any reported phase
  uint64_t launches = 0;
  cudaEvent_t ev[20];
  float dev_update_ms = 0, dev_eval_ms = 0, dev_factor_ms = 0;   // BAGPU_TIMING: finer split of the step / solve phases
  bool ev_ok = false;
  bool keep_S_copy = false;

  // ---- multi-GPU ----
  void* comm = nullptr;
  PeerReduce peer;   // small reductions over NVLink peer memory (peer_reduce.h)
  PeerBulk bulk;     // the reduced system (S tiles | b_p | rhs_sc) in one peer-mapped allocation; NCCL when it cannot be mapped
  std::vector<int> pcg_row_ptr; std::vector<int3> pcg_ent; DevBuf<int> d_pcg_row_ptr; DevBuf<int3> d_pcg_ent; DevBuf<double> d_pcg_work;
  bool pcg_uploaded = false; int pcg_iterations = 0;
  std::vector<int> red_slots; DevBuf<int> d_red_slots;   // tile slots of the symbolic pattern (fill tiles are zero on every rank)
  int rank = 0, world = 1;

  // =======================================================================================
  // every phase ends here: execution faults come back from the synchronise, launch-configuration errors (too much
  // shared memory, an empty grid, ...) of any launch since the last call from cudaGetLastError
  void sync() { CU(cudaStreamSynchronize(st)); CU(cudaGetLastError()); }

  void ensure_device() {
    if (has_device) return;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) throw std::string("no CUDA device (the product path has no CPU fallback)");
    int dev = 0;
    CU(cudaGetDevice(&dev));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));   // (cudaGetDeviceProperties takes milliseconds)
    set_sm_count(sms);
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& e : ev) CU(cudaEventCreate(&e));
    ev_ok = true;
    has_device = true;
  }

  // ---- set-up, host part: opt ids and the membership flags of the few unary / binary / inertial residuals ----
  void setup() {
    n_cams = (int)cam_params.size() / 4;
    n_poses = (int)pose_t7.size() / 7;
    n_lms = (int)n_lm_set;
    n_obs = n_obs_total = (int)n_raw;
    if (n_cams == 0) { n_cams = 1; cam_params = {1, 1, 0, 0}; cam_tvs = {0, 0, 0, 1, 0, 0, 0}; }
    pose_opt.assign(n_poses, -1); n_ap = 0;
    for (int i = 0; i < n_poses; ++i) if (pose_active[i]) pose_opt[i] = n_ap++;
    // opt ids of the landmarks: with every landmark active (the usual case) opt id = id, written on the device; the host
    // vector is built only when somebody walks it (pattern queries, bagpu_get_system, the J dump)
    n_al = (int)std::count_if(lm_active.begin(), lm_active.begin() + n_lms, [](uint8_t a) { return a != 0; });
    lm_opt.clear(); lm_opt_valid = false;
    // residual membership per pose (reference BundleAdjuster.h:490-497, :405, :452-453, :537-538); the projection
    // residuals add theirs on the device (device_symbolic) or in host_order()
    pose_has_res.assign(n_poses, 0); pose_has_imu.assign(n_poses, 0);
    for (uint32_t p : un_pose) pose_has_res[p] = 1;
    for (size_t i = 0; i < bin_p1.size(); ++i) { pose_has_res[bin_p1[i]] = 1; pose_has_res[bin_p2[i]] = 1; }
    for (size_t i = 0; i < imu_p1.size(); ++i) { pose_has_res[imu_p1[i]] = 1; pose_has_res[imu_p2[i]] = 1; pose_has_imu[imu_p1[i]] = 1; pose_has_imu[imu_p2[i]] = 1; }
    T = kTilePoses * D;
    nT = (n_ap + kTilePoses - 1) / kTilePoses;
    host_order_valid = false;
  }

  // landmark-major order of the residual ids on the host: only the pattern queries (bagpu_get_pattern) walk it;
  // the solver derives the same order on the device
  void host_ids() {
    if (ids_on_host) return;
    proj_pose.resize(n_raw); proj_lm.resize(n_raw);
    d_raw_pose.download(proj_pose.data(), n_raw, st); d_raw_lm.download(proj_lm.data(), n_raw, st);
    if (raw_has_cam) { proj_cam.resize(n_raw); d_raw_cam.download(proj_cam.data(), n_raw, st); }
    sync();
    ids_on_host = true;
  }
  void host_order() {
    host_ids();
    if (host_order_valid) return;
    const int n = (int)proj_pose.size();
    lm_obs_start.assign(n_lms + 1, 0);
    for (int i = 0; i < n; ++i) lm_obs_start[proj_lm[i] + 1]++;
    for (int l = 0; l < n_lms; ++l) lm_obs_start[l + 1] += lm_obs_start[l];
    obs_perm.resize(n);
    std::vector<uint32_t> cur(lm_obs_start.begin(), lm_obs_start.end() - 1);
    for (int i = 0; i < n; ++i) obs_perm[cur[proj_lm[i]]++] = i;
    host_order_valid = true;
  }

  // ---- set-up, device part (setup_device.cu): landmark-major observation stream, record slots, tile pattern ----
  void device_symbolic(std::vector<uint8_t>& present) {
    PhaseClock pc;
    std::string err;
    const uint32_t n = (uint32_t)n_obs_total;
    present.assign((size_t)nT * nT, 0);
    for (int k = 0; k < nT; ++k) present[(size_t)k * nT + k] = 1;
    pose_rec_start.assign(n_ap + 1, 0); pose_lmrec_start.assign(n_ap + 1, 0);
    n_rec = 0; n_lmrec = 0; n_cond_proj = 0; s_lo = 0; s_hi = n;
    lm_owner_lo = 0; lm_owner_hi = n_lms;
    G = 4; n_tasks = 0; n_long = 0;
    d_lm_obs_start.alloc((size_t)n_lms + 1); d_lm_obs_start.zero(st);
    if (n == 0 || n_lms == 0) { n_obs = 0; return; }
    if (!raw_on_device) {   // the setters ran before a device was usable: take the host copies
      d_raw_pose.upload(proj_pose, st); d_raw_lm.upload(proj_lm, st);
      if (raw_has_cam) d_raw_cam.upload(proj_cam, st);
      d_raw_z.upload(reinterpret_cast<const double2*>(proj_z.data()), n, st);
      if (raw_has_w0) d_raw_w0.upload(proj_w0, st);
      raw_on_device = true;
    }
    // 1. landmark-major order
    d_sorted_lm.alloc(n); d_perm_global.alloc(n); d_gstarts.alloc((size_t)n_lms + 1);
    if (sort_positions_by_key(d_raw_lm.p, d_sorted_lm.p, d_perm_global.p, n, (uint32_t)n_lms, st, err)) throw err;
    launch_key_starts(d_sorted_lm.p, n, (uint32_t)n_lms, d_gstarts.p, st);
    // 2. membership flags, conditioning count, tile pattern (global)
    DevBuf<uint8_t> d_has, d_present; DevBuf<unsigned int> d_cnt;
    d_has.upload(pose_has_res, st); d_cnt.alloc(1); d_cnt.zero(st);
    launch_obs_flags(d_raw_pose.p, d_raw_lm.p, n, L, d_lm_ref_pose.p, d_pose_opt.p, d_has.p, d_cnt.p, st);
    d_present.upload(present, st);
    if (n_ap > 0) launch_tile_pattern(d_gstarts.p, d_perm_global.p, d_raw_pose.p, d_lm_ref_pose.p, d_pose_opt.p, (uint32_t)n_lms, L, nT, d_present.p, st);
    std::vector<uint32_t> gstarts((size_t)n_lms + 1);
    d_gstarts.download(gstarts.data(), gstarts.size(), st);
    d_has.download(pose_has_res.data(), pose_has_res.size(), st);
    d_present.download(present.data(), present.size(), st);
    d_cnt.download(&n_cond_proj, 1, st);
    sync();
    pc.lap("  symbolic: sort + flags + pattern");
    // 3. landmark sharding: contiguous landmark ranges balanced by observation count; in the
    // landmark-major stream a shard is one contiguous slice
    if (world > 1) {
      const double per = (double)n / world;
      int lo = 0, hi = n_lms;
      for (int l = 0; l <= n_lms; ++l) {
        if ((double)gstarts[l] <= per * rank) lo = l;
        if ((double)gstarts[l] <= per * (rank + 1)) hi = l;
      }
      if (rank == world - 1) hi = n_lms;
      if (rank == 0) lo = 0;
      lm_owner_lo = lo; lm_owner_hi = hi;
      s_lo = gstarts[lo]; s_hi = gstarts[hi];
    }
    n_obs = (int)(s_hi - s_lo);
    if (pc.on) { sync(); pc.lap("    sharding"); }
    launch_local_starts(d_gstarts.p, (uint32_t)n_lms, s_lo, s_hi, d_lm_obs_start.p, st);
    // warp tasks of the landmark-major kernels: runs of whole landmarks with at most 32 observations in total
    // (first observation, count); landmarks with longer tracks are listed apart (one warp each).  Greedy packing from
    // the CSR bounds on a host thread that runs beside the rest of the set-up; finish_task_packing() joins it.
    join_packer();
    pk_gstarts.swap(gstarts);
    pk_tasks.clear(); pk_longs.clear();
    {
      const int lo = lm_owner_lo, hi = lm_owner_hi;
      const uint32_t slo = s_lo, shi = s_hi;
      packer = std::thread([this, lo, hi, slo, shi]() { pack_warp_tasks(pk_gstarts.data(), lo, hi, slo, shi, pk_tasks, pk_longs); });
    }
    // 4. the local stream: gathered values, pose-major record slots
    const uint32_t nl = (uint32_t)n_obs;
    d_obs_z.alloc(nl); d_obs_pose.alloc(nl); d_obs_w0.alloc(nl); d_obs_G.alloc((size_t)nl * kObsG); d_obs_h.alloc((size_t)nl * kObsH); if (L == 1) d_obs_gv.alloc((size_t)nl * kObsGv);
    if (n_cams > 1) d_obs_cam.alloc(nl);
    launch_gather_obs(d_perm_global.p + s_lo, nl, d_raw_z.p, d_raw_pose.p, raw_has_cam ? d_raw_cam.p : nullptr, raw_has_w0 ? d_raw_w0.p : nullptr,
                      d_obs_z.p, d_obs_pose.p, n_cams > 1 ? d_obs_cam.p : nullptr, d_obs_w0.p, nullptr, st);
    if (nl) {
      DevBuf<uint32_t> d_keys, d_keys_s, d_vals_s, d_starts; DevBuf<uint8_t> d_any;
      d_keys.alloc(std::max<uint32_t>(nl, (uint32_t)n_lms)); d_keys_s.alloc(std::max<uint32_t>(nl, (uint32_t)n_lms));
      d_vals_s.alloc(std::max<uint32_t>(nl, (uint32_t)n_lms)); d_starts.alloc((size_t)n_ap + 1);
      d_any.alloc((size_t)n_lms); d_any.zero(st);
      launch_slot_keys(d_obs_pose.p, d_sorted_lm.p + s_lo, nl, d_lm_ref_pose.p, d_pose_opt.p, (uint32_t)n_ap, d_keys.p, d_any.p, st);
      if (sort_positions_by_key(d_keys.p, d_keys_s.p, d_vals_s.p, nl, (uint32_t)n_ap, st, err)) throw err;
      launch_key_starts(d_keys_s.p, nl, (uint32_t)n_ap, d_starts.p, st);
      d_starts.download(pose_rec_start.data(), (size_t)n_ap + 1, st);
      sync();
      n_rec = pose_rec_start[n_ap];
      // the pose-major lists: the first n_rec positions of the pose-sorted order belong to optimised poses
      d_pose_obs.alloc(n_rec); d_pose_obs_lm.alloc(n_rec);
      launch_pose_lists(d_vals_s.p, n_rec, d_sorted_lm.p + s_lo, d_pose_obs.p, d_pose_obs_lm.p, st);
      if (L == 1) {   // reference-pose record slots of the landmarks
        launch_rslot_keys(d_any.p, d_lm_ref_pose.p, d_pose_opt.p, (uint32_t)n_lms, (uint32_t)n_ap, d_keys.p, st);
        if (sort_positions_by_key(d_keys.p, d_keys_s.p, d_vals_s.p, (uint32_t)n_lms, (uint32_t)n_ap, st, err)) throw err;
        launch_key_starts(d_keys_s.p, (uint32_t)n_lms, (uint32_t)n_ap, d_starts.p, st);
        d_starts.download(pose_lmrec_start.data(), (size_t)n_ap + 1, st);
        sync();
        n_lmrec = pose_lmrec_start[n_ap];
        d_pose_lms.alloc(n_lmrec);
        if (n_lmrec) CU(cudaMemcpyAsync(d_pose_lms.p, d_vals_s.p, sizeof(uint32_t) * n_lmrec, cudaMemcpyDeviceToDevice, st));
        sync();   // d_vals_s goes out of scope
      }
    }
    pc.lap("  symbolic: shard + gather + slots");
  }

  // ---- tile structure of S: binary / inertial couplings, elimination order, symbolic fill, tile table ----
  // `flat` is the nT x nT upper pattern from the observations (device_symbolic).  Everything after the one scan of it
  // works on sorted column lists per row, so the cost follows the number of tiles, not nT^2 (a 20k-pose graph has
  // nT = 2500).
  int tile_slot(int a, int b) const {   // storage slot of tile (a <= b) in permuted tile ids, -1 if structurally empty
    const int base = a + row_ptr[a];
    if (a == b) return base;
    const int* lo = row_col.data() + row_ptr[a];
    const int* hi = row_col.data() + row_ptr[a + 1];
    const int* it = std::lower_bound(lo, hi, b);
    return (it != hi && *it == b) ? base + 1 + (int)(it - lo) : -1;
  }
  void tile_structure(std::vector<uint8_t>& flat) {
    std::vector<std::vector<int>> rows(nT);   // columns > row of the upper pattern, natural tile ids
    for (int a = 0; a < nT; ++a) {
      const uint8_t* r = flat.data() + (size_t)a * nT;
      int b = a + 1;
      while (b < nT) {
        const void* hit = memchr(r + b, 1, (size_t)(nT - b));
        if (!hit) break;
        b = (int)(static_cast<const uint8_t*>(hit) - r);
        rows[a].push_back(b);
        ++b;
      }
    }
    bool marked = false;
    auto mark = [&](uint32_t p1, uint32_t p2) {
      const int o1 = pose_opt[p1], o2 = pose_opt[p2];
      if (o1 < 0 || o2 < 0) return;
      const int a = std::min(o1, o2) / kTilePoses, b = std::max(o1, o2) / kTilePoses;
      if (a != b) { rows[a].push_back(b); marked = true; }
    };
    for (size_t i = 0; i < bin_p1.size(); ++i) mark(bin_p1[i], bin_p2[i]);
    for (size_t i = 0; i < imu_p1.size(); ++i) mark(imu_p1[i], imu_p2[i]);
    if (marked) for (auto& r : rows) { std::sort(r.begin(), r.end()); r.erase(std::unique(r.begin(), r.end()), r.end()); }
    // ---- elimination order of the tiles: nested dissection (a banded S is a chain of nT dependent tile rows;
    // dissection turns it into a tree of depth ~log nT whose branches the dataflow solver runs concurrently) ----
    nested_dissection(rows);
    {
      std::vector<std::vector<int>> pp(nT);
      for (int a = 0; a < nT; ++a) for (int b : rows[a]) {
        const int x = tile_perm[a], y = tile_perm[b];
        pp[std::min(x, y)].push_back(std::max(x, y));
      }
      for (auto& r : pp) std::sort(r.begin(), r.end());
      rows.swap(pp);
    }
    const std::vector<std::vector<int>> pattern_rows = rows;   // before fill: the tiles the build can write
    // fill: eliminating tile row k couples every pair of its columns.  It is enough to merge the columns behind the
    // first one into the row of the first one (its parent in the elimination tree): the parent passes them on.
    {
      std::vector<int> merged;
      for (int k = 0; k < nT; ++k) {
        const std::vector<int>& cols = rows[k];
        if (cols.size() < 2) continue;
        std::vector<int>& par = rows[cols[0]];
        merged.clear(); merged.reserve(par.size() + cols.size());
        std::set_union(par.begin(), par.end(), cols.begin() + 1, cols.end(), std::back_inserter(merged));
        par.assign(merged.begin(), merged.end());
      }
    }
    // tiles are numbered row by row, the diagonal tile first (tile_slot()); the dense nT x nT table the kernels look
    // tiles up in is filled on the device (launch_fill_tile_map)
    row_ptr.assign(nT + 1, 0); row_col.clear();
    n_tiles = 0; max_row = 0;
    for (int k = 0; k < nT; ++k) {
      n_tiles += 1 + (int)rows[k].size();
      row_col.insert(row_col.end(), rows[k].begin(), rows[k].end());
      row_ptr[k + 1] = (int)row_col.size();
      max_row = std::max(max_row, row_ptr[k + 1] - row_ptr[k]);
    }
    red_slots.clear();
    for (int k = 0; k < nT; ++k) {
      red_slots.push_back(tile_slot(k, k));
      for (int b : pattern_rows[k]) red_slots.push_back(tile_slot(k, b));
    }
    // rows of y = S x for the PCG (pcg.cu): per tile row its diagonal tile, the tiles of its row of the upper storage and,
    // transposed, the tiles of its column, in ascending order of the other tile
    {
      std::vector<std::vector<int3>> re(nT);
      for (int a = 0; a < nT; ++a) re[a].push_back(make_int3(tile_slot(a, a), a, 2));
      for (int a = 0; a < nT; ++a) for (int b : pattern_rows[a]) {
        const int slot = tile_slot(a, b);
        re[a].push_back(make_int3(slot, b, 0));
        re[b].push_back(make_int3(slot, a, 1));
      }
      pcg_row_ptr.assign(nT + 1, 0); pcg_ent.clear();
      for (int a = 0; a < nT; ++a) {
        std::sort(re[a].begin(), re[a].end(), [](const int3& x, const int3& y) { return x.y < y.y; });
        pcg_ent.insert(pcg_ent.end(), re[a].begin(), re[a].end());
        pcg_row_ptr[a + 1] = (int)pcg_ent.size();
      }
      pcg_uploaded = false;
    }
  }

  // Nested dissection on the tile graph (upper pattern `present`, natural tile ids), generic recursive
  // bisection of index-sorted node lists: the separator is the smaller set of nodes on one side of the
  // split that touch the other side.  Falls back to the natural order where the separator is not small
  // (dense S).  Output: tile_perm[natural tile] = position in the elimination order.
  void nested_dissection(const std::vector<std::vector<int>>& upper) {
    tile_perm.assign(nT, 0);
    for (int i = 0; i < nT; ++i) tile_perm[i] = i;
    if (nT < 8 || getenv("BAGPU_NO_ND")) return;
    // a node list is split while its separator is smaller than 2 / nd_ratio of it (development knob BAGPU_ND_RATIO)
    const int nd_ratio = getenv("BAGPU_ND_RATIO") ? std::max(2, atoi(getenv("BAGPU_ND_RATIO"))) : 6;
    std::vector<std::vector<int>> adj(nT);
    for (int a = 0; a < nT; ++a) for (int b : upper[a]) { adj[a].push_back(b); adj[b].push_back(a); }
    for (auto& r : adj) std::sort(r.begin(), r.end());   // neighbours in ascending order, as a dense scan would list them
    std::vector<int> order; order.reserve(nT);
    std::vector<int> side(nT, 0);   // scratch: 1 = left, 2 = right of the current split
    struct Frame { std::vector<int> nodes; bool emit; };
    std::vector<Frame> stack;
    { Frame f; f.nodes.resize(nT); for (int i = 0; i < nT; ++i) f.nodes[i] = i; f.emit = false; stack.push_back(std::move(f)); }
    while (!stack.empty()) {
      Frame f = std::move(stack.back()); stack.pop_back();
      const int n = (int)f.nodes.size();
      if (f.emit || n <= 4) { for (int v : f.nodes) order.push_back(v); continue; }
      const int mid = n / 2;
      for (int k = 0; k < n; ++k) side[f.nodes[k]] = k < mid ? 1 : 2;
      std::vector<int> sl, sr;
      for (int k = 0; k < n; ++k) {
        const int v = f.nodes[k], other = k < mid ? 2 : 1;
        bool touch = false;
        for (int u : adj[v]) if (side[u] == other) { touch = true; break; }
        if (touch) (k < mid ? sl : sr).push_back(v);
      }
      const bool use_left = sl.size() <= sr.size();
      const std::vector<int>& sep = use_left ? sl : sr;
      for (int v : f.nodes) side[v] = 0;
      if (sep.empty() || (int)sep.size() * nd_ratio >= n * 2) { for (int v : f.nodes) order.push_back(v); continue; }
      std::vector<uint8_t> in_sep_mark;   // small: mark through `side`
      for (int v : sep) side[v] = 3;
      Frame left, right, sp;
      for (int k = 0; k < n; ++k) { const int v = f.nodes[k]; if (side[v] == 3) continue; (k < mid ? left : right).nodes.push_back(v); }
      for (int v : sep) side[v] = 0;
      left.emit = right.emit = false;
      sp.nodes = sep; sp.emit = true;
      // post-order: left subtree, right subtree, separator (stack pops in reverse)
      stack.push_back(std::move(sp)); stack.push_back(std::move(right)); stack.push_back(std::move(left));
    }
    for (int pos = 0; pos < nT; ++pos) tile_perm[order[pos]] = pos;
  }

  void upload_all() {
    PhaseClock pc;
    // cameras
    std::vector<double> Rvs(9 * n_cams), tvs(3 * n_cams);
    for (int c = 0; c < n_cams; ++c) {
      quat_normalize(&cam_tvs[7 * c]);   // Sophus::SO3Group's constructor normalises
      quat_to_R(&cam_tvs[7 * c], &Rvs[9 * c]);
      for (int k = 0; k < 3; ++k) tvs[3 * c + k] = cam_tvs[7 * c + 4 + k];
    }
    d_cam_params.upload(cam_params, st); d_cam_Rvs.upload(Rvs, st); d_cam_tvs.upload(tvs, st);
    // poses
    std::vector<double> q(4 * n_poses), t(3 * n_poses);
    for (int i = 0; i < n_poses; ++i) { quat_normalize(&pose_t7[7 * i]); for (int k = 0; k < 4; ++k) q[4 * i + k] = pose_t7[7 * i + k]; for (int k = 0; k < 3; ++k) t[3 * i + k] = pose_t7[7 * i + 4 + k]; }
    for (size_t i = 0; i < un_pose.size(); ++i) quat_normalize(&un_t[7 * i]);
    for (size_t i = 0; i < bin_p1.size(); ++i) quat_normalize(&bin_t[7 * i]);
    d_pose_q.upload(q, st); d_pose_t.upload(t, st); d_pose_v.upload(pose_v, st); d_pose_b.upload(pose_b, st);
    d_bk_q.alloc(q.size()); d_bk_t.alloc(t.size()); d_bk_v.alloc(pose_v.size()); d_bk_b.alloc(pose_b.size());
    d_pose_opt.upload(pose_opt, st); d_pose_free.alloc(n_poses); d_pose_mask_used.alloc(n_poses);
    d_Tsw.alloc((size_t)n_poses * n_cams * 12); d_Tws.alloc((size_t)n_poses * n_cams * 12); d_Rwp.alloc((size_t)n_poses * 9); d_kd_pack.alloc((size_t)n_poses * 12);
    // landmarks
    if (lm_host_fresh) d_lm_xw.upload(lm_xw, st);
    d_lm_x.alloc(4 * n_lm_set); d_bk_lm_x.alloc(4 * n_lm_set);
    if (!lm_ids_on_device) { d_lm_ref_pose.upload(lm_ref_pose, st); d_lm_ref_cam.upload(lm_ref_cam, st); lm_ids_on_device = true; }
    if (n_al == n_lms) { d_lm_opt.alloc((size_t)n_lms); launch_iota_i32(d_lm_opt.p, (uint32_t)n_lms, st); }
    else { host_lm_opt(); d_lm_opt.upload(lm_opt, st); }
    if (!lm_rel_on_device) {   // every landmark starts reliable (host copies of these flags exist only after a fetch)
      d_lm_reliable.alloc(n_lm_set);
      if (n_lm_set) CU(cudaMemsetAsync(d_lm_reliable.p, 1, n_lm_set, st));
      lm_rel_on_device = true;
    }
    d_bk_rel.alloc(n_lm_set); d_lm_outliers.alloc(n_lms); d_bk_outliers.alloc(n_lms);
    d_lm_outliers.zero(st);
    pc.lap("  upload: cameras, poses, landmarks");
    // observations: landmark-major stream, record slots and the tile pattern, derived on the device
    std::vector<uint8_t> present;
    device_symbolic(present);
    pc.lap("  device symbolic");
    // The tile structure of S (elimination order, symbolic fill) and the task plan of the factorisation are host-only
    // work (3 - 4 ms for 10k poses): they run on a helper thread beside the pair-list set-up below, which needs
    // nothing of S.
    std::vector<CholTask> plan_tasks; std::vector<CholTerm> plan_terms;
    std::string tiler_err;
    std::thread tiler([&]() {
      try { tile_structure(present); chol_build_plan(T, nT, row_ptr, row_col, plan_tasks, plan_terms); }
      catch (const std::exception& e) { tiler_err = e.what(); }
    });
    struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } tiler_guard{tiler};
    const bool multi_cam = n_cams > 1;
    d_obs_maha0.alloc(n_obs); d_obs_w.alloc(n_obs);
    const int Ls = std::max(1, L);
    d_lm_rec.alloc((size_t)n_lms * kLmRec); d_lm_rec.zero(st);
    if (L == 1) d_lm_recC.alloc((size_t)n_lms * kLmRecC);
    d_lm_bl.alloc((size_t)n_lms * Ls);
    d_lm_bl.zero(st);
    d_red_partial.alloc((size_t)kRedMaxGrid * 8);
    if (!d_red_counter.p) { d_red_counter.alloc(4); d_red_counter.zero(st); }
    {
      std::vector<int> o2p((size_t)std::max(1, n_ap), 0);
      for (int i = 0; i < n_poses; ++i) if (pose_opt[i] >= 0) o2p[pose_opt[i]] = i;
      d_opt_to_pose.upload(o2p, st);
      sync();
    }
    pc.lap("  upload: allocs");
    d_sc.alloc(1); d_sc.zero(st); d_sel.alloc(1);
    d_bl_compact.alloc((size_t)n_al * Ls); d_gn_l.alloc((size_t)n_al * Ls);
    d_dl_l.alloc((size_t)n_al * Ls); d_tmp_l.alloc((size_t)n_al * Ls);
    d_gn_l.zero(st); d_dl_l.zero(st); d_bl_compact.zero(st);
    // other residuals
    d_un_pose.upload(un_pose, st); d_un_t.upload(un_t, st); d_un_cov_inv.upload(un_cov_inv, st); d_un_rot.upload(un_rot, st); d_un_maha.alloc(un_pose.size());
    d_bin_p1.upload(bin_p1, st); d_bin_p2.upload(bin_p2, st); d_bin_t.upload(bin_t, st); d_bin_cov_inv.upload(bin_cov_inv, st);
    d_bin_w.upload(bin_w, st); d_bin_rot.upload(bin_rot, st);
    d_imu_p1.upload(imu_p1, st); d_imu_p2.upload(imu_p2, st); d_imu_ms.upload(imu_ms, st); d_imu_meas.upload(imu_meas, st); d_imu_w.upload(imu_w, st);
    const size_t ni = imu_p1.size();
    d_imu_cov_inv.alloc(ni * D * D); d_imu_J1.alloc(ni * D * D); d_imu_J2.alloc(ni * D * D); d_imu_res.alloc(ni * D);
    d_un_out.alloc(7 * un_pose.size()); d_bin_out.alloc(7 * bin_p1.size()); d_imu_out.alloc(ni * (D + 1));
    d_un_out.zero(st); d_bin_out.zero(st); d_imu_out.zero(st);
    d_imu_once.alloc(ni * 160); d_imu_once_done.alloc(ni); d_imu_once_done.zero(st);

    memset(&P, 0, sizeof(P));
    P.L = L; P.D = D; P.n_cams = n_cams; P.n_poses = n_poses; P.n_active_poses = n_ap; P.n_lms = n_lms; P.n_active_lms = n_al; P.n_obs = n_obs;
    P.n_unary = (int)un_pose.size(); P.n_binary = (int)bin_p1.size(); P.n_imu = (int)imu_p1.size();
    P.cam_params = d_cam_params.p; P.cam_Rvs = d_cam_Rvs.p; P.cam_tvs = d_cam_tvs.p;
    P.pose_q = d_pose_q.p; P.pose_t = d_pose_t.p; P.pose_v = d_pose_v.p; P.pose_b = d_pose_b.p; P.pose_opt = d_pose_opt.p; P.pose_free = d_pose_free.p;
    P.Tsw = d_Tsw.p; P.Tws = d_Tws.p; P.Rwp = d_Rwp.p; P.kd_pack = d_kd_pack.p;
    P.lm_x = d_lm_x.p; P.lm_ref_pose = d_lm_ref_pose.p; P.lm_ref_cam = d_lm_ref_cam.p; P.lm_opt = d_lm_opt.p; P.lm_reliable = d_lm_reliable.p;
    P.lm_outliers = d_lm_outliers.p; P.lm_obs_start = d_lm_obs_start.p;
    P.obs_z = d_obs_z.p; P.obs_pose = d_obs_pose.p; P.obs_cam = multi_cam ? d_obs_cam.p : nullptr; P.obs_w0 = d_obs_w0.p; P.obs_perm = d_perm_global.p ? d_perm_global.p + s_lo : nullptr;
    P.obs_lm = d_sorted_lm.p ? d_sorted_lm.p + s_lo : nullptr;
    P.obs_maha0 = d_obs_maha0.p; P.obs_w = d_obs_w.p;
    P.obs_G = d_obs_G.p; P.obs_h = d_obs_h.p; P.obs_gv = d_obs_gv.p; P.pose_obs = d_pose_obs.p; P.pose_obs_lm = d_pose_obs_lm.p; P.pose_lms = d_pose_lms.p; P.opt_to_pose = d_opt_to_pose.p;
    P.lm_rec = d_lm_rec.p; P.lm_recC = (L == 1) ? d_lm_recC.p : nullptr; P.lm_bl = d_lm_bl.p;
    P.red_partial = d_red_partial.p; P.red_counter = d_red_counter.p;
    P.cam_identity = 1;
    for (int c = 0; c < n_cams; ++c) {
      const double* t7 = &cam_tvs[7 * c];
      if (!(t7[0] == 0.0 && t7[1] == 0.0 && t7[2] == 0.0 && t7[3] == 1.0 && t7[4] == 0.0 && t7[5] == 0.0 && t7[6] == 0.0)) P.cam_identity = 0;
    }
    P.sc = d_sc.p; P.T = T;
    P.c_huber_factor = 1.2107;

    memset(&R, 0, sizeof(R));
    R.n_unary = P.n_unary; R.n_binary = P.n_binary; R.n_imu = P.n_imu;
    R.un_pose = d_un_pose.p; R.un_t = d_un_t.p; R.un_cov_inv = d_un_cov_inv.p; R.un_use_rot = d_un_rot.p; R.un_maha = d_un_maha.p;
    R.bin_p1 = d_bin_p1.p; R.bin_p2 = d_bin_p2.p; R.bin_t12 = d_bin_t.p; R.bin_cov_inv = d_bin_cov_inv.p; R.bin_weight = d_bin_w.p; R.bin_use_rot = d_bin_rot.p;
    R.imu_p1 = d_imu_p1.p; R.imu_p2 = d_imu_p2.p; R.imu_meas_start = d_imu_ms.p; R.imu_meas = d_imu_meas.p; R.imu_weight = d_imu_w.p;
    R.imu_cov_inv = d_imu_cov_inv.p; R.imu_J1 = d_imu_J1.p; R.imu_J2 = d_imu_J2.p; R.imu_res = d_imu_res.p;
    R.un_out = d_un_out.p; R.bin_out = d_bin_out.p; R.imu_out = d_imu_out.p; R.imu_once = d_imu_once.p; R.imu_once_done = d_imu_once_done.p;
    apply_options();
    sync(); pc.lap("  upload: rest");
    build_reduce_items();
    pc.lap("  build_reduce_items");
    tiler.join();
    if (!tiler_err.empty()) throw std::string("tile structure: ") + tiler_err;
    pc.lap("  tile structure + factorisation plan (join)");
    // reduced system
    const size_t tsz = (size_t)T * T;
    d_uinv.alloc((size_t)nT * tsz); d_scratch.alloc(3 * (size_t)nT * T + 2 * (size_t)n_tiles * T);
    d_rhs_work.alloc((size_t)nT * T);
    {
      // the system the ranks sum: S tiles | b_p | rhs_sc, in one peer-mapped allocation when the windows work
      const size_t nS = (size_t)n_tiles * tsz, nV = (size_t)nT * T;
      ensure_bulk_window((nS + 2 * nV) * sizeof(double));
      if (bulk.ok) {
        double* b = bulk.base[rank];
        d_S.adopt(b, nS); d_bp.adopt(b + nS, nV); d_rhs_sc.adopt(b + nS + nV, nV);
        d_red_slots.upload(red_slots, st);
      } else {
        d_S.alloc(nS); d_bp.alloc(nV); d_rhs_sc.alloc(nV);
      }
    }
    d_row_ptr.upload(row_ptr, st); d_row_col.upload(row_col, st); d_tile_perm.upload(tile_perm, st);
    d_tile_map.alloc((size_t)nT * nT); launch_fill_tile_map(d_row_ptr.p, d_row_col.p, nT, d_tile_map.p, st);
    {
      n_chol_tasks = (int)plan_tasks.size();
      if (getenv("BAGPU_CHOL_TRACE")) h_chol_tasks = plan_tasks;
      d_chol_tasks.upload(plan_tasks, st); d_chol_terms.upload(plan_terms, st); d_chol_sync.alloc(chol_sync_ints(n_tiles, nT));
      sync();
    }
    pc.lap("  upload: S + chol plan");
    d_gn_p.alloc((size_t)nT * T); d_dl_p.alloc((size_t)nT * T); d_tmp_p.alloc((size_t)nT * T);
    d_gn_p.zero(st); d_dl_p.zero(st);
    P.S_tiles = d_S.p; P.tile_map = d_tile_map.p; P.tile_perm = d_tile_perm.p; P.nT = nT; P.b_p = d_bp.p; P.rhs_sc = d_rhs_sc.p;
    sync();
    finish_task_packing();
    pc.lap("  task packing (join)");
    dirty = false;
  }

  // (Re)allocate the peer-mapped system window; every rank calls this with the same size at the same point of Solve().
  void ensure_bulk_window(size_t bytes) {
    if (world <= 1 || !peer.ok || getenv("BAGPU_NO_BULK")) return;
    if (bulk.ok && bulk.cap_bytes >= bytes) return;
    DevBuf<unsigned int> d_tab; d_tab.alloc((size_t)world * 16 + 1);
    if (bulk.base[rank] || bulk.ok) {   // unmap, meet, free (a window may be freed only after every importer has unmapped it)
      d_S.release(); d_bp.release(); d_rhs_sc.release();
      bulk.close_peers();
      d_tab.zero(st); allreduce_u32(d_tab.p, 1); sync();
      bulk.destroy();
    }
    const size_t cap = ((bytes + bytes / 4) + ((size_t)2 << 20) - 1) / ((size_t)2 << 20) * ((size_t)2 << 20);
    std::vector<unsigned int> tab((size_t)world * 16 + 1, 0u);
    bool mine = bulk.create_local(rank, world, cap, &tab[(size_t)rank * 16]) == cudaSuccess;
    if (!mine) { cudaGetLastError(); memset(&tab[(size_t)rank * 16], 0, 64); }
    CU(cudaMemcpyAsync(d_tab.p, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, st));
    nccl_allreduce_u32(d_tab.p, tab.size());
    CU(cudaMemcpyAsync(tab.data(), d_tab.p, tab.size() * 4, cudaMemcpyDeviceToHost, st));
    sync();
    const bool opened = mine && bulk.open_peers(tab.data());
    unsigned int vote = opened ? 1u : 0u;
    CU(cudaMemcpyAsync(d_tab.p, &vote, 4, cudaMemcpyHostToDevice, st));
    nccl_allreduce_u32(d_tab.p, 1);
    CU(cudaMemcpyAsync(&vote, d_tab.p, 4, cudaMemcpyDeviceToHost, st));
    sync();
    if (vote == (unsigned int)world) bulk.ok = true;
    else { bulk.close_peers(); nccl_allreduce_u32(d_tab.p, 1); sync(); bulk.destroy(); }
    if (getenv("BAGPU_TIMING")) fprintf(stderr, "[bagpu] rank %d: peer-mapped system window %s (%.1f MB)\n", rank, bulk.ok ? "on" : "off (NCCL)", cap / 1e6);
  }
  void nccl_allreduce_u32(unsigned int* buf, size_t n) {
    int rc = g_nccl.allReduce(buf, buf, n, kNcclUint32, kNcclSum, comm, st);
    if (rc != 0) throw std::string("ncclAllReduce(u32) failed");
  }

  // ---- work items of the pose-pair-major reduction (schur_kernels.cu): the device builds the pair
  // lists (symbolic.cu), the host cuts every S block into chunks of at most kChunk entries ----
  void build_reduce_items() {
    if (d_pairs) { dev_free(d_pairs); d_pairs = nullptr; }
    n_pairs = 0; n_items = 0; n_items_sym = 0;
    if (L == 0 || n_obs == 0 || n_ap == 0) return;
    std::vector<unsigned long long> bkey; std::vector<uint32_t> bcnt; std::string err;
    PhaseClock pcr;
    if (build_pair_lists(P, st, &d_pairs, &n_pairs, bkey, bcnt, err) != 0) throw std::string("pair lists: ") + err;
    pcr.lap("    pair lists (device)");
    std::vector<ReduceItem> items, sym;
    uint32_t n_partial = 0, n_groups = 0;
    build_work_items(n_ap, L, pose_rec_start.data(), pose_lmrec_start.data(), bkey, bcnt, n_pairs, items, sym, n_partial, n_groups);
    n_items = (int)items.size(); n_items_sym = (int)sym.size();
    pcr.lap("    work items (host)");
    d_items.upload(items, st); d_items_sym.upload(sym, st);
    d_partial.alloc((size_t)std::max<uint32_t>(1, n_partial) * 64); d_counters.alloc((size_t)n_groups + 2); d_counters.zero(st);   // [0] = item counter of the running launch
    sync();   // the host vectors above go out of scope
  }

  // Options that the kernels read (Init(), BundleAdjuster.h:192-218)
  void apply_options() {
    P.use_robust = (opt.use_robust_norm_for_proj_residuals ? 1 : 0) | (opt.use_robust_norm_for_inertial_residuals ? 2 : 0);
    P.outlier_threshold = opt.projection_outlier_threshold;
    R.cov_once = opt.calculate_inertial_covariance_once ? 1 : 0;
    for (int k = 0; k < 3; ++k) R.gravity[k] = opt.gravity[k];
    for (int k = 0; k < 3; ++k) {
      R.imu_r[k] = opt.gyro_sigma * opt.gyro_sigma; R.imu_r[3 + k] = opt.accel_sigma * opt.accel_sigma;
      R.imu_rb[k] = opt.gyro_bias_sigma * opt.gyro_bias_sigma; R.imu_rb[3 + k] = opt.accel_bias_sigma * opt.accel_bias_sigma;
    }
  }

  // ---- parameter masks, literal from src/BundleAdjuster.cpp:1237-1330 ------------------------
  uint32_t gravity_reg_dim(const double* q) {   // BundleAdjuster.h:634-652
    double Rm[9];
    quat_to_R(q, Rm);
    double max_dot = 0; uint32_t max_dim = 0;
    for (uint32_t ii = 0; ii < 3; ++ii) {
      const double d = std::fabs(Rm[ii] * opt.gravity[0] + Rm[3 + ii] * opt.gravity[1] + Rm[6 + ii] * opt.gravity[2]);
      if (d > max_dot) { max_dot = d; max_dim = ii; }
    }
    return max_dim + 3;
  }

  void compute_masks() {
    const bool vel = D >= 9, bias = D >= 15;
    bool are_all_active = true;
    for (int i = 0; i < n_poses; ++i) {
      if (!pose_active[i]) { are_all_active = false; break; }      // quirk Q8
      if (!pose_has_res[i]) { pose_mask_used[i] = 1; for (int k = 0; k < D; ++k) pose_mask[(size_t)i * D + k] = 0; }
    }
    if (vel) {
      for (int i = 0; i < n_poses; ++i) {
        if (!pose_has_imu[i] && pose_active[i]) {
          pose_mask_used[i] = 1;
          for (int k = 0; k < D; ++k) pose_mask[(size_t)i * D + k] = 1;
          pose_mask[(size_t)i * D + 6] = pose_mask[(size_t)i * D + 7] = pose_mask[(size_t)i * D + 8] = 0;
          if (bias) for (int k = 9; k < 15; ++k) pose_mask[(size_t)i * D + k] = 0;
        }
      }
    }
    if (are_all_active && un_pose.empty() && opt.enable_auto_regularization && n_poses > 0) {
      const int r = (int)opt.root_pose_id;
      uint8_t* m = &pose_mask[(size_t)r * D];
      pose_mask_used[r] = 1;
      for (int k = 0; k < D; ++k) m[k] = 1;
      m[0] = m[1] = m[2] = 0;
      if (bias && opt.regularize_biases_in_batch) for (int k = 9; k < 15; ++k) m[k] = 0;
      if (!vel) {
        m[3] = m[4] = m[5] = 0;
      } else {
        double q[4];
        CU(cudaMemcpyAsync(q, d_pose_q.p + 4 * (size_t)r, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
        sync();
        m[gravity_reg_dim(q)] = 0;
      }
    }
    std::vector<uint32_t> fr(n_poses, 0);
    is_param_mask_used = false;
    for (int i = 0; i < n_poses; ++i) {
      uint32_t bits = (D >= 32) ? 0xffffffffu : ((1u << D) - 1u);
      if (pose_mask_used[i]) {
        bits = 0;
        for (int k = 0; k < D; ++k) if (pose_mask[(size_t)i * D + k]) bits |= (1u << k);
        if (pose_active[i] && pose_has_res[i]) is_param_mask_used = true;
      }
      fr[i] = bits;
    }
    d_pose_free.upload(fr, st);
    d_pose_mask_used.upload(pose_mask_used, st);
  }

  // ---- median of maha0 (exact) -------------------------------------------------------------
  // `distributed`: x is one rank's shard of n_total values (histograms summed over ranks); false: x is the whole,
  // replicated list and no collective is issued (unary residuals, evaluated on rank 0 only)
  void run_select(const double* x, size_t n_local, unsigned long long n_total, double* out, bool distributed = true) {
    if (n_total == 0) return;
    const bool multi = world > 1 && distributed;
    launch_select_init(d_sel.p, n_total / 2, st); ++launches;
    int pass = 0;
    // Two 12-bit passes (sign + exponent + 13 mantissa bits), then — when the bucket of the k-th value is small enough for one
    // CTA, which it is unless thousands of values tie — its members are collected, gathered over the ranks and the selection is
    // finished by one CTA: 2 histogram exchanges + 1 gather instead of 6 exchanges, 3 passes over the data instead of 6.
    // The branch is taken on the GLOBAL bucket count, which every rank reads: all ranks issue the same collectives.
    const bool cand_path = (!multi || peer.ok) && !getenv("BAGPU_SELECT_RADIX");
    if (cand_path) {
      for (; pass < 2; ++pass) {
        launch_select_hist(x, n_local, d_sel.p, pass, sms, st); ++launches;
        if (multi) allreduce_u32(d_sel.p->hist, 4096);
        launch_select_scan(d_sel.p, pass, out, st); ++launches;
      }
      unsigned long long cnt = 0;
      CU(cudaMemcpyAsync(&cnt, &d_sel.p->bucket_count, sizeof(cnt), cudaMemcpyDeviceToHost, st));
      sync();
      if (cnt <= (unsigned long long)kSelCandCap) {
        d_sel_cand.alloc((size_t)kSelCandCap + 1); d_sel_all.alloc((size_t)kSelCandCap * (multi ? world : 1) + 1);
        launch_select_compact(x, n_local, d_sel.p, 2, d_sel_cand.p, sms, st); ++launches;
        const unsigned long long* all = d_sel_cand.p; const unsigned int* all_n = &d_sel.p->cand_n;
        if (multi) { peer.gather_u64(d_sel_cand.p, &d_sel.p->cand_n, d_sel_all.p, &d_sel.p->all_n, st); ++launches; all = d_sel_all.p; all_n = &d_sel.p->all_n; }
        launch_select_final(all, all_n, d_sel.p, out, st); ++launches;
        return;
      }
    }
    for (; pass < select_num_passes(); ++pass) {
      launch_select_hist(x, n_local, d_sel.p, pass, sms, st); ++launches;
      if (multi) allreduce_u32(d_sel.p->hist, 4096);
      launch_select_scan(d_sel.p, pass, out, st); ++launches;
    }
  }

  void allreduce_f64(double* buf, size_t n) {
    if (world <= 1 || n == 0) return;
    if (peer.fits(n * sizeof(double))) { peer.allreduce_f64(buf, (int)n, st); ++launches; return; }
    int rc = g_nccl.allReduce(buf, buf, n, kNcclFloat64, kNcclSum, comm, st);
    if (rc != 0) throw std::string("ncclAllReduce(f64) failed: ") + (g_nccl.getErrorString ? g_nccl.getErrorString(rc) : "?");
  }
  void allreduce_u32(unsigned int* buf, size_t n) {
    if (world <= 1 || n == 0) return;
    if (peer.fits(n * sizeof(unsigned int))) { peer.allreduce_u32(buf, (int)n, st); ++launches; return; }
    int rc = g_nccl.allReduce(buf, buf, n, kNcclUint32, kNcclSum, comm, st);
    if (rc != 0) throw std::string("ncclAllReduce(u32) failed");
  }

  // ---- BuildProblem + Schur assembly ----------------------------------------------------------
  void build() {
    compute_masks();
    CU(cudaEventRecord(ev[0], st));
    d_S.zero(st, (size_t)n_tiles * T * T); d_bp.zero(st); d_sc.zero(st);
    { const double flag = n_obs_total > 0 ? 1.0 : 0.0; CU(cudaMemcpyAsync(&d_sc.p->dot[7], &flag, sizeof(double), cudaMemcpyHostToDevice, st)); }
    launch_pose_frames(P, st); ++launches;
    // pass 1 (residuals -> median input) is skipped when the accepted EvaluateResiduals of the previous iteration
    // already left |r|^2 w0 of this very state in obs_maha0
    if (!maha0_valid) { launch_residual(P, G, st); if (n_obs) ++launches; }
    maha0_valid = false;
    CU(cudaEventRecord(ev[1], st));
    const unsigned long long n_total = (unsigned long long)n_obs_total;   // ranks hold disjoint landmark shards
    run_select(d_obs_maha0.p, (size_t)n_obs, n_total, &d_sc.p->median);
    CU(cudaEventRecord(ev[2], st));
    launch_linearize(P, d_tasks.p, n_tasks, d_long_lms.p, n_long, st); if (n_lms && L) launches += 1 + (n_tasks > 0) + (n_long > 0);
    CU(cudaEventRecord(ev[3], st));
    if (rank == 0) {   // replicated residuals are added once
      if (R.n_unary) {
        launch_unary_maha(P, R, st); ++launches;
        run_select(d_un_maha.p, R.n_unary, R.n_unary, &d_sc.p->unary_median, false);
      }
      launch_other_build(P, R, st, &launches);
    }
    CU(cudaEventRecord(ev[4], st));
    CU(cudaMemcpyAsync(d_rhs_sc.p, d_bp.p, sizeof(double) * (size_t)nT * T, cudaMemcpyDeviceToDevice, st));
    if (n_items) { launch_reduce_blocks(P, d_items.p, n_items, d_pairs, d_pairs + n_pairs, d_pairs + 2 * n_pairs, d_partial.p, d_counters.p + 1, d_counters.p, st); ++launches; }
    if (n_items_sym) { launch_reduce_blocks(P, d_items_sym.p, n_items_sym, d_pairs, d_pairs + n_pairs, d_pairs + 2 * n_pairs, d_partial.p, d_counters.p + 1, d_counters.p, st); ++launches; }
    CU(cudaEventRecord(ev[5], st));
    if (world > 1 && bulk.ok) {
      // the error sums over the small windows double as the barrier "every rank has built its partial system"; then
      // the fused reduce-scatter + all-gather over peer memory, then the barrier "every owner has delivered"
      allreduce_f64(&d_sc.p->proj_error, 8);
      bulk.allreduce(d_red_slots.p, (int)red_slots.size(), T * T, (size_t)n_tiles * T * T, 2 * (size_t)nT * T, sms, st); ++launches;
      allreduce_u32(&d_sel.p->hist[0], 1);
    } else if (world > 1) {
      // one NCCL group: the four reductions share a single launch
      const bool grp = g_nccl.groupStart && g_nccl.groupEnd;
      if (grp) g_nccl.groupStart();
      allreduce_f64(d_S.p, (size_t)n_tiles * T * T);
      allreduce_f64(d_bp.p, (size_t)nT * T);
      allreduce_f64(d_rhs_sc.p, (size_t)nT * T);
      allreduce_f64(&d_sc.p->proj_error, 8);   // proj_error .. inertial_error
      if (grp && g_nccl.groupEnd() != 0) throw std::string("ncclGroupEnd failed");
    }
    launch_mask_diagonal(P, d_pose_mask_used.p, is_param_mask_used ? 1 : 0, st); launches += is_param_mask_used ? 2 : 1;
    CU(cudaEventRecord(ev[6], st));
    if (keep_S_copy) {
      d_S_copy.alloc((size_t)n_tiles * T * T);
      CU(cudaMemcpyAsync(d_S_copy.p, d_S.p, sizeof(double) * (size_t)n_tiles * T * T, cudaMemcpyDeviceToDevice, st));
    }
  }

  void collect_build_times() {
    float ms;
    cudaEventElapsedTime(&ms, ev[0], ev[1]); phase_ms[BAGPU_PH_RESIDUAL] += ms;
    cudaEventElapsedTime(&ms, ev[1], ev[2]); phase_ms[BAGPU_PH_MEDIAN] += ms;
    cudaEventElapsedTime(&ms, ev[2], ev[3]); phase_ms[BAGPU_PH_ACCUMULATE] += ms;
    cudaEventElapsedTime(&ms, ev[3], ev[4]); phase_ms[BAGPU_PH_OTHER_RES] += ms;
    cudaEventElapsedTime(&ms, ev[4], ev[5]); phase_ms[BAGPU_PH_SCHUR] += ms;
    cudaEventElapsedTime(&ms, ev[5], ev[6]); phase_ms[BAGPU_PH_ALLREDUCE] += ms;
  }

  Scalars read_scalars() {
    Scalars s;
    d_sc.download(&s, 1, st);
    sync();
    return s;
  }

  double dot(const double* a, const double* b, size_t n, int slot) {
    (void)slot;
    CU(cudaMemsetAsync(&d_sc.p->dot[0], 0, sizeof(double), st));
    launch_dot(P, a, b, n, &d_sc.p->dot[0], st); if (n) ++launches;
    double v;
    CU(cudaMemcpyAsync(&v, &d_sc.p->dot[0], sizeof(double), cudaMemcpyDeviceToHost, st));
    sync();
    return v;
  }
  // (a1 . b1, a2 . b2) with one read-back; the second product is over landmark-sharded vectors (summed over ranks)
  void dot_pair(const double* a1, const double* b1, size_t n1, const double* a2, const double* b2, size_t n2, double* r1, double* r2,
                bool wait = true) {
    CU(cudaMemsetAsync(&d_sc.p->dot[0], 0, 2 * sizeof(double), st));
    if (n1) { launch_dot(P, a1, b1, n1, &d_sc.p->dot[0], st); ++launches; }
    if (n2) { launch_dot(P, a2, b2, n2, &d_sc.p->dot[1], st); ++launches; }
    if (world > 1) allreduce_f64(&d_sc.p->dot[1], 1);
    if (!wait) return;
    double v[2];
    CU(cudaMemcpyAsync(v, &d_sc.p->dot[0], 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    sync();
    *r1 = v[0]; *r2 = v[1];
  }
  // sum over ranks of a landmark-sharded quantity
  double shard_sum(double v) {
    if (world <= 1) return v;
    CU(cudaMemcpyAsync(&d_sc.p->dot[6], &v, sizeof(double), cudaMemcpyHostToDevice, st));
    allreduce_f64(&d_sc.p->dot[6], 1);
    CU(cudaMemcpyAsync(&v, &d_sc.p->dot[6], sizeof(double), cudaMemcpyDeviceToHost, st));
    sync();
    return v;
  }

  bool outliers_cleared = false;   // k_update_landmarks zeroed lm_outliers for the evaluation that follows
  void backup_state(bool poses = true, bool landmarks = true) {
    if (poses) {
      CU(cudaMemcpyAsync(d_bk_q.p, d_pose_q.p, sizeof(double) * 4 * n_poses, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_bk_t.p, d_pose_t.p, sizeof(double) * 3 * n_poses, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_bk_v.p, d_pose_v.p, sizeof(double) * 3 * n_poses, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_bk_b.p, d_pose_b.p, sizeof(double) * 6 * n_poses, cudaMemcpyDeviceToDevice, st));
    }
    if (n_lms && landmarks) {
      CU(cudaMemcpyAsync(d_bk_lm_x.p, d_lm_x.p, sizeof(double) * 4 * n_lms, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_bk_rel.p, d_lm_reliable.p, n_lms, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_bk_outliers.p, d_lm_outliers.p, sizeof(uint32_t) * n_lms, cudaMemcpyDeviceToDevice, st));   // landmarks_copy carries num_outlier_residuals (cpp:1022,1062)
    }
  }
  void restore_state() {
    CU(cudaMemcpyAsync(d_pose_q.p, d_bk_q.p, sizeof(double) * 4 * n_poses, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(d_pose_t.p, d_bk_t.p, sizeof(double) * 3 * n_poses, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(d_pose_v.p, d_bk_v.p, sizeof(double) * 3 * n_poses, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(d_pose_b.p, d_bk_b.p, sizeof(double) * 6 * n_poses, cudaMemcpyDeviceToDevice, st));
    if (n_lms) {
      CU(cudaMemcpyAsync(d_lm_x.p, d_bk_lm_x.p, sizeof(double) * 4 * n_lms, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_lm_reliable.p, d_bk_rel.p, n_lms, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d_lm_outliers.p, d_bk_outliers.p, sizeof(uint32_t) * n_lms, cudaMemcpyDeviceToDevice, st));
    }
    launch_pose_frames(P, st); ++launches;
  }

  // EvaluateResiduals at the current state (cpp:144-274); returns the four sums
  void evaluate(double* e_proj, double* e_bin, double* e_un, double* e_in, cudaEvent_t mark = nullptr) {
    CU(cudaMemsetAsync(&d_sc.p->eval_proj, 0, 4 * sizeof(double), st));
    launch_evaluate(P, G, &d_sc.p->eval_proj, nullptr, nullptr, 1, outliers_cleared ? 2 : 1, st); if (n_obs) ++launches;
    outliers_cleared = false;
    if (mark) CU(cudaEventRecord(ev[17], st));
    maha0_valid = false;   // overwritten with this state's values; valid for the next build only if the step is kept
    if (rank == 0) launch_other_evaluate(P, R, st, &launches);
    if (world > 1) allreduce_f64(&d_sc.p->eval_proj, 4);
    if (mark) CU(cudaEventRecord(mark, st));   // end of the step phase on the device: read with the same round trip
    last_eval = read_scalars();   // the whole block: the dogleg loop parks other sums in it (dot[4..5])
    *e_proj = last_eval.eval_proj; *e_bin = last_eval.eval_binary; *e_un = last_eval.eval_unary; *e_in = last_eval.eval_inertial;
  }
  Scalars last_eval;

  // development aid: BAGPU_CHOL_TRACE=<file> dumps the task table and 4 globaltimer stamps per task of the last solve
  DevBuf<unsigned long long> d_chol_trace;
  std::vector<CholTask> h_chol_tasks;
  unsigned long long* chol_trace_buf() {
    if (!getenv("BAGPU_CHOL_TRACE")) return nullptr;
    d_chol_trace.alloc((size_t)n_chol_tasks * 4); d_chol_trace.zero(st);
    return d_chol_trace.p;
  }
  void chol_trace_dump() {
    const char* path = getenv("BAGPU_CHOL_TRACE");
    if (!path || !d_chol_trace.p) return;
    std::vector<unsigned long long> tr((size_t)n_chol_tasks * 4);
    d_chol_trace.download(tr.data(), tr.size(), st); sync();
    FILE* f = fopen(path, "wb");
    if (!f) return;
    fwrite(&n_chol_tasks, sizeof(int), 1, f);
    fwrite(h_chol_tasks.data(), sizeof(CholTask), n_chol_tasks, f);
    fwrite(tr.data(), sizeof(unsigned long long), tr.size(), f);
    fclose(f);
  }

  // CalculateGn + GetLandmarkDelta -> d_gn_p, d_gn_l
  bool gauss_newton() {
    CU(cudaEventRecord(ev[8], st));
    // options.reduced_solver: 0 = tile LDL^T, 1 = block-Jacobi PCG, 2 = PCG for small systems (at most 64 poses), LDL^T otherwise
    const bool use_pcg = n_ap > 0 && (opt.reduced_solver == 1 || (opt.reduced_solver == 2 && n_ap <= 64)) && nT <= 9000;
    if (use_pcg) {
      if (!pcg_uploaded) { d_pcg_row_ptr.upload(pcg_row_ptr, st); d_pcg_ent.upload(pcg_ent, st); pcg_uploaded = true; }
      d_pcg_work.alloc(pcg_work_doubles(nT, T, D));
      launch_permute_tiles(d_rhs_work.p, d_rhs_sc.p, d_tile_perm.p, T, nT, 1, st); ++launches;
      const int rc = launch_pcg_solve(D, d_S.p, d_tile_map.p, d_pcg_row_ptr.p, d_pcg_ent.p, nT, d_rhs_work.p, d_rhs_work.p, d_pcg_work.p,
                                      d_red_partial.p, d_red_counter.p, opt.pcg_tolerance, (int)opt.pcg_max_iterations, &pcg_iterations, &launches, st);
      if (rc < 0) throw std::string("pcg: CUDA error");
      launch_permute_tiles(d_gn_p.p, d_rhs_work.p, d_tile_perm.p, T, nT, 0, st); ++launches;
      // not positive definite -> the factorisation of the reference would have failed; no convergence -> its solve step
      if (rc == 2) { summary.result = BAGPU_FactorizationError; return false; }
      if (rc == 1) { summary.result = BAGPU_SolverError; return false; }
    } else if (n_ap > 0) {
      launch_permute_tiles(d_rhs_work.p, d_rhs_sc.p, d_tile_perm.p, T, nT, 1, st); ++launches;
      const int rc = launch_cholesky_solve(T, d_S.p, d_uinv.p, d_scratch.p, d_rhs_work.p, d_chol_tasks.p, d_chol_terms.p, d_row_ptr.p, d_row_col.p, n_chol_tasks,
                                           d_chol_sync.p, n_tiles, nT, sms, st, chol_trace_buf());
      if (rc != 0) throw std::string("cholesky launch: ") + cudaGetErrorString((cudaError_t)rc);
      ++launches;
      launch_permute_tiles(d_gn_p.p, d_rhs_work.p, d_tile_perm.p, T, nT, 0, st); ++launches;
      chol_flag_pending = true;
    }
    summary.result = BAGPU_Success;
    CU(cudaEventRecord(ev[18], st));
    if (n_al > 0) { launch_back_substitute(P, d_tasks.p, n_tasks, d_long_lms.p, n_long, d_gn_p.p, d_gn_l.p, st); launches += 1 + (n_tasks > 0) + (n_long > 0); }
    CU(cudaEventRecord(ev[9], st));
    // ONE host round trip for the whole Gauss-Newton step: the factorisation's failure flag and the two norms of the step come
    // back together (on a failed factorisation the back-substitution and the norms ran on garbage and are thrown away)
    const size_t np = (size_t)n_ap * D, nl = (size_t)n_al * std::max(1, L);
    dot_pair(d_gn_p.p, d_gn_p.p, np, d_gn_l.p, d_gn_l.p, nl, nullptr, nullptr, false);
    int f = 0;
    double v[2] = {0.0, 0.0};
    if (chol_flag_pending) CU(cudaMemcpyAsync(&f, d_chol_sync.p + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(v, &d_sc.p->dot[0], 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    sync();
    chol_flag_pending = false;
    chol_trace_dump();
    gn_norm2_p = v[0]; gn_norm2_l = v[1];
    float ms; cudaEventElapsedTime(&ms, ev[8], ev[9]); phase_ms[BAGPU_PH_SOLVE] += ms;
    cudaEventElapsedTime(&ms, ev[8], ev[18]); dev_factor_ms += ms;
    if (f != 0) { summary.result = BAGPU_FactorizationError; return false; }
    return true;
  }
  bool chol_flag_pending = false;
  double gn_norm2_p = 0, gn_norm2_l = 0;   // |delta_p|^2, |delta_l|^2 of the last Gauss-Newton step

  // the steepest-descent sums |b_p|^2, |b_l|^2 and |J_p b_p + J_l b_l|^2 (cpp:858-925), enqueued behind the build
  void enqueue_sd() {
    const size_t np = (size_t)n_ap * D, nl = (size_t)n_al * std::max(1, L);
    CU(cudaEventRecord(ev[12], st));
    if (L > 0 && n_al > 0) { launch_compact_bl(P, d_bl_compact.p, st); ++launches; }
    dot_pair(d_bp.p, d_bp.p, np, d_bl_compact.p, d_bl_compact.p, (L > 0) ? nl : 0, nullptr, nullptr, false);
    CU(cudaMemsetAsync(&d_sc.p->sd_den_proj, 0, 2 * sizeof(double), st));
    if (n_ap > 0 || n_al > 0) { launch_sd_denominator(P, G, st); if (n_obs) launches += 2; }
    if (n_ap > 0 && rank == 0) launch_other_sd(P, R, st, &launches);
    if (world > 1) allreduce_f64(&d_sc.p->sd_den_proj, 2);
    CU(cudaEventRecord(ev[13], st));
  }

  // SolveInternal (cpp:838-1161)
  bool solve_internal(double gn_damping, bool error_increase_allowed, bagpu_iter_trace& tr, const Scalars& sb) {
    const size_t np = (size_t)n_ap * D, nl = (size_t)n_al * std::max(1, L);
    const double pre_proj = sb.proj_error, pre_un = sb.unary_error, pre_in = sb.inertial_error;
    bool gn_computed = false;
    double e_proj, e_bin, e_un, e_in;
    if (opt.use_dogleg) {
      // steepest descent step (cpp:857-929): its kernels were enqueued right behind the build (enqueue_sd) and their sums came
      // back with the build scalars in one round trip
      const double n2_bp = sb.dot[0], n2_bl = sb.dot[1];
      const double den[2] = {sb.sd_den_proj, sb.sd_den_other};
      const double numerator = n2_bp + n2_bl;
      const double denominator = den[0] + den[1];
      const double factor = numerator / denominator;
      const double delta_sd_norm = std::sqrt(factor * factor * n2_bp + factor * factor * n2_bl);
      tr.sd_numerator = numerator; tr.sd_denominator = denominator; tr.sd_norm = delta_sd_norm;
      uint32_t iteration_count = 0;
      while (true) {
        iteration_count++;
        if (iteration_count > opt.dogleg_max_inner_iterations) break;
        tr.inner_iterations = iteration_count;
        double n2_dl_p = 0, n2_dl_l = 0;
        bool norms_pending = false;
        if (delta_sd_norm > trust_region && trust_region != -1.0) {
          const double f = trust_region / delta_sd_norm * factor;
          launch_axpby(d_dl_p.p, f, d_bp.p, 0, nullptr, np, st); if (np) ++launches;
          if (nl) { launch_axpby(d_dl_l.p, f, d_bl_compact.p, 0, nullptr, nl, st); ++launches; }
          n2_dl_p = f * f * n2_bp; n2_dl_l = f * f * n2_bl;
        } else {
          if (!gn_computed) {
            if (!gauss_newton()) return false;
            gn_computed = true;
          }
          const double delta_gn_norm = std::sqrt(gn_norm2_p + gn_norm2_l);
          tr.gn_norm = delta_gn_norm;
          const bool good = !std::isnan(delta_gn_norm) && !std::isinf(delta_gn_norm);
          if (good && trust_region == -1.0) trust_region = delta_gn_norm;
          if (good && delta_gn_norm <= trust_region) {
            CU(cudaMemcpyAsync(d_dl_p.p, d_gn_p.p, sizeof(double) * np, cudaMemcpyDeviceToDevice, st));
            if (nl) CU(cudaMemcpyAsync(d_dl_l.p, d_gn_l.p, sizeof(double) * nl, cudaMemcpyDeviceToDevice, st));
            n2_dl_p = gn_norm2_p; n2_dl_l = gn_norm2_l;
          } else {
            // diff = gn - sd, formed explicitly like the reference (cpp:991-1017); its two dot products in the same pass
            launch_blend_diff(P, d_tmp_p.p, d_gn_p.p, d_bp.p, np, d_tmp_l.p, d_gn_l.p, d_bl_compact.p, nl, factor, &d_sc.p->dot[0], st); ++launches;
            if (world > 1) allreduce_f64(&d_sc.p->dot[2], 2);   // the landmark parts are sums over the shards
            double v4[4];
            CU(cudaMemcpyAsync(v4, &d_sc.p->dot[0], 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
            sync();
            const double ap = v4[0], bpp = v4[1], al = v4[2], bll = v4[3];
            const double a = ap + al;
            const double b = 2 * (factor * bpp + factor * bll);
            const double c = (factor * factor * n2_bp + factor * factor * n2_bl) - trust_region * trust_region;
            double beta = 0;
            if (b * b > 4 * a * c && a > 1e-10) beta = (-(b * b) + std::sqrt(b * b - 4 * a * c)) / (2 * a);   // quirk Q4
            // the step and its two norms; they come back with the scalars of the evaluation below
            launch_blend_step(P, d_dl_p.p, d_tmp_p.p, d_bp.p, np, d_dl_l.p, d_tmp_l.p, d_bl_compact.p, nl, beta, factor, &d_sc.p->dot[4], st); ++launches;
            if (world > 1) allreduce_f64(&d_sc.p->dot[5], 1);
            norms_pending = true;
          }
        }
        CU(cudaEventRecord(ev[10], st));
        // with apply_results the copy of the state is taken by the update kernels themselves
        const bool lm_fused = n_lms > 0 && L > 0;
        if (!opt.apply_results) backup_state();
        // "pre" evaluation: same state and weights as the build, so the build sums are reused;
        // only the binary error uses a different formula in EvaluateResiduals (quirk Q7).
        double pre_bin = 0;
        if (R.n_binary) {   // on every rank (binary residuals are replicated): all ranks must take the same accept / reject branch
          CU(cudaMemsetAsync(&d_sc.p->eval_binary, 0, sizeof(double), st));
          OtherResiduals Rb = R; Rb.n_unary = 0; Rb.n_imu = 0;
          launch_other_evaluate(P, Rb, st, &launches);
          CU(cudaMemcpyAsync(&pre_bin, &d_sc.p->eval_binary, sizeof(double), cudaMemcpyDeviceToHost, st));
          sync();
        }
        summary.pre_solve_norm = pre_proj + pre_in + pre_bin + pre_un;
        if (opt.apply_results) {
          const StateBackup bk = {d_bk_q.p, d_bk_t.p, d_bk_v.p, d_bk_b.p, d_bk_lm_x.p, d_bk_rel.p, d_bk_outliers.p};
          launch_update_poses(P, d_dl_p.p, 1.0, st, &bk); ++launches;
          if (lm_fused) { launch_update_landmarks(P, d_dl_l.p, 1.0, st, &bk); ++launches; outliers_cleared = true; }
          else backup_state(false, true);
          launch_pose_frames(P, st); ++launches;
        }
        CU(cudaEventRecord(ev[16], st));
        evaluate(&e_proj, &e_bin, &e_un, &e_in, ev[11]);
        { float ms; cudaEventElapsedTime(&ms, ev[10], ev[16]); dev_update_ms += ms; cudaEventElapsedTime(&ms, ev[16], ev[17]); dev_eval_ms += ms; }
        if (norms_pending) { n2_dl_p = last_eval.dot[4]; n2_dl_l = last_eval.dot[5]; }
        if (opt.apply_results) summary.delta_norm = std::sqrt(n2_dl_l) + std::sqrt(n2_dl_p);
        summary.post_solve_norm = e_proj + e_in + e_bin + e_un;
        tr.delta_p_norm = std::sqrt(n2_dl_p); tr.delta_l_norm = std::sqrt(n2_dl_l);
        { float ms; cudaEventElapsedTime(&ms, ev[10], ev[11]); phase_ms[BAGPU_PH_STEP] += ms; }
        if (summary.post_solve_norm > summary.pre_solve_norm) {
          if (opt.apply_results) restore_state();
          trust_region /= 2;
        } else {
          proj_error = e_proj; unary_error = e_un; binary_error = e_bin; inertial_error = e_in;
          trust_region *= 2;
          maha0_valid = opt.apply_results != 0;
          break;
        }
      }
    } else {
      // damped Gauss-Newton (cpp:1084-1159)
      if (L > 0 && n_al > 0) { launch_compact_bl(P, d_bl_compact.p, st); ++launches; }
      if (!gauss_newton()) return false;
      backup_state();
      launch_axpby(d_dl_p.p, gn_damping, d_gn_p.p, 0, nullptr, np, st); if (np) ++launches;
      if (nl) { launch_axpby(d_dl_l.p, gn_damping, d_gn_l.p, 0, nullptr, nl, st); ++launches; }
      double n2p = 0, n2l = 0;
      dot_pair(d_dl_p.p, d_dl_p.p, np, d_dl_l.p, d_dl_l.p, nl, &n2p, &n2l);
      tr.gn_norm = std::sqrt(n2p + n2l); tr.delta_p_norm = std::sqrt(n2p); tr.delta_l_norm = std::sqrt(n2l);
      evaluate(&e_proj, &e_bin, &e_un, &e_in);
      const double prev_error = e_proj + e_in + e_bin + e_un;
      if (opt.apply_results) {
        summary.delta_norm = std::sqrt(n2l) + std::sqrt(n2p);
        launch_update_poses(P, d_dl_p.p, 1.0, st); ++launches;
        launch_update_landmarks(P, d_dl_l.p, 1.0, st); if (n_lms && L) ++launches;
        launch_pose_frames(P, st); ++launches;
      }
      evaluate(&e_proj, &e_bin, &e_un, &e_in);
      const double post_error = e_proj + e_in + e_bin + e_un;
      tr.inner_iterations = 1; tr.pre_solve_norm = prev_error; tr.post_solve_norm = post_error;
      if (post_error > prev_error && !error_increase_allowed) {
        if (opt.apply_results) restore_state();
        summary.result = BAGPU_ErrorIncreased;
        return false;
      }
      proj_error = e_proj; unary_error = e_un; binary_error = e_bin; inertial_error = e_in;
      maha0_valid = opt.apply_results != 0;
    }
    return true;
  }

  void prepare_solve() {
    PhaseClock pc;
    ensure_device(); pc.lap("ensure_device");
    maha0_valid = false;
    if (dirty) { setup(); pc.lap("setup (host symbolic)"); upload_all(); sync(); pc.lap("upload_all"); solved_once = false; }
    else apply_options();
    // Solve() prologue (cpp:287-296)
    launch_pose_frames(P, st); ++launches;
    if (L == 1) { launch_prepare_xs(P, d_lm_xw.p, st); ++launches; }
    else if (n_lms) CU(cudaMemcpyAsync(d_lm_x.p, d_lm_xw.p, sizeof(double) * 4 * n_lms, cudaMemcpyDeviceToDevice, st));
  }

  void download_state() {
    // poses, landmarks back into the host vectors so a second Solve() continues from here
    std::vector<double> q(4 * n_poses), t(3 * n_poses);
    d_pose_q.download(q.data(), q.size(), st); d_pose_t.download(t.data(), t.size(), st);
    d_pose_v.download(pose_v.data(), pose_v.size(), st); d_pose_b.download(pose_b.data(), pose_b.size(), st);
    if (n_lms) {
      if (world > 1) {   // every landmark is owned by exactly one rank: zero the others and sum
        if (lm_owner_lo > 0) CU(cudaMemsetAsync(d_lm_x.p, 0, sizeof(double) * 4 * (size_t)lm_owner_lo, st));
        if (lm_owner_hi < n_lms) CU(cudaMemsetAsync(d_lm_x.p + 4 * (size_t)lm_owner_hi, 0, sizeof(double) * 4 * (size_t)(n_lms - lm_owner_hi), st));
        allreduce_f64(d_lm_x.p, 4 * (size_t)n_lms);
        // reliable flags (set by the owner's ApplyUpdate, cpp:127-134) and outlier counts (counted over the owner's
        // observations, cpp:184-186): AND / sum over ranks, so every rank reports what a single GPU would
        DevBuf<unsigned int> d_unrel; d_unrel.alloc((size_t)n_lms);
        launch_reliable_merge(d_lm_reliable.p, d_unrel.p, (uint32_t)n_lms, 0, st); ++launches;
        allreduce_u32(d_unrel.p, (size_t)n_lms);
        launch_reliable_merge(d_lm_reliable.p, d_unrel.p, (uint32_t)n_lms, 1, st); ++launches;
        allreduce_u32(d_lm_outliers.p, (size_t)n_lms);
        sync();   // d_unrel goes out of scope
      }
      if (L == 1) { launch_finish_xw(P, d_lm_xw.p, st); ++launches; }
      else CU(cudaMemcpyAsync(d_lm_xw.p, d_lm_x.p, sizeof(double) * 4 * n_lms, cudaMemcpyDeviceToDevice, st));
      lm_host_fresh = false;   // bagpu_get_landmarks reads them straight into the caller's buffers
    }
    sync();
    for (int i = 0; i < n_poses; ++i) { for (int k = 0; k < 4; ++k) pose_t7[7 * i + k] = q[4 * i + k]; for (int k = 0; k < 3; ++k) pose_t7[7 * i + 4 + k] = t[3 * i + k]; }
  }

  // the host copies of the landmark state, needed only when the problem is set up again from the host vectors
  void fetch_landmarks_host() {
    if (lm_host_fresh || !n_lms || !d_lm_xw.p) { lm_host_fresh = true; return; }
    lm_xw.resize(4 * n_lm_set);
    d_lm_xw.download(lm_xw.data(), lm_xw.size(), st);
    lm_xs.resize(lm_xw.size());
    d_lm_x.download(lm_xs.data(), lm_xs.size(), st);
    lm_reliable.resize(n_lms);
    d_lm_reliable.download(lm_reliable.data(), lm_reliable.size(), st);
    lm_outliers.resize(n_lms);
    d_lm_outliers.download(lm_outliers.data(), n_lms, st);
    sync();
    lm_host_fresh = true;
  }
  void copy_landmarks_out(double* xw, double* xs, uint8_t* rel, uint32_t* nout) {
    if (xw) d_lm_xw.download(xw, 4 * (size_t)n_lms, st);
    if (xs) d_lm_x.download(xs, 4 * (size_t)n_lms, st);
    if (rel) d_lm_reliable.download(rel, n_lms, st);
    if (nout) d_lm_outliers.download(nout, n_lms, st);
    sync();
  }

  // Solve (cpp:278-705)
  void solve(uint32_t max_iter, double gn_damping, bool error_increase_allowed) {
    trace.clear(); rec_dp.clear(); rec_dl.clear();
    for (float& m : phase_ms) m = 0;
    dev_sd_ms = 0; dev_update_ms = 0; dev_eval_ms = 0; dev_factor_ms = 0;
    launches = 0;
    summary.iterations = 0;
    if (n_raw == 0 && bin_p1.empty() && un_pose.empty() && imu_p1.empty()) return;
    PhaseClock pcs;
    prepare_solve();
    CU(cudaEventRecord(ev[14], st));
    for (uint32_t kk = 0; kk < max_iter; ++kk) {
      build();
      if (opt.use_dogleg) enqueue_sd();
      const Scalars sb = read_scalars();   // one round trip: build scalars + steepest-descent sums
      collect_build_times();
      if (opt.use_dogleg) { float ms; cudaEventElapsedTime(&ms, ev[12], ev[13]); dev_sd_ms += ms; }
      // options.write_reduced_camera_matrix (cpp:600-606): s.txt / rhs.txt of this iteration's system into the working
      // directory, overwritten every iteration like the reference's std::ios_base::trunc streams
      if (opt.write_reduced_camera_matrix && bagpu_dump(this, ".") != BAGPU_OK) throw std::string("write_reduced_camera_matrix: cannot write s.txt / rhs.txt");
      bagpu_iter_trace tr; memset(&tr, 0, sizeof(tr));
      tr.proj_error_build = sb.proj_error; tr.binary_error_build = sb.binary_error;
      tr.unary_error_build = sb.unary_error; tr.inertial_error_build = sb.inertial_error; tr.median_maha = sb.median;
      proj_error = sb.proj_error; binary_error = sb.binary_error; unary_error = sb.unary_error; inertial_error = sb.inertial_error;
      const bool ok = solve_internal(gn_damping, error_increase_allowed, tr, sb);
      summary.iterations = kk + 1;
      if (opt.use_dogleg) { tr.pre_solve_norm = summary.pre_solve_norm; tr.post_solve_norm = summary.post_solve_norm; }
      tr.delta_norm = summary.delta_norm; tr.trust_region_after = trust_region;
      if (opt.record_deltas) {
        std::vector<double> dp((size_t)n_ap * D), dl((size_t)n_al * std::max(1, L));
        d_dl_p.download(dp.data(), dp.size(), st); d_dl_l.download(dl.data(), dl.size(), st); sync();
        rec_dp.push_back(dp); rec_dl.push_back(dl);
      }
      if (!ok) { tr.result = summary.result; trace.push_back(tr); break; }
      if ((std::fabs(summary.post_solve_norm - summary.pre_solve_norm) / summary.pre_solve_norm) < opt.error_change_threshold) {
        summary.result = BAGPU_ErrorChangeBelowThreshold; tr.result = summary.result; trace.push_back(tr); break;
      }
      if (summary.delta_norm < opt.param_change_threshold) {
        summary.result = BAGPU_ParamChangeBelowThreshold; tr.result = summary.result; trace.push_back(tr); break;
      }
      tr.result = summary.result; trace.push_back(tr);
    }
    CU(cudaEventRecord(ev[15], st));
    sync();
    { float ms; cudaEventElapsedTime(&ms, ev[14], ev[15]); phase_ms[BAGPU_PH_TOTAL] = ms; }
    pcs.lap("solve: prepare + iterations");
    if (pcs.on) {
      float sum = 0; for (int i = 0; i < BAGPU_PH_TOTAL; ++i) sum += phase_ms[i];
      fprintf(stderr, "[bagpu timing]   device: total %.3f ms, phases %.3f, steepest-descent section %.3f, rest (host round trips, vector kernels) %.3f over %u iterations\n",
              phase_ms[BAGPU_PH_TOTAL], sum, dev_sd_ms, phase_ms[BAGPU_PH_TOTAL] - sum - dev_sd_ms, summary.iterations);
      fprintf(stderr, "[bagpu timing]   device: step = backup + update %.3f ms, evaluation kernel %.3f ms; solve = factorisation %.3f ms + back-substitution\n",
              dev_update_ms, dev_eval_ms, dev_factor_ms);
    }
    if (!opt.keep_on_device) download_state();
    pcs.lap("download_state");
    solved_once = true;
    // summary (cpp:680-704)
    summary.num_proj_residuals = (uint32_t)n_obs_total; summary.num_inertial_residuals = (uint32_t)imu_p1.size();
    summary.proj_error = proj_error; summary.inertial_error = inertial_error; summary.unary_error = unary_error; summary.binary_error = binary_error;
    summary.trust_region_size = trust_region;
    // conditioning residuals (cpp:682-703): counted at set-up, their maha / weight summed on the device
    summary.num_cond_proj_residuals = n_cond_proj;
    uint32_t nci = 0;
    for (size_t i = 0; i < imu_p1.size(); ++i) if (!pose_active[imu_p1[i]] && pose_active[imu_p2[i]]) ++nci;
    summary.num_cond_inertial_residuals = nci;
    summary.cond_proj_error = 0; summary.cond_inertial_error = 0;
    if (nci) {   // cpp:688-691: sum of res.mahalanobis_distance (as left by the last EvaluateResiduals) over the conditioning residuals
      std::vector<double> io = other_outputs(d_imu_out, imu_p1.size() * (D + 1));
      for (size_t i = 0; i < imu_p1.size(); ++i)
        if (!pose_active[imu_p1[i]] && pose_active[imu_p2[i]]) summary.cond_inertial_error += io[i * (D + 1) + D];
    }
    if (n_cond_proj) {
      evaluate_by_id();
      CU(cudaMemsetAsync(&d_sc.p->dot[4], 0, sizeof(double), st));
      launch_cond_error(d_obs_pose.p, d_sorted_lm.p + s_lo, P.obs_perm, (uint32_t)n_obs, d_lm_ref_pose.p, d_pose_opt.p, d_maha_out.p, d_obs_w.p,
                        &d_sc.p->dot[4], st); ++launches;
      double v = 0;
      CU(cudaMemcpyAsync(&v, &d_sc.p->dot[4], sizeof(double), cudaMemcpyDeviceToHost, st));
      sync();
      summary.cond_proj_error = shard_sum(v);
    }
    pcs.lap("summary");
  }

  // Per-residual outputs of the replicated residual types live on rank 0 (the rank that evaluates them); the other ranks
  // get them through one sum with their own copy zeroed.
  std::vector<double> other_outputs(DevBuf<double>& buf, size_t n) {
    std::vector<double> out(n, 0.0);
    if (n == 0) return out;
    if (world > 1) {
      if (rank != 0) buf.zero(st, n);
      allreduce_f64(buf.p, n);
    }
    buf.download(out.data(), n, st);
    sync();
    return out;
  }

  // residual and maha of every local observation, written at its residual id (d_res_out, d_maha_out)
  void evaluate_by_id() {
    const size_t nt = (size_t)n_obs_total;
    d_res_out.alloc(2 * nt); d_maha_out.alloc(nt);
    if (world > 1) { d_res_out.zero(st); d_maha_out.zero(st); }
    CU(cudaMemsetAsync(&d_sc.p->dot[5], 0, sizeof(double), st));
    launch_evaluate(P, G, &d_sc.p->dot[5], d_res_out.p, d_maha_out.p, 0, 0, st); ++launches;
  }

  void export_residuals(double* res2, double* weight, double* maha) {
    if (n_obs_total == 0) return;
    const size_t nt = (size_t)n_obs_total;
    evaluate_by_id();
    DevBuf<double> d_w_id;
    if (weight) {
      d_w_id.alloc(nt);
      if (world > 1) d_w_id.zero(st);
      launch_scatter_by_perm(P.obs_perm, d_obs_w.p, (uint32_t)n_obs, d_w_id.p, st); ++launches;
      d_w_id.download(weight, nt, st);
    }
    if (res2) d_res_out.download(res2, 2 * nt, st);
    if (maha) d_maha_out.download(maha, nt, st);
    sync();
  }
};

// ============================================================================================
// for the entry points that live in other translation units (select_static.cu)
namespace bagpu { void set_last_error(const std::string& msg) { g_err = msg; } }

extern "C" {

#define GUARD(...)                                                                \
  try { __VA_ARGS__ } catch (const std::string& e) {                              \
    const bool nodev = e.find("no CUDA device") != std::string::npos;             \
    return fail(nodev ? BAGPU_ERR_NO_DEVICE : BAGPU_ERR_CUDA, e);                 \
  } catch (const std::exception& e) {   /* bad_alloc, system_error: never across the C boundary */ \
    return fail(BAGPU_ERR_CUDA, std::string("host exception: ") + e.what());      \
  }

const char* bagpu_last_error(void) { return g_err.c_str(); }

// Page-locked host buffers for the arrays that cross the ABI (copies from / to them are direct DMA).  Without a
// device the buffer is ordinary memory, so host code can use the same allocator everywhere.
int bagpu_host_alloc(size_t bytes, void** out) {
  if (!out) return fail(BAGPU_ERR_INVALID, "null out pointer");
  *out = nullptr;
  if (bytes == 0) bytes = 1;
  int n = 0;
  void* p = nullptr;
  if (cudaGetDeviceCount(&n) == cudaSuccess && n > 0 && cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess) {
    std::lock_guard<std::mutex> g(g_pinned_mu);
    g_pinned.insert(p);
  } else {
    cudaGetLastError();
    p = std::malloc(bytes);
    if (!p) return fail(BAGPU_ERR_INVALID, "out of host memory");
  }
  *out = p;
  return BAGPU_OK;
}
int bagpu_host_free(void* p) {
  if (!p) return BAGPU_OK;
  bool pinned = false;
  { std::lock_guard<std::mutex> g(g_pinned_mu); pinned = g_pinned.erase(p) > 0; }
  if (pinned) cudaFreeHost(p); else std::free(p);
  return BAGPU_OK;
}

void bagpu_default_options(bagpu_options* o) {
  memset(o, 0, sizeof(*o));
  o->trust_region_size = -1;            // reference BundleAdjuster.h:75 (kTrustRegionAuto)
  o->gyro_sigma = 5.3088444e-5; o->accel_sigma = 0.001883649;           // Types.h:33,35
  o->gyro_bias_sigma = 1.4125375e-4; o->accel_bias_sigma = 1.2589254e-2;  // Types.h:34,36
  o->projection_outlier_threshold = 1.0; o->error_change_threshold = 0.01; o->param_change_threshold = 1e-3;
  o->dogleg_max_inner_iterations = 100; o->apply_results = 1; o->use_dogleg = 1; o->use_triangular_matrices = 1;
  o->use_sparse_solver = 1; o->regularize_biases_in_batch = 1; o->enable_auto_regularization = 1;
  o->use_robust_norm_for_proj_residuals = 1;
  o->pcg_tolerance = 1e-12; o->pcg_max_iterations = 1000;
}

int bagpu_create(const bagpu_dims* d, bagpu_handle* out) {
  if (!d || !out) return fail(BAGPU_ERR_INVALID, "null argument");
  // scalar_bytes = sizeof(REAL_TYPE) of the caller (reference CMakeLists.txt:16-20).  4 = the float instantiation: its host
  // class converts at the boundary (every array of this ABI is double), the device computes in FP64 either way -- the
  // build is latency / L2 bound with the FP64 pipe under 30 % busy (profiles/), so FP32 arithmetic would buy nothing and
  // the float caller gets results well inside its 1e-4 bar
  if (d->scalar_bytes != 8 && d->scalar_bytes != 4) return fail(BAGPU_ERR_UNSUPPORTED, "REAL_TYPE must be double (8) or float (4)");
  if (d->calib_dim != 0 || d->do_tvs != 0) return fail(BAGPU_ERR_UNSUPPORTED, "CalibSize>0 / DoTvs not built (SURVEY.md §8f rank 3)");
  if (!(d->lm_dim == 0 || d->lm_dim == 1 || d->lm_dim == 3)) return fail(BAGPU_ERR_UNSUPPORTED, "LmSize must be 0, 1 or 3");
  if (!(d->pose_dim == 6 || d->pose_dim == 9 || d->pose_dim == 15)) return fail(BAGPU_ERR_UNSUPPORTED, "PoseSize must be 6, 9 or 15");
  bagpu_handle h = new bagpu_handle_s();
  h->dims = *d; h->L = d->lm_dim; h->D = d->pose_dim;
  bagpu_default_options(&h->opt);
  h->trust_region = h->opt.trust_region_size;
  memset(&h->summary, 0, sizeof(h->summary));
  *out = h;
  return BAGPU_OK;
}

void bagpu_destroy(bagpu_handle h) {
  if (!h) return;
  if (h->st) cudaStreamSynchronize(h->st);
  bool comm_stuck = false;
  if (h->comm && h->peer.w.world > 1 && h->peer.w.win[h->peer.w.rank]) {
    // CUDA IPC: a window may be freed only after every importer has unmapped it.  Unmap the peers' windows, then meet
    // the other ranks in one tiny NCCL all-reduce before freeing.  The wait is bounded: if the others never come (a
    // rank died) the window and the communicator are abandoned instead of hanging the caller.
    h->peer.close_peers();
    h->d_S.release(); h->d_bp.release(); h->d_rhs_sc.release();   // views into the bulk window
    h->bulk.close_peers();
    try {
      DevBuf<unsigned int> one; one.alloc(1); one.zero(h->st);
      h->allreduce_u32(one.p, 1);
      const auto t0 = std::chrono::steady_clock::now();
      while (cudaStreamQuery(h->st) == cudaErrorNotReady) {
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 10.0) { comm_stuck = true; break; }
        std::this_thread::sleep_for(std::chrono::microseconds(50));
      }
      if (comm_stuck) one.p = nullptr;   // still in use by the pending collective
    } catch (const std::string&) {}
  }
  h->join_packer();
  if (comm_stuck) return;   // abandon the handle (freeing its buffers would wait for the pending collective): the job is failing anyway
  h->peer.destroy();
  h->bulk.destroy();
  if (h->comm && g_nccl.commDestroy) g_nccl.commDestroy(h->comm);
  if (h->ev_ok) for (auto& e : h->ev) cudaEventDestroy(e);
  if (h->d_pairs) dev_free(h->d_pairs);
  if (h->st) cudaStreamDestroy(h->st);
  delete h;
}

int bagpu_set_options(bagpu_handle h, const bagpu_options* o) {
  if (!h || !o) return fail(BAGPU_ERR_INVALID, "null argument");
  // options this path cannot honour are refused, not ignored: per-pose camera parameters (parallel_algos.h:54-57,138-140:
  // the ABI carries one parameter set per camera), calibration marginals (cpp:771-784: CalibSize > 0 only)
  if (o->use_per_pose_cam_params) return fail(BAGPU_ERR_UNSUPPORTED, "use_per_pose_cam_params is not supported (one parameter set per camera)");
  if (o->calculate_calibration_marginals) return fail(BAGPU_ERR_UNSUPPORTED, "calculate_calibration_marginals needs CalibSize > 0, which this build does not have");
  if (o->reduced_solver < 0 || o->reduced_solver > 2) return fail(BAGPU_ERR_INVALID, "reduced_solver must be 0 (tile LDL^T), 1 (PCG) or 2 (PCG below 65 poses)");
  h->opt = *o;
  // Init() sets trust_region_size_ = options_.trust_region_size (BundleAdjuster.h:193); later changes through
  // options() (BundleAdjuster.h:605) leave the running trust region alone
  if (!h->solved_once) h->trust_region = o->trust_region_size;
  return BAGPU_OK;
}

int bagpu_set_cameras(bagpu_handle h, uint32_t n, const double* p, const double* t) {
  h->cam_params.assign(p, p + 4 * (size_t)n); h->cam_tvs.assign(t, t + 7 * (size_t)n); h->dirty = true; return BAGPU_OK;
}
int bagpu_set_poses(bagpu_handle h, uint32_t n, const double* t, const double* v, const double* b, const uint8_t* a,
                    const uint8_t* mu, const uint8_t* m) {
  ApiClock api_clock("bagpu_set_poses");
  const int D = h->D;
  h->pose_t7.assign(t, t + 7 * (size_t)n);
  if (v) h->pose_v.assign(v, v + 3 * (size_t)n); else h->pose_v.assign(3 * (size_t)n, 0.0);
  if (b) h->pose_b.assign(b, b + 6 * (size_t)n); else h->pose_b.assign(6 * (size_t)n, 0.0);
  if (a) h->pose_active.assign(a, a + n); else h->pose_active.assign(n, 1);
  if (mu) h->pose_mask_used.assign(mu, mu + n); else h->pose_mask_used.assign(n, 0);
  if (m) h->pose_mask.assign(m, m + (size_t)n * D); else h->pose_mask.assign((size_t)n * D, 1);
  h->dirty = true;
  return BAGPU_OK;
}
int bagpu_set_landmarks(bagpu_handle h, uint32_t n, const double* x, const uint32_t* rp, const uint32_t* rc, const uint8_t* a) {
  ApiClock api_clock("bagpu_set_landmarks");
  const size_t np = h->pose_t7.size() / 7;
  for (uint32_t i = 0; i < n; ++i) if (rp[i] >= np) return fail(BAGPU_ERR_INVALID, "landmark ref_pose_id out of range");
  // with a usable device the positions go straight into HBM (the host copy is made only if somebody asks for it)
  bool dev = true;
  try { h->ensure_device(); } catch (const std::string&) { dev = false; }
  h->n_lm_set = n; h->lm_xs.clear();   // x_s is filled by the first Solve()
  h->lm_rel_on_device = false; h->lm_ids_on_device = false;
  if (dev && n > 0) {
    GUARD(h->d_lm_xw.upload(x, 4 * (size_t)n, h->st);)   // DMA in flight while the id arrays are copied below
    h->lm_xw.clear(); h->lm_host_fresh = false;
  } else {
    h->lm_xw.assign(x, x + 4 * (size_t)n); h->lm_host_fresh = true;
  }
  h->lm_ref_pose.assign(rp, rp + n);
  if (rc) h->lm_ref_cam.assign(rc, rc + n); else h->lm_ref_cam.assign(n, 0);
  if (a) h->lm_active.assign(a, a + n); else h->lm_active.assign(n, 1);
  if (dev && n > 0) { GUARD(h->sync();) }   // the caller's buffer is free again on return
  h->lm_reliable.clear(); h->lm_outliers.clear();   // all reliable, no outliers until a Solve() says otherwise
  h->dirty = true;
  return BAGPU_OK;
}
int bagpu_set_proj_residuals(bagpu_handle h, uint32_t n, const double* z, const uint32_t* mp, const uint32_t* lm, const uint32_t* cam,
                             const double* w) {
  ApiClock api_clock("bagpu_set_proj_residuals");
  const size_t np = h->pose_t7.size() / 7, nl = h->n_lm_set;
  bool dev = true;
  try { h->ensure_device(); } catch (const std::string&) { dev = false; }
  h->raw_on_device = false; h->raw_has_cam = cam != nullptr; h->raw_has_w0 = w != nullptr; h->n_raw = n;
  if (dev && n > 0) {
    // the observations go straight into HBM (the copies below are staged by the driver: the caller's buffers are
    // free again on return); ids are checked there
    GUARD(
      h->d_raw_z.upload(reinterpret_cast<const double2*>(z), n, h->st); h->d_raw_pose.upload(mp, n, h->st); h->d_raw_lm.upload(lm, n, h->st);
      if (cam) h->d_raw_cam.upload(cam, n, h->st);
      if (w) h->d_raw_w0.upload(w, n, h->st);
      if (api_clock.pc.on) { h->sync(); api_clock.pc.lap("  raw observation upload"); }
      h->d_lm_ref_pose.upload(h->lm_ref_pose, h->st); h->d_lm_ref_cam.upload(h->lm_ref_cam, h->st); h->lm_ids_on_device = true;
      DevBuf<int> flag; flag.alloc(2); flag.zero(h->st);
      launch_validate_obs(h->d_raw_pose.p, h->d_raw_lm.p, cam ? h->d_raw_cam.p : nullptr, n, (uint32_t)np, (uint32_t)nl, h->L, h->d_lm_ref_pose.p,
                          h->d_lm_ref_cam.p, flag.p, h->st);
      int f[2] = {0, 0};
      flag.download(f, 2, h->st); h->sync();
      if (f[0]) { h->n_raw = 0; return fail(BAGPU_ERR_INVALID, "projection residual id out of range"); }
      if (f[1]) { h->n_raw = 0; return fail(BAGPU_ERR_INVALID, "residual of a landmark in its own reference frame (AddProjectionResidual returns -1 for it)"); }
      h->raw_on_device = true;
    )
  } else {
    for (uint32_t i = 0; i < n; ++i) {
      if (mp[i] >= np || lm[i] >= nl) return fail(BAGPU_ERR_INVALID, "projection residual id out of range");
      const uint32_t c = cam ? cam[i] : 0;
      if (h->L == 1 && mp[i] == h->lm_ref_pose[lm[i]] && c == h->lm_ref_cam[lm[i]])
        return fail(BAGPU_ERR_INVALID, "residual of a landmark in its own reference frame (AddProjectionResidual returns -1 for it)");
    }
    h->proj_z.assign(z, z + 2 * (size_t)n);
    if (w) h->proj_w0.assign(w, w + n); else h->proj_w0.clear();
  }
  // host copies of the ids: only the pattern queries (bagpu_get_pattern) walk them; with the observations in HBM
  // they are fetched back on demand (host_ids())
  h->ids_on_host = !h->raw_on_device;
  if (h->ids_on_host) {
    h->proj_pose.assign(mp, mp + n); h->proj_lm.assign(lm, lm + n);
    if (cam) h->proj_cam.assign(cam, cam + n); else h->proj_cam.clear();
  } else {
    h->proj_pose.clear(); h->proj_lm.clear(); h->proj_cam.clear();
  }
  h->dirty = true;
  return BAGPU_OK;
}
int bagpu_set_unary_residuals(bagpu_handle h, uint32_t n, const uint32_t* pid, const double* t, const double* ci, const uint8_t* ur) {
  const size_t np = h->pose_t7.size() / 7;
  for (uint32_t i = 0; i < n; ++i) if (pid[i] >= np) return fail(BAGPU_ERR_INVALID, "unary pose id out of range");
  h->un_pose.assign(pid, pid + n); h->un_t.assign(t, t + 7 * (size_t)n); h->un_cov_inv.assign(ci, ci + 36 * (size_t)n);
  if (ur) h->un_rot.assign(ur, ur + n); else h->un_rot.assign(n, 1);
  h->dirty = true;
  return BAGPU_OK;
}
int bagpu_set_binary_residuals(bagpu_handle h, uint32_t n, const uint32_t* p1, const uint32_t* p2, const double* t, const double* ci,
                               const double* w, const uint8_t* ur) {
  const size_t np = h->pose_t7.size() / 7;
  for (uint32_t i = 0; i < n; ++i) if (p1[i] >= np || p2[i] >= np) return fail(BAGPU_ERR_INVALID, "binary pose id out of range");
  h->bin_p1.assign(p1, p1 + n); h->bin_p2.assign(p2, p2 + n); h->bin_t.assign(t, t + 7 * (size_t)n); h->bin_cov_inv.assign(ci, ci + 36 * (size_t)n);
  if (w) h->bin_w.assign(w, w + n); else h->bin_w.assign(n, 1.0);
  if (ur) h->bin_rot.assign(ur, ur + n); else h->bin_rot.assign(n, 1);
  h->dirty = true;
  return BAGPU_OK;
}
int bagpu_set_imu_residuals(bagpu_handle h, uint32_t n, const uint32_t* p1, const uint32_t* p2, const uint32_t* ms, const double* m7,
                            const double* w) {
  if (n > 0 && h->D < 9) return fail(BAGPU_ERR_INVALID, "inertial residuals need PoseSize >= 9");
  const size_t np = h->pose_t7.size() / 7;
  for (uint32_t i = 0; i < n; ++i) if (p1[i] >= np || p2[i] >= np) return fail(BAGPU_ERR_INVALID, "imu pose id out of range");
  h->imu_p1.assign(p1, p1 + n); h->imu_p2.assign(p2, p2 + n);
  if (n) { h->imu_ms.assign(ms, ms + n + 1); h->imu_meas.assign(m7, m7 + 7 * (size_t)ms[n]); } else { h->imu_ms.assign(1, 0); h->imu_meas.clear(); }
  if (w) h->imu_w.assign(w, w + n); else h->imu_w.assign(n, 1.0);
  h->dirty = true;
  return BAGPU_OK;
}

int bagpu_solve(bagpu_handle h, uint32_t max_iter, double gn_damping, int allow, bagpu_summary* out) {
  if (!h) return fail(BAGPU_ERR_INVALID, "null handle");
  GUARD(h->solve(max_iter, gn_damping, allow != 0);)
  if (h->peer.ok && h->peer.error(h->st))
    return fail(BAGPU_ERR_NCCL, "a peer-memory reduction timed out: some rank did not reach the collective");
  if (out) *out = h->summary;
  return BAGPU_OK;
}

int bagpu_build(bagpu_handle h, float* ms_out) {
  if (!h) return fail(BAGPU_ERR_INVALID, "null handle");
  GUARD(
    if (!h->solved_once || h->dirty) h->prepare_solve();
    h->solved_once = true;
    h->build();
    h->sync();
    h->collect_build_times();
    if (ms_out) { float ms; cudaEventElapsedTime(&ms, h->ev[0], h->ev[6]); *ms_out = ms; }
  )
  return BAGPU_OK;
}

int bagpu_get_poses(bagpu_handle h, double* t, double* v, double* b, uint8_t* mu, uint8_t* m) {
  ApiClock api_clock("bagpu_get_poses");
  const size_t n = h->pose_t7.size() / 7;
  if (t) memcpy(t, h->pose_t7.data(), sizeof(double) * 7 * n);
  if (v) memcpy(v, h->pose_v.data(), sizeof(double) * 3 * n);
  if (b) memcpy(b, h->pose_b.data(), sizeof(double) * 6 * n);
  if (mu) memcpy(mu, h->pose_mask_used.data(), n);
  if (m) memcpy(m, h->pose_mask.data(), n * h->D);
  return BAGPU_OK;
}
int bagpu_get_landmarks(bagpu_handle h, double* xw, double* xs, uint8_t* rel, uint32_t* nout) {
  ApiClock api_clock("bagpu_get_landmarks");
  const size_t n = h->n_lm_set;
  if (!h->lm_host_fresh && h->lm_rel_on_device) { GUARD(h->copy_landmarks_out(xw, xs, rel, nout);) return BAGPU_OK; }
  if (!h->lm_host_fresh) {   // set but not solved yet: positions as sent, x_s not known
    GUARD(if (xw) { h->d_lm_xw.download(xw, 4 * n, h->st); h->sync(); })
    if (xs) memset(xs, 0, sizeof(double) * 4 * n);
    if (rel) memset(rel, 1, n);
    if (nout) memset(nout, 0, sizeof(uint32_t) * n);
    return BAGPU_OK;
  }
  if (xw) memcpy(xw, h->lm_xw.data(), sizeof(double) * 4 * n);
  if (xs) { if (h->lm_xs.size() == 4 * n) memcpy(xs, h->lm_xs.data(), sizeof(double) * 4 * n); else memset(xs, 0, sizeof(double) * 4 * n); }
  if (rel) { if (h->lm_reliable.size() == n) memcpy(rel, h->lm_reliable.data(), n); else memset(rel, 1, n); }
  if (nout) { if (h->lm_outliers.size() == n) memcpy(nout, h->lm_outliers.data(), sizeof(uint32_t) * n); else memset(nout, 0, sizeof(uint32_t) * n); }
  return BAGPU_OK;
}
int bagpu_get_proj_residuals(bagpu_handle h, double* r, double* w, double* m) {
  GUARD(h->ensure_device(); if (h->dirty) return fail(BAGPU_ERR_INVALID, "no solve yet"); h->export_residuals(r, w, m);)
  return BAGPU_OK;
}
int bagpu_get_other_residuals(bagpu_handle h, double* un_residual, double* un_maha, double* un_cov_inv, double* bin_residual, double* bin_maha,
                              double* imu_residual, double* imu_maha, double* imu_cov_inv) {
  if (!h) return fail(BAGPU_ERR_INVALID, "null handle");
  GUARD(
    h->ensure_device();
    if (h->dirty) return fail(BAGPU_ERR_INVALID, "no solve yet");
    const int D = h->D;
    const size_t nu = h->un_pose.size(), nb = h->bin_p1.size(), ni = h->imu_p1.size();
    if (nu && (un_residual || un_maha)) {
      const std::vector<double> o = h->other_outputs(h->d_un_out, 7 * nu);
      for (size_t i = 0; i < nu; ++i) { if (un_residual) memcpy(un_residual + 6 * i, &o[7 * i], 48); if (un_maha) un_maha[i] = o[7 * i + 6]; }
    }
    if (nu && un_cov_inv) { const std::vector<double> o = h->other_outputs(h->d_un_cov_inv, 36 * nu); memcpy(un_cov_inv, o.data(), o.size() * 8); }
    if (nb && (bin_residual || bin_maha)) {
      const std::vector<double> o = h->other_outputs(h->d_bin_out, 7 * nb);
      for (size_t i = 0; i < nb; ++i) { if (bin_residual) memcpy(bin_residual + 6 * i, &o[7 * i], 48); if (bin_maha) bin_maha[i] = o[7 * i + 6]; }
    }
    if (ni && (imu_residual || imu_maha)) {
      const std::vector<double> o = h->other_outputs(h->d_imu_out, (size_t)(D + 1) * ni);
      for (size_t i = 0; i < ni; ++i) { if (imu_residual) memcpy(imu_residual + (size_t)D * i, &o[(size_t)(D + 1) * i], sizeof(double) * D); if (imu_maha) imu_maha[i] = o[(size_t)(D + 1) * i + D]; }
    }
    if (ni && imu_cov_inv) { const std::vector<double> o = h->other_outputs(h->d_imu_cov_inv, (size_t)D * D * ni); memcpy(imu_cov_inv, o.data(), o.size() * 8); }
  )
  return BAGPU_OK;
}

int bagpu_get_trace(bagpu_handle h, uint32_t maxr, bagpu_iter_trace* out, uint32_t* n) {
  const uint32_t c = std::min<uint32_t>(maxr, (uint32_t)h->trace.size());
  if (out && c) memcpy(out, h->trace.data(), c * sizeof(bagpu_iter_trace));
  if (n) *n = (uint32_t)h->trace.size();
  return BAGPU_OK;
}
int bagpu_get_deltas(bagpu_handle h, uint32_t iter, double* dp, double* dl) {
  if (iter >= h->rec_dp.size()) return fail(BAGPU_ERR_INVALID, "delta not recorded (options.record_deltas)");
  if (dp) memcpy(dp, h->rec_dp[iter].data(), h->rec_dp[iter].size() * 8);
  if (dl) memcpy(dl, h->rec_dl[iter].data(), h->rec_dl[iter].size() * 8);
  return BAGPU_OK;
}

// Block patterns exactly as the reference's sparse products produce them (host, from the ids):
//   U  = pattern of J^T J (upper, SparseBlockMatrixOps.h:236-238) over all residual types
//   W  = jt_pr_j_l_ (P x N_l)   S = U  union  pattern of (W V^-1) W^T (upper)
int bagpu_get_pattern(bagpu_handle h, int which, uint32_t* ncols, uint32_t* nnzb, uint32_t* colptr, uint32_t* rowidx) {
  if (h->dirty) { h->setup(); }
  GUARD(h->host_order();)
  h->host_lm_opt();
  const int P = h->n_ap, NL = h->n_al, L = h->L;
  std::vector<std::vector<uint32_t>> cols;
  auto lm_poses = [&](int l, std::vector<uint32_t>& out) {   // pose rows of W column l (sorted unique)
    out.clear();
    const int rp = (int)h->lm_ref_pose[l];
    bool any = false;
    for (uint32_t e = h->lm_obs_start[l]; e < h->lm_obs_start[l + 1]; ++e) {
      const int m = (int)h->proj_pose[h->obs_perm[e]];
      if (m != rp || L != 1) {
        if (m != rp) any = true;
        if (h->pose_opt[m] >= 0 && (m != rp || L != 1)) out.push_back(h->pose_opt[m]);
      }
    }
    if (L == 1 && any && h->pose_opt[rp] >= 0) out.push_back(h->pose_opt[rp]);
    std::sort(out.begin(), out.end()); out.erase(std::unique(out.begin(), out.end()), out.end());
  };
  if (which == BAGPU_PATTERN_W) {
    cols.assign(NL, {});
    for (int l = 0; l < h->n_lms; ++l) if (h->lm_opt[l] >= 0) lm_poses(l, cols[h->lm_opt[l]]);
  } else if (which == BAGPU_PATTERN_U || which == BAGPU_PATTERN_S) {
    cols.assign(P, {});
    auto add = [&](int a, int b) { if (a < 0 || b < 0) return; cols[std::max(a, b)].push_back(std::min(a, b)); };
    for (size_t i = 0; i < h->proj_pose.size(); ++i) {
      const int m = (int)h->proj_pose[i], rp = (int)h->lm_ref_pose[h->proj_lm[i]];
      if (m != rp || L != 1) {
        add(h->pose_opt[m], h->pose_opt[m]);
        if (L == 1) { add(h->pose_opt[rp], h->pose_opt[rp]); add(h->pose_opt[m], h->pose_opt[rp]); }
      }
    }
    for (size_t i = 0; i < h->bin_p1.size(); ++i) {
      const int a = h->pose_opt[h->bin_p1[i]], b = h->pose_opt[h->bin_p2[i]];
      add(a, a); add(b, b); add(a, b);
    }
    for (uint32_t p : h->un_pose) add(h->pose_opt[p], h->pose_opt[p]);
    for (size_t i = 0; i < h->imu_p1.size(); ++i) {
      const int a = h->pose_opt[h->imu_p1[i]], b = h->pose_opt[h->imu_p2[i]];
      add(a, a); add(b, b); add(a, b);
    }
    if (which == BAGPU_PATTERN_S && L > 0 && NL > 0) {
      std::vector<uint32_t> ps;
      for (int l = 0; l < h->n_lms; ++l) {
        if (h->lm_opt[l] < 0) continue;
        lm_poses(l, ps);
        for (size_t a = 0; a < ps.size(); ++a) for (size_t b = a; b < ps.size(); ++b) cols[ps[b]].push_back(ps[a]);
      }
    }
    for (auto& c : cols) { std::sort(c.begin(), c.end()); c.erase(std::unique(c.begin(), c.end()), c.end()); }
  } else {
    return fail(BAGPU_ERR_INVALID, "unknown pattern id");
  }
  size_t nz = 0;
  for (auto& c : cols) nz += c.size();
  if (ncols) *ncols = (uint32_t)cols.size();
  if (nnzb) *nnzb = (uint32_t)nz;
  if (colptr) { colptr[0] = 0; for (size_t j = 0; j < cols.size(); ++j) colptr[j + 1] = colptr[j] + (uint32_t)cols[j].size(); }
  if (rowidx) { size_t k = 0; for (auto& c : cols) for (uint32_t r : c) rowidx[k++] = r; }
  return BAGPU_OK;
}

int bagpu_get_system(bagpu_handle h, double* s_dense, double* rhs_sc, double* rhs_p, double* rhs_l, double* v_inv) {
  GUARD(
    h->ensure_device();
    if (h->dirty) return fail(BAGPU_ERR_INVALID, "no build yet");
    const int n = h->n_ap * h->D, T = h->T, nT = h->nT;
    if (s_dense) {
      std::vector<double> tiles((size_t)h->n_tiles * T * T);
      h->d_S.download(tiles.data(), tiles.size(), h->st);
      h->sync();
      memset(s_dense, 0, sizeof(double) * (size_t)n * n);
      for (int tI = 0; tI < nT; ++tI) for (int tJ = tI; tJ < nT; ++tJ) {   // natural tile ids
        const int a = h->tile_perm[tI], b = h->tile_perm[tJ];
        const int slot = h->tile_slot(std::min(a, b), std::max(a, b));
        if (slot < 0) continue;
        for (int c = 0; c < T; ++c) for (int r = 0; r < T; ++r) {
          const int gi = tI * T + r, gj = tJ * T + c;
          const size_t off = (a <= b) ? (size_t)c * T + r : (size_t)r * T + c;   // stored transposed when the order swaps the tiles
          if (gi < n && gj < n && gi <= gj) s_dense[(size_t)gj * n + gi] = tiles[(size_t)slot * T * T + off];
        }
      }
    }
    if (rhs_sc) h->d_rhs_sc.download(rhs_sc, n, h->st);
    if (rhs_p) h->d_bp.download(rhs_p, n, h->st);
    const int Ls = std::max(1, h->L);
    if ((rhs_l || v_inv) && h->L > 0 && h->n_lms) {
      h->host_lm_opt();
      std::vector<double> bl((size_t)h->n_lms * Ls), rec((size_t)h->n_lms * kLmRec);
      h->d_lm_bl.download(bl.data(), bl.size(), h->st); h->d_lm_rec.download(rec.data(), rec.size(), h->st);
      h->sync();
      for (int l = 0; l < h->n_lms; ++l) {
        const int o = h->lm_opt[l];
        if (o < 0) continue;
        if (rhs_l) for (int k = 0; k < Ls; ++k) rhs_l[(size_t)o * Ls + k] = bl[(size_t)l * Ls + k];
        if (v_inv) {   // the landmark record holds V^-1 (LmSize 3: its upper triangle)
          const double* r = &rec[(size_t)l * kLmRec];
          double* v = v_inv + (size_t)o * Ls * Ls;
          if (Ls == 1) v[0] = r[7];
          else { v[0] = r[4]; v[1] = r[5]; v[2] = r[6]; v[3] = r[5]; v[4] = r[7]; v[5] = r[8]; v[6] = r[6]; v[7] = r[8]; v[8] = r[9]; }
        }
      }
    }
    h->sync();
  )
  return BAGPU_OK;
}

int bagpu_dump(bagpu_handle h, const char* dir) {
  // s.txt / rhs.txt with Eigen's FullPrecision ", " separated rows (reference Utils.h:66, cpp:604-605)
  if (!h) return fail(BAGPU_ERR_INVALID, "null handle");
  const int n = h->n_ap * h->D;
  // the reference writes the dense (P d)^2 matrix as text (cpp:604): refuse sizes nobody can hold or read (C4 = 28.8 GB)
  if ((size_t)n * n > ((size_t)1 << 28)) return fail(BAGPU_ERR_UNSUPPORTED, "dense s.txt dump refused above 2 GB (use bagpu_get_pattern / the tile storage)");
  std::vector<double> s, r;
  try { s.resize((size_t)n * n); r.resize(n); } catch (const std::exception&) { return fail(BAGPU_ERR_INVALID, "out of host memory for the dense dump"); }
  int rc = bagpu_get_system(h, s.data(), r.data(), nullptr, nullptr, nullptr);
  if (rc) return rc;
  const std::string d = dir ? dir : ".";
  FILE* f = fopen((d + "/s.txt").c_str(), "w");
  if (!f) return fail(BAGPU_ERR_INVALID, "cannot open s.txt");
  // Eigen's kLongCsvFmt puts "\n" between rows, none after the last
  for (int i = 0; i < n; ++i) { if (i) fputc('\n', f); for (int j = 0; j < n; ++j) fprintf(f, j ? ", %.17g" : "%.17g", s[(size_t)j * n + i]); }
  fclose(f);
  f = fopen((d + "/rhs.txt").c_str(), "w");
  if (!f) return fail(BAGPU_ERR_INVALID, "cannot open rhs.txt");
  for (int i = 0; i < n; ++i) fprintf(f, i ? "\n%.17g" : "%.17g", r[i]);
  fclose(f);
  // j_pr.txt / r_pr.txt / j_l.txt (cpp:607-620): the dense images of the block-sparse Jacobians, rows in residual-id
  // order.  The build is matrix-free, so the rows are evaluated for the dump only (k_jacobian_rows), with the Huber
  // weights of the last build.  Single GPU only (a rank holds the observations of its landmarks).
  const int L = h->L;
  const size_t no = (size_t)h->n_obs_total;
  if (L == 0 || no == 0 || h->world > 1) return BAGPU_OK;
  const size_t cols_p = (size_t)h->n_ap * 6, cols_l = (size_t)h->n_al * L;
  if (2 * no * std::max(cols_p, cols_l) > ((size_t)1 << 27)) return BAGPU_OK;   // like the oracle: no dense J beyond 1 GB
  const int W = 28 + 2 * L;
  std::vector<double> rows;
  GUARD(
    DevBuf<double> d_rows; d_rows.alloc(no * W);
    launch_jacobian_rows(h->P, d_rows.p, h->st);
    rows.resize(no * W);
    d_rows.download(rows.data(), rows.size(), h->st);
    h->sync();
  )
  GUARD(h->host_ids();)
  h->host_lm_opt();
  FILE* fj = fopen((d + "/j_pr.txt").c_str(), "w");
  FILE* fr = fopen((d + "/r_pr.txt").c_str(), "w");
  FILE* fl = fopen((d + "/j_l.txt").c_str(), "w");
  if (!fj || !fr || !fl) { if (fj) fclose(fj); if (fr) fclose(fr); if (fl) fclose(fl); return fail(BAGPU_ERR_INVALID, "cannot open j_pr.txt / r_pr.txt / j_l.txt"); }
  std::vector<double> line(std::max(cols_p, cols_l));
  for (size_t i = 0; i < no; ++i) {
    const double* q = &rows[i * W];
    const int om = (int)q[26 + 2 * L], orf = (int)q[27 + 2 * L];
    const int ol = h->lm_opt[h->proj_lm[i]];
    for (int rr = 0; rr < 2; ++rr) {
      std::fill(line.begin(), line.begin() + cols_p, 0.0);
      if (om >= 0) for (int k = 0; k < 6; ++k) line[(size_t)om * 6 + k] = q[2 + 6 * rr + k];
      if (orf >= 0) for (int k = 0; k < 6; ++k) line[(size_t)orf * 6 + k] = q[14 + 6 * rr + k];
      if (i || rr) fputc('\n', fj);
      for (size_t j = 0; j < cols_p; ++j) fprintf(fj, j ? ", %.17g" : "%.17g", line[j]);
      fprintf(fr, (i || rr) ? "\n%.17g" : "%.17g", q[rr]);
      std::fill(line.begin(), line.begin() + cols_l, 0.0);
      if (ol >= 0) for (int k = 0; k < L; ++k) line[(size_t)ol * L + k] = q[26 + L * rr + k];
      if (i || rr) fputc('\n', fl);
      for (size_t j = 0; j < cols_l; ++j) fprintf(fl, j ? ", %.17g" : "%.17g", line[j]);
    }
  }
  fclose(fj); fclose(fr); fclose(fl);
  return BAGPU_OK;
}

int bagpu_nccl_unique_id(void* out128) {
  if (!g_nccl.load()) return fail(BAGPU_ERR_NCCL, "libnccl not found (set BAGPU_NCCL_LIB)");
  UniqueId id;
  int rc = g_nccl.getUniqueId(&id);
  if (rc) return fail(BAGPU_ERR_NCCL, "ncclGetUniqueId failed");
  memcpy(out128, &id, 128);
  return BAGPU_OK;
}

int bagpu_comm_init(bagpu_handle h, int rank, int world, const void* id128) {
  if (world <= 1) { h->rank = 0; h->world = 1; return BAGPU_OK; }
  if (!g_nccl.load()) return fail(BAGPU_ERR_NCCL, "libnccl not found (set BAGPU_NCCL_LIB)");
  GUARD(h->ensure_device();)
  UniqueId id;
  memcpy(&id, id128, 128);
  int rc = g_nccl.commInitRank(&h->comm, world, id, rank);
  if (rc) return fail(BAGPU_ERR_NCCL, std::string("ncclCommInitRank failed: ") + (g_nccl.getErrorString ? g_nccl.getErrorString(rc) : "?"));
  h->rank = rank; h->world = world;
  // NCCL connects its channels lazily, at the first collective of each size class (hundreds of ms): do that here, as
  // part of communicator set-up, with one small and one bulk all-reduce, so that the first Solve() does not pay for it.
  GUARD({
    const size_t n_big = (size_t)4 << 20;  // 32 MB of doubles: the order of an S all-reduce
    DevBuf<double> warm; warm.alloc(n_big);
    warm.zero(h->st);
    h->allreduce_u32(reinterpret_cast<unsigned int*>(warm.p), 4096);
    h->allreduce_f64(warm.p, 64);
    h->allreduce_f64(warm.p, n_big);
    CU(cudaStreamSynchronize(h->st));
    // windows for the small reductions: IPC handles gathered with one all-reduce of a zero-padded table, then a vote
    // (every rank must have mapped every peer, otherwise all of them stay on NCCL)
    const char* pe = getenv("BAGPU_P2P");
    if (world <= kPeerMaxWorld && !(pe && pe[0] == '0')) {
      std::vector<unsigned int> tab((size_t)world * 16 + 1, 0u);
      bool mine = h->peer.create_local(rank, world, &tab[(size_t)rank * 16]) == cudaSuccess;
      if (!mine) { cudaGetLastError(); memset(&tab[(size_t)rank * 16], 0, 64); }
      unsigned int* d_tab = reinterpret_cast<unsigned int*>(warm.p);
      CU(cudaMemcpyAsync(d_tab, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, h->st));
      h->allreduce_u32(d_tab, tab.size());
      CU(cudaMemcpyAsync(tab.data(), d_tab, tab.size() * 4, cudaMemcpyDeviceToHost, h->st));
      CU(cudaStreamSynchronize(h->st));
      const bool opened = mine && h->peer.open_peers(tab.data());
      tab.assign(1, opened ? 1u : 0u);
      CU(cudaMemcpyAsync(d_tab, tab.data(), 4, cudaMemcpyHostToDevice, h->st));
      h->allreduce_u32(d_tab, 1);
      CU(cudaMemcpyAsync(tab.data(), d_tab, 4, cudaMemcpyDeviceToHost, h->st));
      CU(cudaStreamSynchronize(h->st));
      if (tab[0] == (unsigned int)world) h->peer.ok = true;
      else {   // stay on NCCL: unmap, meet the others, then free (same order as in bagpu_destroy)
        h->peer.close_peers();
        h->allreduce_u32(d_tab, 1);
        CU(cudaStreamSynchronize(h->st));
        h->peer.destroy();
      }
      if (getenv("BAGPU_TIMING")) fprintf(stderr, "[bagpu] rank %d: peer-memory reductions %s\n", rank, h->peer.ok ? "on" : "off (NCCL only)");
    }
  })
  return BAGPU_OK;
}

int bagpu_build_work_items(uint32_t n_active_poses, int lm_dim, const uint32_t* pose_rec_start, const uint32_t* pose_lmrec_start,
                           uint32_t n_blocks, const uint64_t* block_key, const uint32_t* block_count, uint32_t* items10,
                           uint32_t* n_items, uint32_t* sym10, uint32_t* n_sym, uint32_t* n_partial, uint32_t* n_groups) {
  if (!pose_rec_start || !n_items || !n_sym || (lm_dim == 1 && !pose_lmrec_start) || (n_blocks && (!block_key || !block_count)))
    return fail(BAGPU_ERR_INVALID, "null argument");
  static_assert(sizeof(ReduceItem) == 10 * sizeof(uint32_t), "ReduceItem is ten uint32");
  std::vector<unsigned long long> bkey(block_key, block_key + n_blocks);
  std::vector<uint32_t> bcnt(block_count, block_count + n_blocks);
  size_t n_pairs = 0;
  for (uint32_t c : bcnt) n_pairs += c;
  std::vector<ReduceItem> items, sym;
  uint32_t np = 0, ng = 0;
  build_work_items((int)n_active_poses, lm_dim, pose_rec_start, pose_lmrec_start, bkey, bcnt, n_pairs, items, sym, np, ng);
  if (items.size() > *n_items || sym.size() > *n_sym) return fail(BAGPU_ERR_INVALID, "output does not fit");
  if (items10 && !items.empty()) memcpy(items10, items.data(), items.size() * sizeof(ReduceItem));
  if (sym10 && !sym.empty()) memcpy(sym10, sym.data(), sym.size() * sizeof(ReduceItem));
  *n_items = (uint32_t)items.size(); *n_sym = (uint32_t)sym.size();
  if (n_partial) *n_partial = np;
  if (n_groups) *n_groups = ng;
  return BAGPU_OK;
}

int bagpu_pack_warp_tasks(uint32_t n_lms, const uint32_t* starts, uint32_t lm_lo, uint32_t lm_hi, uint32_t* task_first,
                          uint32_t* task_count, uint32_t* n_tasks, uint32_t* long_lms, uint32_t* n_long) {
  if (!starts || !n_tasks || !n_long || lm_lo > lm_hi || lm_hi > n_lms) return fail(BAGPU_ERR_INVALID, "bad argument");
  std::vector<uint2> tasks; std::vector<uint32_t> longs;
  pack_warp_tasks(starts, (int)lm_lo, (int)lm_hi, starts[lm_lo], starts[lm_hi], tasks, longs);
  if (tasks.size() > *n_tasks || longs.size() > *n_long) return fail(BAGPU_ERR_INVALID, "output does not fit");
  for (size_t i = 0; i < tasks.size(); ++i) { if (task_first) task_first[i] = tasks[i].x; if (task_count) task_count[i] = tasks[i].y; }
  for (size_t i = 0; i < longs.size(); ++i) if (long_lms) long_lms[i] = longs[i];
  *n_tasks = (uint32_t)tasks.size(); *n_long = (uint32_t)longs.size();
  return BAGPU_OK;
}

int bagpu_symbolic_tile_fill(uint32_t nT, uint32_t n_edges, const uint32_t* ea, const uint32_t* eb, uint32_t* perm, uint32_t* row_ptr,
                             uint32_t* row_col, uint32_t cap) {
  if (!perm || !row_ptr || (n_edges && (!ea || !eb))) return fail(BAGPU_ERR_INVALID, "null argument");
  bagpu_handle_s h;
  h.nT = (int)nT;
  std::vector<uint8_t> flat((size_t)nT * nT, 0);
  for (uint32_t e = 0; e < n_edges; ++e) {
    if (ea[e] >= nT || eb[e] >= nT) return fail(BAGPU_ERR_INVALID, "tile id out of range");
    flat[(size_t)std::min(ea[e], eb[e]) * nT + std::max(ea[e], eb[e])] = 1;
  }
  h.tile_structure(flat);
  if (h.row_col.size() > cap) return fail(BAGPU_ERR_INVALID, "row_col does not fit");
  for (uint32_t i = 0; i < nT; ++i) perm[i] = (uint32_t)h.tile_perm[i];
  for (uint32_t i = 0; i <= nT; ++i) row_ptr[i] = (uint32_t)h.row_ptr[i];
  for (size_t i = 0; i < h.row_col.size(); ++i) row_col[i] = (uint32_t)h.row_col[i];
  return BAGPU_OK;
}

int bagpu_chol_task_plan(int tile_dim, uint32_t nT, const uint32_t* row_ptr, const uint32_t* row_col, int32_t* tasks8, uint32_t* n_tasks,
                         int32_t* terms3, uint32_t* n_terms) {
  if (!row_ptr || !n_tasks || !n_terms || (nT && row_ptr[nT] && !row_col)) return fail(BAGPU_ERR_INVALID, "null argument");
  if (tile_dim != 48 && tile_dim != 72 && tile_dim != 120) return fail(BAGPU_ERR_INVALID, "tile_dim must be 48, 72 or 120");
  GUARD(
    std::vector<int> rp(row_ptr, row_ptr + nT + 1), rc(row_col, row_col + row_ptr[nT]);
    for (uint32_t k = 0; k < nT; ++k)
      for (int e = rp[k]; e < rp[k + 1]; ++e)
        if (rc[e] <= (int)k || rc[e] >= (int)nT || (e > rp[k] && rc[e] <= rc[e - 1])) return fail(BAGPU_ERR_INVALID, "rows must list ascending columns above the diagonal");
    std::vector<CholTask> tasks; std::vector<CholTerm> terms;
    chol_build_plan(tile_dim, (int)nT, rp, rc, tasks, terms);
    const bool fits = tasks.size() <= *n_tasks && terms.size() <= *n_terms && (tasks8 || tasks.empty()) && (terms3 || terms.empty());
    *n_tasks = (uint32_t)tasks.size(); *n_terms = (uint32_t)terms.size();
    if (!fits) return fail(BAGPU_ERR_INVALID, "task plan does not fit");
    static_assert(sizeof(CholTask) == 8 * sizeof(int32_t) && sizeof(CholTerm) == 3 * sizeof(int32_t), "plan records are plain ints");
    if (!tasks.empty()) memcpy(tasks8, tasks.data(), tasks.size() * sizeof(CholTask));
    if (!terms.empty()) memcpy(terms3, terms.data(), terms.size() * sizeof(CholTerm));
  )
  return BAGPU_OK;
}

int bagpu_sync_state(bagpu_handle h) {
  if (!h) return fail(BAGPU_ERR_INVALID, "null handle");
  GUARD(h->ensure_device(); if (h->dirty) return fail(BAGPU_ERR_INVALID, "no solve yet"); h->download_state(); h->fetch_landmarks_host();)
  return BAGPU_OK;
}

int bagpu_get_phase_ms(bagpu_handle h, float* ms, uint64_t* launches) {
  if (ms) for (int i = 0; i < BAGPU_PH_COUNT; ++i) ms[i] = h->phase_ms[i];
  if (launches) *launches = h->launches;
  return BAGPU_OK;
}

}  // extern "C"
