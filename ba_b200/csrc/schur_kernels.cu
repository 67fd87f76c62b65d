// schur_kernels.cu — atomic-free, record-free assembly of the reduced camera system.
//
// Replaces, per outer iteration (reference src/BundleAdjuster.cpp):
//   U += Jpr^T Jpr, rhs_p += Jpr^T r                                    cpp:337-354
//   V = sum Jl^T w Jl, b_l, V^-1                                        cpp:409-443
//   W = Jpr^T Jl, W V^-1, S = U - W V^-1 W^T, rhs_sc = rhs_p - W V^-1 b_l   cpp:448-485
// (the four sparse block products of include/ba/SparseBlockMatrixOps.h:159-380).
//
// The reference scatters: per landmark it adds k(k+1)/2 outer products into S.  On the GPU that is ~36 FP64
// reductions per product into L2, bound by the L2 atomic units (profiles/r1a_*: 7 ms for 5M observations).  Round 1
// turned the product around with a 224/320-byte record per observation (write once, gather once per pose pair): no
// atomics, but 1.8x - 3.7x the algorithmic DRAM bytes and an L2-bound gather (profiles/r1f, r1i).  Here nothing per
// observation is stored at all:
//
//   pass A  k_linearize   (landmark-major, one observation per lane)
//       Jacobians, Huber weight, V, b_l, V^-1 by segmented warp shuffles; writes the weight (8 B per observation)
//       and ONE 128-byte record per landmark: the point, V^-1 and dhat = V^-1 b_l (LmSize 1: also the three sums
//       from which the reference pose's blocks follow).
//   pass B  k_reduce_blocks (pose-pair-major, one warp per S block chunk, one list entry per lane)
//       S_ij -= w_a w_b Jm_a^T (Jl_a V^-1 Jl_b^T) Jm_b over the block's pair list; both Jacobians of a pair are
//       re-evaluated from two 32-byte observations and the landmark record (5 sectors through L2 instead of 9-10)
//       against the two pose frames of the block, which are uniform over the warp.  Diagonal blocks, b_p and rhs_sc
//       come from the pose's own observation list the same way.  Each block is owned by one warp (or a fixed
//       chain of chunks combined in index order): NO atomics, bit-reproducible run to run.
// The pair lists depend only on the ids; they are built once per Solve() on the device (symbolic.cu).
#include <algorithm>
#include <cstdlib>
#include "proj_kernels.cuh"
#include "launch.h"

namespace bagpu {

BA_D void st2(double* p, double a, double b) { *reinterpret_cast<double2*>(p) = make_double2(a, b); }
BA_D void ld2(const double* p, double& a, double& b) { const double2 v = *reinterpret_cast<const double2*>(p); a = v.x; b = v.y; }

// ---------------------------------------------------------------------------------------
// pass A
// Running sums of one landmark over its observations.  LmSize 1: the reference-pose Jacobian of every observation of
// the landmark is Bp M with the same 3 x 6 M = [-rho I | Ar], so H_rr = M^T (sum w Bp^T Bp) M, g_r = M^T sum w Bp^T r,
// W_r = M^T sum w Bp^T J_l: 12 running sums instead of 33.
template <int L> struct LmSums {
  static constexpr int NV = L * (L + 1) / 2;
  static constexpr int N = NV + L + (L == 1 ? 12 : 0);
  double v[N];   // V (upper, NV) | b_l (L) | LmSize 1: SBB (6) | SBr (3) | SBl (3)
};

template <int L>
BA_D void sums_add(LmSums<L>& S, const ObsJac<L>& o, double w, bool lm_on) {
  constexpr int NV = LmSums<L>::NV;
  const double wr0 = w * o.r[0], wr1 = w * o.r[1];
  if (lm_on) {
    int k = 0;
#pragma unroll
    for (int a = 0; a < L; ++a) {
#pragma unroll
      for (int b = a; b < L; ++b) { S.v[k] += w * (o.Jl[a] * o.Jl[b] + o.Jl[L + a] * o.Jl[L + b]); ++k; }
      S.v[NV + a] += o.Jl[a] * wr0 + o.Jl[L + a] * wr1;
    }
  }
  if (L == 1 && o.ref_on) {
    double* SBB = S.v + NV + L; double* SBr = SBB + 6; double* SBl = SBr + 3;
    int k = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
      for (int b = a; b < 3; ++b) { SBB[k] += w * (o.Bp[a] * o.Bp[b] + o.Bp[3 + a] * o.Bp[3 + b]); ++k; }
      SBr[a] += o.Bp[a] * wr0 + o.Bp[3 + a] * wr1;
      if (lm_on) SBl[a] += w * (o.Bp[a] * o.Jl[0] + o.Bp[3 + a] * o.Jl[1]);
    }
  }
}

// V^-1 of a landmark from its sums (cpp:431-442); Vi = upper triangle (LmSize 3: 00 01 02 11 12 22), dhat = V^-1 b_l
template <int L>
BA_D void lm_invert(const LmSums<L>& S, bool lm_on, double* Vi, double* bl, double* dhat) {
  constexpr int NV = LmSums<L>::NV;
  const double* V = S.v;
#pragma unroll
  for (int i = 0; i < L; ++i) bl[i] = lm_on ? S.v[NV + i] : 0.0;
  if (lm_on) {
    if (L == 1) {
      double v = V[0];
      if (fabs(v) < 1e-6) v += 1e-6;
      Vi[0] = 1.0 / v;
    } else {
      double M[9] = {V[0], V[1], V[2], V[1], V[3], V[4], V[2], V[4], V[5]};
      double n2 = 0.0;
#pragma unroll
      for (int i = 0; i < 9; ++i) n2 += M[i] * M[i];
      if (sqrt(n2) < 1e-6) { M[0] += 1e-6; M[4] += 1e-6; M[8] += 1e-6; }
      double Mi[9];
      inv3(M, Mi);
      Vi[0] = Mi[0]; Vi[1] = Mi[1]; Vi[2] = Mi[2]; Vi[3] = Mi[4]; Vi[4] = Mi[5]; Vi[5] = Mi[8];
    }
  } else {
#pragma unroll
    for (int i = 0; i < NV; ++i) Vi[i] = 0.0;
  }
  if (L == 1) {
    dhat[0] = Vi[0] * bl[0];
  } else {
    dhat[0] = Vi[0] * bl[0] + Vi[1] * bl[1] + Vi[2] * bl[2];
    dhat[1] = Vi[1] * bl[0] + Vi[3] * bl[1] + Vi[4] * bl[2];
    dhat[2] = Vi[2] * bl[0] + Vi[4] * bl[1] + Vi[5] * bl[2];
  }
}

// the landmark's record (ba_gpu.h), written by one lane: eight 16-byte stores into one 128-byte line
template <int L>
BA_D void lm_write(const DeviceProblem& P, int lm, const LmCtx<L>& c, bool lm_on, const LmSums<L>& S) {
  constexpr int NV = LmSums<L>::NV;
  double Vi[NV], bl[L], dhat[L];
  lm_invert<L>(S, lm_on, Vi, bl, dhat);
  double* o = P.lm_rec + (size_t)lm * kLmRec;
  st2(o, c.X[0], c.X[1]); st2(o + 2, c.X[2], c.rho);
  if (L == 3) {
    st2(o + 4, Vi[0], Vi[1]); st2(o + 6, Vi[2], Vi[3]); st2(o + 8, Vi[4], Vi[5]);
    st2(o + 10, dhat[0], dhat[1]); st2(o + 12, dhat[2], 0.0); st2(o + 14, 0.0, 0.0);
  } else {
    const double* SBB = S.v + NV + L; const double* SBr = SBB + 6; const double* SBl = SBr + 3;
    const double* xs = P.lm_x + 4 * (size_t)lm;
    st2(o + 4, c.tws_r[0], c.tws_r[1]); st2(o + 6, c.tws_r[2], Vi[0]);
    st2(o + 8, dhat[0], xs[0]); st2(o + 10, xs[1], xs[2]);
    st2(o + 12, SBl[0], SBl[1]); st2(o + 14, SBl[2], 0.0);
    double* oc = P.lm_recC + (size_t)lm * kLmRecC;
    st2(oc, SBB[0], SBB[1]); st2(oc + 2, SBB[2], SBB[3]); st2(oc + 4, SBB[4], SBB[5]);
    st2(oc + 6, SBr[0], SBr[1]); st2(oc + 8, SBr[2], 0.0);
  }
#pragma unroll
  for (int i = 0; i < L; ++i) P.lm_bl[(size_t)lm * L + i] = bl[i];
}

// Landmarks with more than 32 observations (lm_list): one warp per landmark, lanes stride over the track.
template <int L>
__global__ void __launch_bounds__(128) k_linearize_long(DeviceProblem P, const uint32_t* __restrict__ lm_list, int n_list) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median), c2 = c_huber * c_huber;
  double err[1] = {0.0};
  for (int q = warp; q < n_list; q += n_warps) {
    const int lm = (int)lm_list[q];
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    const bool lm_on = c.lm_opt >= 0;
    LmSums<L> S;
#pragma unroll
    for (int i = 0; i < LmSums<L>::N; ++i) S.v[i] = 0.0;
    for (int i = s + lane; i < e; i += 32) {
      ObsJac<L> o;
      obs_jacobians<L>(P, c, i, o);
      const double w0 = P.obs_w0[i];
      const double n2 = o.r[0] * o.r[0] + o.r[1] * o.r[1];
      const double w = huber_weight(n2 * w0, w0, c_huber, c2, P.use_robust);
      P.obs_w[i] = w;
      err[0] += n2 * w;
      sums_add<L>(S, o, w, lm_on);
    }
#pragma unroll
    for (int i = 0; i < LmSums<L>::N; ++i) S.v[i] = warp_sum(S.v[i]);
    if (lane == 0) lm_write<L>(P, lm, c, lm_on, S);
  }
  grid_sum_fixed<1>(err, P.red_partial, P.red_counter, &P.sc->proj_error, true);
}

// Packed tasks: a task is a run of consecutive landmarks with at most 32 observations in total (packed on the host from
// the CSR bounds, symbolic set-up).  One warp per task, ONE OBSERVATION PER LANE: the Jacobians are evaluated once,
// the per-landmark sums are segmented warp reductions over the lanes of the landmark, the head lane of a landmark
// writes its record.
template <int L>
__global__ void __launch_bounds__(128) k_linearize_packed(DeviceProblem P, const uint2* __restrict__ tasks, int n_tasks) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median), c2 = c_huber * c_huber;
  double err[1] = {0.0};
  // two levels of the dependent chain (task -> landmark / pose ids -> landmark point / pose frame) are requested one
  // task ahead: the next task's bounds, and its lanes' landmark and pose ids
  uint2 tk_n = warp < n_tasks ? tasks[warp] : make_uint2(0, 0);
  uint2 tk_nn = warp + n_warps < n_tasks ? tasks[warp + n_warps] : make_uint2(0, 0);
  int lm_n = 0, m_n = 0;
  if (warp < n_tasks) { const int i0 = (int)tk_n.x + (lane < (int)tk_n.y ? lane : 0); lm_n = (int)P.obs_lm[i0]; m_n = (int)P.obs_pose[i0]; }
  for (int t = warp; t < n_tasks; t += n_warps) {
    const uint2 tk = tk_n;
    const int lm = lm_n, m_pre = m_n;
    tk_n = tk_nn;
    if (t + 2 * n_warps < n_tasks) tk_nn = tasks[t + 2 * n_warps];
    if (t + n_warps < n_tasks) { const int i0 = (int)tk_n.x + (lane < (int)tk_n.y ? lane : 0); lm_n = (int)P.obs_lm[i0]; m_n = (int)P.obs_pose[i0]; }
    const int s = (int)tk.x, n = (int)tk.y;            // first observation, number of observations (<= 32)
    const bool on = lane < n;
    const int i = s + (on ? lane : 0);
    // segments = landmarks: head lane of mine, end of mine
    const int lm_prev = __shfl_up_sync(0xffffffffu, lm, 1);
    const bool head = on && (lane == 0 || lm != lm_prev);
    const unsigned heads = __ballot_sync(0xffffffffu, head);
    const unsigned above = (lane == 31) ? 0u : (heads & (0xffffffffu << (lane + 1)));
    const int seg_end = above ? (__ffs(above) - 1) : n;
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    const bool lm_on = c.lm_opt >= 0;
    LmSums<L> S;
#pragma unroll
    for (int k = 0; k < LmSums<L>::N; ++k) S.v[k] = 0.0;
    if (on) {
      ObsJac<L> o;
      obs_jacobians<L>(P, c, i, o, m_pre);
      const double w0 = P.obs_w0[i];
      const double n2 = o.r[0] * o.r[0] + o.r[1] * o.r[1];
      const double w = huber_weight(n2 * w0, w0, c_huber, c2, P.use_robust);
      P.obs_w[i] = w;
      err[0] += n2 * w;
      sums_add<L>(S, o, w, lm_on);
    }
    // segmented reduction to the head lane (a lane adds its neighbour at distance d only inside its own landmark)
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const bool take = on && (lane + d < seg_end);
#pragma unroll
      for (int k = 0; k < LmSums<L>::N; ++k) {
        const double other = __shfl_down_sync(0xffffffffu, S.v[k], d);
        if (take) S.v[k] += other;
      }
    }
    if (head) lm_write<L>(P, lm, c, lm_on, S);
  }
  grid_sum_fixed<1>(err, P.red_partial, P.red_counter, &P.sc->proj_error, true);
}

// landmarks without a (local) observation: the outputs the kernels above would have written for an empty track
template <int L>
__global__ void k_lm_defaults(DeviceProblem P) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  if (P.lm_obs_start[lm] != P.lm_obs_start[lm + 1]) return;
  LmSums<L> S;
#pragma unroll
  for (int k = 0; k < LmSums<L>::N; ++k) S.v[k] = 0.0;
  LmCtx<L> c;
  load_lm<L>(P, lm, c);
  lm_write<L>(P, lm, c, c.lm_opt >= 0, S);
}

// ---------------------------------------------------------------------------------------
// pass B: one warp per work item (a chunk of one S block), items claimed in order from one counter so that the warps
// in flight work on neighbouring blocks (their observations and landmark records stay in L2).  One list entry per
// lane.  The 5 sectors of an entry are brought into shared memory by cp.async with CHUNK-wise lane mapping —
// consecutive lanes copy consecutive 16-byte pieces of one record, so an instruction touches 4-16 cache lines instead
// of 32 (L1 serves one line per cycle) — in a multi-stage pipeline; the lane then reads its own entry back (stride of
// an odd number of 16-byte words: conflict-free).  Accumulators: the 36 elements of the block (or 21 + 6 + 6 for a
// diagonal block with b_p and rhs_sc) per lane, transposed through shared memory at the end of the item so that lane e
// owns element e.
constexpr int kPartial = 64;       // doubles per partial-sum slot
constexpr int kRedWarps = 4;
constexpr int kNAcc = 36;
template <int L> struct RedCfg {
  static constexpr int STRIDE = (L == 3) ? 22 : 30;   // doubles per staged entry: 176 / 240 bytes
  static constexpr int STAGES = (L == 3) ? 3 : 2;
  static constexpr int DATA = 32 * STRIDE;
  static constexpr int BUF = DATA + 32;               // + 32 (a, b) index pairs
  static constexpr size_t smem = sizeof(double) * STAGES * BUF * kRedWarps;
};
static_assert(RedCfg<3>::STAGES * RedCfg<3>::BUF >= kNAcc * 33 && RedCfg<1>::STAGES * RedCfg<1>::BUF >= kNAcc * 33, "transpose buffer");

BA_D void cp_async16(double* smem_dst, const void* gsrc) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
BA_D void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> BA_D void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// stage NC 16-byte chunks of the record `base + id * rec_bytes` of every lane's entry (id = kNoSlot: none) at doubles
// offset `off` of the entry's stage slot
template <int NC, int STRIDE>
BA_D void stage_chunks(double* buf, int off, const char* __restrict__ base, int rec_bytes, uint32_t id, int lane) {
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const int idx = k * 32 + lane;
    const int p = idx / NC, ch = idx - p * NC;
    const uint32_t src = __shfl_sync(0xffffffffu, id, p);
    if (src != kNoSlot) cp_async16(buf + p * STRIDE + off + 2 * ch, base + (size_t)src * rec_bytes + 16 * ch);
  }
}

struct PairLists { const uint32_t *a, *b, *lm; };   // off-diagonal entries: observation of the row pose, of the column pose (or tag), landmark

// upper-triangle packing of the diagonal accumulators: k = c (c + 1) / 2 + r, r <= c
BA_D void add_upper(double* acc, const double* x0, const double* x1, const double* y0, const double* y1) {
  int k = 0;
#pragma unroll
  for (int c = 0; c < 6; ++c)
#pragma unroll
    for (int r = 0; r <= c; ++r) { acc[k] += x0[r] * y0[c] + x1[r] * y1[c]; ++k; }
}

template <int L>
__global__ void __launch_bounds__(kRedWarps * 32, (L == 3) ? 3 : 3) k_reduce_blocks(DeviceProblem P, const ReduceItem* __restrict__ items, int n_items, PairLists pl,
                                                                     double* __restrict__ partial, unsigned int* __restrict__ counters,
                                                                     unsigned int* __restrict__ next_item) {
  typedef RedCfg<L> Cfg;
  constexpr int ST = Cfg::STRIDE, NS = Cfg::STAGES;
  extern __shared__ double sdyn[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  double* stage = sdyn + (size_t)wib * NS * Cfg::BUF;   // the transpose-reduce reuses the stage buffers
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median), c2 = c_huber * c_huber;
  const char* obs_base = reinterpret_cast<const char*>(P.obs_rec);
  const char* lm_base = reinterpret_cast<const char*>(P.lm_rec);
  const char* lmc_base = reinterpret_cast<const char*>(P.lm_recC);
  for (;;) {
    int it = 0;
    if (lane == 0) it = (int)atomicAdd(next_item, 1u);
    it = __shfl_sync(0xffffffffu, it, 0);
    if (it >= n_items) break;
    const ReduceItem w = items[it];
    const int kind = (int)w.kind;
    const bool diag = (kind == kItemDiagObs || kind == kItemDiagLm);
    const int row_pose = P.opt_to_pose[w.row], col_pose = P.opt_to_pose[w.col];
    double acc[kNAcc];
#pragma unroll
    for (int k = 0; k < kNAcc; ++k) acc[k] = 0.0;
    const int nb = (int)((w.end - w.begin + 31) / 32);
    // list entries of batch b (one per lane): e0 = observation of the row pose (or of the pose, diagonal),
    // e1 = observation of the column pose / tag, e2 = landmark
    auto entry = [&](int b, uint32_t& e0, uint32_t& e1, uint32_t& e2) {
      const uint32_t p = w.begin + (uint32_t)b * 32 + lane;
      e0 = kNoSlot; e1 = kNoSlot; e2 = kNoSlot;
      if (b >= nb || p >= w.end) return;
      if (kind == kItemDiagObs) { e0 = P.pose_obs[p]; e2 = P.pose_obs_lm[p]; }
      else if (kind == kItemDiagLm) { e2 = P.pose_lms[p]; }
      else { e0 = pl.a[p]; e1 = pl.b[p]; e2 = pl.lm[p]; }
    };
    uint32_t n0, n1, n2;
    entry(0, n0, n1, n2);
    for (int bt = -(NS - 1); bt < nb; ++bt) {
      const int ld = bt + NS - 1;
      if (ld < nb) {
        double* dst = stage + (ld % NS) * Cfg::BUF;
        uint32_t e0 = n0, e1 = n1, e2 = n2;
        entry(ld + 1, n0, n1, n2);
        reinterpret_cast<uint2*>(dst + Cfg::DATA)[lane] = make_uint2(e0, (kind == kItemDiagLm) ? e2 : e1);
        if (kind == kItemOff || kind == kItemSym) {
          stage_chunks<2, ST>(dst, 0, obs_base, 32, e0, lane);
          stage_chunks<2, ST>(dst, 4, obs_base, 32, e1, lane);
          if (L == 3) stage_chunks<5, ST>(dst, 8, lm_base, kLmRec * 8, e2, lane);
          else stage_chunks<4, ST>(dst, 8, lm_base, kLmRec * 8, e2, lane);
        } else if (kind == kItemDiagObs) {
          stage_chunks<2, ST>(dst, 0, obs_base, 32, e0, lane);
          stage_chunks<8, ST>(dst, 4, lm_base, kLmRec * 8, e2, lane);
        } else if (L == 1 && kind == kItemOffRef) {
          stage_chunks<2, ST>(dst, 0, obs_base, 32, e0, lane);
          stage_chunks<8, ST>(dst, 4, lm_base, kLmRec * 8, e2, lane);
        } else if (L == 1) {   // kItemDiagLm
          stage_chunks<5, ST>(dst, 0, lmc_base, kLmRecC * 8, e2, lane);
          stage_chunks<8, ST>(dst, 12, lm_base, kLmRec * 8, e2, lane);
        }
      }
      cp_async_commit();
      if (bt < 0) continue;
      cp_async_wait<NS - 1>();
      __syncwarp();
      const double* cur = stage + (bt % NS) * Cfg::BUF;
      const double* d = cur + lane * ST;
      const uint2 ids = reinterpret_cast<const uint2*>(cur + Cfg::DATA)[lane];
      const bool valid = (kind == kItemDiagLm) ? (ids.y != kNoSlot) : (ids.x != kNoSlot);
      if (valid) {
        if (kind == kItemOff || kind == kItemSym) {
          const double* lm = d + 8;
          const int cam_a = P.obs_cam ? (int)P.obs_cam[ids.x] : 0, cam_b = P.obs_cam ? (int)P.obs_cam[ids.y] : 0;
          Side<L> sa, sb;
          side_eval<L>(P, lm, d[0], d[1], d[2], row_pose, cam_a, c_huber, c2, sa);
          side_eval<L>(P, lm, d[4], d[5], d[6], col_pose, cam_b, c_huber, c2, sb);
          if (L == 3) {
            // M = w_a w_b B_a V^-1 B_b^T  (J_l = -B on both sides)
            const double v00 = lm[4], v01 = lm[5], v02 = lm[6], v11 = lm[7], v12 = lm[8], v22 = lm[9];
            double BV[6];
#pragma unroll
            for (int x = 0; x < 2; ++x) {
              const double* b = sa.B + 3 * x;
              BV[3 * x] = b[0] * v00 + b[1] * v01 + b[2] * v02;
              BV[3 * x + 1] = b[0] * v01 + b[1] * v11 + b[2] * v12;
              BV[3 * x + 2] = b[0] * v02 + b[1] * v12 + b[2] * v22;
            }
            const double ww = sa.w * sb.w;
            const double m00 = ww * (BV[0] * sb.B[0] + BV[1] * sb.B[1] + BV[2] * sb.B[2]);
            const double m01 = ww * (BV[0] * sb.B[3] + BV[1] * sb.B[4] + BV[2] * sb.B[5]);
            const double m10 = ww * (BV[3] * sb.B[0] + BV[4] * sb.B[1] + BV[5] * sb.B[2]);
            const double m11 = ww * (BV[3] * sb.B[3] + BV[4] * sb.B[4] + BV[5] * sb.B[5]);
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              const double t0 = m00 * sb.Jm[c] + m01 * sb.Jm[6 + c], t1 = m10 * sb.Jm[c] + m11 * sb.Jm[6 + c];
#pragma unroll
              for (int r = 0; r < 6; ++r) acc[r * 6 + c] -= sa.Jm[r] * t0 + sa.Jm[6 + r] * t1;
            }
          } else {
            // - V^-1 W_a W_b^T,  W = w Jm^T J_l
            const double vi = lm[7];
            double Wa[6], Wb[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) {
              Wa[k] = vi * sa.w * (sa.Jm[k] * sa.jl[0] + sa.Jm[6 + k] * sa.jl[1]);
              Wb[k] = sb.w * (sb.Jm[k] * sb.jl[0] + sb.Jm[6 + k] * sb.jl[1]);
            }
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
              for (int c = 0; c < 6; ++c) acc[r * 6 + c] -= Wa[r] * Wb[c];
          }
        } else if (kind == kItemDiagObs) {
          const double* lm = d + 4;
          const int cam_a = P.obs_cam ? (int)P.obs_cam[ids.x] : 0;
          Side<L> s;
          side_eval<L>(P, lm, d[0], d[1], d[2], row_pose, cam_a, c_huber, c2, s);
          double T0[6], T1[6];
          double q0, q1;   // r - J_l dhat
          if (L == 3) {
            const double v00 = lm[4], v01 = lm[5], v02 = lm[6], v11 = lm[7], v12 = lm[8], v22 = lm[9];
            double BV[6];
#pragma unroll
            for (int x = 0; x < 2; ++x) {
              const double* b = s.B + 3 * x;
              BV[3 * x] = b[0] * v00 + b[1] * v01 + b[2] * v02;
              BV[3 * x + 1] = b[0] * v01 + b[1] * v11 + b[2] * v12;
              BV[3 * x + 2] = b[0] * v02 + b[1] * v12 + b[2] * v22;
            }
            // N = I - w B V^-1 B^T
            const double n00 = 1.0 - s.w * (BV[0] * s.B[0] + BV[1] * s.B[1] + BV[2] * s.B[2]);
            const double n01 = -s.w * (BV[0] * s.B[3] + BV[1] * s.B[4] + BV[2] * s.B[5]);
            const double n10 = -s.w * (BV[3] * s.B[0] + BV[4] * s.B[1] + BV[5] * s.B[2]);
            const double n11 = 1.0 - s.w * (BV[3] * s.B[3] + BV[4] * s.B[4] + BV[5] * s.B[5]);
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              T0[c] = s.w * (n00 * s.Jm[c] + n01 * s.Jm[6 + c]);
              T1[c] = s.w * (n10 * s.Jm[c] + n11 * s.Jm[6 + c]);
            }
            q0 = s.r[0] + (s.B[0] * lm[10] + s.B[1] * lm[11] + s.B[2] * lm[12]);
            q1 = s.r[1] + (s.B[3] * lm[10] + s.B[4] * lm[11] + s.B[5] * lm[12]);
          } else {
            // w Jm^T Jm - V^-1 W W^T = Jm^T (w I - w^2 V^-1 jl jl^T) Jm
            const double vi = lm[7], ww = s.w * s.w * vi;
            const double n00 = s.w - ww * s.jl[0] * s.jl[0], n01 = -ww * s.jl[0] * s.jl[1], n11 = s.w - ww * s.jl[1] * s.jl[1];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              T0[c] = n00 * s.Jm[c] + n01 * s.Jm[6 + c];
              T1[c] = n01 * s.Jm[c] + n11 * s.Jm[6 + c];
            }
            q0 = s.r[0] - s.jl[0] * lm[8];
            q1 = s.r[1] - s.jl[1] * lm[8];
          }
          add_upper(acc, s.Jm, s.Jm + 6, T0, T1);
          const double wr0 = s.w * s.r[0], wr1 = s.w * s.r[1], wq0 = s.w * q0, wq1 = s.w * q1;
#pragma unroll
          for (int k = 0; k < 6; ++k) {
            acc[21 + k] += s.Jm[k] * wr0 + s.Jm[6 + k] * wr1;
            acc[27 + k] += s.Jm[k] * wq0 + s.Jm[6 + k] * wq1;
          }
        } else if (L == 1 && kind == kItemOffRef) {
          // block between the measurement pose and the reference pose of one observation:
          //   X = w Jm^T Jr - V^-1 W_a W_ref^T   (rows: measurement pose, columns: reference pose)
          const double* lm = d + 4;
          const bool meas_row = (ids.y == kPairTagMeasRow);
          const int mp = meas_row ? row_pose : col_pose, rp = meas_row ? col_pose : row_pose;
          const int cam_a = P.obs_cam ? (int)P.obs_cam[ids.x] : 0;
          Side<L> s;
          side_eval<L>(P, lm, d[0], d[1], d[2], mp, cam_a, c_huber, c2, s);
          double M[18];
          ref_M(P, rp, P.n_cams > 1 ? (int)P.lm_ref_cam[(uint32_t)((unsigned long long)__double_as_longlong(d[3]) >> 32)] : 0, lm[3], lm + 9, M);
          double J0[6], J1[6], Wr[6], Wa[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) {
            J0[k] = s.w * (s.B[0] * M[k] + s.B[1] * M[6 + k] + s.B[2] * M[12 + k]);
            J1[k] = s.w * (s.B[3] * M[k] + s.B[4] * M[6 + k] + s.B[5] * M[12 + k]);
            Wr[k] = M[k] * lm[12] + M[6 + k] * lm[13] + M[12 + k] * lm[14];
            Wa[k] = lm[7] * s.w * (s.Jm[k] * s.jl[0] + s.Jm[6 + k] * s.jl[1]);
          }
          if (meas_row) {
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
              for (int c = 0; c < 6; ++c) acc[r * 6 + c] += s.Jm[r] * J0[c] + s.Jm[6 + r] * J1[c] - Wa[r] * Wr[c];
          } else {
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
              for (int c = 0; c < 6; ++c) acc[r * 6 + c] += J0[r] * s.Jm[c] + J1[r] * s.Jm[6 + c] - Wr[r] * Wa[c];
          }
        } else if (L == 1) {   // kItemDiagLm: the reference pose's own blocks of one landmark
          const double* lc = d;         // SBB (6) | SBr (3)
          const double* lm = d + 12;
          double M[18];
          ref_M(P, row_pose, P.n_cams > 1 ? (int)P.lm_ref_cam[ids.y] : 0, lm[3], lm + 9, M);
          const double S3[9] = {lc[0], lc[1], lc[2], lc[1], lc[3], lc[4], lc[2], lc[4], lc[5]};
          double TM[18];
#pragma unroll
          for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int k = 0; k < 6; ++k) TM[6 * j + k] = S3[3 * j] * M[k] + S3[3 * j + 1] * M[6 + k] + S3[3 * j + 2] * M[12 + k];
          double gr[6], Wr[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) {
            gr[k] = M[k] * lc[6] + M[6 + k] * lc[7] + M[12 + k] * lc[8];
            Wr[k] = M[k] * lm[12] + M[6 + k] * lm[13] + M[12 + k] * lm[14];
          }
          const double vi = lm[7], dh = lm[8];
          int k = 0;
#pragma unroll
          for (int c = 0; c < 6; ++c)
#pragma unroll
            for (int r = 0; r <= c; ++r) { acc[k] += (M[r] * TM[c] + M[6 + r] * TM[6 + c] + M[12 + r] * TM[12 + c]) - vi * Wr[r] * Wr[c]; ++k; }
#pragma unroll
          for (int a = 0; a < 6; ++a) { acc[21 + a] += gr[a]; acc[27 + a] += gr[a] - Wr[a] * dh; }
        }
      }
      __syncwarp();
    }
    cp_async_wait<0>();
    __syncwarp();
    // ---- transpose-reduce: lane e takes element e (and 32 + e) over the 32 lanes, in lane order ----
    double (*buf)[33] = reinterpret_cast<double (*)[33]>(stage);
#pragma unroll
    for (int k = 0; k < kNAcc; ++k) buf[k][lane] = acc[k];
    __syncwarp();
    double v0 = 0.0, v1 = 0.0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) v0 += buf[lane][j];
    if (lane < kNAcc - 32) {
#pragma unroll 8
      for (int j = 0; j < 32; ++j) v1 += buf[32 + lane][j];
    }
    __syncwarp();
    // ---- chunks of one block are combined in index order by the last one to arrive --------
    if (w.mcount > 1) {
      double* mine = partial + (size_t)w.mself * kPartial;
      mine[lane] = v0;
      if (lane < kNAcc - 32) mine[32 + lane] = v1;
      __threadfence();
      __syncwarp();
      unsigned int prev = 0;
      if (lane == 0) prev = atomicAdd(&counters[w.mgroup], 1u);
      prev = __shfl_sync(0xffffffffu, prev, 0);
      if (prev + 1 != w.mcount) continue;
      __threadfence();
      if (lane == 0) counters[w.mgroup] = 0;   // ready for the next build
      v0 = 0.0; v1 = 0.0;
      for (uint32_t ch = 0; ch < w.mcount; ++ch) {
        const volatile double* q = partial + (size_t)(w.mfirst + ch) * kPartial;
        v0 += q[lane];
        if (lane < kNAcc - 32) v1 += q[32 + lane];
      }
    }
    // ---- write-out (exclusive owner: plain read-modify-write) ------------------------------
    const SBlock blk = s_block(P, (int)w.row, (int)w.col);
    if (kind == kItemSym) {   // same pose observed twice: X + X^T lands in the (upper) diagonal block
      stage[lane] = v0;
      if (lane < kNAcc - 32) stage[32 + lane] = v1;
      __syncwarp();
    }
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int e = half * 32 + lane;
      if (half == 1 && lane >= kNAcc - 32) break;
      double v = half ? v1 : v0;
      if (diag) {
        if (e < 21) {
          const int c = (e >= 15) ? 5 : (e >= 10) ? 4 : (e >= 6) ? 3 : (e >= 3) ? 2 : (e >= 1) ? 1 : 0;
          const int r = e - c * (c + 1) / 2;
          blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
        } else if (e < 33) {
          double* dst = (e < 27) ? P.b_p : P.rhs_sc;
          dst[(size_t)w.row * P.D + (e < 27 ? e - 21 : e - 27)] += v;
        }
      } else {
        const int r = e / 6, c = e - r * 6;
        if (kind == kItemSym) { v += stage[c * 6 + r]; if (r <= c) blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v; }
        else blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
      }
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------
// launchers
static int sm_count_cached() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

void launch_linearize(const DeviceProblem& P, const uint2* tasks, int n_tasks, const uint32_t* long_lms, int n_long, cudaStream_t st) {
  if (P.n_lms == 0 || P.L == 0) return;
  const int cap = std::min(kRedMaxGrid, sm_count_cached() * 12);
  if (P.L == 1) k_lm_defaults<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P); else k_lm_defaults<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P);
  if (n_tasks > 0) {
    const int grid = std::max(1, std::min(cap, (n_tasks + 3) / 4));
    if (P.L == 1) k_linearize_packed<1><<<grid, 128, 0, st>>>(P, tasks, n_tasks); else k_linearize_packed<3><<<grid, 128, 0, st>>>(P, tasks, n_tasks);
  }
  if (n_long > 0) {
    const int grid = std::max(1, std::min(cap, (n_long + 3) / 4));
    if (P.L == 1) k_linearize_long<1><<<grid, 128, 0, st>>>(P, long_lms, n_long); else k_linearize_long<3><<<grid, 128, 0, st>>>(P, long_lms, n_long);
  }
}

void launch_reduce_blocks(const DeviceProblem& P, const ReduceItem* items, int n_items, const uint32_t* pair_a, const uint32_t* pair_b,
                          const uint32_t* pair_lm, double* partial, unsigned int* counters, unsigned int* next_item, cudaStream_t st) {
  if (n_items <= 0) return;
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(k_reduce_blocks<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RedCfg<1>::smem);
    cudaFuncSetAttribute(k_reduce_blocks<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RedCfg<3>::smem);
    attr_done = true;
  }
  int grid = (n_items + kRedWarps - 1) / kRedWarps;
  const int cap = sm_count_cached() * 3;
  if (grid > cap) grid = cap;
  cudaMemsetAsync(next_item, 0, sizeof(unsigned int), st);
  const PairLists pl = {pair_a, pair_b, pair_lm};
  if (P.L == 1) k_reduce_blocks<1><<<grid, kRedWarps * 32, RedCfg<1>::smem, st>>>(P, items, n_items, pl, partial, counters, next_item);
  else k_reduce_blocks<3><<<grid, kRedWarps * 32, RedCfg<3>::smem, st>>>(P, items, n_items, pl, partial, counters, next_item);
}

}  // namespace bagpu
