// schur_kernels.cu — atomic-free, Jacobian-free assembly of the reduced camera system.
//
// Replaces, per outer iteration (reference src/BundleAdjuster.cpp):
//   U += Jpr^T Jpr, rhs_p += Jpr^T r                                    cpp:337-354
//   V = sum Jl^T w Jl, b_l, V^-1                                        cpp:409-443
//   W = Jpr^T Jl, W V^-1, S = U - W V^-1 W^T, rhs_sc = rhs_p - W V^-1 b_l   cpp:448-485
// (the four sparse block products of include/ba/SparseBlockMatrixOps.h:159-380).
//
// The reference scatters: per landmark it adds k(k+1)/2 outer products into S.  On the GPU that is ~36 FP64
// reductions per product into L2, bound by the L2 atomic units (profiles/r1a_*: 7 ms for 5M observations).  Round 1
// turned the product around with a 224/320-byte record per observation (1.8x - 3.7x the algorithmic DRAM bytes); the
// first round-2 version stored nothing and re-evaluated both Jacobians of every pair (profiles/r2a_*: instruction- and
// latency-bound, 26 warp instructions per list entry).  This version uses the factorisation of ba_gpu.h: with
// B = d project / d X_world every projection Jacobian is B K_pose (or -B, -B u for the landmark), so
//
//   pass A  k_linearize   (landmark-major, one observation per lane)
//       residual, Huber weight, G = w B^T B (64-byte record) and h = w B^T r (32-byte record) per observation;
//       V = sum G (LmSize 1: sum u^T G u), b_l, V^-1 by segmented warp shuffles; ONE 128-byte record per landmark.
//   pass B  k_reduce_blocks (pose-pair-major, one warp per S block chunk, one list entry per lane)
//       S_ij += sum_e K~_i^T M_e K~_j,  K~_p = [rho I | -[X - rho t_p]x],  M_e = -G_a V^-1 G_b (off-diagonal, LmSize 3),
//       G_a - G_a V^-1 G_a (diagonal), -V^-1 g_a g_b^T (LmSize 1), ...: no camera model, no division, no weights.
//       The pose's rotation and parameter mask (D_i^T . D_j) are applied once per block at write-out.  Each block is
//       owned by one warp (or a fixed chain of chunks combined in index order): NO atomics, bit-reproducible.
// The pair lists depend only on the ids; they are built once per Solve() on the device (symbolic.cu).
#include <algorithm>
#include <cstdlib>
#include "proj_kernels.cuh"
#include "launch.h"

namespace bagpu {

// 256-bit store (sm_100: STG.E.ENL2.256): one whole 32-byte sector per lane and instruction
BA_D void st4(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

// ---------------------------------------------------------------------------------------
// pass A
// Running sums of one landmark over its observations.
//   LmSize 3: V = sum G (6) | b_l = -sum h (3)
//   LmSize 1: V = sum w j.j | b_l = -sum w j.r | Gbar = sum G (6) | hbar = sum h (3) | gbar = sum w B^T j (3)
//             (the last three over the observations whose reference pose takes part: from them follow the reference
//             pose's own blocks, W_ref = K_ref^T gbar, U_ref = K_ref^T Gbar K_ref, b_ref = -K_ref^T hbar)
template <int L> struct LmSums {
  static constexpr int NV = L * (L + 1) / 2;
  static constexpr int N = NV + L + (L == 1 ? 12 : 0);
  double v[N];
};

template <int L>
BA_D void sums_add(LmSums<L>& S, const ObsLin<L>& o, bool lm_on) {
  if constexpr (L == 3) {
    if (lm_on) {
#pragma unroll
      for (int k = 0; k < 6; ++k) S.v[k] += o.G[k];
#pragma unroll
      for (int k = 0; k < 3; ++k) S.v[6 + k] -= o.h[k];
    }
  } else {
    if (lm_on) { S.v[0] += o.w * (o.j[0] * o.j[0] + o.j[1] * o.j[1]); S.v[1] -= o.wjr; }
    if (o.ref_on) {
#pragma unroll
      for (int k = 0; k < 6; ++k) S.v[2 + k] += o.G[k];
#pragma unroll
      for (int k = 0; k < 3; ++k) S.v[8 + k] += o.h[k];
      if (lm_on) {
#pragma unroll
        for (int k = 0; k < 3; ++k) S.v[11 + k] += o.gj[k];
      }
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

// the landmark's record (ba_gpu.h), written by one lane: four 32-byte stores into one 128-byte line
template <int L>
BA_D double lm_write(const DeviceProblem& P, int lm, const LmCtx<L>& c, bool lm_on, const LmSums<L>& S) {
  constexpr int NV = LmSums<L>::NV;
  double Vi[NV], bl[L], dhat[L];
  lm_invert<L>(S, lm_on, Vi, bl, dhat);
  double* o = P.lm_rec + (size_t)lm * kLmRec;
  st4(o, c.X[0], c.X[1], c.X[2], c.rho);
  if constexpr (L == 3) {
    st4(o + 4, Vi[0], Vi[1], Vi[2], Vi[3]); st4(o + 8, Vi[4], Vi[5], dhat[0], dhat[1]); st4(o + 12, dhat[2], 0.0, 0.0, 0.0);
  } else {
    const double* Gb = S.v + 2; const double* hb = S.v + 8; const double* gb = S.v + 11;
    st4(o + 4, c.tws_r[0], c.tws_r[1], c.tws_r[2], Vi[0]);
    st4(o + 8, dhat[0], gb[0], gb[1], gb[2]);
    st4(o + 12, 0.0, 0.0, 0.0, 0.0);
    double* oc = P.lm_recC + (size_t)lm * kLmRecC;
    st4(oc, Gb[0], Gb[1], Gb[2], Gb[3]); st4(oc + 4, Gb[4], Gb[5], hb[0], hb[1]); st4(oc + 8, hb[2], 0.0, 0.0, 0.0);
  }
#pragma unroll
  for (int i = 0; i < L; ++i) P.lm_bl[(size_t)lm * L + i] = bl[i];
  return Vi[0];   // LmSize 1: V^-1, which the gv records of the landmark's observations carry as well
}

// one observation of pass A: linearise, write weight and records, add into the landmark's sums
template <int L>
BA_D void linearize_one(const DeviceProblem& P, const LmCtx<L>& c, int i, int m, bool lm_on, double c_huber, double c2, LmSums<L>& S,
                        double& err, double* gu) {
  ObsLin<L> o;
  obs_linearize<L>(P, c, i, m, c_huber, c2, o);
  P.obs_w[i] = o.w;
  err += o.n2 * o.w;
  // G for every observation (the steepest-descent denominator and the back-substitution read it as well); h only where the
  // pose-major lists will ask for it
  double* g = P.obs_G + (size_t)i * kObsG;
  st4(g, o.G[0], o.G[1], o.G[2], o.G[3]); st4(g + 4, o.G[4], o.G[5], 0.0, 0.0);   // whole sectors
  if (o.pm_on) st4(P.obs_h + (size_t)i * kObsH, o.h[0], o.h[1], o.h[2], 0.0);
  sums_add<L>(S, o, lm_on);
  if constexpr (L == 1) { gu[0] = o.gj[0]; gu[1] = o.gj[1]; gu[2] = o.gj[2]; }
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
      double gu[3];
      linearize_one<L>(P, c, i, (int)P.obs_pose[i], lm_on, c_huber, c2, S, err[0], gu);
      if constexpr (L == 1) st4(P.obs_gv + (size_t)i * kObsGv, gu[0], gu[1], gu[2], 0.0);   // V^-1 follows below
    }
#pragma unroll
    for (int i = 0; i < LmSums<L>::N; ++i) S.v[i] = warp_sum(S.v[i]);
    double vi = 0.0;
    if (lane == 0) vi = lm_write<L>(P, lm, c, lm_on, S);
    if constexpr (L == 1) {
      vi = __shfl_sync(0xffffffffu, vi, 0);
      for (int i = s + lane; i < e; i += 32) P.obs_gv[(size_t)i * kObsGv + 3] = vi;   // the lane that wrote the record
    }
  }
  grid_sum_fixed<1>(err, P.red_partial, P.red_counter, &P.sc->proj_error, true);
}

// Packed tasks: a task is a run of consecutive landmarks with at most 32 observations in total (packed on the host from
// the CSR bounds, symbolic set-up).  One warp per task, ONE OBSERVATION PER LANE: the observation is linearised once,
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
    double gu[3] = {0.0, 0.0, 0.0};
    if (on) linearize_one<L>(P, c, i, m_pre, lm_on, c_huber, c2, S, err[0], gu);
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
    double vi = 0.0;
    if (head) vi = lm_write<L>(P, lm, c, lm_on, S);
    if constexpr (L == 1) {
      // V^-1 from the head lane of my landmark, then the observation's gv record
      const int hl = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
      vi = __shfl_sync(0xffffffffu, vi, on ? hl : 0);
      if (on) st4(P.obs_gv + (size_t)i * kObsGv, gu[0], gu[1], gu[2], vi);
    }
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
// in flight work on neighbouring blocks (their records stay in L2).  One list entry per lane.  The records of an entry
// (two 64-byte G records and 64-80 bytes of the landmark record off the diagonal; G, h and the landmark record on
// it) are brought into shared memory by cp.async with CHUNK-wise lane mapping — consecutive lanes copy consecutive
// 16-byte pieces of one record, so an instruction touches 6-11 cache lines instead of 32 (L1 serves one line per
// cycle) — double buffered; the lane then reads its own entry back (stride of an odd number of 16-byte words:
// conflict-free).  Accumulators: the 36 elements of the block (or 21 + 6 + 6 for a diagonal block with b_p and rhs_sc)
// per lane, transposed through shared memory at the end of the item so that lane e owns element e.
constexpr int kPartial = 64;       // doubles per partial-sum slot
constexpr int kRedWarps = 4;
constexpr int kNAcc = 36;
template <int L> struct RedCfg {
  static constexpr int STRIDE = (L == 3) ? 26 : 22;   // doubles per staged entry
  static constexpr int STAGES = 2;
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

// stage the first NC 16-byte chunks of the record `base + id * rec_bytes` of every lane's entry (id = kNoSlot: none)
// at doubles offset `off` of the entry's stage slot
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

BA_D void sym3_full(const double* G, double* F) {
  F[0] = G[0]; F[1] = G[1]; F[2] = G[2]; F[3] = G[1]; F[4] = G[3]; F[5] = G[4]; F[6] = G[2]; F[7] = G[4]; F[8] = G[5];
}
// Y = X - rho t
BA_D void y_of(const double* X, double rho, const double* t, double* Y) { Y[0] = X[0] - rho * t[0]; Y[1] = X[1] - rho * t[1]; Y[2] = X[2] - rho * t[2]; }
// diagonal accumulators: 21 (block, packed upper) | 6 (b_p) | 6 (rhs_sc)
BA_D void acc_diag(double* acc, double rho, const double* Y, const double* M, const double* v1, const double* v2) {
  acc_kmk_sym(acc, rho, Y, M);
  double p[6];
  kt_vec(rho, Y, v1, p);
#pragma unroll
  for (int k = 0; k < 6; ++k) acc[21 + k] += p[k];
  kt_vec(rho, Y, v2, p);
#pragma unroll
  for (int k = 0; k < 6; ++k) acc[27 + k] += p[k];
}

template <int L>
__global__ void __launch_bounds__(kRedWarps * 32, 4) k_reduce_blocks(DeviceProblem P, const ReduceItem* __restrict__ items, int n_items, PairLists pl,
                                                                     double* __restrict__ partial, unsigned int* __restrict__ counters,
                                                                     unsigned int* __restrict__ next_item) {
  typedef RedCfg<L> Cfg;
  constexpr int ST = Cfg::STRIDE, NS = Cfg::STAGES;
  extern __shared__ double sdyn[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  double* stage = sdyn + (size_t)wib * NS * Cfg::BUF;   // the transpose-reduce and the write-out reuse the stage buffers
  const char* g_base = reinterpret_cast<const char*>(P.obs_G);
  const char* h_base = reinterpret_cast<const char*>(P.obs_h);
  const char* lm_base = reinterpret_cast<const char*>(P.lm_rec);
  const char* lmc_base = reinterpret_cast<const char*>(P.lm_recC);
  const char* gv_base = reinterpret_cast<const char*>(P.obs_gv);
  constexpr int kGB = kObsG * 8, kHB = kObsH * 8, kLB = kLmRec * 8, kCB = kLmRecC * 8, kVB = kObsGv * 8;
  for (;;) {
    int it = 0;
    if (lane == 0) it = (int)atomicAdd(next_item, 1u);
    it = __shfl_sync(0xffffffffu, it, 0);
    if (it >= n_items) break;
    const ReduceItem w = items[it];
    const int kind = (int)w.kind;
    const bool diag = (kind == kItemDiagObs || kind == kItemDiagLm);
    const int row_pose = P.opt_to_pose[w.row], col_pose = P.opt_to_pose[w.col];
    // positions of the two poses, once per item (the cp.async fences would make the compiler reload them every batch)
    const double ti[3] = {P.pose_t[3 * (size_t)row_pose], P.pose_t[3 * (size_t)row_pose + 1], P.pose_t[3 * (size_t)row_pose + 2]};
    const double tj[3] = {P.pose_t[3 * (size_t)col_pose], P.pose_t[3 * (size_t)col_pose + 1], P.pose_t[3 * (size_t)col_pose + 2]};
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
        if (L == 1 && (kind == kItemOff || kind == kItemSym)) {
          // three sectors per entry: gv of both observations, X | rho of the landmark
          stage_chunks<2, ST>(dst, 0, gv_base, kVB, e0, lane);
          stage_chunks<2, ST>(dst, 4, gv_base, kVB, e1, lane);
          stage_chunks<2, ST>(dst, 8, lm_base, kLB, e2, lane);
        } else if (kind == kItemOff || kind == kItemSym) {
          stage_chunks<3, ST>(dst, 0, g_base, kGB, e0, lane);
          stage_chunks<3, ST>(dst, 6, g_base, kGB, e1, lane);
          stage_chunks<5, ST>(dst, 12, lm_base, kLB, e2, lane);
        } else if (kind == kItemDiagObs) {
          stage_chunks<3, ST>(dst, 0, g_base, kGB, e0, lane);
          stage_chunks<2, ST>(dst, 6, h_base, kHB, e0, lane);
          if (L == 3) stage_chunks<7, ST>(dst, 10, lm_base, kLB, e2, lane);
          else stage_chunks<5, ST>(dst, 10, lm_base, kLB, e2, lane);
        } else if (L == 1 && kind == kItemOffRef) {
          stage_chunks<3, ST>(dst, 0, g_base, kGB, e0, lane);
          stage_chunks<6, ST>(dst, 6, lm_base, kLB, e2, lane);
        } else if (L == 1) {   // kItemDiagLm
          stage_chunks<5, ST>(dst, 0, lmc_base, kCB, e2, lane);
          stage_chunks<6, ST>(dst, 10, lm_base, kLB, e2, lane);
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
          const double* lm = d + (L == 3 ? 12 : 8);
          const double rho = lm[3];
          double Ya[3], Yb[3];
          y_of(lm, rho, ti, Ya); y_of(lm, rho, tj, Yb);
          if (L == 3) {
            // M = -G_a V^-1 G_b
            double Ga[9], Vi[9], Gb[9], T[9], M[9];
            sym3_full(d, Ga); sym3_full(lm + 4, Vi); sym3_full(d + 6, Gb);
            mat3_mul(Ga, Vi, T);
            mat3_mul(T, Gb, M);
#pragma unroll
            for (int k = 0; k < 9; ++k) M[k] = -M[k];
            acc_kmk(acc, rho, Ya, Yb, M);
          } else {
            // -V^-1 (K~_a^T g_a)(K~_b^T g_b)^T,  g = G u (u = baseline reference sensor -> measuring sensor) and V^-1
            // straight from the gv records
            double pa[6], pb[6];
            kt_vec(rho, Ya, d, pa); kt_vec(rho, Yb, d + 4, pb);
            const double vi = d[3];
#pragma unroll
            for (int r = 0; r < 6; ++r) {
              const double s = -vi * pa[r];
#pragma unroll
              for (int c = 0; c < 6; ++c) acc[r * 6 + c] += s * pb[c];
            }
          }
        } else if (kind == kItemDiagObs) {
          const double* h = d + 6;
          const double* lm = d + 10;
          const double rho = lm[3];
          double Y[3], M[9], v2[3];
          y_of(lm, rho, ti, Y);
          if (L == 3) {
            // M = G - G V^-1 G (exactly symmetric: upper computed, mirrored); v2 = h + G dhat
            double Ga[9], Vi[9], T[9];
            sym3_full(d, Ga); sym3_full(lm + 4, Vi);
            mat3_mul(Ga, Vi, T);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int c = r; c < 3; ++c) {
                const double m = Ga[3 * r + c] - (T[3 * r] * Ga[c] + T[3 * r + 1] * Ga[3 + c] + T[3 * r + 2] * Ga[6 + c]);
                M[3 * r + c] = m; M[3 * c + r] = m;
              }
            double gd[3];
            sym3_vec(d, lm + 10, gd);
            v2[0] = h[0] + gd[0]; v2[1] = h[1] + gd[1]; v2[2] = h[2] + gd[2];
          } else {
            // M = G - V^-1 g g^T; v2 = h + g dhat
            const int cam_a = P.obs_cam ? (int)P.obs_cam[ids.x] : 0;
            const double* twa = P.Tws + ((size_t)row_pose * P.n_cams + cam_a) * 12 + 9;
            const double u[3] = {lm[4] - twa[0], lm[5] - twa[1], lm[6] - twa[2]};
            double g[3];
            sym3_vec(d, u, g);
            const double vi = lm[7], dh = lm[8];
            sym3_full(d, M);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int c = r; c < 3; ++c) { const double m = M[3 * r + c] - vi * g[r] * g[c]; M[3 * r + c] = m; M[3 * c + r] = m; }
            v2[0] = h[0] + g[0] * dh; v2[1] = h[1] + g[1] * dh; v2[2] = h[2] + g[2] * dh;
          }
          acc_diag(acc, rho, Y, M, h, v2);
        } else if (L == 1 && kind == kItemOffRef) {
          // block between the measurement pose and the reference pose of one observation:
          //   K_m^T (-G_a + V^-1 g_a gbar^T) K_r   (rows: measurement pose, columns: reference pose), or its transpose
          const double* lm = d + 6;
          const double rho = lm[3];
          const bool meas_row = (ids.y == kPairTagMeasRow);
          const int mp = meas_row ? row_pose : col_pose;
          const int cam_a = P.obs_cam ? (int)P.obs_cam[ids.x] : 0;
          const double* twa = P.Tws + ((size_t)mp * P.n_cams + cam_a) * 12 + 9;
          const double u[3] = {lm[4] - twa[0], lm[5] - twa[1], lm[6] - twa[2]};
          double g[3], Gf[9], M[9], Ya[3], Yb[3];
          sym3_vec(d, u, g);
          sym3_full(d, Gf);
          const double vi = lm[7];
          const double* gbar = lm + 9;
          y_of(lm, rho, ti, Ya); y_of(lm, rho, tj, Yb);
          if (meas_row) {
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int c = 0; c < 3; ++c) M[3 * r + c] = vi * g[r] * gbar[c] - Gf[3 * r + c];
          } else {
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int c = 0; c < 3; ++c) M[3 * r + c] = vi * gbar[r] * g[c] - Gf[3 * r + c];
          }
          acc_kmk(acc, rho, Ya, Yb, M);
        } else if (L == 1) {   // kItemDiagLm: the reference pose's own blocks of one landmark
          const double* lc = d;         // Gbar (6) | hbar (3)
          const double* lm = d + 10;
          const double rho = lm[3];
          const double vi = lm[7], dh = lm[8];
          const double* gbar = lm + 9;
          double Y[3], M[9], v1[3], v2[3];
          y_of(lm, rho, ti, Y);
          sym3_full(lc, M);
#pragma unroll
          for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = r; c < 3; ++c) { const double m = M[3 * r + c] - vi * gbar[r] * gbar[c]; M[3 * r + c] = m; M[3 * c + r] = m; }
#pragma unroll
          for (int k = 0; k < 3; ++k) { v1[k] = -lc[6 + k]; v2[k] = -(lc[6 + k] + gbar[k] * dh); }
          acc_diag(acc, rho, Y, M, v1, v2);
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
    // ---- the poses' rotation and parameter masks: out = D_i^T A D_j (vectors: D_i^T v), once per block ----
    double* tA = stage;            // 36 (or 21 + 12 packed)
    double* tT = stage + 64;       // A D_j, 6 x 6 row-major
    double* tO = stage + 128;      // out, 6 x 6 row-major (kItemSym needs the transpose)
    tA[lane] = v0;
    if (lane < kNAcc - 32) tA[32 + lane] = v1;
    __syncwarp();
    const double* Ri = P.Rwp + 9 * (size_t)row_pose;
    const double* Rj = P.Rwp + 9 * (size_t)col_pose;
    const uint32_t fi = P.pose_free[row_pose], fj = P.pose_free[col_pose];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int e = half * 32 + lane;
      if (e < 36) {
        const int r = e / 6, c = e - r * 6;
        auto A = [&](int p, int q) { return diag ? tA[p <= q ? q * (q + 1) / 2 + p : p * (p + 1) / 2 + q] : tA[p * 6 + q]; };
        double t = 0.0;
        if ((fj >> c) & 1u) {
          if (c < 3) t = A(r, c);
          else t = A(r, 3) * Rj[c - 3] + A(r, 4) * Rj[3 + c - 3] + A(r, 5) * Rj[6 + c - 3];
        }
        tT[e] = t;
      }
    }
    __syncwarp();
    double o0 = 0.0, o1 = 0.0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int e = half * 32 + lane;
      if (e < 36) {
        const int r = e / 6, c = e - r * 6;
        double t = 0.0;
        if ((fi >> r) & 1u) {
          if (r < 3) t = tT[r * 6 + c];
          else t = Ri[r - 3] * tT[18 + c] + Ri[3 + r - 3] * tT[24 + c] + Ri[6 + r - 3] * tT[30 + c];
        }
        if (half) o1 = t; else o0 = t;
        tO[e] = t;
      }
    }
    double bv = 0.0;   // diagonal kinds: lanes 0..11 transform the two vectors (b_p: 0..5, rhs_sc: 6..11)
    if (diag && lane < 12) {
      const int r = lane % 6;
      const double* v = tA + 21 + (lane / 6) * 6;
      if ((fi >> r) & 1u) bv = (r < 3) ? v[r] : Ri[r - 3] * v[3] + Ri[3 + r - 3] * v[4] + Ri[6 + r - 3] * v[5];
    }
    __syncwarp();
    // ---- write-out (exclusive owner: plain read-modify-write) ------------------------------
    const SBlock blk = s_block(P, (int)w.row, (int)w.col);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int e = half * 32 + lane;
      if (e >= 36) break;
      const int r = e / 6, c = e - r * 6;
      double v = half ? o1 : o0;
      if (diag) {
        if (r <= c) blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
      } else if (kind == kItemSym) {   // same pose observed twice: X + X^T lands in the (upper) diagonal block
        v += tO[c * 6 + r];
        if (r <= c) blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
      } else {
        blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
      }
    }
    if (diag && lane < 12) {
      double* dst = (lane < 6) ? P.b_p : P.rhs_sc;
      dst[(size_t)w.row * P.D + (lane % 6)] += bv;
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
  // one wave of resident CTAs (launch.h, resident_ctas_per_sm): the warps stride over the tasks
  auto cap_of = [&](const void* k) { return std::min(kRedMaxGrid, sm_count_cached() * resident_ctas_per_sm(k, 128)); };
  if (P.L == 1) k_lm_defaults<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P); else k_lm_defaults<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P);
  if (n_tasks > 0) {
    const int cap = cap_of(P.L == 1 ? reinterpret_cast<const void*>(k_linearize_packed<1>) : reinterpret_cast<const void*>(k_linearize_packed<3>));
    const int grid = std::max(1, std::min(cap, (n_tasks + 3) / 4));
    if (P.L == 1) k_linearize_packed<1><<<grid, 128, 0, st>>>(P, tasks, n_tasks); else k_linearize_packed<3><<<grid, 128, 0, st>>>(P, tasks, n_tasks);
  }
  if (n_long > 0) {
    const int cap = cap_of(P.L == 1 ? reinterpret_cast<const void*>(k_linearize_long<1>) : reinterpret_cast<const void*>(k_linearize_long<3>));
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
  const int cap = sm_count_cached() * 4;
  if (grid > cap) grid = cap;
  cudaMemsetAsync(next_item, 0, sizeof(unsigned int), st);
  const PairLists pl = {pair_a, pair_b, pair_lm};
  if (P.L == 1) k_reduce_blocks<1><<<grid, kRedWarps * 32, RedCfg<1>::smem, st>>>(P, items, n_items, pl, partial, counters, next_item);
  else k_reduce_blocks<3><<<grid, kRedWarps * 32, RedCfg<3>::smem, st>>>(P, items, n_items, pl, partial, counters, next_item);
}

}  // namespace bagpu
