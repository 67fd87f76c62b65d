// schur_kernels.cu — atomic-free assembly of the reduced camera system.
//
// Replaces, per outer iteration (reference src/BundleAdjuster.cpp):
//   U += Jpr^T Jpr, rhs_p += Jpr^T r                                    cpp:337-354
//   V = sum Jl^T w Jl, b_l, V^-1                                        cpp:409-443
//   W = Jpr^T Jl, W V^-1, S = U - W V^-1 W^T, rhs_sc = rhs_p - W V^-1 b_l   cpp:448-485
// (the four sparse block products of include/ba/SparseBlockMatrixOps.h:159-380).
//
// The reference scatters: per landmark it adds k(k+1)/2 outer products into S.  On the GPU
// that is ~36 FP64 reductions per product into L2 and is bound by the L2 atomic units
// (profiles/r1a_*: 7 ms for 5M observations).  Here the product is turned around:
//
//   pass A  k_linearize   (landmark-major, one group of G lanes per landmark)
//       Jacobians, Huber weight, V, b_l, V^-1 by warp shuffles, then ONE record per
//       observation written to its POSE-MAJOR slot:  sqrt(w) J_pose, sqrt(w) J_lm V^-1, ...
//       with the landmark already projected out (rho = r~ - B~ V^-1 b_l), see ba_gpu.h.
//   pass B  k_reduce_blocks (pose-pair-major, one warp per S block chunk)
//       S_ii   = sum over the pose's records        A~^T (I - C B~^T) A~      (streaming)
//       S_ij   = sum over the block's pair list    -A~_a^T C_a B~_b^T A~_b    (gather, L2 resident)
//       b_p, rhs_sc from the same records.  Each block is owned by one warp (or a fixed chain
//       of chunks combined in index order), so there are NO atomics and the result is
//       bit-reproducible run to run.
// The pair lists depend only on the ids; they are built once per Solve() on the device
// (symbolic.cu).
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

// From the complete sums of a landmark: V^-1 (cpp:431-442), dhat = V^-1 b_l, LmSize 1 W_r and the reference-pose
// record.  Every lane that holds the sums computes Vi / dhat / Wr (it needs them for its own record); the landmark's
// outputs are written by the lanes with `writer` set, element k of the reference-pose record by the lane with
// k % MOD == rem.
template <int L, int MOD>
BA_D void lm_finalize(const DeviceProblem& P, int lm, const LmCtx<L>& c, bool lm_on, const LmSums<L>& S, double* Vi, double* dhat, double* Wr,
                      bool writer, int rem) {
  constexpr int NV = LmSums<L>::NV;
  const double* V = S.v;
  double bl[L];
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
#pragma unroll
      for (int i = 0; i < L * L; ++i) Vi[i] = Mi[i];
    }
  } else {
#pragma unroll
    for (int i = 0; i < L * L; ++i) Vi[i] = 0.0;
  }
#pragma unroll
  for (int r = 0; r < L; ++r) {
    double v = 0.0;
#pragma unroll
    for (int k = 0; k < L; ++k) v += Vi[r * L + k] * bl[k];
    dhat[r] = v;
  }
  if (L == 1) {
    const uint32_t rslot = P.lm_rslot[lm];
#pragma unroll
    for (int i = 0; i < 6; ++i) Wr[i] = 0.0;
    if (rslot != kNoSlot) {
      const double* SBB = S.v + NV + L; const double* SBr = SBB + 6; const double* SBl = SBr + 3;
      double M[18];   // 3 x 6 row-major, masked parameter columns zeroed
#pragma unroll
      for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int k = 0; k < 3; ++k) { M[6 * j + k] = (j == k) ? -c.rho : 0.0; M[6 * j + 3 + k] = c.Ar[3 * j + k]; }
      if ((c.ref_free & 63u) != 63u) {
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (!((c.ref_free >> k) & 1u)) { M[k] = 0.0; M[6 + k] = 0.0; M[12 + k] = 0.0; }
      }
      const double S3[9] = {SBB[0], SBB[1], SBB[2], SBB[1], SBB[3], SBB[4], SBB[2], SBB[4], SBB[5]};
      double TM[18];  // S3 M
#pragma unroll
      for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int k = 0; k < 6; ++k) TM[6 * j + k] = S3[3 * j] * M[k] + S3[3 * j + 1] * M[6 + k] + S3[3 * j + 2] * M[12 + k];
      double gr[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        gr[k] = M[k] * SBr[0] + M[6 + k] * SBr[1] + M[12 + k] * SBr[2];
        Wr[k] = M[k] * SBl[0] + M[6 + k] * SBl[1] + M[12 + k] * SBl[2];
      }
      // reference-pose record: H_rr - W_r V^-1 W_r^T (upper) | g_r | g_r - W_r dhat
      double* o = P.lm_rec + (size_t)rslot * kLmRec;
      int k = 0;
#pragma unroll
      for (int b = 0; b < 6; ++b)
#pragma unroll
        for (int a = 0; a <= b; ++a) {
          if (writer && (k % MOD) == rem) o[k] = (M[a] * TM[b] + M[6 + a] * TM[6 + b] + M[12 + a] * TM[12 + b]) - Vi[0] * Wr[a] * Wr[b];
          ++k;
        }
#pragma unroll
      for (int a = 0; a < 6; ++a)
        if (writer && ((21 + a) % MOD) == rem) { o[21 + a] = gr[a]; o[27 + a] = gr[a] - Wr[a] * dhat[0]; }
    }
    if (writer && rem == 0) {
#pragma unroll
      for (int a = 0; a < 6; ++a) P.lm_Wref[(size_t)lm * 6 + a] = Wr[a];
    }
  }
  if (writer && rem == 0) {
#pragma unroll
    for (int i = 0; i < L * L; ++i) P.lm_Vinv[(size_t)lm * L * L + i] = Vi[i];
#pragma unroll
    for (int i = 0; i < L; ++i) { P.lm_bl[(size_t)lm * L + i] = bl[i]; P.lm_dhat[(size_t)lm * L + i] = dhat[i]; }
  }
}

// the pose-major record of one observation (ba_gpu.h), written once its landmark's V^-1, dhat (and W_r) are known
template <int L>
BA_D void emit_record(double* out, const LmCtx<L>& c, const ObsJac<L>& o, double w, bool lm_on, const double* Vi, const double* dhat,
                      const double* Wr) {
  const double sw = sqrt(w);
  double A[12], B[2 * L];
#pragma unroll
  for (int k = 0; k < 12; ++k) A[k] = sw * o.Jm[k];
#pragma unroll
  for (int k = 0; k < 2 * L; ++k) B[k] = sw * o.Jl[k];
  const double r0 = sw * o.r[0], r1 = sw * o.r[1];
  if (L == 3) {
    double C[6];
#pragma unroll
    for (int rr = 0; rr < 2; ++rr)
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) C[rr * 3 + cc] = B[rr * 3] * Vi[cc] + B[rr * 3 + 1] * Vi[3 + cc] + B[rr * 3 + 2] * Vi[6 + cc];
    const double p0 = r0 - (B[0] * dhat[0] + B[1] * dhat[1] + B[2] * dhat[2]);
    const double p1 = r1 - (B[3] * dhat[0] + B[4] * dhat[1] + B[5] * dhat[2]);
#pragma unroll
    for (int k = 0; k < 3; ++k) st2(out + 2 * k, C[2 * k], C[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) st2(out + 6 + 2 * k, A[2 * k], A[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 3; ++k) st2(out + 18 + 2 * k, lm_on ? B[2 * k] : 0.0, lm_on ? B[2 * k + 1] : 0.0);
    st2(out + 24, r0, r1);
    st2(out + 26, p0, p1);
  } else {
    const double bv0 = lm_on ? B[0] : 0.0, bv1 = lm_on ? B[1] : 0.0;
    double W[6], Q[12];
#pragma unroll
    for (int k = 0; k < 6; ++k) W[k] = A[k] * bv0 + A[6 + k] * bv1;
    const double c0 = bv0 * Vi[0], c1 = bv1 * Vi[0];
    if (o.ref_on) {
      ref_jacobian<L>(c, o.Bp, Q);
#pragma unroll
      for (int k = 0; k < 6; ++k) { Q[k] = sw * Q[k] - c0 * Wr[k]; Q[6 + k] = sw * Q[6 + k] - c1 * Wr[k]; }
    } else {
#pragma unroll
      for (int k = 0; k < 12; ++k) Q[k] = 0.0;
    }
    const double p0 = r0 - bv0 * dhat[0], p1 = r1 - bv1 * dhat[0];
#pragma unroll
    for (int k = 0; k < 3; ++k) st2(out + 2 * k, W[2 * k] * Vi[0], W[2 * k + 1] * Vi[0]);
#pragma unroll
    for (int k = 0; k < 3; ++k) st2(out + 6 + 2 * k, W[2 * k], W[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) st2(out + 12 + 2 * k, A[2 * k], A[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) st2(out + 24 + 2 * k, Q[2 * k], Q[2 * k + 1]);
    st2(out + 36, r0, r1);
    st2(out + 38, p0, p1);
  }
}

// Landmarks with more than 32 observations (lm_list): one warp per landmark, lanes stride over the track.  Two sweeps:
// weights and sums, then the records (the lane's last observation keeps its Jacobians in registers).
template <int L>
__global__ void __launch_bounds__(128, (L == 1) ? 3 : 4) k_linearize_long(DeviceProblem P, const uint32_t* __restrict__ lm_list, int n_list) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median);
  double err = 0.0;
  for (int q = warp; q < n_list; q += n_warps) {
    const int lm = (int)lm_list[q];
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    const bool lm_on = c.lm_opt >= 0;
    LmSums<L> S;
#pragma unroll
    for (int i = 0; i < LmSums<L>::N; ++i) S.v[i] = 0.0;
    ObsJac<L> o;
    double w_keep = 0.0;
    int i_keep = -1;
    for (int i = s + lane; i < e; i += 32) {
      obs_jacobians<L>(P, c, i, o);
      const double w0 = P.obs_w0[i];
      const double n2 = o.r[0] * o.r[0] + o.r[1] * o.r[1];
      const double w = huber_weight(n2 * w0, w0, c_huber, P.use_robust);
      P.obs_w[i] = w;
      w_keep = w; i_keep = i;
      err += n2 * w;
      sums_add<L>(S, o, w, lm_on);
    }
#pragma unroll
    for (int i = 0; i < LmSums<L>::N; ++i) S.v[i] = warp_sum(S.v[i]);
    double Vi[L * L], dhat[L], Wr[6];
    lm_finalize<L, 32>(P, lm, c, lm_on, S, Vi, dhat, Wr, true, lane);
    for (int i = i_keep; i >= s; i -= 32) {
      const uint32_t slot = P.obs_slot[i];
      if (slot == kNoSlot) continue;
      if (i != i_keep) { obs_jacobians<L>(P, c, i, o); w_keep = P.obs_w[i]; }
      emit_record<L>(P.rec + (size_t)slot * P.rec_stride, c, o, w_keep, lm_on, Vi, dhat, Wr);
    }
  }
  err = warp_sum(err);
  if (lane == 0) red_add(&P.sc->proj_error, err);
}

// Packed tasks: a task is a run of consecutive landmarks with at most 32 observations in total (packed on the host from
// the CSR bounds, symbolic set-up).  One warp per task, ONE OBSERVATION PER LANE: the Jacobians are evaluated once and
// stay in registers, the per-landmark sums are segmented warp reductions over the lanes of the landmark.  (One lane group
// per landmark left 13 of 32 lanes busy on a BAL-like track-length distribution and evaluated long tracks twice.)
// STAGE: the 32 records of a task are assembled in shared memory and written out with consecutive lanes on consecutive
// 16-byte pieces of one record (full 32-byte sectors, a few records per store instruction) instead of every lane
// storing 14-20 separate half sectors of its own record.
template <int L, bool STAGE>
__global__ void __launch_bounds__(128, (L == 1) ? 3 : 4) k_linearize_packed(DeviceProblem P, const uint2* __restrict__ tasks, int n_tasks) {
  constexpr int RS = (L == 3) ? kRec3 : kRec1;
  // stage stride RS + 2 doubles: a lane's 16-byte stores land 240 / 336 bytes apart, which spreads the eight lanes of a
  // store phase over all 32 banks (224 / 320 bytes put them on the same 8 / 16 banks: 2- / 4-way conflicts)
  constexpr int SS = RS + 2;
  __shared__ double2 srec[STAGE ? 4 * 32 * SS / 2 : 1];
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median);
  double err = 0.0;
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
    const int hl = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));          // lane 0 is always a head (n >= 1)
    const unsigned above = (lane == 31) ? 0u : (heads & (0xffffffffu << (lane + 1)));
    const int seg_end = above ? (__ffs(above) - 1) : n;
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    const bool lm_on = c.lm_opt >= 0;
    LmSums<L> S;
#pragma unroll
    for (int k = 0; k < LmSums<L>::N; ++k) S.v[k] = 0.0;
    ObsJac<L> o;
    uint32_t slot = kNoSlot;
    double w = 0.0;
    if (on) {
      slot = P.obs_slot[i];
      obs_jacobians<L>(P, c, i, o, m_pre);
      const double w0 = P.obs_w0[i];
      const double n2 = o.r[0] * o.r[0] + o.r[1] * o.r[1];
      w = huber_weight(n2 * w0, w0, c_huber, P.use_robust);
      P.obs_w[i] = w;
      err += n2 * w;
      sums_add<L>(S, o, w, lm_on);
    }
    // segmented reduction to the head lane (a lane adds its neighbour at distance d only inside its own landmark),
    // then every lane of the landmark reads the totals from its head
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const bool take = on && (lane + d < seg_end);
#pragma unroll
      for (int k = 0; k < LmSums<L>::N; ++k) {
        const double other = __shfl_down_sync(0xffffffffu, S.v[k], d);
        if (take) S.v[k] += other;
      }
    }
#pragma unroll
    for (int k = 0; k < LmSums<L>::N; ++k) S.v[k] = __shfl_sync(0xffffffffu, S.v[k], on ? hl : lane);
    if (on) {
      double Vi[L * L], dhat[L], Wr[6];
      lm_finalize<L, 1>(P, lm, c, lm_on, S, Vi, dhat, Wr, head, 0);
      if (slot != kNoSlot) {
        double* out = STAGE ? reinterpret_cast<double*>(srec) + ((threadIdx.x >> 5) * 32 + lane) * SS : P.rec + (size_t)slot * P.rec_stride;
        emit_record<L>(out, c, o, w, lm_on, Vi, dhat, Wr);
      }
    }
    if (STAGE) {
      __syncwarp();
      const double2* sw2 = srec + (threadIdx.x >> 5) * 32 * (SS / 2);
#pragma unroll 2
      for (int base = 0; base < n * (RS / 2); base += 32) {     // uniform trip count: the shuffle needs every lane
        const int idx = base + lane;
        const bool valid = idx < n * (RS / 2);
        const int rec = valid ? idx / (RS / 2) : 0, ch = idx - rec * (RS / 2);
        const uint32_t sl = __shfl_sync(0xffffffffu, slot, rec);
        if (valid && sl != kNoSlot) reinterpret_cast<double2*>(P.rec + (size_t)sl * RS)[ch] = sw2[rec * (SS / 2) + ch];
      }
      __syncwarp();
    }
  }
  err = warp_sum(err);
  if (lane == 0) red_add(&P.sc->proj_error, err);
}

// landmarks without a (local) observation: the outputs the kernels above would have written for an empty track
template <int L>
__global__ void k_lm_defaults(DeviceProblem P) {
  const int lm = blockIdx.x * blockDim.x + threadIdx.x;
  if (lm >= P.n_lms) return;
  if (P.lm_obs_start[lm] != P.lm_obs_start[lm + 1]) return;
  double Vi[L * L];
#pragma unroll
  for (int i = 0; i < L * L; ++i) Vi[i] = 0.0;
  if (P.lm_opt[lm] >= 0) {
    if (L == 1) {
      Vi[0] = 1.0 / 1e-6;
    } else {
      double M[9] = {1e-6, 0, 0, 0, 1e-6, 0, 0, 0, 1e-6}, Mi[9];
      inv3(M, Mi);
#pragma unroll
      for (int i = 0; i < L * L; ++i) Vi[i] = Mi[i];
    }
  }
#pragma unroll
  for (int i = 0; i < L * L; ++i) P.lm_Vinv[(size_t)lm * L * L + i] = Vi[i];
#pragma unroll
  for (int i = 0; i < L; ++i) { P.lm_bl[(size_t)lm * L + i] = 0.0; P.lm_dhat[(size_t)lm * L + i] = 0.0; }
  if (L == 1) {
#pragma unroll
    for (int a = 0; a < 6; ++a) P.lm_Wref[(size_t)lm * 6 + a] = 0.0;
  }
}

// ---------------------------------------------------------------------------------------
// pass B: one warp per work item (a chunk of one S block), items claimed in order from one counter
// so that the warps in flight always work on neighbouring blocks (their records stay in L2).
// The records of 16 pairs (or 16 consecutive records of a diagonal item) are brought into shared
// memory by cp.async with CHUNK-wise lane mapping — consecutive lanes copy consecutive 16-byte
// pieces of one record, a few cache lines per instruction instead of 32 — in a 3-stage pipeline.
// Two lanes share a pair: each multiplies out three of the six block columns (24 accumulators per
// lane instead of 36+12: registers set the occupancy of this latency-bound gather).  The partial
// sets are transposed through shared memory so that lane k owns element k of the block.
constexpr int kAccL = 24;          // per lane: 6 rows x 3 columns, then 6 entries of b_p (even lane) or rhs_sc (odd lane)
constexpr int kPartial = 64;       // doubles per partial-sum slot (two halves of kAccL, padded)
constexpr int kRedWarps = 4;
constexpr int kPPB = 16;           // pairs (or records) per batch
constexpr int kStages = 3;
template <int L> struct RedCfg {
  static constexpr int PS = (L == 3) ? 18 : 24;                 // doubles per staged pair part
  static constexpr int RS = (L == 3) ? kRec3 : kRec1;           // doubles per record
  static constexpr int DATA = (2 * kPPB * PS > kPPB * RS) ? 2 * kPPB * PS : kPPB * RS;   // >= kPPB * kLmRec
  static constexpr int BUF = DATA + 8;                          // + 16 uint32 tags
  static constexpr size_t smem = sizeof(double) * kStages * BUF * kRedWarps;
};

BA_D void cp_async16(double* smem_dst, const double* gsrc) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
BA_D void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> BA_D void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// diagonal record, lane half h: columns 3h..3h+2 of A^T N A (all six rows) and b_p (h = 0) or rhs_sc (h = 1)
template <int L>
BA_D void diag_record(const double* r, int h, double* acc) {
  double A[12];
#pragma unroll
  for (int k = 0; k < 6; ++k) ld2(r + (L == 3 ? 6 : 12) + 2 * k, A[2 * k], A[2 * k + 1]);
  double q0, q1;
  ld2(r + (L == 3 ? 24 : 36) + 2 * h, q0, q1);     // (r0, r1) or (rho0, rho1)
  if (L == 3) {
    double C[6], B[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) { ld2(r + 2 * k, C[2 * k], C[2 * k + 1]); ld2(r + 18 + 2 * k, B[2 * k], B[2 * k + 1]); }
    const double n00 = 1.0 - (C[0] * B[0] + C[1] * B[1] + C[2] * B[2]);
    const double n01 = -(C[0] * B[3] + C[1] * B[4] + C[2] * B[5]);
    const double n10 = -(C[3] * B[0] + C[4] * B[1] + C[5] * B[2]);
    const double n11 = 1.0 - (C[3] * B[3] + C[4] * B[4] + C[5] * B[5]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const double a0 = h ? A[3 + c] : A[c], a1 = h ? A[9 + c] : A[6 + c];
      const double t0 = n00 * a0 + n01 * a1, t1 = n10 * a0 + n11 * a1;
#pragma unroll
      for (int rr = 0; rr < 6; ++rr) acc[c * 6 + rr] += A[rr] * t0 + A[6 + rr] * t1;
    }
  } else {
    double E[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(r + 2 * k, E[2 * k], E[2 * k + 1]);
    const double* Wh = r + 6 + 3 * h;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const double a0 = h ? A[3 + c] : A[c], a1 = h ? A[9 + c] : A[6 + c], wc = Wh[c];
#pragma unroll
      for (int rr = 0; rr < 6; ++rr) acc[c * 6 + rr] += A[rr] * a0 + A[6 + rr] * a1 - E[rr] * wc;
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) acc[18 + k] += A[k] * q0 + A[6 + k] * q1;
}

// acc += columns 3h..3h+2 of the contribution of one staged pair to its block.  pa: part of the entry in the
// row pose, pb: part of the entry in the column pose.
//   LmSize 3: pa = [C (2x3) | At (2x6)], pb = [At (2x6) | Bt (2x3)]          -At_a^T C_a Bt_b^T At_b
//   LmSize 1: tag < kPairTagRefRow: pa = E (6), pb = W (6)                    -E_a W_b^T
//             else pa = [At (2x6) | Q (2x6)]:  At^T Q (measurement pose is the row) or Q^T At
template <int L>
BA_D void pair_product(const double* pa, const double* pb, uint32_t tag, int h, double* acc) {
  if (L == 3) {
    double C[6], B[6], Aa[12];
#pragma unroll
    for (int k = 0; k < 3; ++k) { ld2(pa + 2 * k, C[2 * k], C[2 * k + 1]); ld2(pb + 12 + 2 * k, B[2 * k], B[2 * k + 1]); }
    const double m00 = -(C[0] * B[0] + C[1] * B[1] + C[2] * B[2]);
    const double m01 = -(C[0] * B[3] + C[1] * B[4] + C[2] * B[5]);
    const double m10 = -(C[3] * B[0] + C[4] * B[1] + C[5] * B[2]);
    const double m11 = -(C[3] * B[3] + C[4] * B[4] + C[5] * B[5]);
#pragma unroll
    for (int k = 0; k < 6; ++k) ld2(pa + 6 + 2 * k, Aa[2 * k], Aa[2 * k + 1]);
    const double* b0 = pb + 3 * h;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const double x0 = b0[c], x1 = b0[6 + c];
      const double t0 = m00 * x0 + m01 * x1, t1 = m10 * x0 + m11 * x1;
#pragma unroll
      for (int r = 0; r < 6; ++r) acc[c * 6 + r] += Aa[r] * t0 + Aa[6 + r] * t1;
    }
  } else {
    if (tag < kPairTagRefRow) {
      double E[6];
#pragma unroll
      for (int k = 0; k < 3; ++k) ld2(pa + 2 * k, E[2 * k], E[2 * k + 1]);
      const double* w = pb + 3 * h;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const double wc = w[c];
#pragma unroll
        for (int r = 0; r < 6; ++r) acc[c * 6 + r] -= E[r] * wc;
      }
    } else {
      double X[12];
      const bool meas_row = (tag == kPairTagMeasRow);
      const double* x = pa + (meas_row ? 0 : 12);
      const double* z = pa + (meas_row ? 12 : 0) + 3 * h;
#pragma unroll
      for (int k = 0; k < 6; ++k) ld2(x + 2 * k, X[2 * k], X[2 * k + 1]);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const double z0 = z[c], z1 = z[6 + c];
#pragma unroll
        for (int r = 0; r < 6; ++r) acc[c * 6 + r] += X[r] * z0 + X[6 + r] * z1;
      }
    }
  }
}

// stage the parts of up to kPPB list entries (batch `bt` of the item) into buf, their tags behind the data.
//   entries between two measurement poses: two parts per entry (row-pose part, column-pose part) of
//     18 doubles (LmSize 3: [C | At] and [At | Bt]) or 6 doubles (LmSize 1: E and W);
//   LmSize 1 measurement/reference entries (kItemOffRef): one part of 24 doubles [At | Q].
template <int L, int KIND> struct PartCfg {
  static constexpr int PC = (KIND == kItemOffRef) ? 12 : (L == 3 ? 9 : 3);     // 16-byte chunks per part
  static constexpr int NPART = (KIND == kItemOffRef) ? kPPB : 2 * kPPB;
  static constexpr int PS = 2 * PC;                                            // doubles per part
};
template <int L, int KIND>
BA_D void stage_pairs(const DeviceProblem& P, const uint2 ab, double* buf, int lane) {
  typedef PartCfg<L, KIND> C;
  if (lane < kPPB) reinterpret_cast<uint32_t*>(buf + RedCfg<L>::DATA)[lane] = (ab.x == kNoSlot) ? kNoSlot - 2 : ab.y;
#pragma unroll
  for (int k = 0; k < (C::NPART * C::PC + 31) / 32; ++k) {
    const int idx = k * 32 + lane;
    const int part = idx / C::PC, chunk = idx - part * C::PC;
    const int pl = ((KIND == kItemOffRef) ? part : (part >> 1)) & (kPPB - 1), side = (KIND == kItemOffRef) ? 0 : (part & 1);
    const uint32_t sa = __shfl_sync(0xffffffffu, ab.x, pl), sb = __shfl_sync(0xffffffffu, ab.y, pl);
    if (sa == kNoSlot || idx >= C::NPART * C::PC) continue;
    const double* src;
    if (KIND == kItemOffRef) src = P.rec + (size_t)sa * kRec1 + 12;
    else src = P.rec + (size_t)(side ? sb : sa) * (L == 3 ? kRec3 : kRec1) + (side ? 6 : 0);
    cp_async16(buf + part * C::PS + chunk * 2, src + chunk * 2);
  }
}

// stage up to kPPB consecutive records (stride RS doubles) starting at slot s0 (s0 may be past the end)
template <int RS>
BA_D void stage_records(const double* __restrict__ base, uint32_t s0, uint32_t end, double* buf, int lane) {
  if (s0 >= end) return;
  const int cnt = (int)min((uint32_t)kPPB, end - s0);
  const double* src = base + (size_t)s0 * RS;
#pragma unroll
  for (int k = 0; k < (kPPB * RS / 2 + 31) / 32; ++k) {
    const int idx = k * 32 + lane;
    if (idx < cnt * (RS / 2)) cp_async16(buf + idx * 2, src + idx * 2);
  }
}

template <int L>
__global__ void __launch_bounds__(kRedWarps * 32, 4) k_reduce_blocks(DeviceProblem P, const ReduceItem* __restrict__ items, int n_items,
                                                                     const uint2* __restrict__ pairs, double* __restrict__ partial,
                                                                     unsigned int* __restrict__ counters, unsigned int* __restrict__ next_item) {
  typedef RedCfg<L> Cfg;
  extern __shared__ double sdyn[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int pl = lane >> 1, h = lane & 1;
  double* stage = sdyn + (size_t)wib * kStages * Cfg::BUF;   // the transpose-reduce reuses the stage buffers
  for (;;) {
    int it = 0;
    if (lane == 0) it = (int)atomicAdd(next_item, 1u);
    it = __shfl_sync(0xffffffffu, it, 0);
    if (it >= n_items) break;
    const ReduceItem w = items[it];
    double acc[kAccL];
#pragma unroll
    for (int k = 0; k < kAccL; ++k) acc[k] = 0.0;
    const int nb = (int)((w.end - w.begin + kPPB - 1) / kPPB);
    const bool diag = (w.kind == kItemDiagObs || w.kind == kItemDiagLm);
    const bool obs = (w.kind == kItemDiagObs);
    // ---- software pipeline: batch bt + kStages - 1 is in flight while batch bt is multiplied out ----
    // the list entries of a batch are requested one batch ahead of the record gather that needs them
    auto entry = [&](int b) {
      const uint32_t p = w.begin + (uint32_t)b * kPPB + (lane & (kPPB - 1));
      return (!diag && b < nb && p < w.end) ? pairs[p] : make_uint2(kNoSlot, kNoSlot);
    };
    uint2 ab_next = entry(0);
    for (int bt = -(kStages - 1); bt < nb; ++bt) {
      const int ld = bt + kStages - 1;
      if (ld < nb) {
        double* dst = stage + (ld % kStages) * Cfg::BUF;
        const uint2 ab = ab_next;
        ab_next = entry(ld + 1);
        if (w.kind == kItemOffRef) { if (L == 1) stage_pairs<1, kItemOffRef>(P, ab, dst, lane); }
        else if (!diag) stage_pairs<L, kItemOff>(P, ab, dst, lane);
        else if (obs) stage_records<Cfg::RS>(P.rec, w.begin + (uint32_t)ld * kPPB, w.end, dst, lane);
        else stage_records<kLmRec>(P.lm_rec, w.begin + (uint32_t)ld * kPPB, w.end, dst, lane);
      }
      cp_async_commit();
      if (bt < 0) continue;
      cp_async_wait<kStages - 1>();
      __syncwarp();
      const double* cur = stage + (bt % kStages) * Cfg::BUF;
      if (!diag) {
        const uint32_t tag = reinterpret_cast<const uint32_t*>(cur + Cfg::DATA)[pl];
        if (tag != kNoSlot - 2) {
          if (w.kind == kItemOffRef) pair_product<L>(cur + pl * 24, cur, tag, h, acc);
          else pair_product<L>(cur + (2 * pl) * PartCfg<L, kItemOff>::PS, cur + (2 * pl + 1) * PartCfg<L, kItemOff>::PS, tag, h, acc);
        }
      } else if (w.begin + (uint32_t)bt * kPPB + pl < w.end) {
        if (obs) {
          diag_record<L>(cur + pl * Cfg::RS, h, acc);
        } else {   // reference-pose records (LmSize 1): upper triangle (21) | g_r (6) | g_r~ (6), plain sums
          const double* r = cur + pl * kLmRec;
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int rr = 0; rr < 6; ++rr) { const int cg = 3 * h + c; if (rr <= cg) acc[c * 6 + rr] += r[cg * (cg + 1) / 2 + rr]; }
#pragma unroll
          for (int k = 0; k < 6; ++k) acc[18 + k] += r[21 + 6 * h + k];
        }
      }
      __syncwarp();
    }
    cp_async_wait<0>();
    // ---- transpose-reduce over the 16 lanes of each half: element e = h * 24 + k, e < 48 ----
    double (*buf)[33] = reinterpret_cast<double (*)[33]>(stage);
#pragma unroll
    for (int k = 0; k < kAccL; ++k) buf[k][lane] = acc[k];
    __syncwarp();
    double v0 = 0.0, v1 = 0.0;
    {
      const int e = lane, eh = e / kAccL, ek = e - eh * kAccL;
#pragma unroll 8
      for (int j = 0; j < 16; ++j) v0 += buf[ek][2 * j + eh];
    }
    if (lane < 2 * kAccL - 32) {
      const int e = 32 + lane, eh = e / kAccL, ek = e - eh * kAccL;
#pragma unroll 8
      for (int j = 0; j < 16; ++j) v1 += buf[ek][2 * j + eh];
    }
    __syncwarp();
    // ---- chunks of one block are combined in index order by the last one to arrive --------
    if (w.mcount > 1) {
      double* mine = partial + (size_t)w.mself * kPartial;
      mine[lane] = v0;
      if (lane < 2 * kAccL - 32) mine[32 + lane] = v1;
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
        if (lane < 2 * kAccL - 32) v1 += q[32 + lane];
      }
    }
    // ---- write-out (exclusive owner: plain read-modify-write) ------------------------------
    const SBlock blk = s_block(P, (int)w.row, (int)w.col);
    if (w.kind == kItemSym) {   // same pose observed twice: X + X^T lands in the (upper) diagonal block
      double* sy = stage;
      sy[lane] = v0;
      if (lane < 2 * kAccL - 32) sy[32 + lane] = v1;
      __syncwarp();
    }
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int e = half * 32 + lane;
      if (half == 1 && lane >= 2 * kAccL - 32) break;
      double v = half ? v1 : v0;
      const int eh = e / kAccL, ek = e - eh * kAccL;
      if (ek < 18) {
        const int c = 3 * eh + ek / 6, r = ek % 6;
        if (w.kind == kItemSym) v += stage[(r / 3) * kAccL + (r % 3) * 6 + c];   // element (c, r)
        if (w.kind == kItemOff || w.kind == kItemOffRef || r <= c) blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
      } else if (diag) {
        double* dst = eh ? P.rhs_sc : P.b_p;
        dst[(size_t)w.row * P.D + (ek - 18)] += v;
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
  const int cap = sm_count_cached() * 12;
  if (P.L == 1) k_lm_defaults<1><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P); else k_lm_defaults<3><<<(P.n_lms + 255) / 256, 256, 0, st>>>(P);
  if (n_tasks > 0) {
    const int grid = std::max(1, std::min(cap, (n_tasks + 3) / 4));
    static const bool stage = getenv("BAGPU_LIN_STAGE") ? atoi(getenv("BAGPU_LIN_STAGE")) != 0 : true;
    if (P.L == 1) { if (stage) k_linearize_packed<1, true><<<grid, 128, 0, st>>>(P, tasks, n_tasks); else k_linearize_packed<1, false><<<grid, 128, 0, st>>>(P, tasks, n_tasks); }
    else { if (stage) k_linearize_packed<3, true><<<grid, 128, 0, st>>>(P, tasks, n_tasks); else k_linearize_packed<3, false><<<grid, 128, 0, st>>>(P, tasks, n_tasks); }
  }
  if (n_long > 0) {
    const int grid = std::max(1, std::min(cap, (n_long + 3) / 4));
    if (P.L == 1) k_linearize_long<1><<<grid, 128, 0, st>>>(P, long_lms, n_long); else k_linearize_long<3><<<grid, 128, 0, st>>>(P, long_lms, n_long);
  }
}

void launch_reduce_blocks(const DeviceProblem& P, const ReduceItem* items, int n_items, const uint2* pairs, double* partial,
                          unsigned int* counters, unsigned int* next_item, cudaStream_t st) {
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
  if (P.L == 1) k_reduce_blocks<1><<<grid, kRedWarps * 32, RedCfg<1>::smem, st>>>(P, items, n_items, pairs, partial, counters, next_item);
  else k_reduce_blocks<3><<<grid, kRedWarps * 32, RedCfg<3>::smem, st>>>(P, items, n_items, pairs, partial, counters, next_item);
}

}  // namespace bagpu
