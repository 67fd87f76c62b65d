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
#include "proj_kernels.cuh"
#include "launch.h"

namespace bagpu {

BA_D void st2(double* p, double a, double b) { *reinterpret_cast<double2*>(p) = make_double2(a, b); }
BA_D void ld2(const double* p, double& a, double& b) { const double2 v = *reinterpret_cast<const double2*>(p); a = v.x; b = v.y; }

// ---------------------------------------------------------------------------------------
// pass A
template <int L, int G>
__global__ void __launch_bounds__(128) k_linearize(DeviceProblem P) {
  constexpr int NV = L * (L + 1) / 2;
  const int lane = threadIdx.x & 31, gl = lane % G;
  const unsigned gmask = group_mask<G>(lane);
  const int group = (blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int n_groups = (gridDim.x * blockDim.x) / G;
  const double c_huber = P.c_huber_factor * sqrt(P.sc->median);
  double err = 0.0;
  for (int lm = group; lm < P.n_lms; lm += n_groups) {
    const int s = P.lm_obs_start[lm], e = P.lm_obs_start[lm + 1];
    LmCtx<L> c;
    load_lm<L>(P, lm, c);
    const bool lm_on = c.lm_opt >= 0;
    double V[NV], bl[L];
#pragma unroll
    for (int i = 0; i < NV; ++i) V[i] = 0.0;
#pragma unroll
    for (int i = 0; i < L; ++i) bl[i] = 0.0;
    double Hrr[L == 1 ? 21 : 1], gr[L == 1 ? 6 : 1], Wr[L == 1 ? 6 : 1];
    if (L == 1) {
#pragma unroll
      for (int i = 0; i < 21; ++i) Hrr[i] = 0.0;
#pragma unroll
      for (int i = 0; i < 6; ++i) { gr[i] = 0.0; Wr[i] = 0.0; }
    }
    // ---- sweep 1: weights and the landmark sums ---------------------------------------
    for (int i = s + gl; i < e; i += G) {
      ObsJac<L> o;
      obs_jacobians<L>(P, c, i, o);
      const double w0 = P.obs_w0[i];
      const double n2 = o.r[0] * o.r[0] + o.r[1] * o.r[1];
      const double w = huber_weight(n2 * w0, w0, c_huber, P.use_robust);
      P.obs_w[i] = w;
      err += n2 * w;
      if (lm_on) {
        const double wr0 = w * o.r[0], wr1 = w * o.r[1];
        int k = 0;
#pragma unroll
        for (int a = 0; a < L; ++a) {
#pragma unroll
          for (int b = a; b < L; ++b) { V[k] += w * (o.Jl[a] * o.Jl[b] + o.Jl[L + a] * o.Jl[L + b]); ++k; }
          bl[a] += o.Jl[a] * wr0 + o.Jl[L + a] * wr1;
        }
      }
      if (L == 1 && o.ref_on) {
        const double wr0 = w * o.r[0], wr1 = w * o.r[1];
        int k = 0;
#pragma unroll
        for (int b = 0; b < 6; ++b) {
#pragma unroll
          for (int a = 0; a <= b; ++a) { Hrr[k] += w * (o.Jr[a] * o.Jr[b] + o.Jr[6 + a] * o.Jr[6 + b]); ++k; }
          gr[b] += o.Jr[b] * wr0 + o.Jr[6 + b] * wr1;
          if (lm_on) Wr[b] += w * (o.Jr[b] * o.Jl[0] + o.Jr[6 + b] * o.Jl[1]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) V[i] = group_sum<G>(V[i], gmask);
#pragma unroll
    for (int i = 0; i < L; ++i) bl[i] = group_sum<G>(bl[i], gmask);
    // V^-1 (cpp:431-442) and dhat = V^-1 b_l, redundantly in every lane of the group
    double Vi[L * L], dhat[L];
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
#pragma unroll
      for (int i = 0; i < L; ++i) bl[i] = 0.0;
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
      if (rslot != kNoSlot) {
#pragma unroll
        for (int i = 0; i < 21; ++i) Hrr[i] = group_sum<G>(Hrr[i], gmask);
#pragma unroll
        for (int i = 0; i < 6; ++i) { gr[i] = group_sum<G>(gr[i], gmask); Wr[i] = group_sum<G>(Wr[i], gmask); }
        // reference-pose record: H_rr - W_r V^-1 W_r^T (upper) | g_r | g_r - W_r dhat, spread over the lanes
        double* o = P.lm_rec + (size_t)rslot * kLmRec;
        int k = 0;
#pragma unroll
        for (int b = 0; b < 6; ++b)
#pragma unroll
          for (int a = 0; a <= b; ++a) { if ((k % G) == gl) o[k] = Hrr[k] - Vi[0] * Wr[a] * Wr[b]; ++k; }
#pragma unroll
        for (int a = 0; a < 6; ++a)
          if (((21 + a) % G) == gl) { o[21 + a] = gr[a]; o[27 + a] = gr[a] - Wr[a] * dhat[0]; }
      } else {
#pragma unroll
        for (int i = 0; i < 6; ++i) Wr[i] = 0.0;
      }
      if (gl == 0) {
#pragma unroll
        for (int a = 0; a < 6; ++a) P.lm_Wref[(size_t)lm * 6 + a] = Wr[a];
      }
    }
    if (gl == 0) {
#pragma unroll
      for (int i = 0; i < L * L; ++i) P.lm_Vinv[(size_t)lm * L * L + i] = Vi[i];
#pragma unroll
      for (int i = 0; i < L; ++i) { P.lm_bl[(size_t)lm * L + i] = bl[i]; P.lm_dhat[(size_t)lm * L + i] = dhat[i]; }
    }
    // ---- sweep 2: one pose-major record per observation --------------------------------
    for (int i = s + gl; i < e; i += G) {
      const uint32_t slot = P.obs_slot[i];
      if (slot == kNoSlot) continue;
      ObsJac<L> o;
      obs_jacobians<L>(P, c, i, o);
      const double sw = sqrt(P.obs_w[i]);
      double A[12], B[2 * L];
#pragma unroll
      for (int k = 0; k < 12; ++k) A[k] = sw * o.Jm[k];
#pragma unroll
      for (int k = 0; k < 2 * L; ++k) B[k] = sw * o.Jl[k];
      const double r0 = sw * o.r[0], r1 = sw * o.r[1];
      double* out = P.rec + (size_t)slot * P.rec_stride;
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
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          Q[k] = o.ref_on ? sw * o.Jr[k] - c0 * Wr[k] : 0.0;
          Q[6 + k] = o.ref_on ? sw * o.Jr[6 + k] - c1 * Wr[k] : 0.0;
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
  }
  err = warp_sum(err);
  if (lane == 0) red_add(&P.sc->proj_error, err);
}

// ---------------------------------------------------------------------------------------
// pass B: one warp per work item (a chunk of one S block).  Lanes stride over the item's
// records / pairs with a full set of accumulators in registers, then the 32 partial sets are
// transposed through shared memory so that lane k owns element k of the block.
constexpr int kAcc = 36;
constexpr int kRedWarps = 4;

template <int L>
BA_D void diag_record(const double* __restrict__ r, double* acc) {
  // acc[0..20] upper triangle (k = c(c+1)/2 + r), acc[21..26] b_p, acc[27..32] rhs_sc
  double A[12];
  double n00, n01, n10, n11, r0, r1, p0, p1;
  if (L == 3) {
    double C[6], B[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(r + 2 * k, C[2 * k], C[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) ld2(r + 6 + 2 * k, A[2 * k], A[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(r + 18 + 2 * k, B[2 * k], B[2 * k + 1]);
    ld2(r + 24, r0, r1);
    ld2(r + 26, p0, p1);
    n00 = 1.0 - (C[0] * B[0] + C[1] * B[1] + C[2] * B[2]);
    n01 = -(C[0] * B[3] + C[1] * B[4] + C[2] * B[5]);
    n10 = -(C[3] * B[0] + C[4] * B[1] + C[5] * B[2]);
    n11 = 1.0 - (C[3] * B[3] + C[4] * B[4] + C[5] * B[5]);
    int k = 0;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      const double t0 = n00 * A[c] + n01 * A[6 + c], t1 = n10 * A[c] + n11 * A[6 + c];
#pragma unroll
      for (int rr = 0; rr <= c; ++rr) { acc[k] += A[rr] * t0 + A[6 + rr] * t1; ++k; }
    }
  } else {
    double E[6], W[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(r + 2 * k, E[2 * k], E[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(r + 6 + 2 * k, W[2 * k], W[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) ld2(r + 12 + 2 * k, A[2 * k], A[2 * k + 1]);
    ld2(r + 36, r0, r1);
    ld2(r + 38, p0, p1);
    int k = 0;
#pragma unroll
    for (int c = 0; c < 6; ++c)
#pragma unroll
      for (int rr = 0; rr <= c; ++rr) { acc[k] += A[rr] * A[c] + A[6 + rr] * A[6 + c] - E[rr] * W[c]; ++k; }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    acc[21 + k] += A[k] * r0 + A[6 + k] * r1;
    acc[27 + k] += A[k] * p0 + A[6 + k] * p1;
  }
}

// Y = contribution of the pair (a in the row pose, b in the column pose) to the block, el = c*6 + r
template <int L>
BA_D void pair_product(const DeviceProblem& P, uint32_t a, uint32_t b, double* Y) {
  const double* ra = P.rec + (size_t)a * P.rec_stride;
  if (L == 3) {
    const double* rb = P.rec + (size_t)b * kRec3;
    double C[6], B[6], Ab[12], Aa[12];
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(ra + 2 * k, C[2 * k], C[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 3; ++k) ld2(rb + 18 + 2 * k, B[2 * k], B[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) ld2(rb + 6 + 2 * k, Ab[2 * k], Ab[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 6; ++k) ld2(ra + 6 + 2 * k, Aa[2 * k], Aa[2 * k + 1]);
    const double m00 = -(C[0] * B[0] + C[1] * B[1] + C[2] * B[2]);
    const double m01 = -(C[0] * B[3] + C[1] * B[4] + C[2] * B[5]);
    const double m10 = -(C[3] * B[0] + C[4] * B[1] + C[5] * B[2]);
    const double m11 = -(C[3] * B[3] + C[4] * B[4] + C[5] * B[5]);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      const double t0 = m00 * Ab[c] + m01 * Ab[6 + c], t1 = m10 * Ab[c] + m11 * Ab[6 + c];
#pragma unroll
      for (int r = 0; r < 6; ++r) Y[c * 6 + r] += Aa[r] * t0 + Aa[6 + r] * t1;
    }
  } else {
    if (b < kPairTagRefRow) {   // measurement x measurement: -E_a W_b^T
      const double* rb = P.rec + (size_t)b * kRec1;
      double E[6], W[6];
#pragma unroll
      for (int k = 0; k < 3; ++k) ld2(ra + 2 * k, E[2 * k], E[2 * k + 1]);
#pragma unroll
      for (int k = 0; k < 3; ++k) ld2(rb + 6 + 2 * k, W[2 * k], W[2 * k + 1]);
#pragma unroll
      for (int c = 0; c < 6; ++c)
#pragma unroll
        for (int r = 0; r < 6; ++r) Y[c * 6 + r] -= E[r] * W[c];
    } else {                    // measurement x reference pose of the same observation: A^T Q or Q^T A
      double X[12], Z[12];
      const bool meas_row = (b == kPairTagMeasRow);
      const double* x = ra + (meas_row ? 12 : 24);
      const double* z = ra + (meas_row ? 24 : 12);
#pragma unroll
      for (int k = 0; k < 6; ++k) { ld2(x + 2 * k, X[2 * k], X[2 * k + 1]); ld2(z + 2 * k, Z[2 * k], Z[2 * k + 1]); }
#pragma unroll
      for (int c = 0; c < 6; ++c)
#pragma unroll
        for (int r = 0; r < 6; ++r) Y[c * 6 + r] += X[r] * Z[c] + X[6 + r] * Z[6 + c];
    }
  }
}

template <int L>
__global__ void __launch_bounds__(kRedWarps * 32) k_reduce_blocks(DeviceProblem P, const ReduceItem* __restrict__ items, int n_items,
                                                                  const uint2* __restrict__ pairs, double* __restrict__ partial,
                                                                  unsigned int* __restrict__ counters) {
  __shared__ double sbuf[kRedWarps][kAcc][33];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int n_warps = (gridDim.x * blockDim.x) >> 5;
  double (*buf)[33] = sbuf[wib];
  for (int it = warp; it < n_items; it += n_warps) {
    const ReduceItem w = items[it];
    double acc[kAcc];
#pragma unroll
    for (int k = 0; k < kAcc; ++k) acc[k] = 0.0;
    if (w.kind == kItemDiagObs) {
      for (uint32_t s = w.begin + lane; s < w.end; s += 32) diag_record<L>(P.rec + (size_t)s * P.rec_stride, acc);
    } else if (w.kind == kItemDiagLm) {
      for (uint32_t s = w.begin + lane; s < w.end; s += 32) {
        const double* r = P.lm_rec + (size_t)s * kLmRec;
#pragma unroll
        for (int k = 0; k < 17; ++k) { double a, b; ld2(r + 2 * k, a, b); acc[2 * k] += a; if (2 * k + 1 < 33) acc[2 * k + 1] += b; }
      }
    } else if (w.kind == kItemOff) {
      for (uint32_t p = w.begin + lane; p < w.end; p += 32) {
        const uint2 ab = pairs[p];
        pair_product<L>(P, ab.x, ab.y, acc);
      }
    } else {   // same pose observed twice (several cameras): both cross terms land in the diagonal block
      for (uint32_t p = w.begin + lane; p < w.end; p += 32) {
        const uint2 ab = pairs[p];
        double Y[kAcc];
#pragma unroll
        for (int k = 0; k < kAcc; ++k) Y[k] = 0.0;
        pair_product<L>(P, ab.x, ab.y, Y);
#pragma unroll
        for (int c = 0; c < 6; ++c)
#pragma unroll
          for (int r = 0; r < 6; ++r) acc[c * 6 + r] += Y[c * 6 + r] + Y[r * 6 + c];
      }
    }
    // ---- transpose-reduce: lane k ends with element k (lanes 0..3 also element 32 + k) ----
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kAcc; ++k) buf[k][lane] = acc[k];
    __syncwarp();
    double v0 = 0.0, v1 = 0.0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) v0 += buf[lane][j];
    if (lane < kAcc - 32) {
#pragma unroll 8
      for (int j = 0; j < 32; ++j) v1 += buf[32 + lane][j];
    }
    // ---- chunks of one block are combined in index order by the last one to arrive --------
    if (w.mcount > 1) {
      double* mine = partial + (size_t)w.mself * kAcc;
      mine[lane] = v0;
      if (lane < kAcc - 32) mine[32 + lane] = v1;
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
        const volatile double* q = partial + (size_t)(w.mfirst + ch) * kAcc;
        v0 += q[lane];
        if (lane < kAcc - 32) v1 += q[32 + lane];
      }
    }
    // ---- write-out (exclusive owner: plain read-modify-write) ------------------------------
    const bool diag = (w.kind == kItemDiagObs || w.kind == kItemDiagLm);
    const SBlock blk = s_block(P, (int)w.row, (int)w.col);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int el = half * 32 + lane;
      if (half == 1 && lane >= kAcc - 32) break;
      const double v = half ? v1 : v0;
      if (diag) {
        if (el < 21) {
          int c = 0;
          while ((c + 1) * (c + 2) / 2 <= el) ++c;
          const int r = el - c * (c + 1) / 2;
          blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
        } else if (el < 27) {
          P.b_p[(size_t)w.row * P.D + (el - 21)] += v;
        } else if (el < 33) {
          P.rhs_sc[(size_t)w.row * P.D + (el - 27)] += v;
        }
      } else {
        const int c = el / 6, r = el - c * 6;
        if (w.kind == kItemOff || r <= c) blk.p[(size_t)r * blk.rs + (size_t)c * blk.cs] += v;
      }
    }
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

void launch_linearize(const DeviceProblem& P, int G, cudaStream_t st) {
  if (P.n_obs == 0) return;
  size_t blocks = ((size_t)P.n_lms * G + 127) / 128;
  const size_t cap = (size_t)sm_count_cached() * 12;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  const int grid = (int)blocks;
#define CALL(LL, GG) k_linearize<LL, GG><<<grid, 128, 0, st>>>(P)
  if (P.L == 1) {
    if (G == 4) { CALL(1, 4); } else if (G == 8) { CALL(1, 8); } else if (G == 16) { CALL(1, 16); } else { CALL(1, 32); }
  } else {
    if (G == 4) { CALL(3, 4); } else if (G == 8) { CALL(3, 8); } else if (G == 16) { CALL(3, 16); } else { CALL(3, 32); }
  }
#undef CALL
}

void launch_reduce_blocks(const DeviceProblem& P, const ReduceItem* items, int n_items, const uint2* pairs, double* partial,
                          unsigned int* counters, cudaStream_t st) {
  if (n_items <= 0) return;
  int grid = (n_items + kRedWarps - 1) / kRedWarps;
  const int cap = sm_count_cached() * 8;
  if (grid > cap) grid = cap;
  if (P.L == 1) k_reduce_blocks<1><<<grid, kRedWarps * 32, 0, st>>>(P, items, n_items, pairs, partial, counters);
  else k_reduce_blocks<3><<<grid, kRedWarps * 32, 0, st>>>(P, items, n_items, pairs, partial, counters);
}

}  // namespace bagpu
