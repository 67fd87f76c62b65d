// device_math.cuh — small fixed-size FP64 helpers for the sm_100a BundleAdjuster kernels.
//
// Conventions match the reference's third-party math: Hamilton quaternions
// in Eigen coefficient order (x,y,z,w); rotation matrices row-major R[3*r+c];
// SO3 exp/log with Sophus' small-angle branches; decoupled SE3 update R <- R exp(w),
// t <- t + dt (reference include/ba/Utils.h:363-369).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bagpu {

#define BA_HD __host__ __device__ __forceinline__
#define BA_D __device__ __forceinline__

BA_HD void quat_to_R(const double* q, double* R) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

BA_HD void quat_mul(const double* a, const double* b, double* o) {  // o = a (x) b, Hamilton
  const double ox = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  const double oy = a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2];
  const double oz = a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0];
  const double ow = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
  o[0] = ox; o[1] = oy; o[2] = oz; o[3] = ow;
}

BA_HD void quat_normalize(double* q) {
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}

// Sophus SO3::exp (quaternion), epsilon 1e-10
BA_HD void so3_exp(const double* w, double* q) {
  const double theta = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  const double half = 0.5 * theta;
  double imag;
  const double real = cos(half);
  if (theta < 1e-10) {
    const double t2 = theta * theta;
    imag = 0.5 - 0.0208333 * t2 + 0.000260417 * t2 * t2;
  } else {
    imag = sin(half) / theta;
  }
  q[0] = imag * w[0]; q[1] = imag * w[1]; q[2] = imag * w[2]; q[3] = real;
  quat_normalize(q);
}

// Sophus SO3::log: (2 atan(|v|/w) / |v|) v with the small-angle branch
BA_HD void so3_log(const double* q, double* w) {
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
  const double n = sqrt(n2);
  const double qw = q[3];
  double f;
  if (n < 1e-10) {
    f = 2.0 / qw - 2.0 * n2 / (qw * qw * qw);
  } else if (fabs(qw) < 1e-10) {
    f = (qw > 0 ? 3.14159265358979323846 : -3.14159265358979323846) / n;
  } else {
    f = 2.0 * atan(n / qw) / n;
  }
  w[0] = f * q[0]; w[1] = f * q[1]; w[2] = f * q[2];
}

BA_HD void mat3_mul(const double* A, const double* B, double* C) {
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}
BA_HD void mat3_mulT(const double* A, const double* B, double* C) {  // C = A^T B
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) C[3 * r + c] = A[r] * B[c] + A[3 + r] * B[3 + c] + A[6 + r] * B[6 + c];
}
BA_HD void mat3_vec(const double* A, const double* v, double* o) {
  const double a = A[0] * v[0] + A[1] * v[1] + A[2] * v[2];
  const double b = A[3] * v[0] + A[4] * v[1] + A[5] * v[2];
  const double c = A[6] * v[0] + A[7] * v[1] + A[8] * v[2];
  o[0] = a; o[1] = b; o[2] = c;
}
BA_HD void mat3T_vec(const double* A, const double* v, double* o) {
  const double a = A[0] * v[0] + A[3] * v[1] + A[6] * v[2];
  const double b = A[1] * v[0] + A[4] * v[1] + A[7] * v[2];
  const double c = A[2] * v[0] + A[5] * v[1] + A[8] * v[2];
  o[0] = a; o[1] = b; o[2] = c;
}
// C = A [v]x   (3x3 times skew)
BA_HD void mat3_skew(const double* A, const double* v, double* C) {
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    C[3 * r + 0] = A[3 * r + 1] * v[2] - A[3 * r + 2] * v[1];
    C[3 * r + 1] = -A[3 * r + 0] * v[2] + A[3 * r + 2] * v[0];
    C[3 * r + 2] = A[3 * r + 0] * v[1] - A[3 * r + 1] * v[0];
  }
}

// Inverse left/right Jacobian of SO3 at rotation vector phi:
//   log(exp(eps) exp(phi)) ~ phi + Jl_inv(phi) eps ;  log(exp(phi) exp(eps)) ~ phi + Jr_inv(phi) eps
//   Jl_inv = I - 1/2 [phi]x + c [phi]x^2,  Jr_inv = I + 1/2 [phi]x + c [phi]x^2,
//   c = 1/theta^2 - (1 + cos theta) / (2 theta sin theta)
BA_HD void so3_jinv(const double* phi, double sign_half, double* J) {
  const double x = phi[0], y = phi[1], z = phi[2];
  const double t2 = x * x + y * y + z * z;
  double c;
  if (t2 < 1e-8) {
    c = 1.0 / 12.0 + t2 / 720.0;
  } else {
    const double t = sqrt(t2);
    c = 1.0 / t2 - (1.0 + cos(t)) / (2.0 * t * sin(t));
  }
  // [phi]x^2 = phi phi^T - t2 I
  const double h = sign_half;  // -0.5 for left, +0.5 for right
  J[0] = 1.0 + c * (x * x - t2); J[1] = -h * z + c * x * y;     J[2] = h * y + c * x * z;
  J[3] = h * z + c * x * y;      J[4] = 1.0 + c * (y * y - t2); J[5] = -h * x + c * y * z;
  J[6] = -h * y + c * x * z;     J[7] = h * x + c * y * z;      J[8] = 1.0 + c * (z * z - t2);
}

// closed-form inverse of a symmetric 3x3 given as full row-major (Eigen cofactor form)
BA_HD void inv3(const double* m, double* o) {
  const double c00 = m[4] * m[8] - m[5] * m[7];
  const double c01 = m[2] * m[7] - m[1] * m[8];
  const double c02 = m[1] * m[5] - m[2] * m[4];
  const double c10 = m[5] * m[6] - m[3] * m[8];
  const double c11 = m[0] * m[8] - m[2] * m[6];
  const double c12 = m[2] * m[3] - m[0] * m[5];
  const double c20 = m[3] * m[7] - m[4] * m[6];
  const double c21 = m[1] * m[6] - m[0] * m[7];
  const double c22 = m[0] * m[4] - m[1] * m[3];
  const double inv_det = 1.0 / (m[0] * c00 + m[1] * c10 + m[2] * c20);
  o[0] = c00 * inv_det; o[1] = c01 * inv_det; o[2] = c02 * inv_det;
  o[3] = c10 * inv_det; o[4] = c11 * inv_det; o[5] = c12 * inv_det;
  o[6] = c20 * inv_det; o[7] = c21 * inv_det; o[8] = c22 * inv_det;
}

BA_D double shfl_xor_d(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
BA_D double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// sum over the G lanes of a group; gmask = the lanes of that group (groups may diverge)
template <int G>
BA_D double group_sum(double v, unsigned gmask) {
#pragma unroll
  for (int m = G / 2; m > 0; m >>= 1) v += __shfl_xor_sync(gmask, v, m);
  return v;
}
BA_D double warp_sum(double v) { return group_sum<32>(v, 0xffffffffu); }
template <int G>
BA_D unsigned group_mask(int lane) { return G == 32 ? 0xffffffffu : (((1u << G) - 1u) << ((lane / G) * G)); }

BA_D void red_add(double* p, double v) { atomicAdd(p, v); }  // REDG.E.ADD.F64 (no return value used)

// Bit-reproducible grid-wide sums of NV per-thread values (the reference is a serial, deterministic program: every scalar
// the step control consumes must not depend on the arrival order of atomics).  Warp shuffles, then the warps of the CTA
// in index order, then the CTAs in index order by the last CTA to arrive.  `partial` holds gridDim.x * NV doubles,
// `counter` is zero before the launch and is left zero.  `out[k]` is overwritten, or added to when `accumulate`.
// The result depends only on the launch geometry, which every launcher derives from the problem size.
constexpr int kRedMaxGrid = 148 * 16;
template <int NV>
BA_D void grid_sum_fixed(const double (&v)[NV], double* __restrict__ partial, unsigned int* __restrict__ counter, double* __restrict__ out,
                         bool accumulate) {
  __shared__ double sh_w[32 * NV];
  __shared__ bool sh_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const double s = warp_sum(v[k]);
    if (lane == 0) sh_w[warp * NV + k] = s;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    double s = 0.0;
    for (int w = 0; w < nw; ++w) s += sh_w[w * NV + threadIdx.x];
    partial[(size_t)blockIdx.x * NV + threadIdx.x] = s;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) sh_last = (atomicAdd(counter, 1u) + 1u == gridDim.x);
  __syncthreads();
  if (!sh_last) return;
  __threadfence();
  // the last CTA: thread t takes CTAs t, t + blockDim.x, ... in order, then the threads are combined in index order
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double s = 0.0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x) s += __ldcg(&partial[(size_t)b * NV + k]);
    s = warp_sum(s);
    __syncthreads();
    if (lane == 0) sh_w[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < nw; ++w) t += sh_w[w];
      out[k] = accumulate ? out[k] + t : t;
    }
  }
  if (threadIdx.x == 0) *counter = 0;
}

}  // namespace bagpu
