// Development aid: FP64 tensor-pipe (mma.sync.m8n8k4.f64) peak of the device, the denominator for the
// "FP64 tensor-pipe utilisation" of the Cholesky (MEASURED_PEAKS.json has no FP64 figure).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/build/dmma_peak tools/dmma_peak.cu && tools/build/dmma_peak
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int CH>
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters) {
  double c0[CH], c1[CH];
#pragma unroll
  for (int i = 0; i < CH; ++i) { c0[i] = 0.0; c1[i] = 0.0; }
  const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CH; ++i) dmma(c0[i], c1[i], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < CH; ++i) s += c0[i] + c1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// plain FP64 FMA peak for comparison
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters) {
  double c[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i] = i;
  const double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = fma(c[i], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    k_dmma<8><<<sms * 8, 256>>>(out, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * 8 * 8 * 4 * 8.0 * iters * (double)(sms * 8 * 256 / 32);
    if (rep) printf("{\"dmma_m8n8k4_tflops\": %.2f, \"ms\": %.3f, \"sms\": %d}\n", flops / ms / 1e9, ms, sms);
    cudaEventRecord(e0);
    k_dfma<<<sms * 8, 256>>>(out, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double f2 = 2.0 * 8.0 * iters * (double)(sms * 8 * 256);
    if (rep) printf("{\"dfma_tflops\": %.2f, \"ms\": %.3f}\n", f2 / ms / 1e9, ms);
  }
  return 0;
}
