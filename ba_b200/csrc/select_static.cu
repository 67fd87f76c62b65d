// select_static.cu — bagpu_select_kth_static: the exact k-th order statistic of select.cu for callers that cannot allocate.
//
// No counterpart in the reference: added for callers that may not call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The handle-based path (capi.cu,
// run_select) keeps its radix-select state — the merged 4096-bin histogram, the key prefix and the remaining rank — in a
// cudaMalloc'ed SelectState; this variant runs the same kernels (k_select_init / k_select_hist / k_select_scan, six 12-bit
// passes, no host round trip) with that state in a file-scope __device__ array.
//
//   * The array (kSelectStaticBytes = 16640 bytes) is part of the module image: every GPU that loads the library holds it
//     until the process exits.
//   * In host code `g_select_scratch` is a host-side stand-in for the symbol, NOT a device address; the address the kernels
//     get comes from cudaGetSymbolAddress, resolved once per device with that device current (scratch_for_current_device).
//   * Every call shares the array and it keeps the last call's contents: k_select_init clears the bins and the scalars on
//     the caller's stream before the first pass.  Calls that overlap on the GPU (two streams, two threads) race.
//   * The bins are 32-bit counters, hence the stated limit of BAGPU_SELECT_STATIC_MAX_N values.
//   * No cudaMalloc / cudaFree / allocating library call on this path, and no synchronisation: the result is in *d_out once
//     the work queued on `stream` has run.  The first call on a device loads the module (cudaGetSymbolAddress) — make it
//     before a graph capture starts.
#include <cstdio>
#include <mutex>
#include <string>
#include "../../include/ba_b200.h"
#include "ba_gpu.h"
#include "launch.h"

namespace bagpu { void set_last_error(const std::string& msg); }

namespace {

constexpr size_t kSelectStaticBytes = 16640;   // 65 x 256
static_assert(sizeof(bagpu::SelectState) <= kSelectStaticBytes, "SelectState must fit the static scratch");

static __device__ __align__(256) unsigned char g_select_scratch[kSelectStaticBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
bagpu::SelectState* g_state[kMaxDevices] = {};
int g_sms[kMaxDevices] = {};

int fail(int code, const std::string& msg) { bagpu::set_last_error(msg); return code; }

// device address of the scratch on the CURRENT device (resolved once per device) and its SM count
int scratch_for_current_device(bagpu::SelectState** st, int* sms) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(BAGPU_ERR_NO_DEVICE, std::string("bagpu_select_kth_static: cudaGetDevice: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail(BAGPU_ERR_UNSUPPORTED, "bagpu_select_kth_static: device ordinal beyond 63");
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_state[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_select_scratch);
    if (e != cudaSuccess || !p) return fail(BAGPU_ERR_CUDA, std::string("bagpu_select_kth_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess || n < 1) return fail(BAGPU_ERR_CUDA, std::string("bagpu_select_kth_static: cudaDeviceGetAttribute: ") + cudaGetErrorString(e));
    g_sms[dev] = n;
    g_state[dev] = static_cast<bagpu::SelectState*>(p);
  }
  *st = g_state[dev];
  *sms = g_sms[dev];
  return BAGPU_OK;
}

}  // namespace

extern "C" int bagpu_select_kth_static(const double* d_x, uint64_t n, uint64_t k, double* d_out, void* stream) {
  if (!d_x || !d_out) return fail(BAGPU_ERR_INVALID, "bagpu_select_kth_static: null device pointer");
  if (n == 0) return fail(BAGPU_ERR_INVALID, "bagpu_select_kth_static: n = 0");
  if (n > (uint64_t)BAGPU_SELECT_STATIC_MAX_N) {
    char msg[192];
    snprintf(msg, sizeof(msg), "bagpu_select_kth_static: n = %llu exceeds BAGPU_SELECT_STATIC_MAX_N = %llu (32-bit histogram bins of the static scratch)",
             (unsigned long long)n, (unsigned long long)BAGPU_SELECT_STATIC_MAX_N);
    return fail(BAGPU_ERR_INVALID, msg);
  }
  if (k >= n) return fail(BAGPU_ERR_INVALID, "bagpu_select_kth_static: k must be below n");
  bagpu::SelectState* st = nullptr;
  int sms = 0;
  const int rc = scratch_for_current_device(&st, &sms);
  if (rc != BAGPU_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  bagpu::launch_select_init(st, (unsigned long long)k, s);          // clears what the last call left behind
  const int passes = bagpu::select_num_passes();
  for (int pass = 0; pass < passes; ++pass) {
    bagpu::launch_select_hist(d_x, (size_t)n, st, pass, sms, s);
    bagpu::launch_select_scan(st, pass, d_out, s);                  // the last pass writes the value
  }
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(BAGPU_ERR_CUDA, std::string("bagpu_select_kth_static: launch: ") + cudaGetErrorString(e));
  return BAGPU_OK;
}
