// dev_alloc.cu — see dev_alloc.h
#include <map>
#include <mutex>
#include <unordered_map>
#include "dev_alloc.h"

namespace bagpu {
namespace {
std::mutex g_mu;
std::unordered_map<void*, size_t> g_live;          // blocks handed out -> rounded size
std::multimap<size_t, void*> g_free;               // cached blocks by size
size_t g_cached = 0;
constexpr size_t kBudget = (size_t)24 << 30;       // cached (unused) bytes kept at most

size_t round_up(size_t b) {
  if (b == 0) b = 1;
  const size_t g = b >= ((size_t)1 << 20) ? ((size_t)1 << 20) : 512;   // 1 MiB granules for large blocks
  return (b + g - 1) / g * g;
}
void trim_locked() {
  for (auto& kv : g_free) cudaFree(kv.second);
  g_free.clear();
  g_cached = 0;
}
}  // namespace

cudaError_t dev_alloc(void** p, size_t bytes) {
  const size_t sz = round_up(bytes);
  std::lock_guard<std::mutex> g(g_mu);
  auto it = g_free.lower_bound(sz);
  if (it != g_free.end() && it->first <= sz + sz / 4) {
    *p = it->second;
    g_live[*p] = it->first;
    g_cached -= it->first;
    g_free.erase(it);
    return cudaSuccess;
  }
  cudaError_t e = cudaMalloc(p, sz);
  if (e != cudaSuccess) {
    cudaGetLastError();
    cudaDeviceSynchronize();
    trim_locked();
    e = cudaMalloc(p, sz);
    if (e != cudaSuccess) return e;
  }
  g_live[*p] = sz;
  return cudaSuccess;
}

void dev_free(void* p) {
  if (!p) return;
  cudaDeviceSynchronize();   // cudaFree's contract: nothing pending touches the block when it changes hands
  std::lock_guard<std::mutex> g(g_mu);
  auto it = g_live.find(p);
  if (it == g_live.end()) { cudaFree(p); return; }
  const size_t sz = it->second;
  g_live.erase(it);
  g_free.emplace(sz, p);
  g_cached += sz;
  if (g_cached > kBudget) trim_locked();
}

void dev_trim() {
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> g(g_mu);
  trim_locked();
}
}  // namespace bagpu
