// dev_alloc.cu — see dev_alloc.h
#include <map>
#include <mutex>
#include <unordered_map>
#include "dev_alloc.h"

namespace bagpu {
namespace {
std::mutex g_mu;
struct Live { size_t size; int dev; };
std::unordered_map<void*, Live> g_live;            // blocks handed out -> rounded size, owning device
std::multimap<std::pair<int, size_t>, void*> g_free;   // cached blocks by (device, size): a block never changes GPUs
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
  int dev = 0;
  cudaGetDevice(&dev);
  auto it = g_free.lower_bound(std::make_pair(dev, sz));
  if (it != g_free.end() && it->first.first == dev && it->first.second <= sz + sz / 4) {
    *p = it->second;
    g_live[*p] = Live{it->first.second, dev};
    g_cached -= it->first.second;
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
  g_live[*p] = Live{sz, dev};
  return cudaSuccess;
}

void dev_free(void* p) {
  if (!p) return;
  cudaDeviceSynchronize();   // cudaFree's contract: nothing pending touches the block when it changes hands
  std::lock_guard<std::mutex> g(g_mu);
  auto it = g_live.find(p);
  if (it == g_live.end()) { cudaFree(p); return; }
  const size_t sz = it->second.size;
  const int dev = it->second.dev;
  g_live.erase(it);
  g_free.emplace(std::make_pair(dev, sz), p);
  g_cached += sz;
  if (g_cached > kBudget) trim_locked();
}

void dev_trim() {
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> g(g_mu);
  trim_locked();
}
}  // namespace bagpu
