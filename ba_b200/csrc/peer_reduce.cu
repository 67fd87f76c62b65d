// peer_reduce.cu — see peer_reduce.h.
#include "peer_reduce.h"
#include <cstring>

namespace {

__device__ __forceinline__ unsigned char* slot_ptr(unsigned char* win, int parity, int r) {
  return win + kPeerFlagBytes + ((size_t)parity * kPeerMaxWorld + r) * kPeerSlotBytes;
}
__device__ __forceinline__ unsigned long long* flag_ptr(unsigned char* win, int parity, int r) {
  return reinterpret_cast<unsigned long long*>(win) + parity * kPeerMaxWorld + r;
}

// Two slot sets, used alternately: a rank can enter reduction s+2 only after finishing s+1, which needs every peer's
// flag for s+1, which a peer stores only after its kernel for s has finished reading the slots of s (stream order).
template <typename T>
__global__ void __launch_bounds__(1024) k_peer_allreduce(PeerWindow w, T* buf, int n, unsigned long long seq, int* err) {
  const int tid = threadIdx.x, parity = (int)(seq & 1ull);
  if (*reinterpret_cast<volatile int*>(err)) return;
  for (int p = 0; p < w.world; ++p) {
    T* dst = reinterpret_cast<T*>(slot_ptr(w.win[p], parity, w.rank));
    for (int i = tid; i < n; i += blockDim.x) dst[i] = buf[i];
  }
  __syncthreads();
  if (tid < w.world) {
    __threadfence_system();
    unsigned long long* f = flag_ptr(w.win[tid], parity, w.rank);
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(seq) : "memory");
    const unsigned long long* mine = flag_ptr(w.win[w.rank], parity, tid);
    const long long t0 = clock64();
    unsigned long long v;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
      if (v < seq && clock64() - t0 > 120000000000ll) { *err = 1; break; }  // ~60 s: a peer is gone
    } while (v < seq);
  }
  __syncthreads();
  unsigned char* local = w.win[w.rank];
  for (int i = tid; i < n; i += blockDim.x) {
    T s = 0;
    for (int q = 0; q < w.world; ++q) s += reinterpret_cast<const volatile T*>(slot_ptr(local, parity, q))[i];
    buf[i] = s;
  }
}

// all-gather of a few u64 items per rank (the candidates of the exact median): same flag protocol as k_peer_allreduce; a slot
// holds [count | items]
__global__ void __launch_bounds__(1024) k_peer_gather(PeerWindow w, const unsigned long long* items, const unsigned int* count,
                                                      unsigned long long* all, unsigned int* all_n, unsigned long long seq, int* err) {
  const int tid = threadIdx.x, parity = (int)(seq & 1ull);
  if (*reinterpret_cast<volatile int*>(err)) return;
  unsigned int n = *count;
  if (n > (unsigned int)kPeerGatherCap) n = (unsigned int)kPeerGatherCap;
  for (int p = 0; p < w.world; ++p) {
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(slot_ptr(w.win[p], parity, w.rank));
    if (tid == 0) dst[0] = n;
    for (unsigned int i = tid; i < n; i += blockDim.x) dst[1 + i] = items[i];
  }
  __syncthreads();
  if (tid < w.world) {
    __threadfence_system();
    unsigned long long* f = flag_ptr(w.win[tid], parity, w.rank);
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(seq) : "memory");
    const unsigned long long* mine = flag_ptr(w.win[w.rank], parity, tid);
    const long long t0 = clock64();
    unsigned long long v;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
      if (v < seq && clock64() - t0 > 120000000000ll) { *err = 1; break; }
    } while (v < seq);
  }
  __syncthreads();
  unsigned char* local = w.win[w.rank];
  unsigned int off = 0;
  for (int q = 0; q < w.world; ++q) {
    const volatile unsigned long long* src = reinterpret_cast<const volatile unsigned long long*>(slot_ptr(local, parity, q));
    const unsigned int nq = (unsigned int)src[0];
    for (unsigned int i = tid; i < nq; i += blockDim.x) all[off + i] = src[1 + i];
    off += nq;
  }
  if (tid == 0) *all_n = off;
}

struct BulkView { double* base[kPeerMaxWorld]; int rank, world; };

// sum of one 16-byte piece over the ranks, in rank order, delivered to every rank.  All N loads are issued before the
// first add (a remote load is ~2 us: eight dependent ones would serialise), then N stores.
__device__ __forceinline__ void bulk_piece(const BulkView& v, size_t off2) {
  double2 x[kPeerMaxWorld];
#pragma unroll
  for (int q = 0; q < kPeerMaxWorld; ++q) {
    x[q] = make_double2(0.0, 0.0);
    if (q < v.world) {
      const double2* src = reinterpret_cast<const double2*>(v.base[q]) + off2;
      asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(x[q].x), "=d"(x[q].y) : "l"(src));
    }
  }
  double2 s = make_double2(0.0, 0.0);
#pragma unroll
  for (int q = 0; q < kPeerMaxWorld; ++q) if (q < v.world) { s.x += x[q].x; s.y += x[q].y; }
#pragma unroll
  for (int q = 0; q < kPeerMaxWorld; ++q) {
    if (q < v.world) {
      double2* dst = reinterpret_cast<double2*>(v.base[q]) + off2;
      asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1, %2};" ::"l"(dst), "d"(s.x), "d"(s.y) : "memory");
    }
  }
}

constexpr int kBulkTailChunk = 2048;   // doubles
__global__ void __launch_bounds__(256) k_peer_bulk_allreduce(BulkView v, const int* __restrict__ slots, int n_slots, int tile_elems,
                                                             size_t tail_off, size_t tail_n) {
  const int n_tail_chunks = (int)((tail_n + kBulkTailChunk - 1) / kBulkTailChunk);
  const int n_units = n_slots + n_tail_chunks;
  // unit u belongs to rank u % world; this rank's units are dealt to its CTAs
  for (int u = v.rank + blockIdx.x * v.world; u < n_units; u += gridDim.x * v.world) {
    size_t off, n;
    if (u < n_slots) { off = (size_t)slots[u] * tile_elems; n = (size_t)tile_elems; }
    else { const size_t c = (size_t)(u - n_slots) * kBulkTailChunk; off = tail_off + c; n = tail_n - c < (size_t)kBulkTailChunk ? tail_n - c : (size_t)kBulkTailChunk; }
    // offsets and lengths are even (tiles of T x T doubles with even T, an even tail offset): 16-byte pieces
    for (size_t i = threadIdx.x; i < n / 2; i += blockDim.x) bulk_piece(v, off / 2 + i);   // independent pieces: the loads of consecutive iterations overlap
  }
  __threadfence_system();
}

}  // namespace

cudaError_t PeerBulk::create_local(int rank_, int world_, size_t bytes, void* handle64) {
  rank = rank_; world = world_;
  for (auto& p : base) p = nullptr;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) return e;
  e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) return e;
  base[rank] = static_cast<double*>(p);
  cap_bytes = bytes;
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) return e;
  memcpy(handle64, &h, 64);
  return cudaDeviceSynchronize();
}
bool PeerBulk::open_peers(const void* handles) {
  bool all = true;
  for (int p = 0; p < world; ++p) {
    if (p == rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(handles) + (size_t)p * 64, 64);
    void* q = nullptr;
    if (cudaIpcOpenMemHandle(&q, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); all = false; q = nullptr; }
    base[p] = static_cast<double*>(q);
  }
  return all;
}
void PeerBulk::close_peers() {
  for (int p = 0; p < world; ++p) {
    if (!base[p] || p == rank) continue;
    cudaIpcCloseMemHandle(base[p]);
    base[p] = nullptr;
  }
  ok = false;
}
void PeerBulk::destroy() {
  close_peers();
  if (world > 0 && base[rank]) { cudaFree(base[rank]); base[rank] = nullptr; }
  cap_bytes = 0;
}
void PeerBulk::allreduce(const int* slots, int n_slots, int tile_elems, size_t tail_off, size_t tail_n, int sms, cudaStream_t st) {
  BulkView v;
  for (int p = 0; p < kPeerMaxWorld; ++p) v.base[p] = base[p];
  v.rank = rank; v.world = world;
  const int n_units = n_slots + (int)((tail_n + kBulkTailChunk - 1) / kBulkTailChunk);
  int grid = (n_units + world - 1) / world;
  if (grid > sms * 8) grid = sms * 8;
  if (grid < 1) grid = 1;
  k_peer_bulk_allreduce<<<grid, 256, 0, st>>>(v, slots, n_slots, tile_elems, tail_off, tail_n);
}

cudaError_t PeerReduce::create_local(int rank, int world, void* handle64) {
  w.rank = rank; w.world = world;
  for (auto& p : w.win) p = nullptr;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, kPeerWindowBytes + 256);
  if (e != cudaSuccess) return e;
  e = cudaMemset(p, 0, kPeerWindowBytes + 256);
  if (e != cudaSuccess) return e;
  w.win[rank] = static_cast<unsigned char*>(p);
  d_err = reinterpret_cast<int*>(w.win[rank] + kPeerWindowBytes);
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) return e;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  memcpy(handle64, &h, 64);
  return cudaDeviceSynchronize();
}

bool PeerReduce::open_peers(const void* handles) {
  bool all = true;
  for (int p = 0; p < w.world; ++p) {
    if (p == w.rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(handles) + (size_t)p * 64, 64);
    void* q = nullptr;
    if (cudaIpcOpenMemHandle(&q, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); all = false; q = nullptr; }
    w.win[p] = static_cast<unsigned char*>(q);
  }
  return all;
}

void PeerReduce::close_peers() {
  for (int p = 0; p < w.world; ++p) {
    if (!w.win[p] || p == w.rank) continue;
    cudaIpcCloseMemHandle(w.win[p]);
    w.win[p] = nullptr;
  }
  ok = false;
}

void PeerReduce::destroy() {
  close_peers();
  if (w.world > 0 && w.win[w.rank]) { cudaFree(w.win[w.rank]); w.win[w.rank] = nullptr; }
  d_err = nullptr;
}

void PeerReduce::allreduce_f64(double* buf, int n, cudaStream_t st) {
  k_peer_allreduce<double><<<1, n >= 1024 ? 1024 : (n > 32 ? ((n + 31) / 32) * 32 : 32), 0, st>>>(w, buf, n, ++seq, d_err);
}
void PeerReduce::allreduce_u32(unsigned int* buf, int n, cudaStream_t st) {
  k_peer_allreduce<unsigned int><<<1, n >= 1024 ? 1024 : (n > 32 ? ((n + 31) / 32) * 32 : 32), 0, st>>>(w, buf, n, ++seq, d_err);
}
void PeerReduce::gather_u64(const unsigned long long* items, const unsigned int* count, unsigned long long* all, unsigned int* all_n, cudaStream_t st) {
  k_peer_gather<<<1, 1024, 0, st>>>(w, items, count, all, all_n, ++seq, d_err);
}
int PeerReduce::error(cudaStream_t st) const {
  if (!d_err) return 0;
  int e = 0;
  cudaMemcpyAsync(&e, d_err, sizeof(int), cudaMemcpyDeviceToHost, st);
  cudaStreamSynchronize(st);
  return e;
}
