// peer_reduce.h — one-shot all-reduce of small vectors over NVLink peer memory.
//
// The solver's inner loop has about ten small reductions per iteration (six 4096-bin histograms of the distributed exact
// median, dot products, error sums).  Through NCCL each one costs a launch plus 20-40 us of protocol latency at 8 GPUs,
// which at 8 GPUs is more than the sharded Schur build itself.  Here every rank owns a window of device memory that
// its peers map through CUDA IPC; one single-CTA kernel stores the rank's vector into its slot of every peer's window
// (plain NVLink stores), publishes a sequence number with a system-scope release, waits for the peers' sequence
// numbers with acquire loads and sums the slots in rank order — so every rank gets bit-identical sums.
// The bulk reduction of the partial reduced systems (S tiles, b_p, rhs_sc) goes over peer memory too (PeerBulk below);
// NCCL is the fall-back when the windows cannot be mapped.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>

constexpr int kPeerMaxWorld = 8;
constexpr size_t kPeerSlotBytes = 65536;  // all-reduce payloads up to kPeerReduceMax; the candidate gather of the exact median uses the whole slot
constexpr size_t kPeerReduceMax = 16384;  // one 4096-bin u32 histogram: larger vectors are not worth a single-CTA kernel
constexpr int kPeerGatherCap = (int)(kPeerSlotBytes / 8) - 1;   // u64 items one rank can contribute to a gather
constexpr size_t kPeerFlagBytes = 256;    // 2 parities x 8 ranks x u64, padded
constexpr size_t kPeerWindowBytes = kPeerFlagBytes + 2 * kPeerMaxWorld * kPeerSlotBytes;

struct PeerWindow {
  unsigned char* win[kPeerMaxWorld];  // win[rank] is the local window
  int rank, world;
};

struct PeerReduce {
  PeerWindow w{};
  unsigned long long seq = 0;
  int* d_err = nullptr;   // sticky: a peer did not show up within the spin limit
  bool ok = false;
  // the handle bytes of this rank's window, to be exchanged by the caller (64 bytes)
  cudaError_t create_local(int rank, int world, void* handle64);
  // open the peers' windows; returns false when peer mapping is not possible (the caller falls back to NCCL)
  bool open_peers(const void* handles /* world x 64 bytes */);
  void close_peers();   // unmap the peers' windows (every rank must have done this before any rank frees its own)
  void destroy();       // close_peers + free the local window
  bool fits(size_t bytes) const { return ok && bytes <= kPeerReduceMax; }
  void allreduce_f64(double* buf, int n, cudaStream_t st);
  void allreduce_u32(unsigned int* buf, int n, cudaStream_t st);
  // all-gather of at most kPeerGatherCap u64 items per rank: every rank ends with the concatenation in rank order in `all`
  // and the total count in *all_n (device pointers; *count is read on the device)
  void gather_u64(const unsigned long long* items, const unsigned int* count, unsigned long long* all, unsigned int* all_n, cudaStream_t st);
  int error(cudaStream_t st) const;  // synchronises the stream
};

// Bulk all-reduce of the reduced system over NVLink peer memory.  Every rank keeps its S tiles / b_p / rhs_sc in one
// allocation that its peers map through CUDA IPC.  One multi-CTA kernel: the tiles (and 16 KB chunks of the vector
// tail) are dealt round-robin to the ranks; the owner of a chunk loads the N partial copies straight from the peers'
// memory, adds them in rank order and stores the sum into every rank's copy (a fused reduce-scatter + all-gather:
// (N-1)/N of the data in and out per rank, every rank ends with bit-identical sums).  Only tiles of the symbolic
// pattern are visited: fill tiles of the factorisation are zero on every rank.  The caller brackets the kernel with two
// barriers of the small-reduction windows (PeerReduce): no rank reads a peer's partial system before that peer has
// finished building it, and no rank uses the summed system before every owner has delivered.
struct PeerBulk {
  double* base[kPeerMaxWorld] = {nullptr};   // base[rank] is the local allocation
  size_t cap_bytes = 0;
  int rank = 0, world = 1;
  bool ok = false;
  cudaError_t create_local(int rank_, int world_, size_t bytes, void* handle64);
  bool open_peers(const void* handles /* world x 64 bytes */);
  void close_peers();
  void destroy();
  // slots: tile slots of the symbolic pattern (device array), tile_elems doubles each; tail: [tail_off, tail_off + tail_n) doubles
  void allreduce(const int* slots, int n_slots, int tile_elems, size_t tail_off, size_t tail_n, int sms, cudaStream_t st);
};
