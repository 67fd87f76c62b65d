// dev_alloc.h — caching device allocator of the library.
//
// One Solve() of a new problem allocates ~40 device buffers and a dozen sort / scan temporaries; cudaMalloc and
// cudaFree cost 0.1 - 3 ms apiece on a B200 with gigabytes mapped (cudaFree also synchronises the device), which was
// a third of the set-up time of an end-to-end Solve().  Freed blocks are kept and handed out again (best fit, at most
// 25 % larger than asked for); they go back to the driver when the cache outgrows its budget, when an allocation
// fails, or on dev_trim().  dev_free keeps cudaFree's contract of synchronising the device first, so a block is never
// handed to another stream while work on it is pending.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace bagpu {
cudaError_t dev_alloc(void** p, size_t bytes);
void dev_free(void* p);
void dev_trim();
template <typename T> inline cudaError_t dev_alloc_t(T** p, size_t count) { return dev_alloc(reinterpret_cast<void**>(p), count * sizeof(T)); }
}  // namespace bagpu
