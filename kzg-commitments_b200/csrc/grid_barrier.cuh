// Grid-wide barrier for cooperatively launched kernels (every CTA resident): a generation counter in device memory.
// One atomic per CTA to arrive, the last arrival bumps the generation, the others poll it with ld.acquire and a short
// sleep; __threadfence on both sides gives the ordering of cooperative_groups' grid.sync() and invalidates the SM's
// L1 before data written by other CTAs is read.  The object is zero-initialised once and reusable across launches.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
namespace kzgb200 {

struct GridBarrier {        // lives in device memory, zero-initialised once; reusable across launches
  unsigned count;
  unsigned pad0[31];
  unsigned gen;
  unsigned pad1[31];
  unsigned sleep_ns;      // pause between two polls of `gen` (0 = poll back to back); "barrier_sleep" option
  unsigned pad2[31];
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// all CTAs of the (cooperatively launched) grid
__device__ __forceinline__ void grid_sync(GridBarrier* b) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned g = ld_acquire_u32(&b->gen);  // cannot change before this CTA has arrived
    __threadfence();
    if (atomicAdd(&b->count, 1u) == gridDim.x - 1) {
      atomicExch(&b->count, 0u);
      __threadfence();
      atomicAdd(&b->gen, 1u);
    } else {
      const unsigned pause = b->sleep_ns;
      while (ld_acquire_u32(&b->gen) == g)
        if (pause) __nanosleep(pause);
    }
    __threadfence();
  }
  __syncthreads();
}

}  // namespace kzgb200
#endif
