// Device-side flow control (N12): the pull limiter as a credit counter in device memory.
// Reference: M/WorkerLogic.scala:176-225 keeps `pullCounter <= pullLimit` on the worker's JVM
// heap and queues the excess; here a CTA acquires `n` credits before it issues the pulls of a
// chunk and returns them when the chunk's pushes have been issued.  Invariant checked by tests:
// flow->max_in_flight <= pull_limit.
#pragma once
#include "common.cuh"

// all threads of the block call these; contains __syncthreads
__device__ __forceinline__ void flow_acquire(FpsFlow* flow, int n, int limit) {
  if (threadIdx.x == 0) {
    // Fast path: one optimistic add.  The hosts size their grids so that the allowances of all CTAs
    // that can be resident fit the limit, so it all but always succeeds (no retry, one round trip).
    const int old = atomicAdd(&flow->in_flight, n);
    if (old + n <= limit) {
      atomicMax(&flow->max_in_flight, old + n);
    } else {
      // Somebody else holds the credits (another launch on the same flow block): undo, then wait with
      // CAS - not add-then-undo in a loop, which livelocks when many CTAs wait at a small limit.
      atomicSub(&flow->in_flight, n);
      unsigned long long stalls = 0;
      int cur = *(volatile int*)&flow->in_flight;
      while (true) {
        if (cur + n <= limit) {
          const int seen = atomicCAS(&flow->in_flight, cur, cur + n);
          if (seen == cur) {
            atomicMax(&flow->max_in_flight, cur + n);
            break;
          }
          cur = seen;
          continue;
        }
        ++stalls;
        __nanosleep(100 + (blockIdx.x & 63) * 4);
        cur = *(volatile int*)&flow->in_flight;
      }
      atomicAdd((unsigned long long*)&flow->limiter_stalls, stalls + 1);
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void flow_release(FpsFlow* flow, int n) {
  __syncthreads();
  if (threadIdx.x == 0) atomicSub(&flow->in_flight, n);
}
