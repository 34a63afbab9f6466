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
    unsigned long long stalls = 0;
    // CAS, not add-then-undo: speculative adds from hundreds of waiting CTAs would keep the
    // counter above the limit forever (livelock) when the limit is small.
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
    if (stalls) atomicAdd((unsigned long long*)&flow->limiter_stalls, stalls);
  }
  __syncthreads();
}

__device__ __forceinline__ void flow_release(FpsFlow* flow, int n) {
  __syncthreads();
  if (threadIdx.x == 0) atomicSub(&flow->in_flight, n);
}
