// Shared device helpers for the fps_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/fps_b200.h"

// ---------------------------------------------------------------------------------------------
// host-side error plumbing
// ---------------------------------------------------------------------------------------------
void fps_set_error(const char* fmt, ...);

#define FPS_CUDA(call)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (call);                                                             \
    if (_e != cudaSuccess) {                                                             \
      fps_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,                 \
                    cudaGetErrorString(_e));                                             \
      return (int)_e;                                                                    \
    }                                                                                    \
  } while (0)

#define FPS_REQUIRE(cond, code, ...)                                                     \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      fps_set_error(__VA_ARGS__);                                                        \
      return (code);                                                                     \
    }                                                                                    \
  } while (0)

#define FPS_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    cudaError_t _e = cudaGetLastError();                                                 \
    if (_e != cudaSuccess) {                                                             \
      fps_set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__,             \
                    cudaGetErrorString(_e));                                             \
      return (int)_e;                                                                    \
    }                                                                                    \
  } while (0)

int fps_num_sms();   // cached per device
int fps_check_view(const FpsTableView* v, const char* who);
int fps_mf_sgd_combine(const FpsTableView*, const FpsTableView*, const int32_t*, const int32_t*,
                       const float*, int64_t, const FpsMfParams*, FpsFlow*, cudaStream_t);

// ---------------------------------------------------------------------------------------------
// id -> (owner, slot)
// ---------------------------------------------------------------------------------------------
struct RowRef {
  int owner;
  long long slot;
};

__device__ __forceinline__ RowRef locate(const FpsTableView& t, long long id) {
  RowRef r;
  if (t.n_shards == 1) {   // one shard (single GPU, replicated table): no arithmetic at all
    r.owner = 0;
    r.slot = id;
    return r;
  }
  // 64-bit division costs ~100 instructions on the GPU and sat on top of the PA kernels' profile;
  // ids and divisors that fit 32 bits (every table below 4G rows) take the 32-bit path
  const bool small = ((unsigned long long)id | (unsigned long long)t.div) <= 0xffffffffull;
  if (t.partition == FPS_PART_HASH) {
    // n_shards is tiny; the compiler turns the common power-of-two cases into shifts only when
    // it can see them, so special-case them by hand.
    const int n = t.n_shards;
    if ((n & (n - 1)) == 0) {
      r.owner = (int)(id & (n - 1));
      r.slot = id >> (31 - __clz(n));
    } else if (small) {
      const unsigned q = (unsigned)id / (unsigned)n;
      r.owner = (int)((unsigned)id - q * (unsigned)n);
      r.slot = q;
    } else {
      r.owner = (int)(id % n);
      r.slot = id / n;
    }
  } else if (small) {
    r.owner = (int)((unsigned)id / (unsigned)t.div);
    r.slot = (long long)((unsigned)id - (unsigned)r.owner * (unsigned)t.div);
  } else {
    r.owner = (int)(id / t.div);
    r.slot = id - (long long)r.owner * t.div;
  }
  return r;
}

__device__ __forceinline__ bool in_table(const FpsTableView& t, long long id, const RowRef& r) {
  // n_ids, not just the shard capacity: with ceil-division padding the slots past the last id exist in
  // memory but are not parameters (they were trained on and counted as valid before)
  return id >= 0 && (t.n_ids <= 0 || id < t.n_ids) && r.owner < t.n_shards && r.slot < t.rows_per_shard;
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, counter-based init (twin of ops/philox.py)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

// the 4 initial values of float4 number `blk` of row `id`
__device__ __forceinline__ float4 init_block(const FpsTableView& t, long long id, int blk) {
  const uint4 w = philox4x32_10(make_uint4((uint32_t)id, (uint32_t)((unsigned long long)id >> 32),
                                           (uint32_t)blk, (uint32_t)t.stream_tag),
                                make_uint2((uint32_t)t.seed, (uint32_t)(t.seed >> 32)));
  const float scale = t.init_hi - t.init_lo, s24 = 5.9604644775390625e-08f;  // 2^-24
  float4 v;
  v.x = fmaf((float)(w.x >> 8) * s24, scale, t.init_lo);
  v.y = fmaf((float)(w.y >> 8) * s24, scale, t.init_lo);
  v.z = fmaf((float)(w.z >> 8) * s24, scale, t.init_lo);
  v.w = fmaf((float)(w.w >> 8) * s24, scale, t.init_lo);
  const int j = blk * 4;
  if (j + 0 >= t.dim) v.x = 0.f;   // padding columns stay zero so deltas on them are zero
  if (j + 1 >= t.dim) v.y = 0.f;
  if (j + 2 >= t.dim) v.z = 0.f;
  if (j + 3 >= t.dim) v.w = 0.f;
  return v;
}

// ---------------------------------------------------------------------------------------------
// memory access flavours used on (possibly peer-mapped) table rows
// ---------------------------------------------------------------------------------------------
// Pull: rows are mutable and may live on another GPU.  Peer lines are cached in the local L1
// only (L2 is bypassed for the peer aperture), and a persistent kernel would keep a stale copy
// forever, so row reads skip L1 (.cg).
__device__ __forceinline__ float4 ld_row4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_row1(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_row4(float* p, float4 v) {
  asm volatile("st.global.cg.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
// Push: param_update = add, applied in place on the owner.  One 16-byte reduction per lane;
// system scope because the owner may be a peer GPU behind NVSwitch.
__device__ __forceinline__ void red_add4(float* p, float4 v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
// Push-and-refresh: the add returns the row as it was, so `old + v` is the row as everybody sees it
// now, own delta included - one round trip instead of a push followed by a pull that might overtake it.
__device__ __forceinline__ float4 atom_add4(float* p, float4 v) {
  float4 o;
  asm volatile("atom.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4], {%5,%6,%7,%8};"
               : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
               : "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
  return o;
}
__device__ __forceinline__ void red_add1(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
// Push into a REPLICATED table: one multimem.red on the multicast alias and NVSwitch applies the
// add to every GPU's replica (NVLS in-switch reduction) - the sender pays for one copy on its link.
__device__ __forceinline__ void mc_red_add4(float* p, float4 v) {
  asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_red_add1(float* p, float v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
// param_update = add on row `off` (in floats from the shard base) of the row's owner - or, for a
// replicated table, on every replica: deferred (local replica + local delta array, exchanged by
// fps_table_flush_delta), or immediately through the multicast alias, or one red.add per replica.
__device__ __forceinline__ void push_add1(const FpsTableView& t, int owner, long long off, float v) {
  if (t.n_replicas <= 1) red_add1(t.rows[owner] + off, v);
  else if (t.delta_rows != nullptr) { red_add1(t.rows[0] + off, v); red_add1(t.delta_rows + off, v); }
  else if (t.mc_rows != nullptr) mc_red_add1(t.mc_rows + off, v);
  else
    for (int r = 0; r < t.n_replicas; ++r) red_add1(t.replica_rows[r] + off, v);
}
__device__ __forceinline__ void push_add4(const FpsTableView& t, int owner, long long off, float4 v) {
  if (t.n_replicas <= 1) red_add4(t.rows[owner] + off, v);
  else if (t.delta_rows != nullptr) { red_add4(t.rows[0] + off, v); red_add4(t.delta_rows + off, v); }
  else if (t.mc_rows != nullptr) mc_red_add4(t.mc_rows + off, v);
  else
    for (int r = 0; r < t.n_replicas; ++r) red_add4(t.replica_rows[r] + off, v);
}

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ---------------------------------------------------------------------------------------------
// lazy row life-cycle (state byte per row).  Byte-wide CAS does not exist, so the transition is a
// CAS on the aligned 32-bit word that contains the byte.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int touch_state(uint8_t* state, long long slot, int want) {
  // returns the row's state after the touch (FPS_ROW_PULLED / FPS_ROW_PUSHED)
  volatile uint8_t* sp = state + slot;
  int cur = *sp;
  if (cur != FPS_ROW_UNTOUCHED) return cur;
  unsigned int* word = (unsigned int*)((uintptr_t)(state + slot) & ~(uintptr_t)3);
  const int shift = (int)(((uintptr_t)(state + slot)) & 3) * 8;
  unsigned int old = *(volatile unsigned int*)word;
  while (true) {
    const int b = (old >> shift) & 0xff;
    if (b != FPS_ROW_UNTOUCHED) return b;
    const unsigned int assumed = old;
    old = atomicCAS_system(word, assumed, assumed | ((unsigned int)want << shift));
    if (old == assumed) return want;
  }
}

// sub-warp sum over `LANES` consecutive lanes (LANES power of two <= 32)
template <int LANES>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = LANES / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float dot4(float4 a, float4 b) {
  return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, a.w * b.w)));
}
