// Shared pieces of the top-K kernels (topk.cu: in-kernel operand staging; topk_packed.cu: pre-packed
// item tiles streamed with bulk async copies): PTX wrappers for tcgen05 / TMEM / mbarrier, the
// SWIZZLE_128B staging routine, candidate pruning, scratch planning and the merge kernel.
#pragma once
#include "common.cuh"

namespace {

constexpr int TK_M = 128;          // queries per CTA (UMMA M, TMEM lanes)
constexpr int TK_N = 64;           // items per tile (UMMA N)
constexpr int TK_KB = 32;          // fp32 elements per K block = one 128-byte swizzle row
constexpr int TK_THREADS = 128;
constexpr int TK_CAP = 256;        // candidate slots per (query, split)
constexpr int TK_MAX_K = 128;      // k <= CAP - N must hold: 128 <= 256 - 64

// ------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
}
// The same wait for call sites where ALL 32 lanes of a warp poll one barrier.  When every operand of the poll
// is provably uniform, ptxas takes the result for warp-uniform and emits no reconvergence point behind the
// loop (seen in SASS: the loop branch is followed directly by uniform-datapath updates of phase bits and
// counters, and a __syncwarp() placed there is compiled to a NOP).  Whether the hardware always answers the
// 32 lanes of one poll alike is not documented, so the exit is made uniform in fact with a vote; a completed
// phase stays completed until the waiters have moved on.  Measured cost: none (config 6 unchanged).
__device__ __forceinline__ void mbar_wait_warp(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!__all_sync(0xffffffffu, done));
}
// D[tmem] (+)= A[smem] * B[smem]^T, tf32 inputs, fp32 accumulate, issued by one thread
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
//   [0,14) start >> 4 | [16,30) LBO >> 4 (unused for swizzled K-major, 1) | [32,46) SBO >> 4 = 1024 B
//   between 8-row groups | [46,48) version = 1 | [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor: D fp32 (1<<4), A/B tf32 (2<<7, 2<<10), both K-major, N>>3 at 17, M>>4 at 24
constexpr uint32_t TK_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TK_N >> 3) << 17) | ((uint32_t)(TK_M >> 4) << 24);

__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// order-preserving float <-> uint key
__device__ __forceinline__ uint32_t f2key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// Stage `rows` x (KBLKS*32) fp32 values (row-major, `stride` floats apart, rows >= valid are zero,
// columns >= dim are zero) into the swizzled K-major layout: K block kb is a [rows x 128 B] slab;
// inside it row r sits at r*128 and its 16-byte chunk c at ((c ^ (r & 7)) * 16).
template <bool SPLIT>
__device__ __forceinline__ void stage_tile(float* __restrict__ s_hi, float* __restrict__ s_lo,
                                           const float* __restrict__ src, int stride, int rows, int valid,
                                           int dim, int kblks) {
  const int chunks_per_row = kblks * 8;
  for (int i = threadIdx.x; i < rows * chunks_per_row; i += TK_THREADS) {
    const int r = i / chunks_per_row, cc = i - r * chunks_per_row;
    const int kb = cc >> 3, c = cc & 7;
    const int col = kb * TK_KB + c * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < valid && col < dim) {
      const float* p = src + (size_t)r * stride + col;
      if (col + 3 < dim) x = *reinterpret_cast<const float4*>(p);
      else {
        x.x = p[0];
        if (col + 1 < dim) x.y = p[1];
        if (col + 2 < dim) x.z = p[2];
      }
    }
    const int off = kb * rows * TK_KB + r * TK_KB + ((c ^ (r & 7)) << 2);   // in floats
    if (SPLIT) {
      float4 h, l;
      h.x = to_tf32(x.x); h.y = to_tf32(x.y); h.z = to_tf32(x.z); h.w = to_tf32(x.w);
      l.x = to_tf32(x.x - h.x); l.y = to_tf32(x.y - h.y); l.z = to_tf32(x.z - h.z); l.w = to_tf32(x.w - h.w);
      *reinterpret_cast<float4*>(s_hi + off) = h;
      *reinterpret_cast<float4*>(s_lo + off) = l;
    } else {
      *reinterpret_cast<float4*>(s_hi + off) = x;
    }
  }
}

struct TopkScratch {
  float* cand_s;          // [n_ctas][CAP][128]
  int* cand_i;            // [n_ctas][CAP][128]
  uint32_t* theta;        // [n_q_tiles*128] shared k-th-best lower bound per query (ordered key)
  float* part_s;          // [n_q_pad][n_splits][k]
  int* part_i;
};

// Warp-cooperative exact prune of one row-major list (c > k entries, c <= TK_CAP = 8 per lane) back to
// its k best; returns the k-th best score.  Everything after the 16 coalesced loads runs in
// registers: common-prefix skip, then one REDUX per remaining key bit, then a ballot compaction.
template <int CAP = TK_CAP>
__device__ __forceinline__ float prune_row(float* rs, int* ri, int c, int k, int lane) {
  constexpr int PER = CAP / 32;
  uint32_t key[PER];
  int id[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int idx = lane + 32 * i;
    const bool valid = idx < c;
    key[i] = valid ? f2key(rs[idx]) : 0u;      // 0 sorts below every real score
    id[i] = valid ? ri[idx] : -1;
  }
  const uint32_t key0 = __shfl_sync(0xffffffffu, key[0], 0);
  uint32_t diff = 0;
#pragma unroll
  for (int i = 0; i < PER; ++i) diff |= (lane + 32 * i < c) ? (key[i] ^ key0) : 0u;
  diff = __reduce_or_sync(0xffffffffu, diff);
  const int hb = diff ? 31 - __clz(diff) : -1;
  uint32_t T = hb >= 0 ? (key0 & ~((2u << hb) - 1u)) : key0;
  for (int b = hb; b >= 0; --b) {
    const uint32_t trial = T | (1u << b);
    int n = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) n += (key[i] >= trial);
    n = __reduce_add_sync(0xffffffffu, n);
    if (n >= k) T = trial;
  }
  int n_above = 0;
#pragma unroll
  for (int i = 0; i < PER; ++i) n_above += (key[i] > T);
  n_above = __reduce_add_sync(0xffffffffu, n_above);
  const int ties_left = k - n_above;
  const uint32_t lt = (1u << lane) - 1u;
  int base_a = 0, base_t = 0;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const bool a = key[i] > T, t = key[i] == T && (lane + 32 * i < c);
    const uint32_t ma = __ballot_sync(0xffffffffu, a), mt = __ballot_sync(0xffffffffu, t);
    const int pa = base_a + __popc(ma & lt), pt = base_t + __popc(mt & lt);
    if (a) { rs[pa] = key2f(key[i]); ri[pa] = id[i]; }
    else if (t && pt < ties_left) { rs[n_above + pt] = key2f(key[i]); ri[n_above + pt] = id[i]; }
    base_a += __popc(ma);
    base_t += __popc(mt);
  }
  return key2f(T);
}

// ---- N7: merge partial lists (one CTA per query, bitonic sort in shared memory) ------------------
__global__ void __launch_bounds__(256)
k_topk_merge(const float* __restrict__ part_s, const int* __restrict__ part_i, int n_parts, long long n_q,
             int k_in, int k_out, const int* __restrict__ seen, int seen_stride,
             float* __restrict__ out_s, int* __restrict__ out_i, int pow2) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  float* s = (float*)smem_raw;
  int* id = (int*)(s + pow2);
  const long long q = blockIdx.x;
  const int n = n_parts * k_in;
  for (int j = threadIdx.x; j < pow2; j += blockDim.x) {
    float v = -INFINITY;
    int ii = -1;
    if (j < n) {
      v = part_s[q * n + j];
      ii = part_i[q * n + j];
      if (ii >= 0 && seen != nullptr) {
        for (int m = 0; m < seen_stride; ++m) {
          const int sv = seen[q * seen_stride + m];
          if (sv < 0) break;
          if (sv == ii) { v = -INFINITY; ii = -1; break; }
        }
      }
      if (ii < 0) v = -INFINITY;
    }
    s[j] = v;
    id[j] = ii;
  }
  __syncthreads();
  // bitonic sort, descending by (score, then smaller id first for determinism)
  for (int size = 2; size <= pow2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int j = threadIdx.x; j < pow2; j += blockDim.x) {
        const int p = j ^ stride;
        if (p > j) {
          const bool desc = ((j & size) == 0);
          const float a = s[j], b = s[p];
          const int ia = id[j], ib = id[p];
          const bool a_first = (a > b) || (a == b && (unsigned)ia <= (unsigned)ib);
          if (desc ? !a_first : a_first) {
            s[j] = b; s[p] = a; id[j] = ib; id[p] = ia;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int j = threadIdx.x; j < k_out; j += blockDim.x) {
    out_s[q * k_out + j] = j < pow2 ? s[j] : -INFINITY;
    out_i[q * k_out + j] = j < pow2 ? id[j] : -1;
  }
}

struct Plan {
  int n_q_tiles, n_splits, n_ctas, kblks;
  size_t off_cand_i, off_theta, off_part_s, off_part_i, total;
};

// `groups` = candidate lists per (CTA, query): 1 for topk.cu, the number of selection groups for
// topk_packed.cu (each group also writes its own partial list)
// `multi_wave`: allow more CTAs than SMs when that fills the machine better (32 query tiles: 4 splits
// = 128 CTAs leave 20 of 148 SMs idle, 9 splits = 288 CTAs run as 1.95 waves); later CTAs start with
// the theta the earlier ones published, so the extra lists cost little.  `forced_splits` > 0 overrides.
Plan make_plan(long long n_q, int k, long long n_items, int dim, int groups = 1, int multi_wave = 0,
               int forced_splits = 0, int max_splits = 16) {
  Plan p;
  p.n_q_tiles = (int)((n_q + TK_M - 1) / TK_M);
  const long long n_tiles = (n_items + TK_N - 1) / TK_N;
  const int sms = fps_num_sms();
  const int qt = p.n_q_tiles > 0 ? p.n_q_tiles : 1;
  long long splits = sms / qt;
  if (splits < 1) splits = 1;
  if (splits > max_splits) splits = max_splits;
  // more than 16 splits only while the merge of their lists stays a 2048-element sort per query: measured on
  // 256 queries x 19K items, k = 75 (ms per learner+generator step): 16 splits 0.565, 24 0.507, 32 (4096-sort)
  // 0.550, 48 0.495, 64 (8192-sort) 0.684
  while (splits > 16 && splits * groups * k > 2048) --splits;
  if (multi_wave) {
    auto eff = [&](long long s) { return (double)(qt * s) / (double)(((qt * s + sms - 1) / sms) * sms); };
    long long best = splits;
    for (long long s = splits + 1; s <= 16; ++s)
      if (s * groups * k <= 8192 && eff(s) > eff(best) + 0.03) best = s;
    splits = best;
  }
  if (forced_splits > 0) splits = forced_splits > max_splits ? max_splits : forced_splits;
  if (splits > n_tiles) splits = n_tiles > 0 ? n_tiles : 1;
  p.n_splits = (int)splits;
  p.n_ctas = p.n_q_tiles * p.n_splits;
  p.kblks = (dim + TK_KB - 1) / TK_KB;
  auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
  size_t o = 0;
  o = align(o + (size_t)p.n_ctas * groups * TK_CAP * TK_M * sizeof(float));
  p.off_cand_i = o;
  o = align(o + (size_t)p.n_ctas * groups * TK_CAP * TK_M * sizeof(int));
  p.off_theta = o;
  o = align(o + (size_t)p.n_q_tiles * TK_M * sizeof(uint32_t));
  p.off_part_s = o;
  o = align(o + (size_t)p.n_q_tiles * TK_M * p.n_splits * groups * k * sizeof(float));
  p.off_part_i = o;
  o = align(o + (size_t)p.n_q_tiles * TK_M * p.n_splits * groups * k * sizeof(int));
  p.total = o;
  return p;
}

int launch_merge(const float* part_s, const int* part_i, int n_parts, long long n_q, int k_in, int k_out,
                 const int* seen, int seen_stride, float* out_s, int* out_i, cudaStream_t st) {
  int pow2 = 1;
  while (pow2 < n_parts * k_in) pow2 <<= 1;
  FPS_REQUIRE(pow2 <= 8192, FPS_E_TOO_LARGE, "fps_topk_merge: %d candidates per query exceed 8192", n_parts * k_in);
  FPS_REQUIRE(k_out <= pow2, FPS_E_INVALID, "fps_topk_merge: k_out %d > candidates %d", k_out, pow2);
  const size_t smem = (size_t)pow2 * 8;
  if (smem > 48 * 1024) FPS_CUDA(cudaFuncSetAttribute(k_topk_merge, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_topk_merge<<<(unsigned)n_q, 256, smem, st>>>(part_s, part_i, n_parts, n_q, k_in, k_out, seen, seen_stride, out_s, out_i, pow2);
  FPS_LAUNCH_CHECK();
  return 0;
}

}  // namespace
