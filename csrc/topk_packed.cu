// LEMP top-K, second generation: pre-packed item tiles streamed by the TMA engine into a
// warp-specialised tcgen05 pipeline.
//
// topk.cu stages every item tile with CUDA-core loads + tf32 splitting inside the scoring loop;
// ncu showed the tensor pipe 4 % busy and the staging arithmetic on top of the stall list.  Items
// change rarely compared with how often they are scored (a recommender answers many query batches
// per model refresh), so here the hi/lo split and the SWIZZLE_128B layout are computed once by
// `fps_topk_pack_items` into an image that is byte-for-byte what the UMMA descriptors expect in
// shared memory.  Scoring then is:
//   prologue (warps 0-3): the CTA's 128 queries are split into tf32 hi / lo once and written to
//       TMEM with tcgen05.st (columns TP_A_COL..), so the A operand never touches shared memory
//       again (TS-mode MMA) and all of shared memory is free for item tiles.
//   producer warp (one elected lane): cp.async.bulk (TMA, UBLKCP) of the next packed tile into a
//       TP_STAGES-deep shared-memory ring, completion on the stage's "full" mbarrier.
//   MMA warp (one elected lane): waits for "full", issues tcgen05.mma (3 x kind::tf32 per K step,
//       A from TMEM, descriptors = base + compile-time constants) into one of four TMEM
//       accumulators, tcgen05.commit to that accumulator's mbarrier and to the stage's "empty"
//       mbarrier.
//   selection warps, G groups of four (G = 1, 2 or 4): group g takes tiles g, g + G, ...; it waits
//       for the accumulator, tcgen05.ld's it (thread = TMEM lane = query), hands the TMEM buffer
//       back, then runs the LEMP selection (group-of-4 max filter, threshold append,
//       warp-synchronous pruning) while the next tiles' copies and MMAs are in flight.  The third
//       ncu capture (profiles/r1_topk) showed the kernel bound by the latency of this branchy code
//       with one warp per scheduler - hence the groups.
// The LEMP bucket test (||q|| * tileMaxLen <= theta for every query of the CTA) is evaluated by the
// selection warps and published per warp; the producer stops fetching tiles once every lane quarter
// has one group that agrees.
#include <cstdlib>

#include "topk_core.cuh"

namespace {

constexpr int TP_MAX_GROUPS = 4;      // selection groups (4 warps each) per CTA
constexpr int TP_DEFAULT_GROUPS = 1;  // measured (config 6): 6.31 / 6.43 / 7.32 ms for 1 / 2 / 4 groups - every extra
                                      // list per query multiplies accepted candidates and prunes
constexpr int TP_MAX_THREADS = (4 * TP_MAX_GROUPS + 2) * 32;   // + TMA producer warp + MMA issuer warp
constexpr int TP_ACC = 4;             // TMEM accumulators of TK_N columns
constexpr int TP_TMEM_COLS = 512;     // 4 x 64 accumulator columns + 2 x 128 query (hi, lo) columns
constexpr int TP_A_COL = TP_ACC * TK_N;   // first TMEM column of the query operand
constexpr int TP_STAGES = 3;          // shared-memory stages of packed item tiles (64 KB each at rank 128)

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done != 0;
}
// one poll by all 32 lanes of a warp, with a warp-uniform answer (see mbar_wait_warp)
__device__ __forceinline__ bool mbar_try_warp(uint64_t* bar, uint32_t parity) {
  return __all_sync(0xffffffffu, mbar_try(bar, parity)) != 0;
}
// global -> shared bulk copy executed by the TMA engine; completes `bytes` on `bar`
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---- one-time packing: fp32 rows -> per-tile [hi | lo] images in the swizzled K-major layout ----
__global__ void __launch_bounds__(256)
k_pack_items(const float* __restrict__ items, int stride, long long n_items, int dim, int kblks,
             float* __restrict__ packed) {
  const long long tile = blockIdx.x;
  const long long i0 = tile * TK_N;
  const int b_floats = TK_N * kblks * TK_KB;
  float* out = packed + tile * 2 * (long long)b_floats;
  const int chunks_per_row = kblks * 8;
  for (int i = threadIdx.x; i < TK_N * chunks_per_row; i += blockDim.x) {
    const int r = i / chunks_per_row, cc = i - r * chunks_per_row;
    const int kb = cc >> 3, c = cc & 7;
    const int col = kb * TK_KB + c * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i0 + r < n_items && col < dim) {
      const float* p = items + (i0 + r) * stride + col;
      if (col + 3 < dim) x = *reinterpret_cast<const float4*>(p);
      else {
        x.x = p[0];
        if (col + 1 < dim) x.y = p[1];
        if (col + 2 < dim) x.z = p[2];
      }
    }
    float4 h, l;
    h.x = to_tf32(x.x); h.y = to_tf32(x.y); h.z = to_tf32(x.z); h.w = to_tf32(x.w);
    l.x = to_tf32(x.x - h.x); l.y = to_tf32(x.y - h.y); l.z = to_tf32(x.z - h.z); l.w = to_tf32(x.w - h.w);
    const int off = kb * TK_N * TK_KB + r * TK_KB + ((c ^ (r & 7)) << 2);
    *reinterpret_cast<float4*>(out + off) = h;
    *reinterpret_cast<float4*>(out + b_floats + off) = l;
  }
}

struct PackedShared {
  uint64_t b_full[TP_STAGES];      // TMA landed the tile in stage s
  uint64_t stage_free[TP_STAGES];  // the MMAs that read stage s are complete
  uint64_t acc_full[TP_ACC];       // MMAs of the tile using accumulator b are complete
  uint64_t acc_empty[TP_ACC];      // the four selection warps of the owning group have read accumulator b
  uint32_t tmem;
  volatile int skip_from[4 * TP_MAX_GROUPS];   // per selection warp: first tile ordinal that may be skipped
  volatile int issued_final;       // number of tiles the producer really fetched (-1 while running)
  float qlen[TK_M];
};

// 32 consecutive TMEM columns of this thread's lane <- registers
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr),
        "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])),
        "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
        "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
        "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])),
        "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T : the query operand never touches shared memory
// Executed by the whole (converged) MMA warp; elect.sync picks the issuing lane inside the asm so
// that the compiler keeps every operand in uniform registers and emits one UTCHMMA per call (an
// `if (lane == 0)` around it costs an ELECT / BRA.U.ANY loop per MMA - seen in the SASS).
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .pred e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, e;\n\t}"
      : "=r"(r));
  return r;
}
template <bool ACC>
__device__ __forceinline__ void umma_tf32_ts(uint32_t leader, uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "setp.ne.b32 e, %5, 0;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(TK_IDESC), "r"(ACC ? 1u : 0u), "r"(leader)
      : "memory");
}
// tcgen05.ld of 32 columns without the wait (the caller waits once for a batch of loads)
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, float (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
        "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15]),
        "=f"(v[16]), "=f"(v[17]), "=f"(v[18]), "=f"(v[19]), "=f"(v[20]), "=f"(v[21]), "=f"(v[22]), "=f"(v[23]),
        "=f"(v[24]), "=f"(v[25]), "=f"(v[26]), "=f"(v[27]), "=f"(v[28]), "=f"(v[29]), "=f"(v[30]), "=f"(v[31])
      : "r"(taddr)
      : "memory");
}
template <bool ACC>
__device__ __forceinline__ void umma_tf32_ts_i(uint32_t leader, uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "setp.ne.b32 e, %5, 0;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(ACC ? 1u : 0u), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void umma_commit_elected(uint32_t leader, uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\t"
      "setp.ne.b32 e, %1, 0;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      ::"r"(smem_u32(bar)), "r"(leader) : "memory");
}

// All MMAs of one tile: 3 x tf32 per 8-wide K step, KBLKS*4 steps.  The shared-memory descriptors of
// the steps differ from the stage's base descriptor by compile-time constants (32 B per K step, one
// [64 x 128 B] slab per K block, the lo image b_bytes after the hi image), so the issuing thread
// executes a couple of uniform adds per MMA instead of rebuilding descriptors.
template <int KBLKS>
__device__ __forceinline__ void issue_tile_mmas(uint32_t leader, uint32_t acc_addr, uint32_t a_hi, uint32_t a_lo,
                                                uint64_t desc_hi0) {
  constexpr uint64_t LO_OFF = ((uint64_t)TK_N * KBLKS * TK_KB * 4) >> 4;      // lo image, in 16-byte units
#pragma unroll
  for (int kb = 0; kb < KBLKS; ++kb) {
#pragma unroll
    for (int ks = 0; ks < TK_KB / 8; ++ks) {
      const uint32_t a_col = (uint32_t)(kb * TK_KB + ks * 8);
      const uint64_t off = ((uint64_t)(kb * TK_N * TK_KB + ks * 8) * 4) >> 4;
      const uint64_t bh = desc_hi0 + off, bl = desc_hi0 + off + LO_OFF;
      if (kb == 0 && ks == 0) umma_tf32_ts<false>(leader, acc_addr, a_lo + a_col, bh);   // small terms first
      else umma_tf32_ts<true>(leader, acc_addr, a_lo + a_col, bh);
      umma_tf32_ts<true>(leader, acc_addr, a_hi + a_col, bl);
      umma_tf32_ts<true>(leader, acc_addr, a_hi + a_col, bh);
    }
  }
}

// Append the entries of v[0..31] that beat theta.  Nearly all entries fail, so groups of four are
// rejected with one max3 + max + compare; only a group with a hit runs the per-element code.
// The candidate lists of the packed kernel are row-major: list of (CTA, group, query) = TK_CAP
// consecutive floats / ints, so that a warp can prune one list cooperatively with coalesced loads.
// The append itself is rare per thread; kept out of line so that the hot filter code stays small
// (the unrolled version was 79 KB of SASS and showed instruction-fetch stalls).
__device__ __noinline__ int append4(float a, float b, float c, float d, float theta, int item0, float* rs, int* ri, int cnt) {
  if (a > theta) { rs[cnt] = a; ri[cnt] = item0; ++cnt; }
  if (b > theta) { rs[cnt] = b; ri[cnt] = item0 + 1; ++cnt; }
  if (c > theta) { rs[cnt] = c; ri[cnt] = item0 + 2; ++cnt; }
  if (d > theta) { rs[cnt] = d; ri[cnt] = item0 + 3; ++cnt; }
  return cnt;
}
// Two-level filter over 32 scores: eight group-of-4 maxima (independent FMNMX3 pairs), their maximum,
// ONE branch when nothing beats theta (the common case even per warp), else a branch per group.
__device__ __forceinline__ void select32(const float (&v)[32], float theta, int item0, float* rs, int* ri, int& cnt) {
  float m[8];
#pragma unroll
  for (int g = 0; g < 8; ++g) m[g] = fmaxf(fmaxf(v[4 * g], v[4 * g + 1]), fmaxf(v[4 * g + 2], v[4 * g + 3]));
  const float mm = fmaxf(fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3])), fmaxf(fmaxf(m[4], m[5]), fmaxf(m[6], m[7])));
  if (mm > theta) {
#pragma unroll
    for (int g = 0; g < 8; ++g)
      if (m[g] > theta) cnt = append4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3], theta, item0 + 4 * g, rs, ri, cnt);
  }
}

__global__ void __launch_bounds__(TP_MAX_THREADS, 1)
k_lemp_topk_packed(const float* __restrict__ queries, int q_stride, long long n_q,
                   const float* __restrict__ packed, const float* __restrict__ item_len,
                   const int* __restrict__ item_ids, long long n_items, int dim, int kblks, int k, int n_splits,
                   int groups, int use_pruning, TopkScratch sc, unsigned long long* __restrict__ stats) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int b_floats = TK_N * kblks * TK_KB;
  float* sB = (float*)base;                 // TP_STAGES x [hi | lo] packed tiles
  __shared__ PackedShared sh;

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int n_sel_warps = 4 * groups;       // warps [0, 4*groups) select, then the producer, then the MMA issuer
  const int qt = blockIdx.x / n_splits, split = blockIdx.x - qt * n_splits;
  const long long q0 = (long long)qt * TK_M;
  const int q_valid = (int)min((long long)TK_M, n_q - q0);
  const long long n_tiles = (n_items + TK_N - 1) / TK_N;
  const int my_tiles = n_tiles > split ? (int)((n_tiles - split + n_splits - 1) / n_splits) : 0;
  const uint32_t tile_bytes = (uint32_t)(2 * b_floats * sizeof(float));
  const int stage_floats = 2 * b_floats;

  if (warp == n_sel_warps) tmem_alloc((uint32_t*)&sh.tmem, TP_TMEM_COLS);
  if (t == 0) {
    for (int s = 0; s < TP_STAGES; ++s) {
      mbar_init(&sh.b_full[s], 1);
      mbar_init(&sh.stage_free[s], 1);
    }
    for (int b = 0; b < TP_ACC; ++b) {
      mbar_init(&sh.acc_full[b], 1);
      mbar_init(&sh.acc_empty[b], 4);
    }
    sh.issued_final = -1;
    for (int w = 0; w < 4 * TP_MAX_GROUPS; ++w) sh.skip_from[w] = 0x7fffffff;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh.tmem;
  // TMEM map: columns [0,256) four accumulators, [256,384) query hi, [384,512) query lo
  const uint32_t tmem_a_hi = tmem_base + TP_A_COL, tmem_a_lo = tmem_base + TP_A_COL + 128;

  // ---- queries -> TMEM (thread t = lane t = query t), split into tf32 hi / lo on the way
  if (warp < 4) {
    const float* qp = queries + (q0 + t) * q_stride;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    float s = 0.f;
    for (int c0 = 0; c0 < kblks * TK_KB; c0 += 32) {
      float hi[32], lo[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int col = c0 + j;
        const float x = (t < q_valid && col < dim) ? qp[col] : 0.f;
        s = fmaf(x, x, s);
        hi[j] = to_tf32(x);
        lo[j] = to_tf32(x - hi[j]);
      }
      tmem_st32(tmem_a_hi + lane_off + c0, hi);
      tmem_st32(tmem_a_lo + lane_off + c0, lo);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    sh.qlen[t] = sqrtf(s);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == n_sel_warps) {
    // ===================== producer: TMA bulk copies of packed tiles =====================
    if (lane == 0) {
      int issued = 0;
      for (int j = 0; j < my_tiles; ++j) {
        const int s = j % TP_STAGES, n = j / TP_STAGES;
        if (n > 0) mbar_wait(&sh.stage_free[s], (n - 1) & 1);
        if (use_pruning) {
          // a lane quarter may skip from tile j on as soon as ONE group says so: each group's theta is a
          // lower bound of the same queries' k-th best score
          bool all = true;
          for (int sub = 0; sub < 4; ++sub) {
            int first = 0x7fffffff;
            for (int g = 0; g < groups; ++g) first = min(first, sh.skip_from[g * 4 + sub]);
            all = all && (first <= j);
          }
          if (all) break;                                                      // LEMP early exit
        }
        const long long tile = split + (long long)j * n_splits;
        const float* src = packed + tile * (long long)stage_floats;
        float* dst = sB + (size_t)s * stage_floats;
        mbar_expect_tx(&sh.b_full[s], tile_bytes);
        for (int part = 0; part < 2 * kblks; ++part)       // 8 KB per copy
          tma_bulk_g2s(dst + part * TK_N * TK_KB, src + part * TK_N * TK_KB, TK_N * TK_KB * 4, &sh.b_full[s]);
        issued = j + 1;
      }
      __threadfence_block();
      sh.issued_final = issued;
    }
  } else if (warp == n_sel_warps + 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs this loop converged (all lanes poll the same barriers); one elected lane
    // issues.  tmem_base goes through a shuffle so that the compiler treats it as warp-uniform.
    const uint32_t leader = elect_one();
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t a_hi = tb + TP_A_COL, a_lo = tb + TP_A_COL + 128;
    for (int j = 0; j < my_tiles; ++j) {
      const int s = j % TP_STAGES, n = j / TP_STAGES;
      bool have = false;
      while (true) {
        if (mbar_try_warp(&sh.b_full[s], n & 1)) { have = true; break; }
        const int fin = sh.issued_final;
        if (__any_sync(0xffffffffu, fin >= 0 && j >= fin)) break;
      }
      have = __shfl_sync(0xffffffffu, have ? 1 : 0, 0) != 0;
      if (!have) break;
      const int b = j % TP_ACC;
      if (j >= TP_ACC) mbar_wait_warp(&sh.acc_empty[b], ((j / TP_ACC) - 1) & 1);      // accumulator drained
      tc_fence_after();
      const uint32_t acc_addr = tb + (uint32_t)b * TK_N;
      const uint64_t d0 = make_desc(smem_u32(sB + (size_t)s * stage_floats));
      switch (kblks) {
        case 1: issue_tile_mmas<1>(leader, acc_addr, a_hi, a_lo, d0); break;
        case 2: issue_tile_mmas<2>(leader, acc_addr, a_hi, a_lo, d0); break;
        case 3: issue_tile_mmas<3>(leader, acc_addr, a_hi, a_lo, d0); break;
        default: issue_tile_mmas<4>(leader, acc_addr, a_hi, a_lo, d0); break;
      }
      umma_commit_elected(leader, &sh.stage_free[s]);
      umma_commit_elected(leader, &sh.acc_full[b]);
    }
  } else {
    // ===================== selection warps =====================
    // group g = warp / 4 takes the tiles j = g, g + groups, ...; inside a group warp (w & 3) owns TMEM
    // lanes 32 (w & 3) .. +31, thread = one query.  More groups = more warps per scheduler to hide
    // the latency of this branchy code; every group keeps its own candidate list per query.
    const int grp = warp >> 2, sub = warp & 3;
    const int tq = sub * 32 + lane;                      // query (TMEM lane) of this thread
    const size_t row0 = (((size_t)blockIdx.x * groups + grp) * TK_M + sub * 32) * TK_CAP;   // this warp's 32 lists
    float* const rs_w = sc.cand_s + row0;
    int* const ri_w = sc.cand_i + row0;
    float* const rs = rs_w + (size_t)lane * TK_CAP;
    int* const ri = ri_w + (size_t)lane * TK_CAP;
    volatile uint32_t* theta_g = sc.theta + q0 + tq;
    const bool q_ok = tq < q_valid;
    const float qlen = sh.qlen[tq];
    int cnt = 0;
    float theta = q_ok ? -INFINITY : INFINITY;           // padding queries never accept anything
    unsigned long long tiles_done = 0, accepted = 0, prunes = 0;
    for (int j = grp; j < my_tiles; j += groups) {
      const int b = j % TP_ACC;
      // the bound published by the other groups / splits: the load is issued before the wait so that
      // its latency hides behind it (it was 7 % of the selection warps' samples after the wait)
      const uint32_t theta_key = *theta_g;
      // wait for tile j, or learn that the producer stopped before it
      bool have = false;
      while (true) {
        if (mbar_try_warp(&sh.acc_full[b], (j / TP_ACC) & 1)) { have = true; break; }
        const int fin = sh.issued_final;
        if (__any_sync(0xffffffffu, fin >= 0 && j >= fin)) break;
      }
      if (!have) break;
      tc_fence_after();
      const long long tile = split + (long long)j * n_splits;
      const long long i0 = tile * TK_N;
      const int i_valid = (int)min((long long)TK_N, n_items - i0);
      float v0[32], v1[32];
      const uint32_t lane_addr = tmem_base + ((uint32_t)(sub * 32) << 16) + (uint32_t)b * TK_N;
      tmem_ld32(lane_addr, v0);
      tmem_ld32(lane_addr + 32, v1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh.acc_empty[b]);
      if (q_ok) theta = fmaxf(theta, key2f(theta_key));
      if (i_valid < TK_N) {                              // ragged last tile
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
          if (jj >= i_valid) v0[jj] = -INFINITY;
          if (32 + jj >= i_valid) v1[jj] = -INFINITY;
        }
      }
      const int before = cnt;
      select32(v0, theta, (int)i0, rs, ri, cnt);
      select32(v1, theta, (int)i0 + 32, rs, ri, cnt);
      accepted += (unsigned)(cnt - before);
      // Prune warp-synchronously and cooperatively: when any lane's list is about to overflow, every
      // list of the warp with more than k entries is pruned, one list at a time by all 32 lanes.
      // (Per-lane pruning was 90 % of the kernel's instructions in the first ncu capture; per-thread
      // pruning in lockstep was bound by 28 dependent passes over L2-resident lists in the third.)
      if (__any_sync(0xffffffffu, cnt > TK_CAP - TK_N)) {
        __syncwarp();
        for (int src = 0; src < 32; ++src) {
          const int c = __shfl_sync(0xffffffffu, cnt, src);
          if (c > k) {
            const float th = prune_row(rs_w + (size_t)src * TK_CAP, ri_w + (size_t)src * TK_CAP, c, k, lane);
            if (lane == src) {
              cnt = k;
              theta = fmaxf(theta, th);
              ++prunes;
              atomicMax((uint32_t*)theta_g, f2key(theta));
            }
          }
        }
        __syncwarp();
      }
      ++tiles_done;
      // LEMP bucket test for the tiles after this one (lengths descend, theta only grows)
      if (use_pruning && j + 1 < my_tiles) {
        const long long nxt = (split + (long long)(j + 1) * n_splits) * TK_N;
        const bool can_skip = !q_ok || (theta > -INFINITY && qlen * item_len[nxt] <= theta);
        if (__all_sync(0xffffffffu, can_skip) && lane == 0 && sh.skip_from[warp] == 0x7fffffff) sh.skip_from[warp] = j + 1;
      }
    }
    // final prune + partial list of (query, split, group), again one list at a time by the whole warp
    __syncwarp();
    for (int src = 0; src < 32; ++src) {
      const int qi = sub * 32 + src;
      if (qi >= q_valid) break;                          // warp-uniform
      int c = __shfl_sync(0xffffffffu, cnt, src);
      float* r_s = rs_w + (size_t)src * TK_CAP;
      int* r_i = ri_w + (size_t)src * TK_CAP;
      if (c > k) {
        prune_row(r_s, r_i, c, k, lane);
        c = k;
        __syncwarp();
      }
      const size_t part = ((size_t)(q0 + qi) * n_splits * groups + (size_t)split * groups + grp) * k;
      for (int jj = lane; jj < k; jj += 32) {
        const bool in = jj < c;
        const int slot = in ? r_i[jj] : -1;
        sc.part_s[part + jj] = in ? r_s[jj] : -INFINITY;
        sc.part_i[part + jj] = in ? (item_ids ? item_ids[slot] : slot) : -1;
      }
    }
    if (stats != nullptr) {
      if (tq == 0) {
        atomicAdd(&stats[0], tiles_done);
        if (grp == 0) atomicAdd(&stats[1], (unsigned long long)my_tiles);
      }
      if (q_ok) atomicAdd(&stats[2], accepted);
      if (prunes) atomicAdd(&stats[3], prunes);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == n_sel_warps) tmem_dealloc(tmem_base, TP_TMEM_COLS);
}

// =================================================================================================
// Wide variant: 128-item tiles (UMMA N = 128), K-block-sized pipeline stages.
//
// ncu of the 64-item kernel (profiles/r2_topk): the selection warps spend 34 % of their samples waiting
// for accumulators, the MMA warp is throttled on issue, yet the tensor pipe is active only 61 % of the
// SM-active cycles - an M128 x N64 x K8 tf32 MMA is about 35 cycles of math, too short against the fixed
// cost of each instruction.  Doubling N doubles the math per instruction.  A 128-item tile with both
// tf32 images is 128 KB at rank 128, so a stage is one K block of a tile ([hi | lo], 32 KB) and the
// MMA warp accumulates over the tile's K-block stages - the K loop of a GEMM - before it commits the
// accumulator.  TMEM: two 128-column accumulators + the 2 x 128 query columns.
// =================================================================================================
constexpr int W_N = 128;              // items per tile = UMMA N (a multiple of 128)
constexpr int W_STAGES = 6;           // shared-memory stages of one K block each (32 KB)
constexpr int W_ACC = 2;              // TMEM accumulators of W_N columns
constexpr int W_CAP = 512;            // candidate slots per (CTA, query): a tile may add W_N
constexpr int W_ISSUERS = 2;           // MMA-issuing warps: ONE issuing thread cannot feed the tensor core (106 cycles per
                                      // M128 x N128 x K8 tf32 instruction against 64 with two, scripts/umma_probe.py)
constexpr int W_THREADS = (5 + W_ISSUERS) * 32;     // 4 selection warps + producer + MMA issuers
constexpr int W_SLAB = W_N * TK_KB;   // floats of one [128 x 128 B] image
constexpr uint32_t W_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(W_N >> 3) << 17) | ((uint32_t)(TK_M >> 4) << 24);

__global__ void __launch_bounds__(256)
k_pack_items_wide(const float* __restrict__ items, int stride, long long n_items, int dim, int kblks,
                  float* __restrict__ packed, const int* __restrict__ live_ids) {
  const long long tile = blockIdx.x;
  const long long i0 = tile * W_N;
  float* out = packed + tile * 2 * (long long)kblks * W_SLAB;
  const int chunks_per_row = kblks * 8;
  for (int i = threadIdx.x; i < W_N * chunks_per_row; i += blockDim.x) {
    const int r = i / chunks_per_row, cc = i - r * chunks_per_row;
    const int kb = cc >> 3, c = cc & 7;
    const int col = kb * TK_KB + c * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i0 + r < n_items && col < dim) {
      const float* p = items + (i0 + r) * stride + col;
      if (col + 3 < dim) x = *reinterpret_cast<const float4*>(p);
      else {
        x.x = p[0];
        if (col + 1 < dim) x.y = p[1];
        if (col + 2 < dim) x.z = p[2];
      }
    }
    // live mask as arithmetic: column `dim` of a dead slot is -1e30, of a live one 0; the caller gives every
    // query a 1 in that column, so a dead slot scores about -1e30 and a live one exactly its inner product
    if (live_ids != nullptr && i0 + r < n_items && dim >= col && dim < col + 4 && live_ids[i0 + r] < 0) {
      const float dead = -1e30f;
      if (dim == col) x.x = dead; else if (dim == col + 1) x.y = dead; else if (dim == col + 2) x.z = dead; else x.w = dead;
    }
    float4 h, l;
    h.x = to_tf32(x.x); h.y = to_tf32(x.y); h.z = to_tf32(x.z); h.w = to_tf32(x.w);
    l.x = to_tf32(x.x - h.x); l.y = to_tf32(x.y - h.y); l.z = to_tf32(x.z - h.z); l.w = to_tf32(x.w - h.w);
    // stage (tile, kb) = [hi slab | lo slab], rows swizzled like every K-major SWIZZLE_128B operand
    const int off = kb * 2 * W_SLAB + r * TK_KB + ((c ^ (r & 7)) << 2);
    *reinterpret_cast<float4*>(out + off) = h;
    *reinterpret_cast<float4*>(out + off + W_SLAB) = l;
  }
}

struct WideShared {
  uint64_t b_full[W_STAGES];
  uint64_t stage_free[W_STAGES];
  uint64_t acc_full[W_ACC];
  uint64_t acc_empty[W_ACC];
  uint32_t tmem;
  volatile int skip_from[4];
  volatile int issued_final;
  float qlen[TK_M];
};

// Work split of the wide kernel.  Every query tile is cut into V "virtual splits" (virtual split v of a
// query tile = item tiles v, v + V, v + 2V, ... - interleaved so that each sees the long items first),
// and CTA c takes the `share` consecutive virtual splits [c * share, (c + 1) * share) of the whole list
// (query tile major).  share = 1, V = n_splits is the classic "one CTA per (query tile, split)" plan;
// V = 37, share = 8 puts 32 query tiles on exactly 148 SMs (37 * 32 = 148 * 8) instead of leaving 20 idle.
// A CTA whose range crosses a query-tile boundary works in SEGMENTS: it finishes its part of one query
// tile (partial list written to its slot), re-stages the next tile's queries in TMEM and goes on; the
// barrier / accumulator / stage ordinals simply keep counting.
struct WideSplit {
  int V, share, n_slots;
};

__global__ void __launch_bounds__(W_THREADS, 1)
k_lemp_topk_wide(const float* __restrict__ queries, int q_stride, long long n_q,
                 const float* __restrict__ packed, const float* __restrict__ item_len,
                 const int* __restrict__ item_ids, long long n_items, int dim, int kblks, int k,
                 const WideSplit ws, int n_q_tiles, int use_pruning, TopkScratch sc,
                 unsigned long long* __restrict__ stats) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  float* sB = (float*)base;                 // W_STAGES x [hi slab | lo slab]
  __shared__ WideShared sh;

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const long long n_tiles = (n_items + W_N - 1) / W_N;
  constexpr uint32_t STAGE_BYTES = 2u * W_SLAB * sizeof(float);
  const long long tile_floats = 2ll * kblks * W_SLAB;
  const long long v_total = (long long)n_q_tiles * ws.V;
  long long v_cur = (long long)blockIdx.x * ws.share;
  const long long v_end = min(v_total, v_cur + ws.share);

  if (warp == 4) tmem_alloc((uint32_t*)&sh.tmem, TP_TMEM_COLS);
  if (t == 0) {
    for (int s = 0; s < W_STAGES; ++s) {
      mbar_init(&sh.b_full[s], 1);
      mbar_init(&sh.stage_free[s], 1);
    }
    for (int b = 0; b < W_ACC; ++b) {
      mbar_init(&sh.acc_full[b], 1);
      mbar_init(&sh.acc_empty[b], 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh.tmem;
  const uint32_t tmem_a_hi = tmem_base + TP_A_COL, tmem_a_lo = tmem_base + TP_A_COL + 128;
  long long g_base = 0;                     // stages issued by the segments before this one
  long long a_base = 0;                     // tiles (accumulator uses) of the segments before this one
  unsigned long long tiles_done = 0, accepted = 0, prunes = 0, tiles_planned = 0;

  while (v_cur < v_end) {
    // ---- this segment: virtual splits [lv0, lv0 + nv) of query tile qt -----------------------------
    const int qt = (int)(v_cur / ws.V);
    const int lv0 = (int)(v_cur - (long long)qt * ws.V);
    const int nv = (int)min((long long)ws.V - lv0, v_end - v_cur);
    const long long q0 = (long long)qt * TK_M;
    const int q_valid = (int)min((long long)TK_M, n_q - q0);
    const int rows = (int)((n_tiles + ws.V - 1) / ws.V);            // item tiles of the longest virtual split
    int my_tiles = 0;
    if (rows > 0) {
      const long long last = n_tiles - (long long)(rows - 1) * ws.V - lv0;     // valid ones in the last row
      my_tiles = (rows - 1) * nv + (int)max(0ll, min((long long)nv, last));
    }
    // ordinal m of the segment -> item tile (row major: all of the segment's virtual splits, then the next row)
    auto tile_of = [&](int m) -> long long { return lv0 + (m % nv) + (long long)(m / nv) * ws.V; };
    const int slot = (int)(blockIdx.x - ((long long)qt * ws.V) / ws.share);     // which partial list of the query tile
    const bool last_of_qt = lv0 + nv == ws.V;

    if (t == 0) {
      sh.issued_final = -1;
      for (int w = 0; w < 4; ++w) sh.skip_from[w] = 0x7fffffff;
    }
    // queries -> TMEM (thread t = lane t = query t), split into tf32 hi / lo on the way
    if (warp < 4) {
      const float* qp = queries + (q0 + t) * q_stride;
      const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
      float s = 0.f;
      for (int c0 = 0; c0 < kblks * TK_KB; c0 += 32) {
        float hi[32], lo[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int col = c0 + j;
          const float x = (t < q_valid && col < dim) ? qp[col] : 0.f;
          s = fmaf(x, x, s);
          hi[j] = to_tf32(x);
          lo[j] = to_tf32(x - hi[j]);
        }
        tmem_st32(tmem_a_hi + lane_off + c0, hi);
        tmem_st32(tmem_a_lo + lane_off + c0, lo);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      sh.qlen[t] = sqrtf(s);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp == 4) {
      // ===================== producer: one bulk copy pair per (tile, K block) =====================
      if (lane == 0) {
        int issued = 0;
        long long g = g_base;
        for (int j = 0; j < my_tiles; ++j) {
          if (use_pruning) {
            bool all = true;
            for (int sub = 0; sub < 4; ++sub) all = all && (sh.skip_from[sub] <= j);
            if (all) break;                                                      // LEMP early exit
          }
          const float* src = packed + tile_of(j) * tile_floats;
          for (int kb = 0; kb < kblks; ++kb, ++g) {
            const int s = (int)(g % W_STAGES);
            const long long n = g / W_STAGES;
            if (n > 0) mbar_wait(&sh.stage_free[s], (uint32_t)((n - 1) & 1));
            float* dst = sB + (size_t)s * 2 * W_SLAB;
            mbar_expect_tx(&sh.b_full[s], STAGE_BYTES);
            tma_bulk_g2s(dst, src + (size_t)kb * 2 * W_SLAB, W_SLAB * 4, &sh.b_full[s]);
            tma_bulk_g2s(dst + W_SLAB, src + (size_t)kb * 2 * W_SLAB + W_SLAB, W_SLAB * 4, &sh.b_full[s]);
          }
          issued = j + 1;
        }
        __threadfence_block();
        sh.issued_final = issued;
      }
    } else if (warp >= 5) {
      // ===================== MMA issuers (converged warps, one elected lane each issues) ===============
      // Issuer w takes the tiles whose accumulator is w (tile ordinals alternate), so both have instructions
      // in flight at the same time; each commits its own stages and its own accumulator.
      const int issuer = warp - 5;
      const uint32_t leader = elect_one();
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t a_hi = tb + TP_A_COL, a_lo = tb + TP_A_COL + 128;
      long long g = g_base;
      for (int j = 0; j < my_tiles; ++j) {
        const long long a = a_base + j;
        const int b = (int)(a % W_ACC);
        if ((int)(a % W_ISSUERS) != issuer) { g += kblks; continue; }
        bool have = true;
        for (int kb = 0; kb < kblks; ++kb, ++g) {
          const int s = (int)(g % W_STAGES);
          const uint32_t par = (uint32_t)((g / W_STAGES) & 1);
          if (kb == 0) {
            // the first stage of a tile may never come (the producer stopped): watch issued_final too
            have = false;
            while (true) {
              if (mbar_try_warp(&sh.b_full[s], par)) { have = true; break; }
              const int fin = sh.issued_final;
              if (__any_sync(0xffffffffu, fin >= 0 && j >= fin)) break;
            }
            have = __shfl_sync(0xffffffffu, have ? 1 : 0, 0) != 0;
            if (!have) break;
            if (a >= W_ACC) mbar_wait_warp(&sh.acc_empty[b], (uint32_t)(((a / W_ACC) - 1) & 1));   // accumulator drained
          } else {
            mbar_wait_warp(&sh.b_full[s], par);
          }
          tc_fence_after();
          const uint32_t acc_addr = tb + (uint32_t)b * W_N;
          const uint64_t d0 = make_desc(smem_u32(sB + (size_t)s * 2 * W_SLAB));
          constexpr uint64_t LO_OFF = ((uint64_t)W_SLAB * 4) >> 4;
#pragma unroll
          for (int ks = 0; ks < TK_KB / 8; ++ks) {
            const uint32_t a_col = (uint32_t)(kb * TK_KB + ks * 8);
            const uint64_t bh = d0 + (uint64_t)(ks * 2), bl = bh + LO_OFF;
            if (kb == 0 && ks == 0) umma_tf32_ts_i<false>(leader, acc_addr, a_lo + a_col, bh, W_IDESC);   // small terms first
            else umma_tf32_ts_i<true>(leader, acc_addr, a_lo + a_col, bh, W_IDESC);
            umma_tf32_ts_i<true>(leader, acc_addr, a_hi + a_col, bl, W_IDESC);
            umma_tf32_ts_i<true>(leader, acc_addr, a_hi + a_col, bh, W_IDESC);
          }
          umma_commit_elected(leader, &sh.stage_free[s]);
        }
        if (!have) break;
        umma_commit_elected(leader, &sh.acc_full[b]);
      }
    } else {
      // ===================== selection warps: thread = TMEM lane = query =====================
      const int sub = warp & 3;
      const int tq = sub * 32 + lane;
      const size_t row0 = ((size_t)blockIdx.x * TK_M + sub * 32) * W_CAP;       // this warp's 32 lists
      float* const rs_w = sc.cand_s + row0;
      int* const ri_w = sc.cand_i + row0;
      float* const rs = rs_w + (size_t)lane * W_CAP;
      int* const ri = ri_w + (size_t)lane * W_CAP;
      volatile uint32_t* theta_g = sc.theta + q0 + tq;
      const bool q_ok = tq < q_valid;
      const float qlen = sh.qlen[tq];
      int cnt = 0;
      float theta = q_ok ? -INFINITY : INFINITY;           // padding queries never accept anything
      for (int j = 0; j < my_tiles; ++j) {
        const long long a = a_base + j;
        const int b = (int)(a % W_ACC);
        const uint32_t theta_key = *theta_g;
        bool have = false;
        while (true) {
          if (mbar_try_warp(&sh.acc_full[b], (uint32_t)((a / W_ACC) & 1))) { have = true; break; }
          const int fin = sh.issued_final;
          if (__any_sync(0xffffffffu, fin >= 0 && j >= fin)) break;
        }
        if (!have) break;
        tc_fence_after();
        const long long i0 = tile_of(j) * W_N;
        const int i_valid = (int)min((long long)W_N, n_items - i0);
        const uint32_t lane_addr = tmem_base + ((uint32_t)(sub * 32) << 16) + (uint32_t)b * W_N;
        if (q_ok) theta = fmaxf(theta, key2f(theta_key));
        const int before = cnt;
        // the tile's W_N columns in halves of 128 (register budget); the four loads of a half are issued
        // back to back and waited for once; the accumulator is released after the last half's loads
#pragma unroll 1
        for (int half = 0; half < W_N / 128; ++half) {
          float v[4][32];
#pragma unroll
          for (int h = 0; h < 4; ++h) tmem_ld32_nowait(lane_addr + 128 * half + 32 * h, v[h]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (half == W_N / 128 - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sh.acc_empty[b]);       // the next tile's MMAs may overwrite it now
          }
          if (i_valid < W_N) {                                 // ragged last tile
#pragma unroll
            for (int h = 0; h < 4; ++h)
#pragma unroll
              for (int jj = 0; jj < 32; ++jj)
                if (128 * half + 32 * h + jj >= i_valid) v[h][jj] = -INFINITY;
          }
#pragma unroll
          for (int h = 0; h < 4; ++h) select32(v[h], theta, (int)i0 + 128 * half + 32 * h, rs, ri, cnt);
        }
        accepted += (unsigned)(cnt - before);
        if (__any_sync(0xffffffffu, cnt > W_CAP - W_N)) {
          __syncwarp();
          for (int src = 0; src < 32; ++src) {
            const int c = __shfl_sync(0xffffffffu, cnt, src);
            if (c > k) {
              const float th = prune_row<W_CAP>(rs_w + (size_t)src * W_CAP, ri_w + (size_t)src * W_CAP, c, k, lane);
              if (lane == src) {
                cnt = k;
                theta = fmaxf(theta, th);
                ++prunes;
                atomicMax((uint32_t*)theta_g, f2key(theta));
              }
            }
          }
          __syncwarp();
        }
        if (tq == 0) ++tiles_done;
        if (use_pruning && j + 1 < my_tiles) {
          const long long nxt = tile_of(j + 1) * W_N;
          const bool can_skip = !q_ok || (theta > -INFINITY && qlen * item_len[nxt] <= theta);
          if (__all_sync(0xffffffffu, can_skip) && lane == 0 && sh.skip_from[warp] == 0x7fffffff) sh.skip_from[warp] = j + 1;
        }
      }
      // final prune + partial list of (query, slot), one list at a time by the whole warp; the CTA that
      // holds the last part of a query tile also blanks the slots nobody owns
      __syncwarp();
      for (int src = 0; src < 32; ++src) {
        const int qi = sub * 32 + src;
        if (qi >= q_valid) break;                          // warp-uniform
        int c = __shfl_sync(0xffffffffu, cnt, src);
        float* r_s = rs_w + (size_t)src * W_CAP;
        int* r_i = ri_w + (size_t)src * W_CAP;
        if (c > k) {
          prune_row<W_CAP>(r_s, r_i, c, k, lane);
          c = k;
          __syncwarp();
        }
        const size_t part = ((size_t)(q0 + qi) * ws.n_slots + (size_t)slot) * k;
        for (int jj = lane; jj < k; jj += 32) {
          const bool in = jj < c;
          const int it = in ? r_i[jj] : -1;
          sc.part_s[part + jj] = in ? r_s[jj] : -INFINITY;
          sc.part_i[part + jj] = in ? (item_ids ? item_ids[it] : it) : -1;
        }
        if (last_of_qt) {
          for (int e = lane + (slot + 1) * k; e < ws.n_slots * k; e += 32) {
            sc.part_s[(size_t)(q0 + qi) * ws.n_slots * k + e] = -INFINITY;
            sc.part_i[(size_t)(q0 + qi) * ws.n_slots * k + e] = -1;
          }
        }
      }
      if (tq == 0) tiles_planned += (unsigned long long)my_tiles;
    }
    // ---- end of segment: everybody agrees on how far the pipeline got, then the next query tile ----
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const int fin = sh.issued_final;
    g_base += (long long)fin * kblks;
    a_base += fin;
    v_cur += nv;
    __syncthreads();                                       // nobody may reset issued_final before all have read it
  }
  if (stats != nullptr && warp < 4) {
    if (t == 0) {
      atomicAdd(&stats[0], tiles_done);
      atomicAdd(&stats[1], tiles_planned);
    }
    if (accepted) atomicAdd(&stats[2], accepted);
    if (prunes) atomicAdd(&stats[3], prunes);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, TP_TMEM_COLS);
}

// Scratch plan of the wide kernel: 128-item tiles, W_CAP slots per list, one selection group.  With ten
// or more query tiles that do not divide the SMs evenly, the virtual-split scheme above fills every SM.
struct WidePlan {
  Plan p;
  WideSplit ws;
};
WidePlan make_plan_wide(long long n_q, int k, long long n_items, int dim) {
  WidePlan wp;
  Plan& p = wp.p;
  p.n_q_tiles = (int)((n_q + TK_M - 1) / TK_M);
  const long long n_tiles = (n_items + W_N - 1) / W_N;
  const int sms = fps_num_sms();
  const int qt = p.n_q_tiles > 0 ? p.n_q_tiles : 1;
  long long splits = sms / qt;
  if (splits < 1) splits = 1;
  if (splits > 16) splits = 16;
  if (splits > n_tiles) splits = n_tiles > 0 ? n_tiles : 1;
  wp.ws = WideSplit{(int)splits, 1, (int)splits};
  p.n_ctas = qt * (int)splits;
  // opt-in (FPS_TOPK_FILL=1): measured on config 6, 148 CTAs of 8/37 of a query tile take 5.85 ms against
  // 5.85 ms for 128 CTAs of 1/4 (3.45 vs 3.32 ms pruned).  The SM clock sits at 1.77 of 1.97 GHz during the
  // kernel (ncu): the tensor work is power-limited, more busy SMs only lower the clock further.
  static const bool fill = []() { const char* e = getenv("FPS_TOPK_FILL"); return e && e[0] == '1'; }();
  if (fill && qt >= 10 && qt < sms && sms % qt != 0) {
    auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
    const long long V = sms / gcd(sms, qt);               // qt * V is a multiple of the SM count
    const long long share = (long long)qt * V / sms;
    const long long slots = (V + share - 1) / share + 1;  // CTAs that can touch one query tile
    if (V <= n_tiles && slots * k <= 8192 && slots <= 16) {
      wp.ws = WideSplit{(int)V, (int)share, (int)slots};
      p.n_ctas = sms;
    }
  }
  p.n_splits = wp.ws.n_slots;
  p.kblks = (dim + TK_KB - 1) / TK_KB;
  auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
  size_t o = 0;
  o = align(o + (size_t)p.n_ctas * W_CAP * TK_M * sizeof(float));
  p.off_cand_i = o;
  o = align(o + (size_t)p.n_ctas * W_CAP * TK_M * sizeof(int));
  p.off_theta = o;
  o = align(o + (size_t)p.n_q_tiles * TK_M * sizeof(uint32_t));
  p.off_part_s = o;
  o = align(o + (size_t)p.n_q_tiles * TK_M * p.n_splits * k * sizeof(float));
  p.off_part_i = o;
  o = align(o + (size_t)p.n_q_tiles * TK_M * p.n_splits * k * sizeof(int));
  p.total = o;
  return wp;
}

int packed_groups() {
  static int g = -1;
  if (g < 0) {
    const char* e = getenv("FPS_TOPK_GROUPS");
    int v = e ? atoi(e) : TP_DEFAULT_GROUPS;
    g = (v == 1 || v == 2 || v == 4) ? v : TP_DEFAULT_GROUPS;
  }
  return g;
}
int packed_forced_splits() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FPS_TOPK_SPLITS");
    v = e ? atoi(e) : 0;
    if (v < 0) v = 0;
  }
  return v;
}
// the richest plan that fits the caller's scratch: all selection groups + multi-wave splits first, then
// single-wave splits, then fewer groups (fps_topk_scratch_bytes covers one group, single wave)
Plan packed_plan(long long n_q, int k, long long n_items, int dim, size_t scratch_bytes, int* groups_out) {
  int groups = packed_groups();
  Plan p;
  while (true) {
    // multi-wave splits are opt-in (FPS_TOPK_SPLITS): measured on config 6, 9 splits (1.95 waves, all SMs
    // busy) take 6.96 ms against 6.19 ms for 4 splits (128 of 148 SMs) - each extra list per query
    // adds accepted candidates and prunes that cost more than the idle SMs
    p = make_plan(n_q, k, n_items, dim, groups, 0, packed_forced_splits());
    if (scratch_bytes >= p.total && (long long)p.n_splits * groups * k <= 8192) break;
    p = make_plan(n_q, k, n_items, dim, groups, 0, 0);
    if ((scratch_bytes >= p.total && (long long)p.n_splits * groups * k <= 8192) || groups == 1) break;
    groups >>= 1;
  }
  *groups_out = groups;
  return p;
}

}  // namespace

extern "C" {

size_t fps_topk_packed_bytes(int64_t n_items, int32_t dim) {
  if (n_items <= 0 || dim <= 0 || dim > 128) return 0;
  const int kblks = (dim + TK_KB - 1) / TK_KB;
  const size_t n_tiles = (size_t)((n_items + TK_N - 1) / TK_N);
  return n_tiles * 2 * (size_t)TK_N * kblks * TK_KB * sizeof(float);
}

size_t fps_topk_packed_scratch_bytes(int64_t n_q, int32_t k, int64_t n_items) {
  if (n_q <= 0 || k <= 0 || n_items <= 0) return 0;
  int groups;
  return packed_plan(n_q, k, n_items, 128, (size_t)-1, &groups).total;
}

int fps_topk_pack_items(const float* d_items, int32_t item_stride, int64_t n_items, int32_t dim, void* d_packed,
                        void* stream) {
  FPS_REQUIRE(d_items && d_packed && n_items > 0 && dim > 0 && dim <= 128 && item_stride % 4 == 0, FPS_E_INVALID,
              "fps_topk_pack_items: bad arguments (rank must be <= 128, stride a multiple of 4)");
  const int kblks = (dim + TK_KB - 1) / TK_KB;
  const long long n_tiles = (n_items + TK_N - 1) / TK_N;
  k_pack_items<<<(unsigned)n_tiles, 256, 0, (cudaStream_t)stream>>>(d_items, item_stride, n_items, dim, kblks, (float*)d_packed);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_lemp_topk_packed(const float* d_queries, int32_t q_stride, int64_t n_q, const void* d_packed,
                         const float* d_item_len, const int32_t* d_item_ids, int64_t n_items, int32_t dim,
                         int32_t k, int32_t use_pruning, void* d_scratch, size_t scratch_bytes,
                         float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats, void* stream) {
  FPS_REQUIRE(n_q > 0 && n_items > 0 && dim > 0, FPS_E_INVALID, "fps_lemp_topk_packed: empty problem");
  FPS_REQUIRE(k >= 1 && k <= TK_MAX_K, FPS_E_TOO_LARGE, "fps_lemp_topk_packed: k = %d outside [1, %d]", k, TK_MAX_K);
  FPS_REQUIRE(dim <= 128 && q_stride % 4 == 0, FPS_E_UNSUPPORTED, "fps_lemp_topk_packed: rank <= 128 and q_stride %% 4 == 0 required");
  FPS_REQUIRE(!use_pruning || d_item_len != nullptr, FPS_E_INVALID, "fps_lemp_topk_packed: pruning needs item lengths");
  cudaStream_t st = (cudaStream_t)stream;
  // More selection groups hide more latency but every group keeps its own candidate list; fall back to
  // fewer groups when the caller's scratch was sized for fewer (fps_topk_scratch_bytes = one group).
  int groups;
  const Plan p = packed_plan(n_q, k, n_items, dim, scratch_bytes, &groups);
  FPS_REQUIRE(scratch_bytes >= p.total, FPS_E_INVALID, "fps_lemp_topk_packed: scratch too small (%zu < %zu)", scratch_bytes, p.total);
  uint8_t* sb = (uint8_t*)d_scratch;
  TopkScratch sc;
  sc.cand_s = (float*)sb;
  sc.cand_i = (int*)(sb + p.off_cand_i);
  sc.theta = (uint32_t*)(sb + p.off_theta);
  sc.part_s = (float*)(sb + p.off_part_s);
  sc.part_i = (int*)(sb + p.off_part_i);
  FPS_CUDA(cudaMemsetAsync(sc.theta, 0, (size_t)p.n_q_tiles * TK_M * sizeof(uint32_t), st));
  const size_t b_bytes = (size_t)TK_N * p.kblks * TK_KB * 4;
  const size_t smem = 1024 + (size_t)TP_STAGES * 2 * b_bytes;
  FPS_CUDA(cudaFuncSetAttribute(k_lemp_topk_packed, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_lemp_topk_packed<<<p.n_ctas, (4 * groups + 2) * 32, smem, st>>>(d_queries, q_stride, n_q, (const float*)d_packed,
                                                                    d_item_len, d_item_ids, n_items, dim, p.kblks, k,
                                                                    p.n_splits, groups, use_pruning, sc,
                                                                    (unsigned long long*)d_stats);
  FPS_LAUNCH_CHECK();
  return launch_merge(sc.part_s, sc.part_i, p.n_splits * groups, n_q, k, k, nullptr, 0, d_out_scores, d_out_ids, st);
}

/* ---- wide variant (128-item tiles) ------------------------------------------------------------- */
size_t fps_topk_wide_packed_bytes(int64_t n_items, int32_t dim) {
  if (n_items <= 0 || dim <= 0 || dim > 128) return 0;
  const int kblks = (dim + TK_KB - 1) / TK_KB;
  const size_t n_tiles = (size_t)((n_items + W_N - 1) / W_N);
  return n_tiles * 2 * (size_t)kblks * W_SLAB * sizeof(float);
}

size_t fps_topk_wide_scratch_bytes(int64_t n_q, int32_t k, int64_t n_items) {
  if (n_q <= 0 || k <= 0 || n_items <= 0) return 0;
  return make_plan_wide(n_q, k, n_items, 128).p.total;
}

int fps_topk_pack_items_wide(const float* d_items, int32_t item_stride, int64_t n_items, int32_t dim, void* d_packed,
                             void* stream) {
  FPS_REQUIRE(d_items && d_packed && n_items > 0 && dim > 0 && dim <= 128 && item_stride % 4 == 0, FPS_E_INVALID,
              "fps_topk_pack_items_wide: bad arguments (rank must be <= 128, stride a multiple of 4)");
  const int kblks = (dim + TK_KB - 1) / TK_KB;
  const long long n_tiles = (n_items + W_N - 1) / W_N;
  k_pack_items_wide<<<(unsigned)n_tiles, 256, 0, (cudaStream_t)stream>>>(d_items, item_stride, n_items, dim, kblks, (float*)d_packed, nullptr);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_topk_pack_items_wide_live(const float* d_items, int32_t item_stride, int64_t n_items, int32_t dim,
                                  const int32_t* d_live_ids, void* d_packed, void* stream) {
  FPS_REQUIRE(d_items && d_packed && d_live_ids && n_items > 0 && dim > 0 && dim < 128 && item_stride % 4 == 0, FPS_E_INVALID,
              "fps_topk_pack_items_wide_live: bad arguments (rank must be < 128: one more column carries the mask)");
  const int kblks = (dim + 1 + TK_KB - 1) / TK_KB;
  const long long n_tiles = (n_items + W_N - 1) / W_N;
  k_pack_items_wide<<<(unsigned)n_tiles, 256, 0, (cudaStream_t)stream>>>(d_items, item_stride, n_items, dim, kblks, (float*)d_packed,
                                                                         d_live_ids);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_lemp_topk_wide(const float* d_queries, int32_t q_stride, int64_t n_q, const void* d_packed,
                       const float* d_item_len, const int32_t* d_item_ids, int64_t n_items, int32_t dim,
                       int32_t k, int32_t use_pruning, void* d_scratch, size_t scratch_bytes,
                       float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats, void* stream) {
  FPS_REQUIRE(n_q > 0 && n_items > 0 && dim > 0, FPS_E_INVALID, "fps_lemp_topk_wide: empty problem");
  FPS_REQUIRE(k >= 1 && k <= TK_MAX_K, FPS_E_TOO_LARGE, "fps_lemp_topk_wide: k = %d outside [1, %d]", k, TK_MAX_K);
  FPS_REQUIRE(dim <= 128 && q_stride % 4 == 0, FPS_E_UNSUPPORTED, "fps_lemp_topk_wide: rank <= 128 and q_stride %% 4 == 0 required");
  FPS_REQUIRE(!use_pruning || d_item_len != nullptr, FPS_E_INVALID, "fps_lemp_topk_wide: pruning needs item lengths");
  cudaStream_t st = (cudaStream_t)stream;
  const WidePlan wp = make_plan_wide(n_q, k, n_items, dim);
  const Plan& p = wp.p;
  FPS_REQUIRE(scratch_bytes >= p.total, FPS_E_INVALID, "fps_lemp_topk_wide: scratch too small (%zu < %zu)", scratch_bytes, p.total);
  uint8_t* sb = (uint8_t*)d_scratch;
  TopkScratch sc;
  sc.cand_s = (float*)sb;
  sc.cand_i = (int*)(sb + p.off_cand_i);
  sc.theta = (uint32_t*)(sb + p.off_theta);
  sc.part_s = (float*)(sb + p.off_part_s);
  sc.part_i = (int*)(sb + p.off_part_i);
  FPS_CUDA(cudaMemsetAsync(sc.theta, 0, (size_t)p.n_q_tiles * TK_M * sizeof(uint32_t), st));
  const size_t smem = 1024 + (size_t)W_STAGES * 2 * W_SLAB * sizeof(float);
  FPS_CUDA(cudaFuncSetAttribute(k_lemp_topk_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_lemp_topk_wide<<<p.n_ctas, W_THREADS, smem, st>>>(d_queries, q_stride, n_q, (const float*)d_packed, d_item_len,
                                                      d_item_ids, n_items, dim, p.kblks, k, wp.ws, p.n_q_tiles,
                                                      use_pruning, sc, (unsigned long long*)d_stats);
  FPS_LAUNCH_CHECK();
  return launch_merge(sc.part_s, sc.part_i, p.n_splits, n_q, k, k, nullptr, 0, d_out_scores, d_out_ids, st);
}

}  // extern "C"
