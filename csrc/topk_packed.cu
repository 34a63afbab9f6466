// LEMP top-K, second generation: pre-packed item tiles streamed by the TMA engine into a
// warp-specialised tcgen05 pipeline.
//
// topk.cu stages every item tile with CUDA-core loads + tf32 splitting inside the scoring loop;
// ncu showed the tensor pipe 4 % busy and the staging arithmetic on top of the stall list.  Items
// change rarely compared with how often they are scored (a recommender answers many query batches
// per model refresh), so here the hi/lo split and the SWIZZLE_128B layout are computed once by
// `fps_topk_pack_items` into an image that is byte-for-byte what the UMMA descriptors expect in
// shared memory.  Scoring then is:
//   warp 4 (one elected lane): cp.async.bulk (TMA, UBLKCP) of the next packed tile into shared
//       memory, completion on an mbarrier; tcgen05.mma (3 x kind::tf32 per K step) into one of two
//       TMEM accumulators; tcgen05.commit to that accumulator's mbarrier.
//   warps 0-3: wait for the accumulator, tcgen05.ld it (thread t = TMEM lane t = query t), hand the
//       TMEM buffer back, then run the LEMP selection (threshold append + pruning) while the
//       next tile's copy and MMAs are already in flight.
// The LEMP bucket test (||q|| * tileMaxLen <= theta for every query of the CTA) is evaluated by the
// selection warps and published per warp; the issuer stops fetching tiles once all four agree.
#include "topk_core.cuh"

namespace {

constexpr int TP_THREADS = 192;       // 4 selection warps + TMA producer warp + MMA issuer warp
constexpr int TP_TMEM_COLS = 512;     // 2 x 64 accumulator columns + 2 x 128 query (hi, lo) columns
constexpr int TP_A_COL = 128;         // first TMEM column of the query operand
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
  uint64_t acc_full[2];            // MMAs of the tile using accumulator b are complete
  uint64_t acc_empty[2];           // all four selection warps have read accumulator b
  uint32_t tmem;
  volatile int skip_from[4];       // per selection warp: first tile ordinal that may be skipped
  volatile int issued_final;       // number of tiles the producer really fetched (-1 while running)
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
__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__global__ void __launch_bounds__(TP_THREADS, 1)
k_lemp_topk_packed(const float* __restrict__ queries, int q_stride, long long n_q,
                   const float* __restrict__ packed, const float* __restrict__ item_len,
                   const int* __restrict__ item_ids, long long n_items, int dim, int kblks, int k, int n_splits,
                   int use_pruning, TopkScratch sc, unsigned long long* __restrict__ stats) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int b_floats = TK_N * kblks * TK_KB;
  float* sB = (float*)base;                 // TP_STAGES x [hi | lo] packed tiles
  __shared__ PackedShared sh;

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int qt = blockIdx.x / n_splits, split = blockIdx.x - qt * n_splits;
  const long long q0 = (long long)qt * TK_M;
  const int q_valid = (int)min((long long)TK_M, n_q - q0);
  const long long n_tiles = (n_items + TK_N - 1) / TK_N;
  const int my_tiles = n_tiles > split ? (int)((n_tiles - split + n_splits - 1) / n_splits) : 0;
  const uint32_t tile_bytes = (uint32_t)(2 * b_floats * sizeof(float));
  const int stage_floats = 2 * b_floats;

  if (warp == 4) tmem_alloc((uint32_t*)&sh.tmem, TP_TMEM_COLS);
  if (t == 0) {
    for (int s = 0; s < TP_STAGES; ++s) {
      mbar_init(&sh.b_full[s], 1);
      mbar_init(&sh.stage_free[s], 1);
    }
    mbar_init(&sh.acc_full[0], 1);
    mbar_init(&sh.acc_full[1], 1);
    mbar_init(&sh.acc_empty[0], 4);
    mbar_init(&sh.acc_empty[1], 4);
    sh.issued_final = -1;
    for (int w = 0; w < 4; ++w) sh.skip_from[w] = 0x7fffffff;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh.tmem;
  // TMEM map: columns [0,128) two accumulators, [128,256) query hi, [256,384) query lo
  const uint32_t tmem_a_hi = tmem_base + TP_A_COL, tmem_a_lo = tmem_base + TP_A_COL + 128;

  // ---- queries -> TMEM (thread t = lane t = query t), split into tf32 hi / lo on the way
  float qlen = 0.f;
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
    qlen = sqrtf(s);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 4) {
    // ===================== producer: TMA bulk copies of packed tiles =====================
    if (lane == 0) {
      int issued = 0;
      for (int j = 0; j < my_tiles; ++j) {
        const int s = j % TP_STAGES, n = j / TP_STAGES;
        if (n > 0) mbar_wait(&sh.stage_free[s], (n - 1) & 1);
        if (use_pruning) {
          const int s0 = sh.skip_from[0], s1 = sh.skip_from[1], s2 = sh.skip_from[2], s3 = sh.skip_from[3];
          if (s0 <= j && s1 <= j && s2 <= j && s3 <= j) break;                // LEMP early exit
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
  } else if (warp == 5) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      for (int j = 0; j < my_tiles; ++j) {
        const int s = j % TP_STAGES, n = j / TP_STAGES;
        bool have = false;
        while (true) {
          if (mbar_try(&sh.b_full[s], n & 1)) { have = true; break; }
          const int fin = sh.issued_final;
          if (fin >= 0 && j >= fin) break;
        }
        if (!have) break;
        if (j >= 2) mbar_wait(&sh.acc_empty[j & 1], ((j >> 1) - 1) & 1);      // accumulator drained
        tc_fence_after();
        const uint32_t acc_addr = tmem_base + (uint32_t)(j & 1) * TK_N;
        const float* bh_s = sB + (size_t)s * stage_floats;
        uint32_t acc = 0;
        for (int kb = 0; kb < kblks; ++kb) {
#pragma unroll
          for (int ks = 0; ks < TK_KB / 8; ++ks) {
            const uint32_t a_col = (uint32_t)(kb * TK_KB + ks * 8);
            const uint32_t b_off = (uint32_t)(kb * TK_N * TK_KB + ks * 8) * 4;
            const uint64_t bh = make_desc(smem_u32(bh_s) + b_off), bl = make_desc(smem_u32(bh_s + b_floats) + b_off);
            umma_tf32_ts(acc_addr, tmem_a_lo + a_col, bh, TK_IDESC, acc);   // small terms first
            umma_tf32_ts(acc_addr, tmem_a_hi + a_col, bl, TK_IDESC, 1);
            umma_tf32_ts(acc_addr, tmem_a_hi + a_col, bh, TK_IDESC, 1);
            acc = 1;
          }
        }
        umma_commit(&sh.stage_free[s]);
        umma_commit(&sh.acc_full[j & 1]);
      }
    }
  } else {
    // ===================== selection warps =====================
    float* cs = sc.cand_s + (size_t)blockIdx.x * TK_CAP * TK_M + t;
    int* ci = sc.cand_i + (size_t)blockIdx.x * TK_CAP * TK_M + t;
    volatile uint32_t* theta_g = sc.theta + q0 + t;
    int cnt = 0;
    float theta = -INFINITY;
    unsigned long long tiles_done = 0, accepted = 0, prunes = 0;
    for (int j = 0; j < my_tiles; ++j) {
      const int b = j & 1;
      // wait for tile j, or learn that the producer stopped before it
      bool have = false;
      while (true) {
        if (mbar_try(&sh.acc_full[b], (j >> 1) & 1)) { have = true; break; }
        const int fin = sh.issued_final;
        if (fin >= 0 && j >= fin) break;
      }
      if (!have) break;
      tc_fence_after();
      const long long tile = split + (long long)j * n_splits;
      const long long i0 = tile * TK_N;
      const int i_valid = (int)min((long long)TK_N, n_items - i0);
      float v0[32], v1[32];
      const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)b * TK_N;
      tmem_ld32(lane_addr, v0);
      tmem_ld32(lane_addr + 32, v1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh.acc_empty[b]);
      if (use_pruning) theta = fmaxf(theta, key2f(*theta_g));
      if (t < q_valid) {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
          if (jj < i_valid && v0[jj] > theta) { cs[cnt * TK_M] = v0[jj]; ci[cnt * TK_M] = (int)(i0 + jj); ++cnt; ++accepted; }
        }
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
          if (32 + jj < i_valid && v1[jj] > theta) { cs[cnt * TK_M] = v1[jj]; ci[cnt * TK_M] = (int)(i0 + 32 + jj); ++cnt; ++accepted; }
        }
      }
      // Prune warp-synchronously: the selection loop costs the same for 1 lane or 32, so when any
      // lane's buffer is about to overflow every lane of the warp with more than k candidates
      // prunes now.  (Per-lane pruning made the warp run this loop on ~40 % of the tiles and was
      // 90 % of the kernel's instructions in the first ncu capture.)
      if (__any_sync(0xffffffffu, cnt > TK_CAP - TK_N)) {
        if (t < q_valid && cnt > k) {
          cnt = prune_candidates(cs, ci, cnt, k, theta);
          ++prunes;
          if (use_pruning) atomicMax((uint32_t*)theta_g, f2key(theta));
        }
      }
      ++tiles_done;
      // LEMP bucket test for the tiles after this one (lengths descend, theta only grows)
      if (use_pruning && j + 1 < my_tiles) {
        const long long nxt = (split + (long long)(j + 1) * n_splits) * TK_N;
        const bool can_skip = (t >= q_valid) || (theta > -INFINITY && qlen * item_len[nxt] <= theta);
        if (__all_sync(0xffffffffu, can_skip) && lane == 0 && sh.skip_from[warp] == 0x7fffffff) sh.skip_from[warp] = j + 1;
      }
    }
    if (t < q_valid) {
      cnt = prune_candidates(cs, ci, cnt, k, theta);
      float* ps = sc.part_s + ((size_t)(q0 + t) * n_splits + split) * k;
      int* pi = sc.part_i + ((size_t)(q0 + t) * n_splits + split) * k;
      for (int j = 0; j < k; ++j) {
        if (j < cnt) {
          const int slot = ci[j * TK_M];
          ps[j] = cs[j * TK_M];
          pi[j] = item_ids ? item_ids[slot] : slot;
        } else {
          ps[j] = -INFINITY;
          pi[j] = -1;
        }
      }
    }
    if (stats != nullptr) {
      if (t == 0) {
        atomicAdd(&stats[0], tiles_done);
        atomicAdd(&stats[1], (unsigned long long)my_tiles);
      }
      atomicAdd(&stats[2], accepted);
      if (prunes) atomicAdd(&stats[3], prunes);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, TP_TMEM_COLS);
}

}  // namespace

extern "C" {

size_t fps_topk_packed_bytes(int64_t n_items, int32_t dim) {
  if (n_items <= 0 || dim <= 0 || dim > 128) return 0;
  const int kblks = (dim + TK_KB - 1) / TK_KB;
  const size_t n_tiles = (size_t)((n_items + TK_N - 1) / TK_N);
  return n_tiles * 2 * (size_t)TK_N * kblks * TK_KB * sizeof(float);
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
  const Plan p = make_plan(n_q, k, n_items, dim);
  FPS_REQUIRE(scratch_bytes >= p.total, FPS_E_INVALID, "fps_lemp_topk_packed: scratch too small (%zu < %zu)", scratch_bytes, p.total);
  uint8_t* sb = (uint8_t*)d_scratch;
  TopkScratch sc;
  sc.cand_s = (float*)sb;
  sc.cand_i = (int*)(sb + p.off_cand_i);
  sc.theta = (uint32_t*)(sb + p.off_theta);
  sc.part_s = (float*)(sb + p.off_part_s);
  sc.part_i = (int*)(sb + p.off_part_i);
  FPS_CUDA(cudaMemsetAsync(sc.theta, 0, (size_t)p.n_q_tiles * TK_M * sizeof(uint32_t), st));
  const size_t a_bytes = (size_t)TK_M * p.kblks * TK_KB * 4, b_bytes = (size_t)TK_N * p.kblks * TK_KB * 4;
  (void)a_bytes;
  const size_t smem = 1024 + (size_t)TP_STAGES * 2 * b_bytes;
  FPS_CUDA(cudaFuncSetAttribute(k_lemp_topk_packed, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_lemp_topk_packed<<<p.n_ctas, TP_THREADS, smem, st>>>(d_queries, q_stride, n_q, (const float*)d_packed, d_item_len,
                                                         d_item_ids, n_items, dim, p.kblks, k, p.n_splits,
                                                         use_pruning, sc, (unsigned long long*)d_stats);
  FPS_LAUNCH_CHECK();
  return launch_merge(sc.part_s, sc.part_i, p.n_splits, n_q, k, k, nullptr, 0, d_out_scores, d_out_ids, st);
}

}  // extern "C"
