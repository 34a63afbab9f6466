// LEMP top-K retrieval on the 5th-gen tensor cores (N6) + partial-list merge (N7).
//
// Reference behaviour: PSTopKGeneratorWorker.scala:49-111 walks items in buckets of decreasing
// length, stops when ||q|| * bucketMaxLen <= theta (theta = current k-th best score), scores the
// surviving items with scalar dot products and keeps a min-heap of workerK entries - per query.
//
// Here a CTA owns a tile of 128 queries and walks 64-item tiles (= buckets) of the length-sorted
// item matrix.  Scores S = Q * P^T are tcgen05.mma tiles (M=128, N=64, kind::tf32) with the
// accumulator in TMEM.  The tables are fp32 and must stay fp32, so both operands are split
// x = hi + lo (hi = tf32(x), lo = tf32(x - hi)) and three MMAs accumulate hi*hi + hi*lo + lo*hi
// (3xTF32, error ~2^-21 relative to the operand magnitudes).  Operands are staged in shared memory
// in the K-major SWIZZLE_128B layout the UMMA descriptors expect.  The LEMP LENGTH test is applied
// per tile: when ||q|| * tileMaxLen <= theta holds for every query of the CTA the walk stops
// (exact, because lengths are sorted).  Selection: thread t owns query t (TMEM lane t); scores
// above its running theta are appended to a per-query candidate buffer in scratch memory; when a
// buffer fills it is pruned back to the k best by a bitwise selection of the k-th largest key,
// which also raises theta.  Item tiles are dealt round-robin to `n_splits` CTAs per query tile;
// their partial lists are merged by fps_topk_merge (X8: CollectTopKFromEachWorker.scala:40-48).
#include "topk_core.cuh"

namespace {

template <bool TC>
__global__ void __launch_bounds__(TK_THREADS, 1)
k_lemp_topk(const float* __restrict__ queries, int q_stride, long long n_q,
            const float* __restrict__ items, int item_stride, const float* __restrict__ item_len,
            const int* __restrict__ item_ids, long long n_items, int dim, int kblks, int k, int n_splits,
            int flags, TopkScratch sc, unsigned long long* __restrict__ stats) {
  // flags: bit 0 = LEMP LENGTH pruning (items sorted by length), bit 1 = items whose id is negative are not
  // candidates (a live-item mask for tables scored in place)
  const int use_pruning = flags & 1;
  const bool mask_dead = (flags & 2) != 0 && item_ids != nullptr;
  __shared__ int s_live[TK_N];
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by SWIZZLE_128B; dynamic smem base is only 16-byte aligned
  uint8_t* base = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int a_floats = TK_M * kblks * TK_KB, b_floats = TK_N * kblks * TK_KB;
  float* sA_hi = (float*)base;
  float* sA_lo = sA_hi + a_floats;
  float* sB_hi = sA_lo + (TC ? a_floats : 0);
  float* sB_lo = sB_hi + b_floats;
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_tmem;

  const int t = threadIdx.x, warp = t >> 5;
  const int qt = blockIdx.x / n_splits, split = blockIdx.x - qt * n_splits;
  const long long q0 = (long long)qt * TK_M;
  const int q_valid = (int)min((long long)TK_M, n_q - q0);

  if (TC) {
    if (warp == 0) tmem_alloc(&s_tmem, 64);
    if (t == 0) {
      mbar_init(&s_bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  stage_tile<TC>(sA_hi, sA_lo, queries + q0 * q_stride, q_stride, TK_M, q_valid, dim, kblks);

  // ||q|| for the LENGTH bound
  float qlen = 0.f;
  if (t < q_valid) {
    const float* qp = queries + (q0 + t) * q_stride;
    float s = 0.f;
    for (int j = 0; j < dim; ++j) s = fmaf(qp[j], qp[j], s);
    qlen = sqrtf(s);
  }
  if (TC) {
    tc_fence_before();
    fence_async_smem();
  }
  __syncthreads();
  uint32_t tmem_base = 0;
  if (TC) {
    tc_fence_after();
    tmem_base = s_tmem;
  }

  // candidate lists: one row of TK_CAP slots per query (row-major, so that a warp can prune one list together)
  const int lane = t & 31;
  float* const cs_w = sc.cand_s + ((size_t)blockIdx.x * TK_M + (size_t)warp * 32) * TK_CAP;    // this warp's 32 lists
  int* const ci_w = sc.cand_i + ((size_t)blockIdx.x * TK_M + (size_t)warp * 32) * TK_CAP;
  float* cs = cs_w + (size_t)lane * TK_CAP;
  int* ci = ci_w + (size_t)lane * TK_CAP;
  volatile uint32_t* theta_g = sc.theta + q0 + t;
  int cnt = 0;
  float theta = -INFINITY;
  uint32_t phase = 0;
  unsigned long long tiles_done = 0, accepted = 0, prunes = 0;

  const long long n_tiles = (n_items + TK_N - 1) / TK_N;
  for (long long tile = split; tile < n_tiles; tile += n_splits) {
    const long long i0 = tile * TK_N;
    const int i_valid = (int)min((long long)TK_N, n_items - i0);
    // ---- LEMP LENGTH test for the whole bucket (exact early exit: lengths are descending)
    if (use_pruning) {
      theta = fmaxf(theta, key2f(*theta_g));
      const float bound = qlen * item_len[i0];
      const int can_skip = (t >= q_valid) || (theta > -INFINITY && bound <= theta);
      if (__syncthreads_and(can_skip)) break;
    }
    // ---- stage the item tile
    if (mask_dead && t < TK_N) s_live[t] = t < i_valid ? item_ids[i0 + t] : -1;
    stage_tile<TC>(sB_hi, sB_lo, items + i0 * item_stride, item_stride, TK_N, i_valid, dim, kblks);
    float sc_row[TK_N];
    if (TC) {
      fence_async_smem();
      __syncthreads();
      if (t == 0) {
        tc_fence_after();
        uint32_t acc = 0;
        for (int kb = 0; kb < kblks; ++kb) {
#pragma unroll
          for (int ks = 0; ks < TK_KB / 8; ++ks) {
            const uint32_t a_off = (uint32_t)(kb * TK_M * TK_KB + ks * 8) * 4, b_off = (uint32_t)(kb * TK_N * TK_KB + ks * 8) * 4;
            const uint64_t ah = make_desc(smem_u32(sA_hi) + a_off), al = make_desc(smem_u32(sA_lo) + a_off);
            const uint64_t bh = make_desc(smem_u32(sB_hi) + b_off), bl = make_desc(smem_u32(sB_lo) + b_off);
            umma_tf32(tmem_base, al, bh, TK_IDESC, acc);   // small terms first
            umma_tf32(tmem_base, ah, bl, TK_IDESC, 1);
            umma_tf32(tmem_base, ah, bh, TK_IDESC, 1);
            acc = 1;
          }
        }
        umma_commit(&s_bar);
      }
      mbar_wait_warp(&s_bar, phase);      // every thread of the CTA waits here
      phase ^= 1;
      tc_fence_after();
      const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
      float v0[32], v1[32];
      tmem_ld32(lane_addr, v0);
      tmem_ld32(lane_addr + 32, v1);
#pragma unroll
      for (int j = 0; j < 32; ++j) { sc_row[j] = v0[j]; sc_row[32 + j] = v1[j]; }
      tc_fence_before();
    } else {
      __syncthreads();
      // CUDA-core fp32 scoring of the same staged tile (baseline / oracle)
#pragma unroll
      for (int j = 0; j < TK_N; ++j) sc_row[j] = 0.f;
      for (int kb = 0; kb < kblks; ++kb) {
        for (int c = 0; c < 8; ++c) {
          const float4 a = *reinterpret_cast<const float4*>(sA_hi + kb * TK_M * TK_KB + t * TK_KB + ((c ^ (t & 7)) << 2));
#pragma unroll
          for (int j = 0; j < TK_N; ++j) {
            const float4 b = *reinterpret_cast<const float4*>(sB_hi + kb * TK_N * TK_KB + j * TK_KB + ((c ^ (j & 7)) << 2));
            sc_row[j] = fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, fmaf(a.w, b.w, sc_row[j]))));
          }
        }
      }
    }
    // ---- selection: append everything that beats the running theta
    if (t < q_valid) {
#pragma unroll
      for (int j = 0; j < TK_N; ++j) {
        if (j < i_valid && sc_row[j] > theta && (!mask_dead || s_live[j] >= 0)) {
          cs[cnt] = sc_row[j];
          ci[cnt] = (int)(i0 + j);
          ++cnt;
          ++accepted;
        }
      }
    }
    ++tiles_done;
    // Prune warp-cooperatively: when any list of the CTA is about to overflow, every list with more than k
    // entries is pruned back to its k best, one list at a time by all 32 lanes of its warp (prune_row).  The
    // first version pruned per thread, each thread making ~30 dependent passes over its own list in global
    // memory: 0.94 ms for 256 queries x 19K items, 87 % of a learner+generator step.
    if (__syncthreads_or(cnt > TK_CAP - TK_N)) {
      for (int src = 0; src < 32; ++src) {
        const int c = __shfl_sync(0xffffffffu, cnt, src);
        if (c > k) {
          const float th = prune_row<TK_CAP>(cs_w + (size_t)src * TK_CAP, ci_w + (size_t)src * TK_CAP, c, k, lane);
          if (lane == src) {
            cnt = k;
            theta = fmaxf(theta, th);
            ++prunes;
            if (use_pruning) atomicMax((uint32_t*)theta_g, f2key(theta));
          }
        }
      }
      __syncwarp();
    }
    __syncthreads();   // the next stage_tile overwrites sB; every thread is done with this tile
  }

  // ---- final: k best of this (query, split), unsorted, padded with (-inf, -1)
  __syncwarp();
  for (int src = 0; src < 32; ++src) {
    const int c = __shfl_sync(0xffffffffu, cnt, src);
    if (c > k) {
      prune_row<TK_CAP>(cs_w + (size_t)src * TK_CAP, ci_w + (size_t)src * TK_CAP, c, k, lane);
      if (lane == src) cnt = k;
    }
  }
  __syncwarp();
  if (t < q_valid) {
    float* ps = sc.part_s + ((size_t)(q0 + t) * n_splits + split) * k;
    int* pi = sc.part_i + ((size_t)(q0 + t) * n_splits + split) * k;
    for (int j = 0; j < k; ++j) {
      if (j < cnt) {
        const int slot = ci[j];
        ps[j] = cs[j];
        pi[j] = item_ids ? item_ids[slot] : slot;
      } else {
        ps[j] = -INFINITY;
        pi[j] = -1;
      }
    }
  }
  if (stats != nullptr && t == 0) {
    atomicAdd(&stats[0], tiles_done);
    atomicAdd(&stats[1], (unsigned long long)((n_tiles - split + n_splits - 1) / n_splits));
  }
  if (stats != nullptr) {
    atomicAdd(&stats[2], accepted);
    if (prunes) atomicAdd(&stats[3], prunes);
  }
  if (TC) {
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 64);
  }
}

template <bool TC>
int run_topk(const float* d_queries, int q_stride, long long n_q, const float* d_items, int item_stride,
             const float* d_item_len, const int* d_item_ids, long long n_items, int dim, int k,
             int use_pruning, void* d_scratch, size_t scratch_bytes, float* d_out_scores, int* d_out_ids,
             unsigned long long* d_stats, cudaStream_t st) {
  FPS_REQUIRE(n_q > 0 && n_items > 0 && dim > 0, FPS_E_INVALID, "topk: empty problem");
  FPS_REQUIRE(k >= 1 && k <= TK_MAX_K, FPS_E_TOO_LARGE, "topk: k = %d outside [1, %d]", k, TK_MAX_K);
  FPS_REQUIRE(dim <= 128, FPS_E_UNSUPPORTED, "topk: rank %d > 128 not supported by the tile kernel", dim);
  FPS_REQUIRE(q_stride % 4 == 0 && item_stride % 4 == 0, FPS_E_INVALID, "topk: strides must be multiples of 4 floats");
  FPS_REQUIRE(!(use_pruning & 1) || d_item_len != nullptr, FPS_E_INVALID, "topk: pruning needs item lengths");
  // as many item splits as the caller's scratch allows (up to 64: a small query batch on few CTAs left most of
  // the SMs idle), never fewer than the plan fps_topk_scratch_bytes sizes for
  Plan p = make_plan(n_q, k, n_items, dim, 1, 0, 0, 64);
  while (p.total > scratch_bytes && p.n_splits > 16) p = make_plan(n_q, k, n_items, dim, 1, 0, p.n_splits - 1, 64);
  if (p.total > scratch_bytes) p = make_plan(n_q, k, n_items, dim);
  FPS_REQUIRE(scratch_bytes >= p.total, FPS_E_INVALID, "topk: scratch too small (%zu < %zu)", scratch_bytes, p.total);
  uint8_t* sb = (uint8_t*)d_scratch;
  TopkScratch sc;
  sc.cand_s = (float*)sb;
  sc.cand_i = (int*)(sb + p.off_cand_i);
  sc.theta = (uint32_t*)(sb + p.off_theta);
  sc.part_s = (float*)(sb + p.off_part_s);
  sc.part_i = (int*)(sb + p.off_part_i);
  // theta keys start at 0 == "below every float"
  FPS_CUDA(cudaMemsetAsync(sc.theta, 0, (size_t)p.n_q_tiles * TK_M * sizeof(uint32_t), st));
  const size_t a_bytes = (size_t)TK_M * p.kblks * TK_KB * 4, b_bytes = (size_t)TK_N * p.kblks * TK_KB * 4;
  const size_t smem = 1024 + (TC ? 2 : 1) * a_bytes + 2 * b_bytes;
  FPS_CUDA(cudaFuncSetAttribute(k_lemp_topk<TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_lemp_topk<TC><<<p.n_ctas, TK_THREADS, smem, st>>>(d_queries, q_stride, n_q, d_items, item_stride, d_item_len,
                                                      d_item_ids, n_items, dim, p.kblks, k, p.n_splits,
                                                      use_pruning, sc, d_stats);
  FPS_LAUNCH_CHECK();
  return launch_merge(sc.part_s, sc.part_i, p.n_splits, n_q, k, k, nullptr, 0, d_out_scores, d_out_ids, st);
}

}  // namespace

extern "C" {

size_t fps_topk_scratch_bytes(int64_t n_q, int32_t k, int64_t n_items) {
  if (n_q <= 0 || n_items <= 0 || k <= 0) return 0;
  return make_plan(n_q, k, n_items, 128).total;
}

size_t fps_topk_scratch_bytes_full(int64_t n_q, int32_t k, int64_t n_items) {
  if (n_q <= 0 || n_items <= 0 || k <= 0) return 0;
  return make_plan(n_q, k, n_items, 128, 1, 0, 0, 64).total;
}

int fps_lemp_topk(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                  int32_t item_stride, const float* d_item_len, const int32_t* d_item_ids, int64_t n_items,
                  int32_t dim, int32_t k, int32_t use_pruning, void* d_scratch, size_t scratch_bytes,
                  float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats, void* stream) {
  return run_topk<true>(d_queries, q_stride, n_q, d_items, item_stride, d_item_len, d_item_ids, n_items, dim, k,
                        use_pruning, d_scratch, scratch_bytes, d_out_scores, d_out_ids,
                        (unsigned long long*)d_stats, (cudaStream_t)stream);
}

int fps_topk_simt(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                  int32_t item_stride, const int32_t* d_item_ids, int64_t n_items, int32_t dim, int32_t k,
                  void* d_scratch, size_t scratch_bytes, float* d_out_scores, int32_t* d_out_ids, void* stream) {
  return run_topk<false>(d_queries, q_stride, n_q, d_items, item_stride, nullptr, d_item_ids, n_items, dim, k, 0,
                         d_scratch, scratch_bytes, d_out_scores, d_out_ids, nullptr, (cudaStream_t)stream);
}

int fps_topk_merge(const float* d_part_scores, const int32_t* d_part_ids, int32_t n_parts, int64_t n_q,
                   int32_t k_in, int32_t k_out, const int32_t* d_seen, int32_t seen_stride,
                   float* d_out_scores, int32_t* d_out_ids, void* stream) {
  FPS_REQUIRE(n_parts >= 1 && k_in >= 1 && k_out >= 1 && n_q >= 0, FPS_E_INVALID, "fps_topk_merge: bad arguments");
  if (n_q == 0) return 0;
  return launch_merge(d_part_scores, d_part_ids, n_parts, n_q, k_in, k_out, d_seen, seen_stride, d_out_scores,
                      d_out_ids, (cudaStream_t)stream);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// fps_lemp_topk_static: the same retrieval for callers that must not allocate.
//
// Who that is: a serving loop that answers queries while a training job (or another framework's
// caching allocator) already owns all the device memory it is going to get, so a cudaMalloc in
// the request path can fail at an arbitrary moment; and latency-sensitive callers for whom the
// device-wide synchronisation that cudaMalloc / cudaFree imply is a stall they cannot afford per
// request.  Such callers bring their inputs and outputs; the candidate buffers, the shared theta
// array and the partial lists (the kernel's scratch) live in the fixed array below instead of a
// caller- or library-allocated buffer.
//
// * The array is part of the module image: every GPU that loads the library carries these
//   FPS_TOPK_STATIC_SCRATCH_BYTES until the process exits, used or not.
// * Problems whose scratch need (fps_topk_scratch_bytes) exceeds the array are refused with
//   FPS_E_TOO_LARGE; nothing is allocated as a fallback.  16 MiB covers up to 256 queries per call.
// * The first call on a device looks the array's device address up (one-time set-up, under a
//   mutex); make that first call outside any stream capture.
// * All calls on a device share the one array and it keeps whatever the previous call left in it.
//   The only part that must start clean - the shared theta keys - is cleared on the call's stream
//   before the kernel.  Two calls whose GPU work overlaps (different streams, or different host
//   threads) would write the same candidate slots: serialise them on one stream.
// * No cudaMalloc, cudaFree, or allocating library routine is called on this path.
// ---------------------------------------------------------------------------------------------
#include <mutex>

static __device__ __align__(256) unsigned char g_topk_static_scratch[FPS_TOPK_STATIC_SCRATCH_BYTES];

namespace {
std::mutex g_static_mu;
void* g_static_ptr[64] = {nullptr};   // device address of the array, per device ordinal
}  // namespace

extern "C" int fps_lemp_topk_static(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                                    int32_t item_stride, const float* d_item_len, const int32_t* d_item_ids,
                                    int64_t n_items, int32_t dim, int32_t k, int32_t use_pruning,
                                    float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats, void* stream) {
  FPS_REQUIRE(n_q > 0 && n_items > 0 && k >= 1, FPS_E_INVALID, "fps_lemp_topk_static: empty problem");
  const size_t need = fps_topk_scratch_bytes(n_q, k, n_items);
  FPS_REQUIRE(need <= (size_t)FPS_TOPK_STATIC_SCRATCH_BYTES, FPS_E_TOO_LARGE,
              "fps_lemp_topk_static: %lld queries x k=%d need %zu bytes of scratch, the built-in array holds %u; "
              "split the query batch or use fps_lemp_topk with your own scratch",
              (long long)n_q, (int)k, need, (unsigned)FPS_TOPK_STATIC_SCRATCH_BYTES);
  int dev = 0;
  FPS_CUDA(cudaGetDevice(&dev));
  FPS_REQUIRE(dev >= 0 && dev < 64, FPS_E_UNSUPPORTED, "fps_lemp_topk_static: device ordinal %d not supported", dev);
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (g_static_ptr[dev] == nullptr) {
      // In host code the array's name is only a stand-in; the address a kernel can use has to be
      // asked for, once per device, with that device current.
      void* p = nullptr;
      FPS_CUDA(cudaGetSymbolAddress((void**)&p, g_topk_static_scratch));
      FPS_REQUIRE(p != nullptr, FPS_E_UNSUPPORTED, "fps_lemp_topk_static: no device address for the scratch array");
      g_static_ptr[dev] = p;
    }
    scratch = g_static_ptr[dev];
  }
  return run_topk<true>(d_queries, q_stride, n_q, d_items, item_stride, d_item_len, d_item_ids, n_items, dim, k,
                        use_pruning, scratch, (size_t)FPS_TOPK_STATIC_SCRATCH_BYTES, d_out_scores, d_out_ids,
                        (unsigned long long*)d_stats, (cudaStream_t)stream);
}
