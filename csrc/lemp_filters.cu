// LEMP candidate filters on the device (E11 / E12): LENGTH, COORD, INCR and the switching strategies LC / LI,
// with the reference's per-bucket loop around them.
//
// Reference: pruning/LEMPPruningFunctions.scala:10-90 (lengthPruning, coordPruning, incrPruning),
// pruning/LEMPPruningStrategy.scala:6-76 (which one a bucket gets) and the retrieval loop
// PSTopKGeneratorWorker.scala:49-111: items sorted by length descending are cut into buckets of
// `bucketSize`; a bucket whose longest item cannot beat the current k-th score ends the search; otherwise
// theta is fixed for the bucket, every item is tested with the bucket's predicate (only once k results
// exist) and the survivors are scored and offered to the queue.
//
// The tensor-core kernels (topk.cu, topk_packed.cu) keep only the bucket-level LENGTH bound, because they
// score whole tiles and a per-candidate filter saves them nothing.  This file is the same algorithm as the
// reference's with the candidate filters executed: one warp per query, lane per candidate, scalar fp32 dot
// products for the survivors.  It is the parity path (its pruned / scored counts are those of the host
// engine's lemp_top_k) and the device-side oracle for the filters; it is not the fast path.
#include "topk_core.cuh"

namespace {

constexpr int LF_WARPS = 4;          // queries per CTA
constexpr int LF_CAP = 256;          // result list slots per query (k <= 128, a chunk adds up to 32)
constexpr int LF_MAX_FOCUS = 16;

enum { LF_LENGTH = 0, LF_COORD = 1, LF_INCR = 2, LF_LC = 3, LF_LI = 4 };

// what a query contributes to the predicates (built once per query by its warp)
struct LfQuery {
  float qlen;
  int f;                        // focus coordinate: argmax q_c^2 (first maximum)
  float qbar_f;                 // q_f / |q|
  int n_focus;                  // |F|
  int F[LF_MAX_FOCUS];          // the n_focus coordinates with the largest q_c^2, ties to the lower index
  float qF[LF_MAX_FOCUS];
  float q_mF_sqr;               // |q|^2 - |q_F|^2
};

// the per-bucket constants of the COORD test: feasible interval [l_f, u_f] of p_f / |p|
struct LfCoord {
  float l_f, u_f, theta_b_q;
};

__device__ __forceinline__ LfCoord lf_coord_bounds(float qbar_f, float theta_b_q) {
  const float t = fminf(fmaxf(theta_b_q, -1.f), 1.f);
  const float a = qbar_f * t;
  const float b = sqrtf(fmaxf(0.f, (1.f - t * t) * (1.f - qbar_f * qbar_f)));
  const float lf_p = a - b, uf_p = a + b;
  LfCoord c;
  c.l_f = (qbar_f >= 0.f || lf_p > t / qbar_f) ? lf_p : -1.f;
  c.u_f = (qbar_f <= 0.f || uf_p < t / qbar_f) ? uf_p : 1.f;
  c.theta_b_q = theta_b_q;
  return c;
}
__device__ __forceinline__ bool lf_coord_keep(const LfCoord& c, float plen, float p_f) {
  if (plen == 0.f) return c.theta_b_q <= 0.f;
  const float pbar = p_f / plen;
  return c.l_f <= pbar && pbar <= c.u_f;
}
__device__ __forceinline__ bool lf_incr_keep(const LfQuery& q, float theta, float plen, const float* __restrict__ p) {
  float qp = 0.f, pp = 0.f;
  for (int i = 0; i < q.n_focus; ++i) {
    const float x = p[q.F[i]];
    qp = fmaf(q.qF[i], x, qp);
    pp = fmaf(x, x, pp);
  }
  const float u = theta - qp;
  return u < 0.f || q.q_mF_sqr * (plen * plen - pp) >= u * u;
}
// which predicate a bucket gets (LEMPPruningStrategy: LC / LI switch on the spread of the bucket's lengths)
__device__ __forceinline__ int lf_bucket_kind(int strategy, float thr, float bmax, float bmin) {
  if (strategy == LF_LC) return bmax > bmin * thr ? LF_LENGTH : LF_COORD;
  if (strategy == LF_LI) return bmax > bmin * thr ? LF_LENGTH : LF_INCR;
  return strategy;
}

// built by the whole warp from the query row in shared memory
__device__ void lf_build_query(LfQuery& q, const float* s_q, int dim, int n_focus, int lane) {
  float s = 0.f;
  for (int c = lane; c < dim; c += 32) s = fmaf(s_q[c], s_q[c], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  q.qlen = sqrtf(s);
  q.n_focus = min(min(n_focus, LF_MAX_FOCUS), dim);
  float taken_sum = 0.f;
  unsigned long long taken_lo = 0, taken_hi = 0;       // coordinates already in F (dim <= 128)
  const int rounds = max(q.n_focus, 1);
  for (int r = 0; r < rounds; ++r) {
    float best = -1.f;
    int arg = 0x7fffffff;
    for (int c = lane; c < dim; c += 32) {
      const bool used = c < 64 ? ((taken_lo >> c) & 1ull) : ((taken_hi >> (c - 64)) & 1ull);
      const float v = s_q[c] * s_q[c];
      if (!used && (v > best || (v == best && c < arg))) { best = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
    }
    if (r == 0) {
      q.f = arg;
      q.qbar_f = q.qlen > 0.f ? s_q[arg] / q.qlen : 0.f;
    }
    if (r < q.n_focus) {
      q.F[r] = arg;
      q.qF[r] = s_q[arg];
      taken_sum = fmaf(s_q[arg], s_q[arg], taken_sum);
      if (arg < 64) taken_lo |= 1ull << arg; else taken_hi |= 1ull << (arg - 64);
    }
  }
  q.q_mF_sqr = q.qlen * q.qlen - taken_sum;
}

// the bucket's predicate for one candidate; `kind` from lf_bucket_kind, theta fixed for the bucket
__device__ __forceinline__ bool lf_keep(int kind, const LfQuery& q, const LfCoord& cb, float theta, float plen,
                                        const float* __restrict__ p) {
  if (kind == LF_COORD) return lf_coord_keep(cb, plen, p[q.f]);
  if (kind == LF_INCR) return lf_incr_keep(q, theta, plen, p);
  return plen >= (q.qlen > 0.f ? theta / q.qlen : 0.f);           // LENGTH
}

// ---- keep-mask of a whole (queries x items) problem under given thetas: the filters in isolation --------
__global__ void __launch_bounds__(LF_WARPS * 32)
k_lemp_keep_mask(const float* __restrict__ queries, int q_stride, long long n_q, const float* __restrict__ items,
                 int item_stride, const float* __restrict__ item_len, long long n_items, int dim,
                 const float* __restrict__ theta, int bucket_size, int strategy, int n_focus, float thr,
                 uint8_t* __restrict__ keep) {
  __shared__ float s_q[LF_WARPS][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long qi = (long long)blockIdx.x * LF_WARPS + warp;
  if (qi >= n_q) return;
  for (int c = lane; c < dim; c += 32) s_q[warp][c] = queries[qi * q_stride + c];
  __syncwarp();
  LfQuery q;
  lf_build_query(q, s_q[warp], dim, n_focus, lane);
  const float th = theta[qi];
  for (long long b0 = 0; b0 < n_items; b0 += bucket_size) {
    const long long b1 = min(n_items, b0 + bucket_size);
    const float bmax = item_len[b0], bmin = item_len[b1 - 1];
    const int kind = lf_bucket_kind(strategy, thr, bmax, bmin);
    const float denom = bmax * q.qlen;
    const LfCoord cb = lf_coord_bounds(q.qbar_f, denom > 0.f ? th / denom : 0.f);
    for (long long i = b0 + lane; i < b1; i += 32)
      keep[qi * n_items + i] = lf_keep(kind, q, cb, th, item_len[i], items + i * item_stride) ? 1 : 0;
  }
}

// ---- the reference's retrieval loop, one warp per query --------------------------------------------------
__global__ void __launch_bounds__(LF_WARPS * 32)
k_lemp_topk_filtered(const float* __restrict__ queries, int q_stride, long long n_q, const float* __restrict__ items,
                     int item_stride, const float* __restrict__ item_len, const int* __restrict__ item_ids,
                     long long n_items, int dim, int k, int bucket_size, int strategy, int n_focus, float thr,
                     float* __restrict__ out_s, int* __restrict__ out_i, unsigned long long* __restrict__ stats) {
  __shared__ float s_q[LF_WARPS][128];
  __shared__ float s_ls[LF_WARPS][LF_CAP];
  __shared__ int s_li[LF_WARPS][LF_CAP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long qi = (long long)blockIdx.x * LF_WARPS + warp;
  if (qi >= n_q) return;
  for (int c = lane; c < dim; c += 32) s_q[warp][c] = queries[qi * q_stride + c];
  __syncwarp();
  LfQuery q;
  lf_build_query(q, s_q[warp], dim, n_focus, lane);
  float* ls = s_ls[warp];
  int* li = s_li[warp];
  int cnt = 0;                         // entries in the list (warp-uniform)
  float kth = -INFINITY;               // k-th best of the list when it holds >= k entries (lower bound in between prunes)
  unsigned long long seen = 0, pruned = 0, scored = 0, buckets_skipped = 0;
  for (long long b0 = 0; b0 < n_items; b0 += bucket_size) {
    const long long b1 = min(n_items, b0 + bucket_size);
    if (cnt > k) {                     // exact k-th score for this bucket's theta
      kth = prune_row<LF_CAP>(ls, li, cnt, k, lane);
      cnt = k;
      __syncwarp();
    } else if (cnt == k && kth == -INFINITY) {
      float m = INFINITY;
      for (int j = lane; j < k; j += 32) m = fminf(m, ls[j]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
      kth = m;
    }
    const bool full = cnt >= k;
    const float bmax = item_len[b0], bmin = item_len[b1 - 1];
    if (full && bmax * q.qlen <= kth) {                 // no later item can enter: lengths descend
      buckets_skipped += (unsigned long long)((n_items - b0 + bucket_size - 1) / bucket_size);
      break;
    }
    const float theta = full ? kth : 0.f;
    const int kind = lf_bucket_kind(strategy, thr, bmax, bmin);
    const float denom = bmax * q.qlen;
    const LfCoord cb = lf_coord_bounds(q.qbar_f, denom > 0.f ? theta / denom : 0.f);
    for (long long c0 = b0; c0 < b1; c0 += 32) {
      const long long i = c0 + lane;
      const bool have = i < b1;
      const float* p = items + (have ? i : b0) * item_stride;
      bool keep = have;
      // the reference tests a candidate only while the queue is full; the queue can fill inside a bucket
      // (then with the bucket's theta = 0): followed here at chunk granularity
      if (have && cnt >= k) keep = lf_keep(kind, q, cb, theta, item_len[i], p);
      if (have) { ++seen; if (!keep) ++pruned; }
      float s = -INFINITY;
      if (keep) {
        ++scored;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        for (int c = 0; c + 3 < dim; c += 4) {
          const float4 x = *reinterpret_cast<const float4*>(p + c);
          a0 = fmaf(s_q[warp][c], x.x, a0); a1 = fmaf(s_q[warp][c + 1], x.y, a1);
          a2 = fmaf(s_q[warp][c + 2], x.z, a2); a3 = fmaf(s_q[warp][c + 3], x.w, a3);
        }
        for (int c = dim & ~3; c < dim; ++c) a0 = fmaf(s_q[warp][c], p[c], a0);
        s = (a0 + a1) + (a2 + a3);
      }
      // offer: what cannot beat the current k-th is dropped, the rest is appended (pruned back when the list fills)
      const bool in = keep && (cnt < k || s > kth);
      const unsigned m = __ballot_sync(0xffffffffu, in);
      if (in) {
        const int pos = cnt + __popc(m & ((1u << lane) - 1u));
        ls[pos] = s;
        li[pos] = (int)i;
      }
      cnt += __popc(m);
      __syncwarp();
      if (cnt > LF_CAP - 32) {
        kth = fmaxf(kth, prune_row<LF_CAP>(ls, li, cnt, k, lane));
        cnt = k;
        __syncwarp();
      }
    }
  }
  if (cnt > k) {
    prune_row<LF_CAP>(ls, li, cnt, k, lane);
    cnt = k;
    __syncwarp();
  }
  // unsorted k best, padded; the caller's merge pass sorts them
  for (int j = lane; j < k; j += 32) {
    const bool in = j < cnt;
    out_s[qi * k + j] = in ? ls[j] : -INFINITY;
    out_i[qi * k + j] = in ? (item_ids ? item_ids[li[j]] : li[j]) : -1;
  }
  if (stats != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      seen += __shfl_xor_sync(0xffffffffu, seen, o);
      pruned += __shfl_xor_sync(0xffffffffu, pruned, o);
      scored += __shfl_xor_sync(0xffffffffu, scored, o);
    }
    if (lane == 0) {
      atomicAdd(&stats[0], seen);
      atomicAdd(&stats[1], pruned);
      atomicAdd(&stats[2], scored);
      atomicAdd(&stats[3], buckets_skipped);
    }
  }
}

}  // namespace

extern "C" {

int fps_lemp_keep_mask(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                       int32_t item_stride, const float* d_item_len, int64_t n_items, int32_t dim,
                       const float* d_theta, int32_t bucket_size, int32_t strategy, int32_t n_focus,
                       float switch_threshold, uint8_t* d_keep, void* stream) {
  FPS_REQUIRE(d_queries && d_items && d_item_len && d_theta && d_keep, FPS_E_INVALID, "fps_lemp_keep_mask: null argument");
  FPS_REQUIRE(dim >= 1 && dim <= 128 && bucket_size >= 1 && strategy >= LF_LENGTH && strategy <= LF_LI &&
                  n_focus >= 0 && n_focus <= LF_MAX_FOCUS,
              FPS_E_INVALID, "fps_lemp_keep_mask: rank in [1, 128], bucket_size >= 1, strategy in [0, 4], focus set <= %d",
              LF_MAX_FOCUS);
  if (n_q <= 0 || n_items <= 0) return 0;
  k_lemp_keep_mask<<<(unsigned)((n_q + LF_WARPS - 1) / LF_WARPS), LF_WARPS * 32, 0, (cudaStream_t)stream>>>(
      d_queries, q_stride, n_q, d_items, item_stride, d_item_len, n_items, dim, d_theta, bucket_size, strategy,
      n_focus, switch_threshold, d_keep);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_lemp_topk_filtered(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                           int32_t item_stride, const float* d_item_len, const int32_t* d_item_ids,
                           int64_t n_items, int32_t dim, int32_t k, int32_t bucket_size, int32_t strategy,
                           int32_t n_focus, float switch_threshold, float* d_part_scores, int32_t* d_part_ids,
                           float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats, void* stream) {
  FPS_REQUIRE(d_queries && d_items && d_item_len && d_part_scores && d_part_ids && d_out_scores && d_out_ids,
              FPS_E_INVALID, "fps_lemp_topk_filtered: null argument");
  FPS_REQUIRE(k >= 1 && k <= TK_MAX_K, FPS_E_TOO_LARGE, "fps_lemp_topk_filtered: k = %d outside [1, %d]", k, TK_MAX_K);
  FPS_REQUIRE(dim >= 1 && dim <= 128 && item_stride % 4 == 0 && bucket_size >= 1 && strategy >= LF_LENGTH &&
                  strategy <= LF_LI && n_focus >= 0 && n_focus <= LF_MAX_FOCUS,
              FPS_E_INVALID, "fps_lemp_topk_filtered: rank in [1, 128], item stride %% 4 == 0, strategy in [0, 4], focus set <= %d",
              LF_MAX_FOCUS);
  if (n_q <= 0) return 0;
  FPS_REQUIRE(n_items > 0, FPS_E_INVALID, "fps_lemp_topk_filtered: no items");
  cudaStream_t st = (cudaStream_t)stream;
  k_lemp_topk_filtered<<<(unsigned)((n_q + LF_WARPS - 1) / LF_WARPS), LF_WARPS * 32, 0, st>>>(
      d_queries, q_stride, n_q, d_items, item_stride, d_item_len, d_item_ids, n_items, dim, k, bucket_size, strategy,
      n_focus, switch_threshold, d_part_scores, d_part_ids, (unsigned long long*)d_stats);
  FPS_LAUNCH_CHECK();
  return launch_merge(d_part_scores, d_part_ids, 1, n_q, k, k, nullptr, 0, d_out_scores, d_out_ids, st);
}

}  // extern "C"
