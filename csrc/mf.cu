// Matrix-factorisation kernels: fused pull-compute-push SGD (N1), streaming ring consumer (N2),
// prediction.  One sub-warp of LANES lanes owns one rating; each lane owns VPL float4 slices of
// the two rows.  Item rows come from the owner shard (a peer GPU's HBM through NVLink when the
// pointer is a peer mapping), the item delta goes back as 16-byte red.add.
#include "common.cuh"
#include "flow.cuh"
#include "mf_core.cuh"

namespace {

template <int LANES, int VPL, bool LAZY, int UNROLL, int MINB = 1, bool REPL = false>
__global__ void __launch_bounds__(MF_THREADS, MINB)
k_mf_sgd(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
         const int* __restrict__ uid, const int* __restrict__ iid, const float* __restrict__ rating,
         long long n, const FpsMfParams p, int chunk, int cta_credits, FpsFlow* flow) {
  constexpr int GPB = MF_THREADS / LANES;   // ratings in flight per block pass
  const int grp = threadIdx.x / LANES;
  MfCounters c;
  c.clear();
  // Pull limiter: the CTA takes its allowance (the pulls one pass of it can have outstanding) from the
  // worker's pool once and keeps it for its lifetime; the host sizes the grid so that all allowances
  // fit the limit.  (The first version acquired per 256-record chunk: every CTA of a 148-SM grid spun
  // on one word, 7x slower at limit 1600.)
  if (p.pull_limit > 0) flow_acquire(flow, cta_credits, p.pull_limit);
  for (long long base = (long long)blockIdx.x * chunk; base < n; base += (long long)gridDim.x * chunk) {
    const int cnt = (int)min((long long)chunk, n - base);
    for (int k0 = 0; k0 < cnt; k0 += GPB * UNROLL) {   // warp-uniform trip count (shuffles inside)
      const int k = k0 + grp * UNROLL;
      MfRec<VPL> rec[UNROLL];
#pragma unroll
      for (int q = 0; q < UNROLL; ++q)
        mf_load<LANES, VPL, LAZY>(rec[q], items, users, uid, iid, rating, base + k + q, k + q < cnt, p.local_shard, c);
#pragma unroll
      for (int q = 0; q < UNROLL; ++q) mf_update<LANES, VPL, LAZY, REPL>(rec[q], items, p, c);
    }
  }
  if (p.pull_limit > 0) flow_release(flow, cta_credits);
  c.flush(flow);
}

// The same step for SMALL pull limits, where a credit is a scarce thing: the ids and the rating of a sub-warp's
// next record are fetched while the current record's rows are in flight, so an allowance is held for about one
// row latency per record instead of two dependent latencies (ids, then rows).  Not used without a limit or
// with a generous one: there thousands of resident warps hide the id loads, and the extra registers cost
// occupancy (the default build runs at 32 registers, 8 CTAs per SM).
template <int LANES, int VPL, bool LAZY>
__global__ void __launch_bounds__(MF_THREADS)
k_mf_sgd_pf(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
            const int* __restrict__ uid, const int* __restrict__ iid, const float* __restrict__ rating,
            long long n, const FpsMfParams p, int chunk, int cta_credits, FpsFlow* flow) {
  constexpr int GPB = MF_THREADS / LANES;
  const int grp = threadIdx.x / LANES;
  MfCounters c;
  c.clear();
  if (p.pull_limit > 0) flow_acquire(flow, cta_credits, p.pull_limit);
  for (long long base = (long long)blockIdx.x * chunk; base < n; base += (long long)gridDim.x * chunk) {
    const int cnt = (int)min((long long)chunk, n - base);
    MfIds nxt = mf_fetch_ids(uid, iid, rating, grp < cnt ? base + grp : -1);
    for (int k0 = 0; k0 < cnt; k0 += GPB) {              // warp-uniform trip count (shuffles inside)
      const MfIds cur = nxt;
      const int kn = k0 + GPB + grp;
      nxt = mf_fetch_ids(uid, iid, rating, kn < cnt ? base + kn : -1);
      MfRec<VPL> rec;
      mf_load_ids<LANES, VPL, LAZY>(rec, items, users, cur, p.local_shard, c);
      mf_update<LANES, VPL, LAZY, false>(rec, items, p, c);
    }
  }
  if (p.pull_limit > 0) flow_release(flow, cta_credits);
  c.flush(flow);
}

template <int LANES, int VPL, bool LAZY>
__global__ void __launch_bounds__(MF_THREADS)
k_mf_predict(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
             const int* __restrict__ uid, const int* __restrict__ iid, long long n, float* __restrict__ out) {
  constexpr int GPB = MF_THREADS / LANES;
  const int lane = threadIdx.x % LANES;
  MfCounters c;
  c.clear();
  for (long long g0 = (long long)blockIdx.x * GPB; g0 < n; g0 += (long long)gridDim.x * GPB) {
    const long long g = g0 + threadIdx.x / LANES;
    MfRec<VPL> rec;
    mf_load<LANES, VPL, LAZY>(rec, items, users, uid, iid, nullptr, g, g < n, -1, c, /*touch=*/false);
    float dot = 0.f;
#pragma unroll
    for (int q = 0; q < VPL; ++q) dot += dot4(rec.u[q], rec.v[q]);
    dot = group_sum<LANES>(dot);
    if (lane == 0 && g < n) out[g] = rec.valid ? dot : __int_as_float(0x7fc00000);
  }
}

// ---- N2: persistent streaming consumer ------------------------------------------------------
// The host (or another stream) appends ratings to a ring and advances *tail.  CTAs claim chunks
// with an atomic on flow->ring_head.  Nothing here waits on another kernel: if the producer
// stops, every CTA leaves after iteration_wait_ms of idleness (device-side `iterate` timeout,
// M/FlinkParameterServer.scala:49-52), on *stop, or at the hard lifetime cap.
template <int LANES, int VPL, bool LAZY>
__global__ void __launch_bounds__(MF_THREADS)
k_mf_stream(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
            const int* ring_u, const int* ring_i, const float* ring_r, unsigned long long capacity,
            const volatile unsigned long long* tail, const volatile int* stop, const FpsMfParams p,
            int chunk, int cta_credits, unsigned long long wait_ns, unsigned long long life_ns, FpsFlow* flow) {
  constexpr int GPB = MF_THREADS / LANES;
  const int grp = threadIdx.x / LANES;
  __shared__ unsigned long long s_base;
  __shared__ int s_cnt;
  MfCounters c;
  c.clear();
  if (p.pull_limit > 0) flow_acquire(flow, cta_credits, p.pull_limit);   // per-CTA allowance, see k_mf_sgd
  const unsigned long long t_start = globaltimer_ns();
  while (true) {
    if (threadIdx.x == 0) {
      int cnt = 0;
      unsigned long long base = 0;
      int reason = 0;
      while (true) {
        const unsigned long long now = globaltimer_ns();
        const unsigned long long t = *tail;
        unsigned long long h = *(volatile unsigned long long*)&flow->ring_head;
        if (h < t) {
          const unsigned long long want = min((unsigned long long)chunk, t - h);
          const unsigned long long got = atomicCAS((unsigned long long*)&flow->ring_head, h, h + want);
          if (got == h) {
            base = h;
            cnt = (int)want;
            *(volatile unsigned long long*)&flow->last_activity_ns = now;
            break;
          }
          continue;   // lost the race, retry immediately
        }
        if (*stop != 0) { reason = 2; break; }
        if (now - t_start > life_ns) { reason = 3; break; }
        if (wait_ns != 0) {
          const unsigned long long last = *(volatile unsigned long long*)&flow->last_activity_ns;
          const unsigned long long ref = last > t_start ? last : t_start;
          // idle only counts once everything claimed has retired
          if (now > ref && now - ref > wait_ns &&
              *(volatile unsigned long long*)&flow->ring_done == h) { reason = 1; break; }
        }
        __nanosleep(2000);
      }
      s_base = base;
      s_cnt = reason ? -reason : cnt;
    }
    __syncthreads();
    const int cnt = s_cnt;
    const unsigned long long base = s_base;
    __syncthreads();
    if (cnt < 0) {
      if (threadIdx.x == 0) atomicMax(&flow->exit_reason, -cnt);
      break;
    }
    for (int k0 = 0; k0 < cnt; k0 += GPB) {
      const int k = k0 + grp;
      MfRec<VPL> rec;
      const long long slot = (long long)((base + k) % capacity);
      // ring slots are rewritten by the copy engine while this kernel runs: read them at L2 (.cg), never
      // through the non-coherent / L1 path a `const __restrict__` load may take
      MfIds ids;
      ids.active = k < cnt;
      ids.u = ids.i = 0;
      ids.r = 0.f;
      if (ids.active) {
        asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(ids.u) : "l"(ring_u + slot));
        asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(ids.i) : "l"(ring_i + slot));
        asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(ids.r) : "l"(ring_r + slot));
        ids.active = ids.u != FPS_SKIP_ID;
      }
      mf_load_ids<LANES, VPL, LAZY>(rec, items, users, ids, p.local_shard, c);
      mf_update<LANES, VPL, LAZY>(rec, items, p, c);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd((unsigned long long*)&flow->ring_done, (unsigned long long)cnt);
      *(volatile unsigned long long*)&flow->last_activity_ns = globaltimer_ns();
    }
  }
  if (p.pull_limit > 0) flow_release(flow, cta_credits);
  c.flush(flow);
}

struct Shape {
  int lanes, vpl;
};

Shape shape_for(int stride) {
  const int nvec = stride / 4;
  Shape s{1, 1};
  while (s.lanes < nvec && s.lanes < 32) s.lanes <<= 1;
  s.vpl = (nvec + s.lanes - 1) / s.lanes;
  return s;
}

int check_pair(const FpsTableView* items, const FpsTableView* users, const char* who) {
  if (int e = fps_check_view(items, who)) return e;
  if (int e = fps_check_view(users, who)) return e;
  FPS_REQUIRE(items->stride == users->stride && items->dim == users->dim, FPS_E_INVALID,
              "%s: item and user tables must have the same dim/stride", who);
  FPS_REQUIRE(items->stride <= 1024, FPS_E_UNSUPPORTED, "%s: rank above 1024 not supported", who);
  return 0;
}

}  // namespace

#define MF_DISPATCH(FN)                                                                      \
  do {                                                                                       \
    const Shape s = shape_for(items->stride);                                                \
    const bool lazy = items->lazy || users->lazy;                                            \
    if (s.vpl == 1) {                                                                        \
      switch (s.lanes) {                                                                     \
        case 1: FN(1, 1, lazy); break;                                                       \
        case 2: FN(2, 1, lazy); break;                                                       \
        case 4: FN(4, 1, lazy); break;                                                       \
        case 8: FN(8, 1, lazy); break;                                                       \
        case 16: FN(16, 1, lazy); break;                                                     \
        default: FN(32, 1, lazy); break;                                                     \
      }                                                                                      \
    } else if (s.vpl == 2) { FN(32, 2, lazy); }                                              \
    else if (s.vpl <= 4) { FN(32, 4, lazy); }                                                \
    else { FN(32, 8, lazy); }                                                                \
  } while (0)

extern "C" {

int fps_mf_sgd_step(const FpsTableView* items, const FpsTableView* users, const int32_t* d_user_ids,
                    const int32_t* d_item_ids, const float* d_ratings, int64_t n,
                    const FpsMfParams* params, FpsFlow* d_flow, void* stream) {
  if (int e = check_pair(items, users, "fps_mf_sgd_step")) return e;
  FPS_REQUIRE(params != nullptr && d_flow != nullptr, FPS_E_INVALID, "fps_mf_sgd_step: params and flow are required");
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const bool repl = items->n_replicas > 1;
  if (repl) {
    FPS_REQUIRE(items->delta_rows != nullptr && !items->lazy && !users->lazy, FPS_E_UNSUPPORTED,
                "fps_mf_sgd_step: a replicated item table must be eager and use deferred pushes (delta_rows)");
    FPS_REQUIRE(params->variant == 0 || params->variant == 2, FPS_E_UNSUPPORTED,
                "fps_mf_sgd_step: replicated items run the direct or the combine variant");
  }
  FPS_REQUIRE(params->variant == 0 || params->variant == 2, FPS_E_INVALID,
              "fps_mf_sgd_step: variant must be 0 (direct) or 2 (combine); the tile path has its own entry points");
  if (params->variant == 2) {
    return fps_mf_sgd_combine(items, users, d_user_ids, d_item_ids, d_ratings, n, params, d_flow, st);
  }
  int chunk = MF_CHUNK;
  if (params->pull_limit > 0 && params->pull_limit < chunk) chunk = params->pull_limit;
  long long blocks = (n + chunk - 1) / chunk;
  // not persistent by default: with 4-8 resident CTAs per SM a grid of exactly 6 per SM ran 1.5
  // waves (1.80 ms vs 1.55 ms); many short CTAs load-balance better than a tuned persistent grid
  const int bps = params->blocks_per_sm > 0 ? params->blocks_per_sm : 64;
  const long long cap = (long long)fps_num_sms() * bps;
  if (blocks > cap) blocks = cap;
  const int unroll = params->unroll;   // 0 = default
  // small limits: the id-prefetching build (see k_mf_sgd_pf); FPS_MF_PREFETCH=0 / 1 forces the choice
  static const int pf_env = []() { const char* e = getenv("FPS_MF_PREFETCH"); return e ? atoi(e) : -1; }();
  const bool use_pf = !repl && (unroll == 0) && params->pull_limit > 0 &&
                      (pf_env == 1 || (pf_env != 0 && params->pull_limit <= 4096));   // measured: 1600: 3.51 -> 3.15 ms, 16384: 1.374 -> 1.416 ms
  int cta_credits = 0;
  if (params->pull_limit > 0) {
    // one pass of a CTA has (sub-warps x ratings per sub-warp) pulls outstanding, or the whole chunk if
    // that is smaller; the grid holds as many CTAs as allowances fit the limit
    const Shape sh = shape_for(items->stride);
    const int per_pass = (MF_THREADS / sh.lanes) * (use_pf ? 1 : (unroll == 2 || unroll == 4 ? unroll : (sh.vpl == 2 ? 2 : 1)));
    cta_credits = per_pass < chunk ? per_pass : chunk;
    // at most SMs x 8 CTAs (the tightest launch bound) are resident and hold credits at a time; only if
    // even those do not fit is the grid cut down to a persistent one
    const long long resident = (long long)fps_num_sms() * 8;
    const long long fit = params->pull_limit / cta_credits > 0 ? params->pull_limit / cta_credits : 1;
    if (fit < resident && blocks > fit) blocks = fit;
  }
#define MF_ARGS <<<(int)blocks, MF_THREADS, 0, st>>>(*items, *users, d_user_ids, d_item_ids, d_ratings, n, *params, chunk, cta_credits, d_flow)
  // Default for eager tables up to rank 128: one rating per sub-warp, 32 registers, 8 CTAs/SM (64
  // resident warps).  Measured on B200 (4.19M ratings, rank 128): 1.37 ms vs 1.55 ms for the
  // 2-ratings-per-sub-warp build at 56 registers - thread-level parallelism hides the row-load
  // latency better than per-warp ILP here.  `unroll` = 2 / 4 select the ILP builds explicitly.
#define FN(L, V, Z)                                                                                   \
  if (use_pf) { if (Z) k_mf_sgd_pf<L, V, true> MF_ARGS; else k_mf_sgd_pf<L, V, false> MF_ARGS; }      \
  else if (repl) k_mf_sgd<L, V, false, 1, 1, true> MF_ARGS;                                           \
  else if (Z) k_mf_sgd<L, V, true, 1> MF_ARGS;                                                        \
  else if (unroll == 4 && V == 1) k_mf_sgd<L, V, false, (V == 1 ? 4 : 1)> MF_ARGS;                    \
  else if (unroll == 2 || V > 1) k_mf_sgd<L, V, false, (V <= 2 ? 2 : 1)> MF_ARGS;                     \
  else if (params->reserved == 6) k_mf_sgd<L, V, false, 1> MF_ARGS;                                   \
  else k_mf_sgd<L, V, false, 1, (V == 1 ? 8 : 1)> MF_ARGS
  MF_DISPATCH(FN);
#undef FN
#undef MF_ARGS
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_mf_predict(const FpsTableView* items, const FpsTableView* users, const int32_t* d_user_ids,
                   const int32_t* d_item_ids, int64_t n, float* d_out, void* stream) {
  if (int e = check_pair(items, users, "fps_mf_predict")) return e;
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const Shape sh = shape_for(items->stride);
  const int gpb = MF_THREADS / sh.lanes;
  long long blocks = (n + gpb - 1) / gpb;
  const long long cap = (long long)fps_num_sms() * MF_BLOCKS_PER_SM;
  if (blocks > cap) blocks = cap;
#define FN(L, V, Z)                                                                                   \
  if (Z) k_mf_predict<L, V, true><<<(int)blocks, MF_THREADS, 0, st>>>(*items, *users, d_user_ids, d_item_ids, n, d_out); \
  else k_mf_predict<L, V, false><<<(int)blocks, MF_THREADS, 0, st>>>(*items, *users, d_user_ids, d_item_ids, n, d_out)
  MF_DISPATCH(FN);
#undef FN
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_mf_online_stream(const FpsTableView* items, const FpsTableView* users, const int32_t* ring_users,
                         const int32_t* ring_items, const float* ring_ratings, uint64_t ring_capacity,
                         const volatile uint64_t* tail, const volatile int32_t* stop,
                         const FpsMfParams* params, uint32_t iteration_wait_ms, uint32_t max_lifetime_ms,
                         FpsFlow* d_flow, void* stream) {
  if (int e = check_pair(items, users, "fps_mf_online_stream")) return e;
  FPS_REQUIRE(params && d_flow && tail && stop && ring_capacity > 0, FPS_E_INVALID, "fps_mf_online_stream: bad arguments");
  FPS_REQUIRE(items->n_replicas <= 1, FPS_E_UNSUPPORTED, "fps_mf_online_stream: replicated item tables are not supported");
  FPS_REQUIRE(max_lifetime_ms > 0 && max_lifetime_ms <= 600000u, FPS_E_INVALID,
              "fps_mf_online_stream: max_lifetime_ms must be in (0, 600000]");
  cudaStream_t st = (cudaStream_t)stream;
  int chunk = 64;
  if (params->pull_limit > 0 && params->pull_limit < chunk) chunk = params->pull_limit;
  // one resident wave only: every CTA of a polling kernel must be co-resident
  int blocks = fps_num_sms() * 2;
  int cta_credits = 0;
  if (params->pull_limit > 0) {
    const Shape sh = shape_for(items->stride);
    const int per_pass = MF_THREADS / sh.lanes;
    cta_credits = per_pass < chunk ? per_pass : chunk;
    const int fit = params->pull_limit / cta_credits > 0 ? params->pull_limit / cta_credits : 1;
    if (blocks > fit) blocks = fit;
  }
  const unsigned long long wait_ns = (unsigned long long)iteration_wait_ms * 1000000ull;
  const unsigned long long life_ns = (unsigned long long)max_lifetime_ms * 1000000ull;
#define FN(L, V, Z)                                                                                   \
  if (Z) k_mf_stream<L, V, true><<<blocks, MF_THREADS, 0, st>>>(*items, *users, ring_users, ring_items, ring_ratings, ring_capacity, (const volatile unsigned long long*)tail, stop, *params, chunk, cta_credits, wait_ns, life_ns, d_flow); \
  else k_mf_stream<L, V, false><<<blocks, MF_THREADS, 0, st>>>(*items, *users, ring_users, ring_items, ring_ratings, ring_capacity, (const volatile unsigned long long*)tail, stop, *params, chunk, cta_credits, wait_ns, life_ns, d_flow)
  MF_DISPATCH(FN);
#undef FN
  FPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
