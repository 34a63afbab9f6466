// MF variant "tiles": hot-key pre-combining in shared memory, pulled rows staged by the TMA engine,
// no library kernel anywhere on the path (north star items 2 and 3).
//
// Step = two phases.
//
//   fps_mf_bucket     groups the step's ratings by (item mod NB) with a counting sort of our own
//                     (k_bucket_count -> k_bucket_scan -> k_bucket_scatter): all records of one item
//                     end up in the same ~1K-record neighbourhood of a packed 16-byte record array.
//                     It only depends on the step's inputs, so the caller may run it on another
//                     stream while the previous step's kernel is still busy (DeviceMF does).
//                     Buckets are small (~256 records); a tile is ~3/4 of the tile size worth of
//                     consecutive records whose ends are moved back to a bucket boundary when one is
//                     near (k_tile_bounds), so that an item's records are not cut in two by a tile
//                     edge (a cut costs a second pull and push of the same row: +8.6 % rows measured
//                     with bucket-sized tiles cut at fixed offsets).
//   fps_mf_sgd_tiles  one CTA per tile of TILE consecutive packed records: load the tile into shared
//                     memory, bitonic-sort it by item there, then every (sub-)warp walks a contiguous
//                     segment of the sorted tile carrying (item, pulled row, accumulated delta) in
//                     registers: a run of equal items costs ONE pull and one combined push per
//                     TL_RUN_FLUSH records instead of one of each per record.
//
// For rows of up to 512 bytes handled by a whole warp (rank 125..128, the headline shape) the rows
// are not loaded through registers: each warp runs a private producer/consumer pipeline in which the
// TMA engine (plain cp.async.bulk, no multicast -> SASS UBLKCP) copies the next RU user rows and the
// rows of the next RI item runs into a shared-memory ring, completion signalled on mbarriers.  That
// is ~3 KB in flight per warp without spending a register on it, against 1 KB for the register path
// - which is what a pull that crosses NVLink (about 3x the latency of local HBM) needs.  The peer
// rows are read by the copy engine of the *requesting* SM straight from the owner's HBM; the delta
// goes back as a 16-byte red.add per lane as in the other variants.
//
// Pull limiter: credits are handed out hierarchically.  A CTA takes its whole allowance from the
// worker's pool (FpsFlow.in_flight) once, when it starts, and gives it back when it leaves; inside
// the CTA the allowance is enforced by construction (a walker never has more item rows outstanding
// than its share: the carried row plus its ring slots).  No CTA spins on the shared word in steady
// state, and the grid is sized so that the allowances of all resident CTAs fit the limit.
#include "common.cuh"
#include "flow.cuh"
#include "mf_core.cuh"

namespace {

constexpr int TL_THREADS = 256;
constexpr int TL_MAX_TILE = 2048;
constexpr int TL_MAX_BUCKETS = 16384;
constexpr int TL_BUCKET_TARGET = 256;   // average records per bucket
constexpr int TL_RUN_FLUSH = 32;   // records of one run after which the combined delta is published (less under a limiter)
constexpr int TL_RU = 4;           // user rows in flight per warp (staged path)
constexpr int TL_RI = 2;           // item rows in flight per warp (staged path)
constexpr size_t TL_HDR_BYTES = 2 * TL_MAX_BUCKETS * sizeof(int);   // counts + cursors (= bucket ends after the scatter)
constexpr int TL_MIN_NOMINAL = 384;      // nominal records per tile at the smallest tile size (512 * 3 / 4)

// scratch layout: [counts | cursors][tile_start: n / TL_MIN_NOMINAL + 2 ints, padded][records]
__host__ __device__ inline size_t tl_bounds_bytes(long long n) {
  return (((size_t)(n / TL_MIN_NOMINAL + 2) * sizeof(int)) + 255) & ~(size_t)255;
}

struct __align__(16) TileRec {
  int user, item;
  float rating;
  int pad;
};
// In shared memory the record carries where its rows live instead of the ids that say so: (owner
// shard << 28 | slot) for both rows, computed once per record by the loading thread rather than by a
// whole warp at every use (id -> (owner, slot) is a division for shard counts that are not powers of 2).
constexpr unsigned int TL_SLOT_MASK = 0x0fffffffu;
__device__ __forceinline__ float* pack_row(const FpsTableView& t, unsigned int pack) {
  return t.rows[pack >> 28] + (size_t)(pack & TL_SLOT_MASK) * t.stride;
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tl_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tl_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tl_mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
  }
}
// global -> shared bulk copy executed by the TMA engine (no multicast); completes `bytes` on `bar`
__device__ __forceinline__ void tl_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_addr(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}

// the same on shared-space addresses computed once per warp (a generic -> shared conversion per call was
// 15 of the kernel's ~360 instructions per record)
__device__ __forceinline__ void tl_expect_and_copy(uint32_t dst_s, const void* gmem_src, uint32_t bytes, uint32_t bar_s) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_s), "l"(gmem_src), "r"(bytes), "r"(bar_s)
               : "memory");
}
__device__ __forceinline__ void tl_wait_s(uint32_t bar_s, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar_s), "r"(parity)
        : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// phase 1: counting sort by (item & mask) into packed records
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
k_bucket_count(const int* __restrict__ iid, long long n, int nb, int* __restrict__ counts) {
  extern __shared__ int s_hist[];
  for (int b = threadIdx.x; b < nb; b += blockDim.x) s_hist[b] = 0;
  __syncthreads();
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(n, lo + per);
  for (long long k = lo + threadIdx.x; k < hi; k += blockDim.x) atomicAdd(&s_hist[iid[k] & (nb - 1)], 1);
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += blockDim.x)
    if (s_hist[b]) atomicAdd(&counts[b], s_hist[b]);
}

// exclusive scan of counts[nb] into cursors[nb]; clears counts for the next step (one CTA).
// Coalesced in and out through shared memory (the first version had every thread walk 16 consecutive
// global words: 23 us for 16K bins).
__global__ void __launch_bounds__(1024)
k_bucket_scan(int* __restrict__ counts, int* __restrict__ cursors, int nb) {
  extern __shared__ int s_bins[];          // nb ints
  __shared__ int s_warp[32];
  for (int b = threadIdx.x; b < nb; b += 1024) {
    s_bins[b] = counts[b];
    counts[b] = 0;
  }
  __syncthreads();
  const int per = (nb + 1023) / 1024;
  const int b0 = threadIdx.x * per;
  int sum = 0;
  for (int q = 0; q < per; ++q)
    if (b0 + q < nb) sum += s_bins[b0 + q];
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((threadIdx.x & 31) >= o) incl += t;
  }
  if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
  __syncthreads();
  if (threadIdx.x < 32) {
    int w = s_warp[threadIdx.x];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (threadIdx.x >= o) w += t;
    }
    s_warp[threadIdx.x] = w;
  }
  __syncthreads();
  int run = incl - sum + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0);
  for (int q = 0; q < per; ++q) {
    if (b0 + q < nb) {
      const int c = s_bins[b0 + q];
      s_bins[b0 + q] = run;
      run += c;
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += 1024) cursors[b] = s_bins[b];
}

// every CTA owns one contiguous range of the input: count its records per bucket in shared memory,
// claim one output range per non-empty bucket with a single global atomic, then place the records
__global__ void __launch_bounds__(1024)
k_bucket_scatter(const int* __restrict__ uid, const int* __restrict__ iid, const float* __restrict__ rating,
                 long long n, int nb, int* __restrict__ cursors, TileRec* __restrict__ out) {
  extern __shared__ int s_mem[];
  int* s_cnt = s_mem;          // [nb] records of this CTA per bucket, then running rank
  int* s_base = s_mem + nb;    // [nb] first output slot of this CTA in the bucket
  for (int b = threadIdx.x; b < nb; b += blockDim.x) s_cnt[b] = 0;
  __syncthreads();
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(n, lo + per);
  for (long long k = lo + threadIdx.x; k < hi; k += blockDim.x) atomicAdd(&s_cnt[iid[k] & (nb - 1)], 1);
  __syncthreads();
  // one global atomic per non-empty (CTA, bucket); four in flight per thread (each is a round trip to L2)
  for (int b = threadIdx.x; b < nb; b += 4 * blockDim.x) {
    int c[4], base[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int bb = b + q * blockDim.x;
      c[q] = bb < nb ? s_cnt[bb] : 0;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) base[q] = c[q] ? atomicAdd(&cursors[b + q * blockDim.x], c[q]) : 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int bb = b + q * blockDim.x;
      if (bb < nb) {
        s_base[bb] = base[q];
        s_cnt[bb] = 0;
      }
    }
  }
  __syncthreads();
  for (long long k = lo + threadIdx.x; k < hi; k += blockDim.x) {
    TileRec r;
    r.item = iid[k];
    r.user = uid[k];
    r.rating = rating ? rating[k] : 0.f;
    r.pad = 0;
    const int b = r.item & (nb - 1);
    const int pos = s_base[b] + atomicAdd(&s_cnt[b], 1);
    out[pos] = r;
  }
}

// Tile t nominally starts at record t * nominal; if a bucket boundary lies at most `snap` records
// before that point the tile starts there instead (no small bucket is ever cut; a bucket larger than
// `snap` - a hot item - is cut at the nominal points).  Tiles are at most nominal + snap records.
// `ends` = the cursors after the scatter = the end offset of every bucket.
__global__ void __launch_bounds__(256)
k_tile_bounds(const int* __restrict__ ends, int nb, long long n, int nominal, int snap, int n_tiles,
              int* __restrict__ tile_start) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > n_tiles) return;
  const long long p = min(n, (long long)t * nominal);
  int s = (int)p;
  if (p > 0 && p < n) {
    int lo = 0, hi = nb;                    // number of buckets whose end is <= p
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (ends[mid] <= p) lo = mid + 1; else hi = mid;
    }
    const int b = lo > 0 ? ends[lo - 1] : 0;
    if (p - b <= snap) s = b;
  }
  tile_start[t] = s;
}

// ---------------------------------------------------------------------------------------------
// phase 2: per-tile shared-memory sort + run-carrying SGD
// ---------------------------------------------------------------------------------------------
struct TileCounters {
  float loss;
  unsigned int records, pulls, pushes, rpulls, rpushes, nan, bad;
  __device__ __forceinline__ void clear() { loss = 0.f; records = pulls = pushes = rpulls = rpushes = nan = bad = 0; }
  __device__ __forceinline__ void flush(FpsFlow* flow) {
    __shared__ float s_loss[TL_THREADS / 32];
    __shared__ unsigned int s_c[TL_THREADS / 32][7];
    float l = loss;
    unsigned int v[7] = {records, pulls, pushes, rpulls, rpushes, nan, bad};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l += __shfl_xor_sync(0xffffffffu, l, o);
#pragma unroll
      for (int q = 0; q < 7; ++q) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
      s_loss[w] = l;
#pragma unroll
      for (int q = 0; q < 7; ++q) s_c[w][q] = v[q];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double L = 0.0;
      unsigned long long t[7] = {0, 0, 0, 0, 0, 0, 0};
      for (int i = 0; i < TL_THREADS / 32; ++i) {
        L += s_loss[i];
        for (int q = 0; q < 7; ++q) t[q] += s_c[i][q];
      }
      if (t[0]) {
        atomicAdd(&flow->loss_sum, L);
        atomicAdd((unsigned long long*)&flow->records, t[0]);
      }
      if (t[1]) {
        atomicAdd((unsigned long long*)&flow->pulls, t[1]);
        atomicAdd((unsigned long long*)&flow->reserved[0], t[1]);     // "combined_rows" of the older variant
      }
      if (t[2]) atomicAdd((unsigned long long*)&flow->pushes, t[2]);
      if (t[3]) atomicAdd((unsigned long long*)&flow->remote_pulls, t[3]);
      if (t[4]) atomicAdd((unsigned long long*)&flow->remote_pushes, t[4]);
      if (t[5]) atomicAdd((unsigned long long*)&flow->nan_count, t[5]);
      if (t[6]) atomicAdd((unsigned long long*)&flow->bad_ids, t[6]);
    }
  }
};

// Load tile `tile` into shared memory, drop records with ids outside the tables (counted as bad),
// sort the rest by item.  Returns the number of valid records; s_key[k] = item << 32 | slot in s_rec.
__device__ __forceinline__ int tile_load_sort(const FpsTableView& items, const FpsTableView& users,
                                              const TileRec* __restrict__ recs, const int* __restrict__ tile_start,
                                              long long tile, TileRec* s_rec, unsigned long long* s_key,
                                              TileCounters& c) {
  const long long base = tile_start[tile];
  const int cnt = tile_start[tile + 1] - (int)base;
  int sort_n = 32;
  while (sort_n < cnt) sort_n <<= 1;
  for (int k = threadIdx.x; k < sort_n; k += TL_THREADS) {
    unsigned long long key = ~0ull;
    if (k < cnt) {
      TileRec r = recs[base + k];
      const RowRef ir = locate(items, r.item), ur = locate(users, r.user);
      if (r.user == FPS_SKIP_ID) {
        // an unused negative-sample slot: neither processed nor counted
      } else if (in_table(items, r.item, ir) && in_table(users, r.user, ur) && users.rows[ur.owner] != nullptr) {
        key = ((unsigned long long)(unsigned int)r.item << 32) | (unsigned int)k;
      } else {
        ++c.bad;
      }
      r.user = (int)(((unsigned int)ur.owner << 28) | ((unsigned int)ur.slot & TL_SLOT_MASK));
      r.pad = (int)(((unsigned int)ir.owner << 28) | ((unsigned int)ir.slot & TL_SLOT_MASK));
      s_rec[k] = r;
    }
    s_key[k] = key;
  }
  __syncthreads();
  // bitonic sort, ascending
  for (int size = 2; size <= sort_n; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < (sort_n >> 1); t += TL_THREADS) {
        const int i = 2 * t - (t & (stride - 1));
        const int j = i + stride;
        const unsigned long long a = s_key[i], b = s_key[j];
        const bool up = (i & size) == 0;
        if ((a > b) == up) {
          s_key[i] = b;
          s_key[j] = a;
        }
      }
      __syncthreads();
    }
  }
  // valid records sort before the ~0 keys: binary search for the first invalid key
  int lo = 0, hi = cnt;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((unsigned int)(s_key[mid] >> 32) == 0xffffffffu) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// Where the walkers' segments of a sorted tile may begin.  Cut points are the run boundaries plus,
// inside a run longer than TL_SERIAL, every TL_SERIAL-th record of it: a run is walked serially by ONE
// walker (exact serial SGD on the carried row, one pull) unless it is so long that this would leave
// the rest of the CTA idle.  Returns the first cut point >= p.  Without this a hot item's run was
// split over every (sub-)warp of the CTA, all starting from the same stale row - the sum of their
// deltas overshoots by the number of walkers (diverged at rank 16, 64 sub-warps per CTA).
constexpr int TL_SERIAL = 128;
__device__ __forceinline__ int tile_cut(const unsigned long long* s_key, int valid, int p) {
  if (p <= 0) return 0;
  if (p >= valid) return valid;
  const unsigned int item = (unsigned int)(s_key[p] >> 32);
  if ((unsigned int)(s_key[p - 1] >> 32) != item) return p;
  int lo = 0, hi = p;                       // first record of the run that contains p
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((unsigned int)(s_key[mid] >> 32) < item) lo = mid + 1; else hi = mid;
  }
  const int start = lo;
  lo = p; hi = valid;                       // first record after the run
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((unsigned int)(s_key[mid] >> 32) <= item) lo = mid + 1; else hi = mid;
  }
  const int next_cut = start + ((p - start + TL_SERIAL - 1) / TL_SERIAL) * TL_SERIAL;
  return min(lo, next_cut);
}

template <int LANES, int VPL>
struct TCarry {
  int item;
  int run_len;
  float* row;
  float4 v[VPL];
  float4 acc[VPL];
  float4 back[VPL];     // long runs: the row as the last push-and-refresh found it (arrives asynchronously)
  float4 sent[VPL];     // ... and the delta that push carried
  bool remote, pending;
};

template <int LANES, int VPL>
__device__ __forceinline__ void tcarry_push(TCarry<LANES, VPL>& cy, int nvec, TileCounters& c) {
  const int lane = threadIdx.x % LANES;
#pragma unroll
  for (int q = 0; q < VPL; ++q) {
    const int b = lane + q * LANES;
    if (b < nvec) red_add4(cy.row + b * 4, cy.acc[q]);
    cy.acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  cy.run_len = 0;
  if (lane == 0) {
    ++c.pushes;
    if (cy.remote) ++c.rpushes;
  }
}

// Long run: every TL_RUN_FLUSH records publish the delta gathered so far AND refresh the carried row
// in the same round trip (the add returns the previous value), so walkers on the same hot item feed
// back into each other - without it the sum of many walkers' deltas, all computed from the same stale
// row, overshoots (diverged on a 6 %-hot item at rank 16).  The returned row is only consumed at the
// NEXT flush point (v = returned + what that push carried + what was gathered since), so the warp
// never waits for the round trip: feedback lags by one flush period instead of stalling the walker
// (the blocking form cost 35 % on the rank-64 online config).
template <int LANES, int VPL>
__device__ __forceinline__ void tcarry_exchange(TCarry<LANES, VPL>& cy, int nvec, TileCounters& c) {
  const int lane = threadIdx.x % LANES;
#pragma unroll
  for (int q = 0; q < VPL; ++q) {
    const int b = lane + q * LANES;
    if (b < nvec) {
      if (cy.pending) {
        cy.v[q].x = cy.back[q].x + cy.sent[q].x + cy.acc[q].x; cy.v[q].y = cy.back[q].y + cy.sent[q].y + cy.acc[q].y;
        cy.v[q].z = cy.back[q].z + cy.sent[q].z + cy.acc[q].z; cy.v[q].w = cy.back[q].w + cy.sent[q].w + cy.acc[q].w;
      }
      cy.back[q] = atom_add4(cy.row + b * 4, cy.acc[q]);
    }
    cy.sent[q] = cy.acc[q];
    cy.acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  cy.pending = true;
  cy.run_len = 0;
  if (lane == 0) {
    ++c.pushes; ++c.pulls;
    if (cy.remote) { ++c.rpushes; ++c.rpulls; }
  }
}

// the SGD step of one record against the carried item row; `us` = user row as loaded
template <int LANES, int VPL>
__device__ __forceinline__ void tile_sgd(TCarry<LANES, VPL>& cy, const float4 (&us)[VPL], float* urow, float r,
                                         bool active, int nvec, const FpsMfParams& p, int flush_every,
                                         TileCounters& c) {
  const int lane = threadIdx.x % LANES;
  float dot = 0.f;
#pragma unroll
  for (int q = 0; q < VPL; ++q) dot += dot4(us[q], cy.v[q]);
  dot = group_sum<LANES>(dot);
  if (!active) return;
  if (lane == 0) ++c.records;
  const float e = r - dot;
  if (!isfinite(e)) {
    if (lane == 0) ++c.nan;
    return;
  }
  const float g = p.lr * e, decay = p.lr * p.lambda;
#pragma unroll
  for (int q = 0; q < VPL; ++q) {
    const int b = lane + q * LANES;
    if (b < nvec) {
      const float4 u = us[q], w = cy.v[q];
      float4 un, dv;
      un.x = u.x + fmaf(g, w.x, -decay * u.x);
      un.y = u.y + fmaf(g, w.y, -decay * u.y);
      un.z = u.z + fmaf(g, w.z, -decay * u.z);
      un.w = u.w + fmaf(g, w.w, -decay * u.w);
      st_row4(urow + b * 4, un);
      dv.x = fmaf(g, u.x, -decay * w.x);
      dv.y = fmaf(g, u.y, -decay * w.y);
      dv.z = fmaf(g, u.z, -decay * w.z);
      dv.w = fmaf(g, u.w, -decay * w.w);
      cy.acc[q].x += dv.x; cy.acc[q].y += dv.y; cy.acc[q].z += dv.z; cy.acc[q].w += dv.w;
      cy.v[q].x += dv.x; cy.v[q].y += dv.y; cy.v[q].z += dv.z; cy.v[q].w += dv.w;   // serial SGD inside a run
    }
  }
  if (lane == 0) c.loss += e * e;
  if (++cy.run_len >= flush_every) tcarry_exchange<LANES, VPL>(cy, nvec, c);
}

// ---- register path: any rank, LANES lanes per record -----------------------------------------
template <int LANES, int VPL>
__global__ void __launch_bounds__(TL_THREADS, (VPL == 1 ? 3 : 1))
k_mf_tiles_reg(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
               const TileRec* __restrict__ recs, const int* __restrict__ tile_start, const FpsMfParams p, int tile_sz,
               long long n_tiles, long long tile_mul, int active_groups, int cta_credits, int flush_every,
               FpsFlow* flow) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  TileRec* s_rec = reinterpret_cast<TileRec*>(s_raw);
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_raw + (size_t)tile_sz * sizeof(TileRec));
  constexpr int GPB = TL_THREADS / LANES;
  const int grp = threadIdx.x / LANES, lane = threadIdx.x % LANES;
  const int nvec = items.stride >> 2;
  const int groups = active_groups > 0 && active_groups < GPB ? active_groups : GPB;   // limiter: fewer walkers
  TileCounters c;
  c.clear();
  if (p.pull_limit > 0) flow_acquire(flow, cta_credits, p.pull_limit);
  for (long long lin = blockIdx.x; lin < n_tiles; lin += gridDim.x) {
    const long long tile = (lin * tile_mul) % n_tiles;
    const int valid = tile_load_sort(items, users, recs, tile_start, tile, s_rec, s_key, c);
    const int nominal = (valid + groups - 1) / groups;
    const int b0 = grp < groups ? tile_cut(s_key, valid, grp * nominal) : valid;
    const int b1 = grp < groups ? tile_cut(s_key, valid, (grp + 1) * nominal) : valid;
    // block-uniform trip count (shuffles inside the loop): the longest segment of the CTA
    __shared__ int s_seg;
    if (threadIdx.x == 0) s_seg = 0;
    __syncthreads();
    if (lane == 0 && b1 > b0) atomicMax(&s_seg, b1 - b0);
    __syncthreads();
    const int seg = s_seg;
    TCarry<LANES, VPL> cy;
    cy.item = -1;
    cy.run_len = 0;
    cy.row = nullptr;
    cy.remote = false;
    cy.pending = false;
#pragma unroll
    for (int q = 0; q < VPL; ++q) cy.acc[q] = cy.v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    // user row of the next record is loaded one iteration ahead (register double buffer)
    float4 nxt_us[VPL];
    float* nxt_urow = nullptr;
    int nxt_user = -1;
    auto load_user = [&](int k, float4 (&dst)[VPL], float*& urow, int& user) {
      urow = nullptr;
      user = -1;
#pragma unroll
      for (int q = 0; q < VPL; ++q) dst[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k < b1) {
        const TileRec r = s_rec[(unsigned int)s_key[k]];
        urow = pack_row(users, (unsigned int)r.user);
        user = r.user;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
          const int b = lane + q * LANES;
          if (b < nvec) dst[q] = ld_row4(urow + b * 4);
        }
      }
    };
    load_user(b0, nxt_us, nxt_urow, nxt_user);
    for (int j = 0; j < seg; ++j) {          // block-uniform trip count: shuffles inside
      const int k = b0 + j;
      const bool active = k < b1;
      float4 us[VPL];
      float* urow = nxt_urow;
      const int user = nxt_user;
#pragma unroll
      for (int q = 0; q < VPL; ++q) us[q] = nxt_us[q];
      float r = 0.f;
      if (active) {
        const unsigned long long key = s_key[k];
        const int item = (int)(key >> 32);
        const TileRec rec = s_rec[(unsigned int)key];
        r = rec.rating;
        if (item != cy.item) {                 // sub-warp-uniform
          if (cy.item >= 0 && cy.run_len > 0) tcarry_push<LANES, VPL>(cy, nvec, c);
          cy.item = item;
          cy.row = pack_row(items, (unsigned int)rec.pad);
          cy.remote = (int)((unsigned int)rec.pad >> 28) != p.local_shard;
          cy.pending = false;
          cy.run_len = 0;
#pragma unroll
          for (int q = 0; q < VPL; ++q) {
            const int b = lane + q * LANES;
            cy.v[q] = b < nvec ? ld_row4(cy.row + b * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
            cy.acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          if (lane == 0) {
            ++c.pulls;
            if (cy.remote) ++c.rpulls;
          }
        }
      }
      load_user(k + 1, nxt_us, nxt_urow, nxt_user);
      tile_sgd<LANES, VPL>(cy, us, urow, r, active, nvec, p, flush_every, c);
      // the row prefetched for the next record was read before this record's store: reload on a clash
      if (active && nxt_user == user && nxt_urow != nullptr) {
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
          const int b = lane + q * LANES;
          if (b < nvec) nxt_us[q] = ld_row4(nxt_urow + b * 4);
        }
      }
    }
    if (cy.item >= 0 && cy.run_len > 0) tcarry_push<LANES, VPL>(cy, nvec, c);
    __syncthreads();    // the next tile overwrites s_rec / s_key
  }
  if (p.pull_limit > 0) flow_release(flow, cta_credits);
  c.flush(flow);
}

// ---- TMA-staged path: one warp per segment, rows of <= 512 bytes -------------------------------
struct WarpRing {
  float rows[(TL_RU + TL_RI) * 128];     // RU user slots then RI item slots, 512 B each
  uint64_t bar[TL_RU + TL_RI];
  float* gptr[TL_RU + TL_RI];            // global address of the row held in each slot
};

__global__ void __launch_bounds__(TL_THREADS, 4)
k_mf_tiles_tma(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
               const TileRec* __restrict__ recs, const int* __restrict__ tile_start, const FpsMfParams p, int tile_sz,
               long long n_tiles, long long tile_mul, int item_depth, int active_warps, int cta_credits,
               int flush_every, FpsFlow* flow) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  TileRec* s_rec = reinterpret_cast<TileRec*>(s_raw);
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_raw + (size_t)tile_sz * sizeof(TileRec));
  WarpRing* rings = reinterpret_cast<WarpRing*>(s_raw + (size_t)tile_sz * (sizeof(TileRec) + 8));
  constexpr int NW = TL_THREADS / 32;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = items.stride >> 2;
  const uint32_t row_bytes = (uint32_t)items.stride * 4u;
  WarpRing& ring = rings[w];
  const uint32_t rows_s = smem_addr(ring.rows), bars_s = smem_addr(ring.bar);   // shared-space addresses of this warp's ring
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < TL_RU + TL_RI; ++s) tl_mbar_init(&ring.bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  TileCounters c;
  c.clear();
  if (p.pull_limit > 0) flow_acquire(flow, cta_credits, p.pull_limit);
  __syncthreads();
  uint32_t phase = 0;                     // bit s = parity the next wait on slot s expects
  item_depth = item_depth < 0 ? 0 : (item_depth > TL_RI ? TL_RI : item_depth);
  const int RI = item_depth < 1 ? 1 : item_depth;     // item ring slots in use
  const int walkers = active_warps > 0 && active_warps < NW ? active_warps : NW;   // limiter: fewer walkers
  for (long long lin = blockIdx.x; lin < n_tiles; lin += gridDim.x) {
    const long long tile = (lin * tile_mul) % n_tiles;
    const int valid = tile_load_sort(items, users, recs, tile_start, tile, s_rec, s_key, c);
    const int nominal = (valid + walkers - 1) / walkers;
    const int b0 = w < walkers ? tile_cut(s_key, valid, w * nominal) : valid;
    const int b1 = w < walkers ? tile_cut(s_key, valid, (w + 1) * nominal) : valid;
    const int len = b1 - b0;
    // ---- producer side ----------------------------------------------------------------------
    auto issue_user = [&](int j) {         // record j of the segment -> user slot j % RU
      if (j < len && lane == 0) {
        const TileRec r = s_rec[(unsigned int)s_key[b0 + j]];
        float* g = pack_row(users, (unsigned int)r.user);
        const int s = j % TL_RU;
        ring.gptr[s] = g;
        tl_expect_and_copy(rows_s + s * 512, g, row_bytes, bars_s + s * 8);
      }
    };
    int pi = 0, runs_issued = 0, runs_consumed = 0, p_item = -1;
    int slot_issue = 0, slot_consume = 0;          // runs_issued % RI, runs_consumed % RI without the division
    auto pump_items = [&](int cap) {       // keep up to `cap` item rows of upcoming runs in flight
      while (runs_issued - runs_consumed < cap && pi < len) {
        const unsigned long long key = s_key[b0 + pi];
        const int item = (int)(key >> 32);
        if (item != p_item) {
          p_item = item;
          if (lane == 0) {
            float* g = pack_row(items, (unsigned int)s_rec[(unsigned int)key].pad);
            const int s = TL_RU + slot_issue;
            ring.gptr[s] = g;
            tl_expect_and_copy(rows_s + s * 512, g, row_bytes, bars_s + s * 8);
          }
          ++runs_issued;
          slot_issue = slot_issue + 1 == RI ? 0 : slot_issue + 1;
        }
        ++pi;
      }
    };
#pragma unroll
    for (int j = 0; j < TL_RU; ++j) issue_user(j);
    pump_items(item_depth);
    // ---- consumer side ----------------------------------------------------------------------
    TCarry<32, 1> cy;
    cy.item = -1;
    cy.run_len = 0;
    cy.row = nullptr;
    cy.remote = false;
    cy.pending = false;
    cy.acc[0] = cy.v[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    int recent[TL_RU];                     // users of the last RU records (their stores may be newer than a prefetch)
#pragma unroll
    for (int q = 0; q < TL_RU; ++q) recent[q] = -1;
    for (int j = 0; j < len; ++j) {
      const unsigned long long key = s_key[b0 + j];
      const int item = (int)(key >> 32);
      const TileRec r = s_rec[(unsigned int)key];
      // user row
      const int su = j % TL_RU;
      tl_wait_s(bars_s + su * 8, (phase >> su) & 1u);
      phase ^= 1u << su;
      float* urow = ring.gptr[su];
      float4 us[1];
      us[0] = lane < nvec ? *reinterpret_cast<const float4*>(&ring.rows[su * 128 + lane * 4])
                          : make_float4(0.f, 0.f, 0.f, 0.f);
      bool clash = false;
#pragma unroll
      for (int q = 0; q < TL_RU; ++q) clash |= recent[q] == r.user;
      if (clash && lane < nvec) us[0] = ld_row4(urow + lane * 4);
#pragma unroll
      for (int q = TL_RU - 1; q > 0; --q) recent[q] = recent[q - 1];
      recent[0] = r.user;
      __syncwarp();
      issue_user(j + TL_RU);
      // item row
      if (item != cy.item) {
        if (cy.item >= 0 && cy.run_len > 0) tcarry_push<32, 1>(cy, nvec, c);
        if (runs_issued == runs_consumed) pump_items(1);     // depth 0 (tight limiter): fetch on demand
        const int si = TL_RU + slot_consume;
        slot_consume = slot_consume + 1 == RI ? 0 : slot_consume + 1;
        tl_wait_s(bars_s + si * 8, (phase >> si) & 1u);
        phase ^= 1u << si;
        cy.item = item;
        cy.row = ring.gptr[si];
        cy.v[0] = lane < nvec ? *reinterpret_cast<const float4*>(&ring.rows[si * 128 + lane * 4])
                              : make_float4(0.f, 0.f, 0.f, 0.f);
        cy.acc[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        cy.run_len = 0;
        cy.pending = false;
        cy.remote = (int)((unsigned int)r.pad >> 28) != p.local_shard;
        if (lane == 0) {
          ++c.pulls;
          if (cy.remote) ++c.rpulls;
        }
        __syncwarp();
        ++runs_consumed;
        pump_items(item_depth);
      }
      tile_sgd<32, 1>(cy, us, urow, r.rating, true, nvec, p, flush_every, c);
    }
    if (cy.item >= 0 && cy.run_len > 0) tcarry_push<32, 1>(cy, nvec, c);
    __syncthreads();
  }
  if (p.pull_limit > 0) flow_release(flow, cta_credits);
  c.flush(flow);
}

int n_buckets_for(long long n) {
  int nb = 1;
  while (nb < TL_MAX_BUCKETS && (long long)nb * TL_BUCKET_TARGET < n) nb <<= 1;
  return nb;
}

}  // namespace

extern "C" {

size_t fps_mf_tiles_scratch_bytes(int64_t n) {
  return TL_HDR_BYTES + tl_bounds_bytes(n > 0 ? n : 0) + (size_t)(n > 0 ? n : 0) * sizeof(TileRec) + 256;
}

int fps_mf_bucket(const int32_t* d_user_ids, const int32_t* d_item_ids, const float* d_ratings, int64_t n,
                  void* d_scratch, size_t scratch_bytes, void* stream) {
  FPS_REQUIRE(d_scratch != nullptr && scratch_bytes >= fps_mf_tiles_scratch_bytes(n), FPS_E_INVALID,
              "fps_mf_bucket: scratch of %zu bytes is smaller than fps_mf_tiles_scratch_bytes(%lld)", scratch_bytes,
              (long long)n);
  FPS_REQUIRE(n <= 0x7fffffffll, FPS_E_TOO_LARGE, "fps_mf_bucket: at most 2^31-1 ratings per step");
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  int* counts = (int*)d_scratch;
  int* cursors = counts + TL_MAX_BUCKETS;
  TileRec* out = (TileRec*)((unsigned char*)d_scratch + TL_HDR_BYTES + tl_bounds_bytes(n));
  const int nb = n_buckets_for(n);
  static bool attr_done[64] = {false};
  int dev = 0;
  FPS_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    FPS_CUDA(cudaFuncSetAttribute(k_bucket_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  2 * TL_MAX_BUCKETS * (int)sizeof(int)));
    FPS_CUDA(cudaFuncSetAttribute(k_bucket_count, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  TL_MAX_BUCKETS * (int)sizeof(int)));
    FPS_CUDA(cudaFuncSetAttribute(k_bucket_scan, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  TL_MAX_BUCKETS * (int)sizeof(int)));
    attr_done[dev] = true;
  }
  FPS_CUDA(cudaMemsetAsync(counts, 0, (size_t)nb * sizeof(int), st));
  // one CTA per SM at most: every CTA walks all nb bins of its shared-memory histogram three times,
  // which only pays when it also owns a long range of records
  long long blocks = (n + 4095) / 4096;
  const long long cap = (long long)fps_num_sms();
  if (blocks > cap) blocks = cap;
  const int threads = n >= (1 << 20) ? 1024 : 256;
  k_bucket_count<<<(int)blocks, threads, nb * sizeof(int), st>>>(d_item_ids, n, nb, counts);
  k_bucket_scan<<<1, 1024, nb * sizeof(int), st>>>(counts, cursors, nb);
  k_bucket_scatter<<<(int)blocks, threads, 2 * nb * sizeof(int), st>>>(d_user_ids, d_item_ids, d_ratings, n, nb, cursors, out);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_mf_sgd_tiles(const FpsTableView* items, const FpsTableView* users, void* d_scratch, int64_t n,
                     const FpsMfParams* params, FpsFlow* d_flow, void* stream) {
  if (int e = fps_check_view(items, "fps_mf_sgd_tiles")) return e;
  if (int e = fps_check_view(users, "fps_mf_sgd_tiles")) return e;
  FPS_REQUIRE(params != nullptr && d_flow != nullptr && d_scratch != nullptr, FPS_E_INVALID,
              "fps_mf_sgd_tiles: params, flow and scratch are required");
  FPS_REQUIRE(items->stride == users->stride && items->dim == users->dim, FPS_E_INVALID,
              "fps_mf_sgd_tiles: item and user tables must have the same dim/stride");
  FPS_REQUIRE(!items->lazy && !users->lazy && items->n_replicas <= 1 && items->stride % 4 == 0 && items->stride <= 512,
              FPS_E_UNSUPPORTED, "fps_mf_sgd_tiles: eager, sharded (not replicated) tables of rank <= 512 only");
  FPS_REQUIRE(items->rows_per_shard <= (int64_t)TL_SLOT_MASK && users->rows_per_shard <= (int64_t)TL_SLOT_MASK,
              FPS_E_UNSUPPORTED, "fps_mf_sgd_tiles: at most 2^28 rows per shard");
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int* ends = (const int*)d_scratch + TL_MAX_BUCKETS;
  int* tile_start = (int*)((unsigned char*)d_scratch + TL_HDR_BYTES);
  const TileRec* recs = (const TileRec*)((const unsigned char*)d_scratch + TL_HDR_BYTES + tl_bounds_bytes(n));
  const int nvec = items->stride / 4;
  int lanes = 1;
  while (lanes < nvec && lanes < 32) lanes <<= 1;
  const int vpl = (nvec + lanes - 1) / lanes;
  const int gpb = TL_THREADS / lanes;
  // params->reserved selects the tile size (512 / 1024 / 2048; 0 = 1024); params->unroll = 1 forces
  // the register path for shapes the staged path would take
  const int tile_sz = params->reserved == 512 ? 512 : params->reserved == 2048 ? 2048 : 1024;
  const int nominal = tile_sz / 4 * 3, snap = tile_sz / 4;
  const long long n_tiles = (n + nominal - 1) / nominal;
  k_tile_bounds<<<(int)((n_tiles + 1 + 255) / 256), 256, 0, st>>>(ends, n_buckets_for(n), n, nominal, snap, (int)n_tiles,
                                                               tile_start);
  long long tile_mul = (long long)((double)n_tiles * 0.6180339887) | 1;
  auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
  while (tile_mul > 1 && gcd(tile_mul, n_tiles) != 1) tile_mul -= 2;
  if (tile_mul < 1) tile_mul = 1;
  const int sms = fps_num_sms();
  const bool staged = lanes == 32 && vpl == 1 && params->unroll != 1;
  long long blocks = n_tiles;
  // Limiter.  What the reference bounds is the number of records whose pull has not been answered and
  // processed yet (M/WorkerLogic.scala:176-225); here a record is "in flight" while the row it was computed
  // against has not been refreshed from the owner and its delta not been published: a walker holds up to
  // `depth` item rows (the carried one + its ring) and, inside a run, up to 2 * flush_every unpublished
  // records (one flush period gathered, one on its way).  The limit is split evenly over the walkers of
  // all CTAs that can be resident; each CTA takes its share from FpsFlow.in_flight once.  A small limit
  // therefore shortens the flush period first (more push-and-refresh round trips on hot items = the
  // feedback that keeps Hogwild stable on small hot tables), then thins the rings, then the grid.
  const int walkers_full = staged ? TL_THREADS / 32 : gpb;   // (sub-)warps per CTA that pull
  const int depth_max = staged ? TL_RI + 1 : 1;
  int active_groups = 0, item_depth = TL_RI, cta_credits = 0, flush_every = TL_RUN_FLUSH;
  if (params->pull_limit > 0) {
    const long long limit = params->pull_limit;
    const long long resident = (long long)sms * 4;    // CTAs per SM the kernels' registers / shared memory allow
    int walkers = walkers_full;
    long long per_walker = 1;
    if (limit < walkers_full) {                       // fewer credits than walkers in one CTA
      active_groups = walkers = (int)limit;
      blocks = 1;
    } else {
      long long ctas = limit / walkers;               // at least one credit per walker
      if (ctas > resident) ctas = resident;
      per_walker = limit / (ctas * walkers);
      if (ctas < resident && blocks > ctas) blocks = ctas;   // else the full grid: only `resident` CTAs hold credits at a time
    }
    if (per_walker > 2 * TL_RUN_FLUSH) per_walker = 2 * TL_RUN_FLUSH;
    const int depth = (int)(per_walker < depth_max ? per_walker : depth_max);
    flush_every = (int)(per_walker / 2 > 0 ? per_walker / 2 : 1);
    item_depth = depth - 1;
    cta_credits = walkers * (int)per_walker;
  }
  const size_t smem_base = (size_t)tile_sz * (sizeof(TileRec) + 8);
  static bool attr_done[64] = {false};
  int dev = 0;
  FPS_CUDA(cudaGetDevice(&dev));
#define SET_ATTR(K) FPS_CUDA(cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                  (int)((size_t)TL_MAX_TILE * (sizeof(TileRec) + 8) + (TL_THREADS / 32) * sizeof(WarpRing))))
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    SET_ATTR(k_mf_tiles_tma);
    SET_ATTR((k_mf_tiles_reg<1, 1>));
    SET_ATTR((k_mf_tiles_reg<2, 1>));
    SET_ATTR((k_mf_tiles_reg<4, 1>));
    SET_ATTR((k_mf_tiles_reg<8, 1>));
    SET_ATTR((k_mf_tiles_reg<16, 1>));
    SET_ATTR((k_mf_tiles_reg<32, 1>));
    SET_ATTR((k_mf_tiles_reg<32, 2>));
    SET_ATTR((k_mf_tiles_reg<32, 4>));
    attr_done[dev] = true;
  }
#undef SET_ATTR
  if (staged) {
    const size_t smem = smem_base + (TL_THREADS / 32) * sizeof(WarpRing);
    k_mf_tiles_tma<<<(int)blocks, TL_THREADS, smem, st>>>(*items, *users, recs, tile_start, *params, tile_sz, n_tiles, tile_mul,
                                                          item_depth, active_groups, cta_credits, flush_every, d_flow);
  } else {
#define LAUNCH(L, V) k_mf_tiles_reg<L, V><<<(int)blocks, TL_THREADS, smem_base, st>>>(*items, *users, recs, tile_start, *params, tile_sz, n_tiles, tile_mul, active_groups, cta_credits, flush_every, d_flow)
    if (vpl == 1) {
      switch (lanes) {
        case 1: LAUNCH(1, 1); break;
        case 2: LAUNCH(2, 1); break;
        case 4: LAUNCH(4, 1); break;
        case 8: LAUNCH(8, 1); break;
        case 16: LAUNCH(16, 1); break;
        default: LAUNCH(32, 1); break;
      }
    } else if (vpl == 2) { LAUNCH(32, 2); }
    else { LAUNCH(32, 4); }
#undef LAUNCH
  }
  FPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
