// MF variant "combine": hot-key pre-combining at batch scope (north star item 3: "hot-key deltas
// are pre-combined" before they cross NVLink).
//
// The step's ratings are grouped by item with a radix sort (cub, keys = item id, payload = record
// index), the (user, rating) payloads are gathered into sorted order, and the fused kernel walks
// the sorted array: every sub-warp owns a contiguous slice and carries (current item, pulled row,
// accumulated delta) across consecutive records.  A run of records with the same item pulls the
// row once and pushes one combined red.add instead of one per record.  Slices are short
// (MF_SLICE records); inside a run the carried row is updated locally after every record (serial
// SGD within the run), chunks are visited in a scattered order so that a hot item's chunks do not
// all start from the same stale row, and the combined delta is pushed at the end of the run.  For MovieLens-shaped
// popularity this removes ~40 % of the item-row traffic (all of it NVLink traffic for remote
// items) and turns the hottest rows' thousands of serialised remote atomics per step into dozens.
//
// Scratch (sorted keys, permutation, gathered payloads, cub temp) is a grow-only per-device cache
// owned by the library (cudaMalloc on first use / growth).
#include <cub/device/device_radix_sort.cuh>

#include <mutex>

#include "common.cuh"
#include "flow.cuh"
#include "mf_core.cuh"

namespace {

constexpr int MF_SLICE = 32;   // records a sub-warp walks before it must flush its carried delta

struct SortScratch {
  void* base = nullptr;
  size_t bytes = 0;
};
std::mutex g_mu;
SortScratch g_scratch[64];

__global__ void __launch_bounds__(256)
k_iota(int* __restrict__ out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = (int)i;
}

__global__ void __launch_bounds__(256)
k_gather_payload(const int* __restrict__ perm, const int* __restrict__ uid, const float* __restrict__ rating,
                 int* __restrict__ su, float* __restrict__ sr, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int p = perm[i];
    su[i] = uid[p];
    sr[i] = rating[p];
  }
}

template <int LANES, int VPL>
struct Carry {
  int item;            // item whose row is held, -1 = none
  float* row;          // owner's row
  float4 v[VPL];       // row as pulled
  float4 acc[VPL];     // combined delta not yet pushed
  bool remote;
};

// REPL: the item table is replicated with deferred pushes - the combined delta also goes to the local
// delta array (fps_table_flush_delta forwards it to the other replicas after the launch).
template <int LANES, int VPL, bool REPL>
__device__ __forceinline__ void carry_flush(Carry<LANES, VPL>& cy, int nvec, const FpsTableView& items) {
  if (cy.item < 0) return;
  const int lane = threadIdx.x % LANES;
#pragma unroll
  for (int q = 0; q < VPL; ++q) {
    const int b = lane + q * LANES;
    if (b < nvec) {
      red_add4(cy.row + b * 4, cy.acc[q]);
      if (REPL) red_add4(items.delta_rows + (cy.row - items.rows[0]) + b * 4, cy.acc[q]);
    }
  }
  cy.item = -1;
}

template <int LANES, int VPL, int UNROLL, bool REPL = false>
__global__ void __launch_bounds__(MF_THREADS, (VPL == 1 ? (UNROLL == 1 ? 4 : UNROLL == 2 ? 3 : 2) : 1))
k_mf_sgd_combine(const __grid_constant__ FpsTableView items, const __grid_constant__ FpsTableView users,
                 const int* __restrict__ suid, const int* __restrict__ siid, const float* __restrict__ srating,
                 long long n, const FpsMfParams p, int chunk, long long chunk_mul, FpsFlow* flow) {
  constexpr int GPB = MF_THREADS / LANES;
  const int grp = threadIdx.x / LANES, lane = threadIdx.x % LANES;
  const int nvec = items.stride >> 2;
  const int slice = chunk / GPB;                  // records per sub-warp per chunk (<= MF_SLICE)
  const int ipc = (slice + UNROLL - 1) / UNROLL;  // iterations per chunk
  const long long n_chunks = (n + chunk - 1) / chunk;
  const long long my_chunks = n_chunks > blockIdx.x ? (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long iters = my_chunks * ipc;
  // Record cursor.  Chunks are visited in a scattered order (multiplicative permutation of the chunk
  // index): in sorted order a hot item's records sit in adjacent chunks, and adjacent chunks running
  // at the same time would all start from the same stale row; scattering spreads them over the
  // launch, which restores the staleness profile of the unsorted kernel.  The permutation needs a
  // 64-bit modulo; the cursor evaluates it once per chunk (the first version recomputed it, and two
  // more 64-bit divisions, for every record: ~250 of the kernel's 340 instructions per record).
  struct Cursor {
    long long ci;      // ordinal of this block's chunk
    int k;             // position inside the sub-warp's slice
    long long base;    // first record of this sub-warp's slice in that chunk, -1 past the end
    long long cbase;   // first record of the chunk
  };
  auto load_chunk = [&](Cursor& cu) {
    if (cu.ci < my_chunks) {
      const long long lin = (long long)blockIdx.x + cu.ci * gridDim.x;
      cu.cbase = ((lin * chunk_mul) % n_chunks) * chunk;
      cu.base = cu.cbase + (long long)grp * slice;
    } else {
      cu.cbase = 0;
      cu.base = -1;
    }
  };
  auto index_at = [&](const Cursor& cu, int q) -> long long {   // record q of the UNROLL records of this iteration
    const int pos = cu.k * UNROLL + q;
    const long long idx = cu.base + pos;
    return (cu.base >= 0 && pos < slice && idx < n) ? idx : -1;
  };
  auto advance = [&](Cursor& cu) {
    if (++cu.k == ipc) {
      cu.k = 0;
      ++cu.ci;
      load_chunk(cu);
    }
  };
  MfCounters c;
  c.clear();
  unsigned int pulls = 0;
  Carry<LANES, VPL> cy;
  cy.item = -1;
  cy.row = nullptr;
  cy.remote = false;
  MfIds cur[UNROLL], nxt[UNROLL];
  Cursor here{0, 0, -1, 0};
  load_chunk(here);
  Cursor ahead = here;
  advance(ahead);
#pragma unroll
  for (int q = 0; q < UNROLL; ++q) cur[q] = mf_fetch_ids(suid, siid, srating, index_at(here, q));
  for (long long it = 0; it < iters; ++it) {
    const int phase = here.k;
    int cnt = 0;
    if (p.pull_limit > 0) {
      cnt = (int)min((long long)chunk, n - here.cbase);
      if (phase == 0) flow_acquire(flow, cnt, p.pull_limit);
    }
#pragma unroll
    for (int q = 0; q < UNROLL; ++q) nxt[q] = mf_fetch_ids(suid, siid, srating, index_at(ahead, q));
    // ---- issue the loads of this iteration: user rows always, item rows only at run starts
    float4 us[UNROLL][VPL], vn[UNROLL][VPL];
    float* urow[UNROLL];
    float* vrow[UNROLL];
    bool ok[UNROLL], fresh[UNROLL], rem[UNROLL];
    int prev_item = cy.item;
#pragma unroll
    for (int q = 0; q < UNROLL; ++q) {
      ok[q] = false;
      fresh[q] = false;
      rem[q] = false;
      urow[q] = vrow[q] = nullptr;
      if (cur[q].active) {
        const long long u_id = cur[q].u, i_id = cur[q].i;
        const RowRef ir = locate(items, i_id), ur = locate(users, u_id);
        if (in_table(items, i_id, ir) && in_table(users, u_id, ur) && users.rows[ur.owner] != nullptr) {
          ok[q] = true;
          urow[q] = users.rows[ur.owner] + ur.slot * users.stride;
          vrow[q] = items.rows[ir.owner] + ir.slot * items.stride;
          rem[q] = ir.owner != p.local_shard;
          fresh[q] = cur[q].i != prev_item;
          prev_item = cur[q].i;
        } else if (lane == 0) {
          ++c.bad;
        }
      }
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int b = lane + v * LANES;
        us[q][v] = (ok[q] && b < nvec) ? ld_row4(urow[q] + b * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        vn[q][v] = (fresh[q] && b < nvec) ? ld_row4(vrow[q] + b * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    // ---- consume them in order
#pragma unroll
    for (int q = 0; q < UNROLL; ++q) {
      if (fresh[q]) {                   // sub-warp-uniform
        carry_flush<LANES, VPL, REPL>(cy, nvec, items);
        cy.item = cur[q].i;
        cy.row = vrow[q];
        cy.remote = rem[q];
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          cy.v[v] = vn[q][v];
          cy.acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (lane == 0) {
          ++pulls;
          if (rem[q]) ++c.remote;
        }
      }
      float dot = 0.f;
#pragma unroll
      for (int v = 0; v < VPL; ++v) dot += dot4(us[q][v], cy.v[v]);
      dot = group_sum<LANES>(dot);
      if (!ok[q]) continue;
      if (lane == 0) ++c.records;
      const float e = cur[q].r - dot;
      if (!isfinite(e)) {
        if (lane == 0) ++c.nan;
        continue;
      }
      const float g = p.lr * e, decay = p.lr * p.lambda;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int b = lane + v * LANES;
        if (b < nvec) {
          const float4 u = us[q][v], w = cy.v[v];
          float4 un;
          un.x = u.x + fmaf(g, w.x, -decay * u.x);
          un.y = u.y + fmaf(g, w.y, -decay * u.y);
          un.z = u.z + fmaf(g, w.z, -decay * u.z);
          un.w = u.w + fmaf(g, w.w, -decay * u.w);
          st_row4(urow[q] + b * 4, un);
          float4 dv;
          dv.x = fmaf(g, u.x, -decay * w.x);
          dv.y = fmaf(g, u.y, -decay * w.y);
          dv.z = fmaf(g, u.z, -decay * w.z);
          dv.w = fmaf(g, u.w, -decay * w.w);
          cy.acc[v].x += dv.x; cy.acc[v].y += dv.y; cy.acc[v].z += dv.z; cy.acc[v].w += dv.w;
          // the next record of the run sees this update: serial SGD inside a run
          cy.v[v].x += dv.x; cy.v[v].y += dv.y; cy.v[v].z += dv.z; cy.v[v].w += dv.w;
        }
      }
      if (lane == 0) c.loss += e * e;
    }
    if (phase == ipc - 1) {
      carry_flush<LANES, VPL, REPL>(cy, nvec, items);      // slice boundary: bounded staleness
      if (p.pull_limit > 0) flow_release(flow, cnt);
    }
#pragma unroll
    for (int q = 0; q < UNROLL; ++q) cur[q] = nxt[q];
    here = ahead;
    advance(ahead);
  }
  carry_flush<LANES, VPL, REPL>(cy, nvec, items);
  // rows actually pulled / pushed (after combining) are reported separately from records
  {
    unsigned int t = pulls;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0 && t) {
      atomicAdd((unsigned long long*)&flow->reserved[0], (unsigned long long)t);
      atomicAdd((unsigned long long*)&flow->pulls, (unsigned long long)t);     // rows, not records
      atomicAdd((unsigned long long*)&flow->pushes, (unsigned long long)t);    // one combined push per pulled run
    }
  }
  c.flush(flow, /*rows_are_records=*/false);
}

int ensure_scratch(size_t need, void** out) {
  int dev = 0;
  FPS_CUDA(cudaGetDevice(&dev));
  FPS_REQUIRE(dev >= 0 && dev < 64, FPS_E_UNSUPPORTED, "device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  SortScratch& s = g_scratch[dev];
  if (s.bytes < need) {
    if (s.base) FPS_CUDA(cudaFree(s.base));
    s.base = nullptr;
    s.bytes = 0;
    const size_t want = need + need / 4;
    FPS_CUDA(cudaMalloc(&s.base, want));
    s.bytes = want;
  }
  *out = s.base;
  return 0;
}

}  // namespace

int fps_mf_sgd_combine(const FpsTableView* items, const FpsTableView* users, const int32_t* d_uid,
                       const int32_t* d_iid, const float* d_rating, int64_t n, const FpsMfParams* params,
                       FpsFlow* flow, cudaStream_t st) {
  const int gpb_min = MF_THREADS / 32;
  if (items->lazy || users->lazy || items->stride > 128 * 4 || items->stride % 4 != 0 ||
      (params->pull_limit > 0 && params->pull_limit < MF_THREADS)) {
    (void)gpb_min;
    // the carried-row scheme is implemented for eager tables up to rank 512; otherwise direct
    FpsMfParams q = *params;
    q.variant = 0;
    return fps_mf_sgd_step(items, users, d_uid, d_iid, d_rating, n, &q, flow, (void*)st);
  }
  // Round-1 path, kept for A/B runs: its sort scratch is one grow-only buffer per device owned by the
  // library (cudaMalloc on first use / growth), so only ONE stream per device may use this variant at a
  // time; the tile path (mf_tiles.cu) takes caller-owned scratch instead.
  FPS_REQUIRE(n <= 0x7fffffffll, FPS_E_TOO_LARGE, "fps_mf_sgd_combine: at most 2^31-1 ratings per step");
  auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t nb = align((size_t)n * 4);
  size_t cub_bytes = 0;
  int end_bit = 1;
  {
    long long max_id = (long long)items->rows_per_shard * items->n_shards;
    while (end_bit < 31 && (1ll << end_bit) < max_id) ++end_bit;
  }
  FPS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const int*)nullptr, (int*)nullptr, (const int*)nullptr,
                                           (int*)nullptr, (int)n, 0, end_bit, st));
  void* base = nullptr;
  if (int e = ensure_scratch(5 * nb + align(cub_bytes), &base)) return e;
  uint8_t* b = (uint8_t*)base;
  int* iota = (int*)b;
  int* keys = (int*)(b + nb);
  int* perm = (int*)(b + 2 * nb);
  int* su = (int*)(b + 3 * nb);
  float* sr = (float*)(b + 4 * nb);
  void* cub_tmp = b + 5 * nb;
  const int sms = fps_num_sms();
  k_iota<<<sms * 4, 256, 0, st>>>(iota, n);
  FPS_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, (const int*)d_iid, keys, (const int*)iota, perm, (int)n, 0,
                                           end_bit, st));
  k_gather_payload<<<sms * 8, 256, 0, st>>>(perm, d_uid, d_rating, su, sr, n);
  FPS_LAUNCH_CHECK();

  const int nvec = items->stride / 4;
  int lanes = 1;
  while (lanes < nvec && lanes < 32) lanes <<= 1;
  const int vpl = (nvec + lanes - 1) / lanes;
  const int gpb = MF_THREADS / lanes;
  int slice = MF_SLICE;
  if (params->pull_limit > 0 && params->pull_limit < gpb * slice) slice = params->pull_limit / gpb > 0 ? params->pull_limit / gpb : 1;
  const int chunk = gpb * slice;
  long long blocks = (n + chunk - 1) / chunk;
  const int bps = params->blocks_per_sm > 0 ? params->blocks_per_sm : 64;
  if (blocks > (long long)sms * bps) blocks = (long long)sms * bps;
  if (params->pull_limit > 0) {
    const long long useful = 2 * ((long long)params->pull_limit / chunk > 0 ? (long long)params->pull_limit / chunk : 1);
    if (blocks > useful) blocks = useful;
  }
  const long long n_chunks = (n + chunk - 1) / chunk;
  long long chunk_mul = (long long)((double)n_chunks * 0.6180339887) | 1;
  auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
  while (chunk_mul > 1 && gcd(chunk_mul, n_chunks) != 1) chunk_mul -= 2;
  if (chunk_mul < 1) chunk_mul = 1;
  // records in flight per sub-warp: 1 is best when the rows are in local HBM (bandwidth-bound), more
  // when they come over NVLink (latency-bound: ~3x the load latency at the same occupancy)
  const int unroll = params->unroll;
  const bool repl = items->n_replicas > 1;
#define LAUNCH_U(L, V, U) k_mf_sgd_combine<L, V, U><<<(int)blocks, MF_THREADS, 0, st>>>(*items, *users, su, keys, sr, n, *params, chunk, chunk_mul, flow)
#define LAUNCH(L, V)                                   \
  do {                                                 \
    if (repl) k_mf_sgd_combine<L, V, 1, true><<<(int)blocks, MF_THREADS, 0, st>>>(*items, *users, su, keys, sr, n, *params, chunk, chunk_mul, flow); \
    else if (V == 1 && unroll == 2) LAUNCH_U(L, V, 2); \
    else if (V == 1 && unroll == 4) LAUNCH_U(L, V, 4); \
    else LAUNCH_U(L, V, 1);                            \
  } while (0)
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
#undef LAUNCH_U
  FPS_LAUNCH_CHECK();
  return 0;
}
