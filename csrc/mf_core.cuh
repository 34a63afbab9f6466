// Per-rating building blocks shared by the MF kernels.
#pragma once
#include "common.cuh"

#define MF_THREADS 256
#define MF_CHUNK 256          // ratings a CTA handles per credit acquisition
#define MF_BLOCKS_PER_SM 6

struct MfCounters {
  float loss;
  unsigned int records, remote, nan, bad;
  __device__ __forceinline__ void clear() { loss = 0.f; records = remote = nan = bad = 0; }
  // block-reduce, then one set of atomics per CTA
  // rows_are_records: the direct kernels pull and push one row per record; the combining kernels count
  // their (fewer) rows themselves
  __device__ __forceinline__ void flush(FpsFlow* flow, bool rows_are_records = true) {
    __shared__ float s_loss[MF_THREADS / 32];
    __shared__ unsigned int s_cnt[MF_THREADS / 32][4];
    float l = loss;
    unsigned int a = records, b = remote, c = nan, d = bad;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l += __shfl_xor_sync(0xffffffffu, l, o);
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
      c += __shfl_xor_sync(0xffffffffu, c, o);
      d += __shfl_xor_sync(0xffffffffu, d, o);
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
      s_loss[w] = l; s_cnt[w][0] = a; s_cnt[w][1] = b; s_cnt[w][2] = c; s_cnt[w][3] = d;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double L = 0.0;
      unsigned long long A = 0, B = 0, C = 0, D = 0;
      for (int i = 0; i < MF_THREADS / 32; ++i) {
        L += s_loss[i]; A += s_cnt[i][0]; B += s_cnt[i][1]; C += s_cnt[i][2]; D += s_cnt[i][3];
      }
      if (A) {
        atomicAdd(&flow->loss_sum, L);
        atomicAdd((unsigned long long*)&flow->records, A);
        if (rows_are_records) {
          atomicAdd((unsigned long long*)&flow->pulls, A);
          atomicAdd((unsigned long long*)&flow->pushes, A - C);
        }
      }
      if (B) {
        atomicAdd((unsigned long long*)&flow->remote_pulls, B);
        atomicAdd((unsigned long long*)&flow->remote_pushes, B);
      }
      if (C) atomicAdd((unsigned long long*)&flow->nan_count, C);
      if (D) atomicAdd((unsigned long long*)&flow->bad_ids, D);
    }
  }
};

template <int VPL>
struct MfRec {
  float4 v[VPL];    // item row value (pulled)
  float4 u[VPL];    // user row value
  float4 us[VPL];   // user row as stored (== u unless the row is in lazy "pulled" state)
  float* vrow;
  float* urow;
  float r;
  bool valid;
};

// The (user, item, rating) triple of one record; fetched one iteration ahead of its row loads so
// the index-load latency is off the critical path (ncu: 14 % of stall samples sat on it).
struct MfIds {
  int u, i;
  float r;
  bool active;
};

__device__ __forceinline__ MfIds mf_fetch_ids(const int* __restrict__ uid, const int* __restrict__ iid,
                                              const float* __restrict__ rating, long long idx) {
  MfIds d;
  d.active = idx >= 0;
  d.u = d.i = 0;
  d.r = 0.f;
  if (d.active) {
    d.u = uid[idx];
    d.i = iid[idx];
    d.r = rating ? rating[idx] : 0.f;
    d.active = d.u != FPS_SKIP_ID;       // an unused negative-sample slot
  }
  return d;
}

// Pull: issue the row loads of one rating.  `touch` = register the access in the lazy life-cycle.
template <int LANES, int VPL, bool LAZY>
__device__ __forceinline__ void mf_load_ids(MfRec<VPL>& rec, const FpsTableView& items, const FpsTableView& users,
                                            const MfIds& ids, int local_shard, MfCounters& c, bool touch = true) {
  const int lane = threadIdx.x % LANES;
  const int nvec = items.stride >> 2;
  rec.valid = false;
#pragma unroll
  for (int q = 0; q < VPL; ++q) rec.v[q] = rec.u[q] = rec.us[q] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (!ids.active) return;
  const long long u_id = ids.u, i_id = ids.i;
  rec.r = ids.r;
  const RowRef ir = locate(items, i_id), ur = locate(users, u_id);
  if (!in_table(items, i_id, ir) || !in_table(users, u_id, ur) || users.rows[ur.owner] == nullptr) {
    if (lane == 0) ++c.bad;
    return;
  }
  rec.valid = true;
  rec.vrow = items.rows[ir.owner] + ir.slot * items.stride;
  rec.urow = users.rows[ur.owner] + ur.slot * users.stride;
#pragma unroll
  for (int q = 0; q < VPL; ++q) {
    const int b = lane + q * LANES;
    if (b < nvec) {
      rec.v[q] = ld_row4(rec.vrow + b * 4);
      rec.us[q] = ld_row4(rec.urow + b * 4);
    } else {
      rec.v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      rec.us[q] = rec.v[q];
    }
    rec.u[q] = rec.us[q];
  }
  if (LAZY) {
    int istate = FPS_ROW_PUSHED, ustate = FPS_ROW_PUSHED;
    if (items.lazy) istate = touch ? touch_state(items.state[ir.owner], ir.slot, FPS_ROW_PULLED)
                                   : (int)items.state[ir.owner][ir.slot];
    if (users.lazy) ustate = touch ? touch_state(users.state[ur.owner], ur.slot, FPS_ROW_PULLED)
                                   : (int)users.state[ur.owner][ur.slot];
    // an untouched row read without touching (predict) still evaluates to init(id)
    if (istate != FPS_ROW_PUSHED && items.lazy) {
#pragma unroll
      for (int q = 0; q < VPL; ++q) {
        const int b = lane + q * LANES;
        if (b < nvec) {
          const float4 i4 = init_block(items, i_id, b);
          rec.v[q].x += i4.x; rec.v[q].y += i4.y; rec.v[q].z += i4.z; rec.v[q].w += i4.w;
        }
      }
    }
    if (ustate != FPS_ROW_PUSHED && users.lazy) {
#pragma unroll
      for (int q = 0; q < VPL; ++q) {
        const int b = lane + q * LANES;
        if (b < nvec) {
          const float4 i4 = init_block(users, u_id, b);
          rec.u[q].x += i4.x; rec.u[q].y += i4.y; rec.u[q].z += i4.z; rec.u[q].w += i4.w;
        }
      }
    }
  }
  if (lane == 0) {
    ++c.records;
    if (ir.owner != local_shard) ++c.remote;
  }
}

template <int LANES, int VPL, bool LAZY>
__device__ __forceinline__ void mf_load(MfRec<VPL>& rec, const FpsTableView& items, const FpsTableView& users,
                                        const int* __restrict__ uid, const int* __restrict__ iid,
                                        const float* __restrict__ rating, long long idx, bool active,
                                        int local_shard, MfCounters& c, bool touch = true) {
  const MfIds ids = mf_fetch_ids(uid, iid, rating, active ? idx : -1);
  mf_load_ids<LANES, VPL, LAZY>(rec, items, users, ids, local_shard, c, touch);
}

// Compute + push: e = r - u.v; user row updated in place (worker-resident half), item delta
// reduced into the owner's row.
// REPL: `items` is a replicated table with deferred pushes - the delta also goes to the local delta
// array, which fps_table_flush_delta forwards to the other replicas after the launch.
template <int LANES, int VPL, bool LAZY, bool REPL = false>
__device__ __forceinline__ void mf_update(MfRec<VPL>& rec, const FpsTableView& items, const FpsMfParams& p,
                                          MfCounters& c) {
  const int lane = threadIdx.x % LANES;
  const int nvec = items.stride >> 2;
  float dot = 0.f;
#pragma unroll
  for (int q = 0; q < VPL; ++q) dot += dot4(rec.u[q], rec.v[q]);
  dot = group_sum<LANES>(dot);          // every lane of the (sub-)warp takes part
  if (!rec.valid) return;
  const float e = rec.r - dot;
  if (!isfinite(e)) {                   // device analogue of FactorIsNotANumberException
    if (lane == 0) ++c.nan;
    return;
  }
  const float g = p.lr * e, decay = p.lr * p.lambda;
#pragma unroll
  for (int q = 0; q < VPL; ++q) {
    const int b = lane + q * LANES;
    if (b < nvec) {
      const float4 u = rec.u[q], v = rec.v[q];
      float4 un, dv;
      un.x = rec.us[q].x + fmaf(g, v.x, -decay * u.x);
      un.y = rec.us[q].y + fmaf(g, v.y, -decay * u.y);
      un.z = rec.us[q].z + fmaf(g, v.z, -decay * u.z);
      un.w = rec.us[q].w + fmaf(g, v.w, -decay * u.w);
      dv.x = fmaf(g, u.x, -decay * v.x);
      dv.y = fmaf(g, u.y, -decay * v.y);
      dv.z = fmaf(g, u.z, -decay * v.z);
      dv.w = fmaf(g, u.w, -decay * v.w);
      st_row4(rec.urow + b * 4, un);
      red_add4(rec.vrow + b * 4, dv);
      if (REPL) red_add4(items.delta_rows + (rec.vrow - items.rows[0]) + b * 4, dv);
    }
  }
  if (lane == 0) c.loss += e * e;
}
