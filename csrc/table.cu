// Sharded parameter table: eager init (N9), gather = pull without compute, scatter = push /
// model load (N10), compaction dump (N11).  See include/fps_b200.h for the contracts.
#include "common.cuh"

namespace {

__device__ __forceinline__ long long slot_to_id(const FpsTableView& t, int shard, long long slot) {
  return t.partition == FPS_PART_HASH ? slot * t.n_shards + shard : (long long)shard * t.div + slot;
}

__global__ void __launch_bounds__(256)
k_init_shard(const __grid_constant__ FpsTableView t, int shard, long long n_rows) {
  const int vec_per_row = t.stride / 4;
  const long long total = n_rows * vec_per_row;
  float* base = t.rows[shard];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long slot = i / vec_per_row;
    const int blk = (int)(i - slot * vec_per_row);
    const float4 v = init_block(t, slot_to_id(t, shard, slot), blk);
    *reinterpret_cast<float4*>(base + slot * t.stride + blk * 4) = v;
  }
}

// value of float4 `blk` of a row, honouring the lazy life-cycle
__device__ __forceinline__ float4 row_value4(const FpsTableView& t, const float* row, long long id,
                                             int blk, int state) {
  float4 v = ld_row4(row + blk * 4);
  if (state == FPS_ROW_PULLED) {
    const float4 b = init_block(t, id, blk);
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
  }
  return v;
}

template <int LANES>
__global__ void __launch_bounds__(256)
k_gather(const __grid_constant__ FpsTableView t, const long long* __restrict__ ids, long long n,
         float* __restrict__ out, int out_stride, FpsFlow* flow, int local_shard) {
  const int lane = threadIdx.x % LANES;
  const long long groups = (long long)gridDim.x * (blockDim.x / LANES);
  const int nvec = t.stride / 4;
  unsigned long long bad = 0, remote = 0, done = 0;
  for (long long g = (long long)blockIdx.x * (blockDim.x / LANES) + threadIdx.x / LANES; g < n; g += groups) {
    const long long id = ids[g];
    const RowRef r = locate(t, id);
    if (!in_table(t, id, r)) {
      if (lane == 0) ++bad;
      continue;
    }
    int state = FPS_ROW_PUSHED;
    if (t.lazy) state = touch_state(t.state[r.owner], r.slot, FPS_ROW_PULLED);
    const float* row = t.rows[r.owner] + r.slot * t.stride;
    for (int b = lane; b < nvec; b += LANES) {
      const float4 v = row_value4(t, row, id, b, state);
      float* o = out + g * out_stride + b * 4;
      const int j = b * 4;
      if (j + 0 < out_stride) o[0] = v.x;
      if (j + 1 < out_stride) o[1] = v.y;
      if (j + 2 < out_stride) o[2] = v.z;
      if (j + 3 < out_stride) o[3] = v.w;
    }
    if (lane == 0) {
      ++done;
      if (r.owner != local_shard) ++remote;
    }
  }
  if (flow != nullptr && lane == 0) {
    if (bad) atomicAdd((unsigned long long*)&flow->bad_ids, bad);
    if (done) atomicAdd((unsigned long long*)&flow->pulls, done);
    if (remote) atomicAdd((unsigned long long*)&flow->remote_pulls, remote);
  }
}

template <int LANES>
__global__ void __launch_bounds__(256)
k_scatter(const __grid_constant__ FpsTableView t, const long long* __restrict__ ids, long long n,
          const float* __restrict__ vals, int val_stride, int mode, FpsFlow* flow, int local_shard) {
  const int lane = threadIdx.x % LANES;
  const long long groups = (long long)gridDim.x * (blockDim.x / LANES);
  const int nvec = t.stride / 4;
  unsigned long long bad = 0, remote = 0, done = 0;
  // warp-uniform trip count: mode 2 shuffles inside the loop
  for (long long g0 = (long long)blockIdx.x * (blockDim.x / LANES); g0 < n; g0 += groups) {
    const long long g = g0 + threadIdx.x / LANES;
    const long long id = g < n ? ids[g] : -1;
    const RowRef r = locate(t, id < 0 ? 0 : id);
    const bool ok = g < n && in_table(t, id, r);
    if (g < n && !ok && lane == 0) ++bad;
    int state = FPS_ROW_PUSHED;
    if (ok && t.lazy) state = touch_state(t.state[r.owner], r.slot, FPS_ROW_PUSHED);
    float* row = ok ? t.rows[r.owner] + r.slot * t.stride : nullptr;
    const float* src = vals + (ok ? g : 0) * val_stride;
    float nrm = 0.f;
    for (int b = lane; ok && b < nvec; b += LANES) {
      const int j = b * 4;
      float4 d;
      d.x = (j + 0 < val_stride && j + 0 < t.dim) ? src[j + 0] : 0.f;
      d.y = (j + 1 < val_stride && j + 1 < t.dim) ? src[j + 1] : 0.f;
      d.z = (j + 2 < val_stride && j + 2 < t.dim) ? src[j + 2] : 0.f;
      d.w = (j + 3 < val_stride && j + 3 < t.dim) ? src[j + 3] : 0.f;
      if (mode == 1) {
        // overwrite: stored := value - base so that base + stored == value
        if (state == FPS_ROW_PULLED) {
          const float4 bse = init_block(t, id, b);
          d.x -= bse.x; d.y -= bse.y; d.z -= bse.z; d.w -= bse.w;
        }
        if (t.stride > t.dim) {   // keep aux columns (e.g. the length column) untouched here
          const float4 old = ld_row4(row + j);
          if (j + 0 >= t.dim) d.x = old.x;
          if (j + 1 >= t.dim) d.y = old.y;
          if (j + 2 >= t.dim) d.z = old.z;
          if (j + 3 >= t.dim) d.w = old.w;
        }
        if (t.n_replicas > 1) {   // model load into a replicated table: every replica gets the value
          for (int rr = 0; rr < t.n_replicas; ++rr) st_row4(t.replica_rows[rr] + r.slot * t.stride + j, d);
        } else {
          st_row4(row + j, d);
        }
      } else {
        push_add4(t, r.owner, r.slot * t.stride + j, d);
      }
    }
    if (mode == 2 && t.stride > t.dim) {
      // add-then-renormalise: the aux column `dim` carries ||row|| (attachLength semantics,
      // PSOnlineMatrixFactorizationAndTopKGenerator.scala:81-84).  Exact when pushes to one row
      // do not overlap; otherwise it reflects one of the interleavings, like any stale read here.
      __threadfence_system();
      for (int b = lane; ok && b < nvec; b += LANES) {
        const float4 v = row_value4(t, row, id, b, state);
        const int j = b * 4;
        if (j + 0 < t.dim) nrm = fmaf(v.x, v.x, nrm);
        if (j + 1 < t.dim) nrm = fmaf(v.y, v.y, nrm);
        if (j + 2 < t.dim) nrm = fmaf(v.z, v.z, nrm);
        if (j + 3 < t.dim) nrm = fmaf(v.w, v.w, nrm);
      }
      nrm = group_sum<LANES>(nrm);
      if (ok && lane == 0) {
        const float len = sqrtf(nrm);   // aux columns are never part of init()
        asm volatile("st.global.cg.f32 [%0], %1;" ::"l"(row + t.dim), "f"(len) : "memory");
      }
    }
    if (ok && lane == 0) {
      ++done;
      if (r.owner != local_shard) ++remote;
    }
  }
  if (flow != nullptr && lane == 0) {
    if (bad) atomicAdd((unsigned long long*)&flow->bad_ids, bad);
    if (done) atomicAdd((unsigned long long*)&flow->pushes, done);
    if (remote) atomicAdd((unsigned long long*)&flow->remote_pushes, remote);
  }
}

__global__ void __launch_bounds__(256)
k_dump_shard(const __grid_constant__ FpsTableView t, int shard, long long row_begin, long long n_rows, long long n_ids_total,
             long long* __restrict__ out_ids, float* __restrict__ out_rows, long long capacity,
             unsigned long long* count) {
  // one warp per row: lane 0 claims an output slot (warp-aggregated per block would save little:
  // the dump runs once per job)
  const int lane = threadIdx.x & 31;
  const long long warps = (long long)gridDim.x * (blockDim.x / 32);
  const int nvec = t.stride / 4;
  const float* base = t.rows[shard];
  for (long long slot = row_begin + (long long)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32; slot < n_rows; slot += warps) {
    const long long id = slot_to_id(t, shard, slot);
    if (id >= n_ids_total) continue;
    int state = FPS_ROW_PUSHED;
    if (t.lazy) {
      state = t.state[shard][slot];
      if (state == FPS_ROW_UNTOUCHED) continue;
    }
    unsigned long long idx = 0;
    if (lane == 0) idx = atomicAdd(count, 1ull);
    idx = __shfl_sync(0xffffffffu, idx, 0);
    if ((long long)idx >= capacity) continue;
    if (lane == 0) out_ids[idx] = id;
    const float* row = base + slot * t.stride;
    for (int b = lane; b < nvec; b += 32) {
      const float4 v = row_value4(t, row, id, b, state);
      *reinterpret_cast<float4*>(out_rows + idx * t.stride + b * 4) = v;
    }
  }
}

// ---- scalar tables (dim == stride == 1: binary PA weights) --------------------------------------
__device__ __forceinline__ float scalar_init_value(const FpsTableView& t, long long id) {
  if (t.init_lo == 0.f && t.init_hi == 0.f) return 0.f;
  return init_block(t, id, 0).x;
}

__global__ void __launch_bounds__(256)
k_init_shard_scalar(const __grid_constant__ FpsTableView t, int shard, long long n_rows) {
  float* base = t.rows[shard];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows;
       i += (long long)gridDim.x * blockDim.x)
    base[i] = scalar_init_value(t, slot_to_id(t, shard, i));
}

__global__ void __launch_bounds__(256)
k_gather_scalar(const __grid_constant__ FpsTableView t, const long long* __restrict__ ids, long long n,
                float* __restrict__ out, int out_stride, FpsFlow* flow) {
  unsigned long long bad = 0, done = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n; g += (long long)gridDim.x * blockDim.x) {
    const long long id = ids[g];
    const RowRef r = locate(t, id);
    if (!in_table(t, id, r)) { ++bad; continue; }
    float v = ld_row1(t.rows[r.owner] + r.slot);
    if (t.lazy && touch_state(t.state[r.owner], r.slot, FPS_ROW_PULLED) == FPS_ROW_PULLED)
      v += scalar_init_value(t, id);
    out[g * out_stride] = v;
    ++done;
  }
  if (flow != nullptr) {
    if (bad) atomicAdd((unsigned long long*)&flow->bad_ids, bad);
    if (done) atomicAdd((unsigned long long*)&flow->pulls, done);
  }
}

__global__ void __launch_bounds__(256)
k_scatter_scalar(const __grid_constant__ FpsTableView t, const long long* __restrict__ ids, long long n,
                 const float* __restrict__ vals, int val_stride, int mode, FpsFlow* flow) {
  unsigned long long bad = 0, done = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n; g += (long long)gridDim.x * blockDim.x) {
    const long long id = ids[g];
    const RowRef r = locate(t, id);
    if (!in_table(t, id, r)) { ++bad; continue; }
    int state = FPS_ROW_PUSHED;
    if (t.lazy) state = touch_state(t.state[r.owner], r.slot, FPS_ROW_PUSHED);
    float* row = t.rows[r.owner] + r.slot;
    float d = vals[g * val_stride];
    if (mode == 1) {
      if (state == FPS_ROW_PULLED) d -= scalar_init_value(t, id);
      if (t.n_replicas > 1) {
        for (int rr = 0; rr < t.n_replicas; ++rr)
          asm volatile("st.global.cg.f32 [%0], %1;" ::"l"(t.replica_rows[rr] + r.slot), "f"(d) : "memory");
      } else {
        asm volatile("st.global.cg.f32 [%0], %1;" ::"l"(row), "f"(d) : "memory");
      }
    } else {
      push_add1(t, r.owner, r.slot, d);
    }
    ++done;
  }
  if (flow != nullptr) {
    if (bad) atomicAdd((unsigned long long*)&flow->bad_ids, bad);
    if (done) atomicAdd((unsigned long long*)&flow->pushes, done);
  }
}

__global__ void __launch_bounds__(256)
k_dump_shard_scalar(const __grid_constant__ FpsTableView t, int shard, long long row_begin, long long n_rows, long long n_ids_total,
                    long long* __restrict__ out_ids, float* __restrict__ out_rows, long long capacity,
                    unsigned long long* count) {
  const float* base = t.rows[shard];
  for (long long slot = row_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; slot < n_rows;
       slot += (long long)gridDim.x * blockDim.x) {
    const long long id = slot_to_id(t, shard, slot);
    if (id >= n_ids_total) continue;
    int state = FPS_ROW_PUSHED;
    if (t.lazy) {
      state = t.state[shard][slot];
      if (state == FPS_ROW_UNTOUCHED) continue;
    }
    const unsigned long long idx = atomicAdd(count, 1ull);
    if ((long long)idx >= capacity) continue;
    out_ids[idx] = id;
    out_rows[idx] = ld_row1(base + slot) + (state == FPS_ROW_PULLED ? scalar_init_value(t, id) : 0.f);
  }
}

int lanes_for(int stride) {
  const int nvec = stride / 4;
  int l = 1;
  while (l < nvec && l < 32) l <<= 1;
  return l;
}

int check_view(const FpsTableView* v, const char* who) {
  FPS_REQUIRE(v != nullptr, FPS_E_INVALID, "%s: null table view", who);
  FPS_REQUIRE(v->n_shards >= 1 && v->n_shards <= FPS_MAX_SHARDS, FPS_E_INVALID, "%s: n_shards %d out of range", who, v->n_shards);
  FPS_REQUIRE((v->stride % 4 == 0 || (v->stride == 1 && v->dim == 1)) && v->stride >= v->dim && v->dim >= 1,
              FPS_E_INVALID, "%s: stride %d must be a multiple of 4 (or 1 for scalar tables) and >= dim %d",
              who, v->stride, v->dim);
  FPS_REQUIRE(v->partition == FPS_PART_HASH || v->div > 0, FPS_E_INVALID, "%s: range partition needs div > 0", who);
  return 0;
}

// ---- replicated tables with deferred pushes: exchange the accumulated deltas ----------------------
// Runs after the step kernels on their stream.  Every 16-byte chunk of the local delta array that is
// not all zero is added to the same place of every OTHER replica (the local one already has it) and
// cleared.  A weight pushed thousands of times during the step crosses NVLink once per peer.
template <bool VEC>
__global__ void __launch_bounds__(256)
k_flush_delta(const __grid_constant__ FpsTableView t, long long n_floats) {
  const long long n_items = VEC ? n_floats / 4 : n_floats;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_items; i += (long long)gridDim.x * blockDim.x) {
    if (VEC) {
      float4* dp = reinterpret_cast<float4*>(t.delta_rows) + i;
      const float4 d = *dp;
      if (d.x == 0.f && d.y == 0.f && d.z == 0.f && d.w == 0.f) continue;
      *dp = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < t.n_replicas; ++r)
        if (r != t.replica_rank) red_add4(t.replica_rows[r] + i * 4, d);
    } else {
      const float d = t.delta_rows[i];
      if (d == 0.f) continue;
      t.delta_rows[i] = 0.f;
      for (int r = 0; r < t.n_replicas; ++r)
        if (r != t.replica_rank) red_add1(t.replica_rows[r] + i, d);
    }
  }
}

// ---- N13: per-row lock word, system-scope CAS (LockPSLogicA.scala:13-46 on the device) ---------------------
// x <- alpha * x + delta as a LOCKED read-modify-write: the reference's lock logic answers a pull only when
// the push of the pull before it has been applied, which makes a worker's pull -> compute -> push an
// exclusive cycle per key.  Here the cycle is one warp's: lane 0 takes the row's lock word (an int32 living
// in an auxiliary column of the row itself, so it is in the owner's memory and reachable over NVLink like
// the row), the warp reads the row, applies the update, writes it back, and lane 0 releases the lock.  A warp
// holds at most one lock, so there is no cycle of waiters; waiters back off with __nanosleep.
// `use_lock == 0` runs the same read-modify-write unprotected (what an overwriting param_update does without
// the lock logic: updates get lost) - kept for the test that shows the difference.
__global__ void __launch_bounds__(256)
k_table_locked_axpby(const __grid_constant__ FpsTableView t, const long long* __restrict__ ids, long long n,
                     float alpha, const float* __restrict__ deltas, int delta_stride, int lock_col, int use_lock,
                     FpsFlow* flow) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  unsigned long long spins = 0, done = 0, bad = 0;
  for (long long k = warp0; k < n; k += n_warps) {
    const long long id = ids[k];
    const RowRef r = locate(t, id);
    if (!in_table(t, id, r)) { if (lane == 0) ++bad; continue; }
    float* row = t.rows[r.owner] + r.slot * t.stride;
    int* lock = reinterpret_cast<int*>(row + lock_col);
    if (use_lock) {
      if (lane == 0) {
        unsigned ns = 32;
        while (atomicCAS_system(lock, 0, 1) != 0) {
          __nanosleep(ns);
          if (ns < 2048) ns <<= 1;
          ++spins;
        }
        __threadfence_system();                 // acquire: the row reads below come after the lock
      }
      __syncwarp();
    }
    if (t.lazy && lane == 0) touch_state(t.state[r.owner], r.slot, FPS_ROW_PUSHED);
    for (int c = lane; c < t.dim; c += 32) {
      float x;
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(x) : "l"(row + c) : "memory");
      x = fmaf(alpha, x, deltas[k * delta_stride + c]);
      asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(row + c), "f"(x) : "memory");
    }
    if (use_lock) {
      __threadfence_system();                   // release: the row is written before the lock opens
      __syncwarp();
      if (lane == 0) atomicExch_system(lock, 0);
    }
    if (lane == 0) ++done;
  }
  if (flow != nullptr && lane == 0) {
    if (done) { atomicAdd((unsigned long long*)&flow->pushes, done); atomicAdd((unsigned long long*)&flow->pulls, done); }
    if (bad) atomicAdd((unsigned long long*)&flow->bad_ids, bad);
    if (spins) atomicAdd((unsigned long long*)&flow->limiter_stalls, spins);
  }
}

}  // namespace

int fps_check_view(const FpsTableView* v, const char* who) { return check_view(v, who); }

extern "C" {

int fps_table_flush_delta(const FpsTableView* view, void* stream) {
  if (int e = check_view(view, "fps_table_flush_delta")) return e;
  if (view->n_replicas <= 1 || view->delta_rows == nullptr) return 0;
  FPS_REQUIRE(view->replica_rank >= 0 && view->replica_rank < view->n_replicas, FPS_E_INVALID,
              "fps_table_flush_delta: replica_rank %d outside [0, %d)", view->replica_rank, view->n_replicas);
  const long long n_floats = (long long)view->rows_per_shard * view->stride;
  const bool vec = (n_floats % 4 == 0) && (((uintptr_t)view->delta_rows) % 16 == 0);
  long long blocks = ((vec ? n_floats / 4 : n_floats) + 255) / 256;
  const long long cap = (long long)fps_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (vec) k_flush_delta<true><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(*view, n_floats);
  else k_flush_delta<false><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(*view, n_floats);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_table_init_shard(const FpsTableView* view, int shard, int64_t n_rows, void* stream) {
  if (int e = check_view(view, "fps_table_init_shard")) return e;
  FPS_REQUIRE(shard >= 0 && shard < view->n_shards && view->rows[shard] != nullptr, FPS_E_INVALID,
              "fps_table_init_shard: shard %d not mapped", shard);
  if (n_rows <= 0) return 0;
  if (view->stride == 1) {
    long long b = (n_rows + 255) / 256;
    if (b > (long long)fps_num_sms() * 16) b = (long long)fps_num_sms() * 16;
    k_init_shard_scalar<<<(int)b, 256, 0, (cudaStream_t)stream>>>(*view, shard, n_rows);
    FPS_LAUNCH_CHECK();
    return 0;
  }
  const long long total = (long long)n_rows * (view->stride / 4);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)fps_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  k_init_shard<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(*view, shard, n_rows);
  FPS_LAUNCH_CHECK();
  return 0;
}

#define FPS_DISPATCH_LANES(lanes, KERNEL, ...)                        \
  switch (lanes) {                                                    \
    case 1: KERNEL<1> __VA_ARGS__; break;                             \
    case 2: KERNEL<2> __VA_ARGS__; break;                             \
    case 4: KERNEL<4> __VA_ARGS__; break;                             \
    case 8: KERNEL<8> __VA_ARGS__; break;                             \
    case 16: KERNEL<16> __VA_ARGS__; break;                           \
    default: KERNEL<32> __VA_ARGS__; break;                           \
  }

int fps_table_gather(const FpsTableView* view, const int64_t* d_ids, int64_t n, float* d_out,
                     int out_stride, FpsFlow* d_flow, void* stream) {
  if (int e = check_view(view, "fps_table_gather")) return e;
  if (n <= 0) return 0;
  if (view->stride == 1) {
    long long b = (n + 255) / 256;
    if (b > (long long)fps_num_sms() * 8) b = (long long)fps_num_sms() * 8;
    k_gather_scalar<<<(int)b, 256, 0, (cudaStream_t)stream>>>(*view, (const long long*)d_ids, n, d_out, out_stride, d_flow);
    FPS_LAUNCH_CHECK();
    return 0;
  }
  const int lanes = lanes_for(view->stride);
  const int gpb = 256 / lanes;
  long long blocks = (n + gpb - 1) / gpb;
  const long long cap = (long long)fps_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = (cudaStream_t)stream;
  FPS_DISPATCH_LANES(lanes, k_gather, <<<(int)blocks, 256, 0, st>>>(*view, (const long long*)d_ids, n, d_out, out_stride, d_flow, -1));
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_table_scatter(const FpsTableView* view, const int64_t* d_ids, int64_t n, const float* d_vals,
                      int val_stride, int mode, FpsFlow* d_flow, void* stream) {
  if (int e = check_view(view, "fps_table_scatter")) return e;
  FPS_REQUIRE(mode >= 0 && mode <= 2, FPS_E_INVALID, "fps_table_scatter: mode %d", mode);
  if (n <= 0) return 0;
  if (view->stride == 1) {
    FPS_REQUIRE(mode != 2, FPS_E_INVALID, "fps_table_scatter: add_norm needs a vector table");
    long long b = (n + 255) / 256;
    if (b > (long long)fps_num_sms() * 8) b = (long long)fps_num_sms() * 8;
    k_scatter_scalar<<<(int)b, 256, 0, (cudaStream_t)stream>>>(*view, (const long long*)d_ids, n, d_vals, val_stride, mode, d_flow);
    FPS_LAUNCH_CHECK();
    return 0;
  }
  const int lanes = lanes_for(view->stride);
  const int gpb = 256 / lanes;
  long long blocks = (n + gpb - 1) / gpb;
  const long long cap = (long long)fps_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = (cudaStream_t)stream;
  FPS_DISPATCH_LANES(lanes, k_scatter, <<<(int)blocks, 256, 0, st>>>(*view, (const long long*)d_ids, n, d_vals, val_stride, mode, d_flow, -1));
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_table_dump_range(const FpsTableView* view, int shard, int64_t row_begin, int64_t row_end, int64_t n_ids_total,
                         int64_t* d_out_ids, float* d_out_rows, int64_t capacity, int64_t* d_count,
                         void* stream) {
  if (int e = check_view(view, "fps_table_dump_range")) return e;
  FPS_REQUIRE(shard >= 0 && shard < view->n_shards && view->rows[shard] != nullptr, FPS_E_INVALID,
              "fps_table_dump_range: shard %d not mapped", shard);
  FPS_REQUIRE(row_begin >= 0 && row_end <= view->rows_per_shard, FPS_E_INVALID, "fps_table_dump_range: bad row range");
  const long long n = row_end - row_begin;
  if (n <= 0) return 0;
  if (view->stride == 1) {
    long long b = (n + 255) / 256;
    if (b > (long long)fps_num_sms() * 8) b = (long long)fps_num_sms() * 8;
    k_dump_shard_scalar<<<(int)b, 256, 0, (cudaStream_t)stream>>>(*view, shard, row_begin, row_end, n_ids_total,
                                                                  (long long*)d_out_ids, d_out_rows, capacity,
                                                                  (unsigned long long*)d_count);
    FPS_LAUNCH_CHECK();
    return 0;
  }
  long long blocks = (n + 7) / 8;
  const long long cap = (long long)fps_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  k_dump_shard<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
      *view, shard, row_begin, row_end, n_ids_total, (long long*)d_out_ids, d_out_rows, capacity,
      (unsigned long long*)d_count);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_table_dump_shard(const FpsTableView* view, int shard, int64_t n_rows, int64_t n_ids_total,
                         int64_t* d_out_ids, float* d_out_rows, int64_t capacity, int64_t* d_count,
                         void* stream) {
  return fps_table_dump_range(view, shard, 0, n_rows, n_ids_total, d_out_ids, d_out_rows, capacity, d_count, stream);
}

int fps_table_locked_axpby(const FpsTableView* view, const int64_t* d_ids, int64_t n, float alpha,
                           const float* d_deltas, int32_t delta_stride, int32_t lock_col, int32_t use_lock,
                           FpsFlow* d_flow, void* stream) {
  if (int e = check_view(view, "fps_table_locked_axpby")) return e;
  FPS_REQUIRE(d_ids && d_deltas && delta_stride >= view->dim, FPS_E_INVALID, "fps_table_locked_axpby: ids, deltas (stride >= dim) required");
  FPS_REQUIRE(view->n_replicas <= 1, FPS_E_UNSUPPORTED, "fps_table_locked_axpby: sharded tables only (a replica has no single owner to lock)");
  FPS_REQUIRE(lock_col >= view->dim && lock_col < view->stride, FPS_E_INVALID,
              "fps_table_locked_axpby: the lock word lives in an auxiliary column: dim <= lock_col < stride (%d, %d, %d)",
              view->dim, lock_col, view->stride);
  FPS_REQUIRE(!view->lazy || (view->init_lo == 0.f && view->init_hi == 0.f), FPS_E_UNSUPPORTED,
              "fps_table_locked_axpby: lazy tables need a zero initialiser");
  if (n <= 0) return 0;
  long long blocks = (n + 7) / 8;
  const long long cap = (long long)fps_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  k_table_locked_axpby<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(*view, (const long long*)d_ids, n, alpha, d_deltas,
                                                                      delta_stride, lock_col, use_lock, d_flow);
  FPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
