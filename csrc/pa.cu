// Passive-aggressive kernels (N3 binary, N4 multiclass OVA / cost-based, N5 predict).
// One CSR mini-batch per launch; per example: gather (pull) every active feature from its
// owner shard, margin / tau, scatter red.add (push) of tau*y*x_i.  The reference does this with
// nnz Pull messages, nnz PullAnswers, a HashMap assembly and nnz Push messages per example
// (PassiveAggressiveParameterServer.scala:289-336).
#include "common.cuh"
#include "flow.cuh"

namespace {

constexpr int PA_THREADS = 256;
// multiclass kernel: one CTA per example; measured on config 4 (no limiter / limit 10 000): 256 threads 0.722 /
// 1.157 ms, 128 threads 0.674 / 0.885 ms, 64 threads 0.728 / 0.879 ms
constexpr int PM_THREADS = 128;

__device__ __forceinline__ float scalar_init(const FpsTableView& t, long long id) {
  if (t.init_lo == 0.f && t.init_hi == 0.f) return 0.f;
  return init_block(t, id, 0).x;
}

// value of a scalar-row (dim == 1) parameter; registers the pull in lazy mode
__device__ __forceinline__ float pull_scalar(const FpsTableView& t, long long id, const RowRef& r) {
  float w = ld_row1(t.rows[r.owner] + r.slot * t.stride);
  if (t.lazy) {
    const int s = touch_state(t.state[r.owner], r.slot, FPS_ROW_PULLED);
    if (s == FPS_ROW_PULLED) w += scalar_init(t, id);
  }
  return w;
}

__device__ __forceinline__ float tau_of(int variant, float C, float loss, float n2) {
  if (n2 <= 0.f) return 0.f;
  if (variant == 0) return loss / n2;
  if (variant == 1) return fminf(C, loss / n2);
  return loss / (n2 + 1.f / (2.f * C));
}

struct PaCounters {
  float loss;
  unsigned int records, pulls, pushes, remote, bad;
};

template <int THREADS = PA_THREADS>
__device__ __forceinline__ void pa_flush(PaCounters c, FpsFlow* flow) {
  __shared__ float s_loss[THREADS / 32];
  __shared__ unsigned int s_cnt[THREADS / 32][5];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c.loss += __shfl_xor_sync(0xffffffffu, c.loss, o);
    c.records += __shfl_xor_sync(0xffffffffu, c.records, o);
    c.pulls += __shfl_xor_sync(0xffffffffu, c.pulls, o);
    c.pushes += __shfl_xor_sync(0xffffffffu, c.pushes, o);
    c.remote += __shfl_xor_sync(0xffffffffu, c.remote, o);
    c.bad += __shfl_xor_sync(0xffffffffu, c.bad, o);
  }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    s_loss[w] = c.loss;
    s_cnt[w][0] = c.records; s_cnt[w][1] = c.pulls; s_cnt[w][2] = c.pushes; s_cnt[w][3] = c.remote; s_cnt[w][4] = c.bad;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double L = 0.0;
    unsigned long long a[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < THREADS / 32; ++i) {
      L += s_loss[i];
      for (int k = 0; k < 5; ++k) a[k] += s_cnt[i][k];
    }
    if (a[0]) {
      atomicAdd(&flow->loss_sum, L);
      atomicAdd((unsigned long long*)&flow->records, a[0]);
    }
    if (a[1]) atomicAdd((unsigned long long*)&flow->pulls, a[1]);
    if (a[2]) atomicAdd((unsigned long long*)&flow->pushes, a[2]);
    if (a[3]) atomicAdd((unsigned long long*)&flow->remote_pulls, a[3]);
    if (a[4]) atomicAdd((unsigned long long*)&flow->bad_ids, a[4]);
  }
}

// ---- N3 / N5 binary: one warp per example -----------------------------------------------------
// measured on config 3 (limit 10 000 / no limit, ms): prefetch depth 4 at 106 registers 0.935 / 0.168, depth 3 at 80
// 0.913 / 0.156, depth 3 at 64 0.920 / 0.153, depth 2 at 64 1.004 / 0.170; the plain loop it replaced 1.43 / 0.160
constexpr int PB_PF = 3;
__global__ void __launch_bounds__(PA_THREADS, 4)
k_pa_binary(const __grid_constant__ FpsTableView t, const long long* __restrict__ indptr,
            const int* __restrict__ indices, const float* __restrict__ values,
            const int* __restrict__ labels, long long n, const FpsPaParams p, int* __restrict__ pred,
            int warps_active, int lanes_active, FpsFlow* flow) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  PaCounters c = {0.f, 0, 0, 0, 0, 0};
  // Limiter (counted in features, like the reference's): an active warp has one weight load per
  // active lane outstanding, so the CTA's allowance is warps_active * lanes_active, taken from the
  // worker's pool once for the CTA's lifetime; the host sizes the grid so that all allowances fit.
  // (The first version took min(limit, nnz of 8 examples) per pass: more loads than credits for long
  // examples, and one CAS round trip per pass: 36x slower with the limiter on.)
  const int WPB = warps_active;
  const int credits = warps_active * lanes_active;
  if (p.pull_limit > 0) flow_acquire(flow, credits, p.pull_limit);
  // Software pipeline over the warp's examples.  What a credit pays for is the weight load; everything that is
  // not a pull - row pointers, feature ids, values, the label - is fetched ahead: the row pointers two examples
  // ahead, ids / values / label one example ahead, so that a warp waits for weights and for nothing else
  // (the first version waited for indptr, then ids, then weights, three times per 75-feature example: a
  // credit was held ~1.45 us per pull, now about the weight latency).  PB_PF chunks of one feature per active
  // lane are prefetched into registers; a longer example takes its tail through a plain loop.
  const bool puller = warp < WPB;
  const long long estride = (long long)gridDim.x * WPB;
  long long e = (long long)blockIdx.x * WPB + warp;
  const bool limited = p.pull_limit > 0;
  long long nb = 0, nend = 0;                              // bounds of the NEXT example (prefetched)
  int nid[PB_PF];
  float nx[PB_PF];
  int nlabel = 1;
  auto fetch_features = [&](long long ee, long long bb, long long eend) {
#pragma unroll
    for (int q = 0; q < PB_PF; ++q) {
      const long long k = bb + lane + (long long)q * lanes_active;
      const bool have = lane < lanes_active && k < eend;
      nid[q] = have ? indices[k] : -1;
      nx[q] = have ? values[k] : 0.f;
    }
    nlabel = p.predict_only ? 1 : labels[ee];
  };
  if (puller && e < n) {
    nb = indptr[e];
    nend = indptr[e + 1];
    fetch_features(e, nb, nend);
  }
  long long b2 = 0, end2 = 0;                              // bounds of the example after the next
  if (puller && e + estride < n) { b2 = indptr[e + estride]; end2 = indptr[e + estride + 1]; }
  for (; puller && e < n; e += estride) {
    // ---- the current example moves out of the prefetch registers
    const long long b = nb, end = nend;
    int id[PB_PF];
    float x[PB_PF];
#pragma unroll
    for (int q = 0; q < PB_PF; ++q) { id[q] = nid[q]; x[q] = nx[q]; }
    const int label = nlabel;
    // ---- prefetch: features of the next example, bounds of the one after it
    nb = b2;
    nend = end2;
    if (e + estride < n) fetch_features(e + estride, nb, nend);
    if (e + 2 * estride < n) { b2 = indptr[e + 2 * estride]; end2 = indptr[e + 2 * estride + 1]; }
    // ---- pull + margin
    float m = 0.f, n2 = 0.f;
#pragma unroll
    for (int q = 0; q < PB_PF; ++q) {
      if (id[q] >= 0) {
        const RowRef r = locate(t, id[q]);
        if (!in_table(t, id[q], r)) { ++c.bad; id[q] = -1; }
        else {
          m = fmaf(x[q], pull_scalar(t, id[q], r), m);
          n2 = fmaf(x[q], x[q], n2);
          ++c.pulls;
          if (r.owner != p.local_shard) ++c.remote;
        }
      }
      // under a limit a lane has ONE weight load outstanding: the next chunk's load may not issue before this one is used
      if (limited) asm volatile("" ::"f"(m) : "memory");
    }
    for (long long k = b + lane + (long long)PB_PF * lanes_active; k < end && lane < lanes_active; k += lanes_active) {
      const long long fid = indices[k];
      const float fx = values[k];
      const RowRef r = locate(t, fid);
      if (!in_table(t, fid, r)) { ++c.bad; continue; }
      m = fmaf(fx, pull_scalar(t, fid, r), m);
      n2 = fmaf(fx, fx, n2);
      ++c.pulls;
      if (r.owner != p.local_shard) ++c.remote;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m += __shfl_xor_sync(0xffffffffu, m, o);
      n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    }
    if (lane == 0) {
      ++c.records;
      if (pred != nullptr) pred[e] = m > 0.f ? 1 : -1;
    }
    if (!p.predict_only) {
      const float y = label > 0 ? 1.f : -1.f;
      const float loss = fmaxf(0.f, 1.f - y * m);
      const float mult = tau_of(p.variant, p.C, loss, n2) * y;
      if (lane == 0) c.loss += loss;
      if (mult != 0.f) {
#pragma unroll
        for (int q = 0; q < PB_PF; ++q) {
          if (id[q] < 0) continue;
          const RowRef r = locate(t, id[q]);
          if (t.lazy) touch_state(t.state[r.owner], r.slot, FPS_ROW_PUSHED);
          push_add1(t, r.owner, r.slot * t.stride, mult * x[q]);
          ++c.pushes;
        }
        for (long long k = b + lane + (long long)PB_PF * lanes_active; k < end && lane < lanes_active; k += lanes_active) {
          const long long fid = indices[k];
          const RowRef r = locate(t, fid);
          if (!in_table(t, fid, r)) continue;
          if (t.lazy) touch_state(t.state[r.owner], r.slot, FPS_ROW_PUSHED);
          push_add1(t, r.owner, r.slot * t.stride, mult * values[k]);
          ++c.pushes;
        }
      }
    }
  }
  if (p.pull_limit > 0) flow_release(flow, credits);
  pa_flush(c, flow);
}

// ---- N4 / N5 multiclass: one CTA per example ---------------------------------------------------
// Warps split the active features, lanes split the class columns (float4 each); the decision
// vector d = W^T x is reduced across warps in shared memory, tau/y per class is computed once,
// then every warp scatters its features' row deltas.
// 64 registers (1024 threads per SM): with 256-thread CTAs measured against 72 and 92 registers: 0.722 / 0.931 / 1.135 ms
__global__ void __launch_bounds__(PM_THREADS, 1024 / PM_THREADS)
k_pa_multi(const __grid_constant__ FpsTableView t, const long long* __restrict__ indptr,
           const int* __restrict__ indices, const float* __restrict__ values,
           const int* __restrict__ labels, long long n, const FpsPaParams p,
           const float* __restrict__ cost, int* __restrict__ pred, int warps_pulling, int pull_u, FpsFlow* flow) {
  constexpr int WPB = PM_THREADS / 32;
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // float4 slices that hold classes; a table padded to line-aligned rows (stride > dim) has more that stay zero
  const int L = p.n_classes, nvec = (L + 3) >> 2;
  float* s_part = smem;                       // [WPB][stride]
  float* s_d = smem + WPB * t.stride;         // [stride] decision, then multiplier vector
  __shared__ float s_n2[WPB];
  __shared__ int s_q;
  __shared__ float s_tau;
  PaCounters c = {0.f, 0, 0, 0, 0, 0};
  // Limiter: a pulling warp has pull_u feature rows outstanding, the CTA's allowance is warps_pulling *
  // pull_u, taken once for the CTA's lifetime (grid sized by the host to fit the limit).
  const int credits = warps_pulling * pull_u;
  if (p.pull_limit > 0) flow_acquire(flow, credits, p.pull_limit);
  for (long long e = blockIdx.x; e < n; e += gridDim.x) {
    const long long b = indptr[e], end = indptr[e + 1];
    // ---- pull + decision vector
    // Features are dealt to the pulling warps round-robin.  Lane u of a warp fetches the id / value of the
    // warp's u-th feature and resolves its row ONCE (32 features per warp and pass), then the warp walks
    // them with shuffles, pull_u (<= PULL_U) rows in flight at a time.  (Before: every lane re-read the id and
    // re-derived the row for each feature, and a dependent id -> row load chain per two features.)
    constexpr int PULL_U = 4;
    float n2 = 0.f;
    for (int v0 = 0; v0 < nvec; v0 += 32) {
      const int vb = v0 + lane;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (long long base = b; base < end && warp < warps_pulling; base += 32ll * warps_pulling) {
        const long long mk = base + warp + (long long)lane * warps_pulling;
        const bool mhave = mk < end;
        const long long mid = mhave ? (long long)indices[mk] : 0;
        const float mx = mhave ? values[mk] : 0.f;
        const RowRef mr = locate(t, mid);
        const bool mok = mhave && in_table(t, mid, mr);
        if (mhave && !mok && v0 == 0) ++c.bad;
        const unsigned okmask = __ballot_sync(0xffffffffu, mok);
        const int rem = (int)min(end - base, 32ll * warps_pulling) - warp;       // features of this warp in this pass
        const int nfeat = rem > 0 ? (rem + warps_pulling - 1) / warps_pulling : 0;
        if (v0 == 0 && mok) {
          n2 = fmaf(mx, mx, n2);
          ++c.pulls;
          if (mr.owner != p.local_shard) ++c.remote;
        }
        for (int u0 = 0; u0 < nfeat; u0 += pull_u) {
          float4 w[PULL_U];
          float x[PULL_U];
#pragma unroll
          for (int u = 0; u < PULL_U; ++u) {
            const int src = min(u0 + u, 31);
            const int owner = __shfl_sync(0xffffffffu, mr.owner, src);
            const long long slot = __shfl_sync(0xffffffffu, mr.slot, src);
            x[u] = __shfl_sync(0xffffffffu, mx, src);
            long long fid = 0;
            if (t.lazy) fid = __shfl_sync(0xffffffffu, mid, src);
            const bool ok = u < pull_u && u0 + u < nfeat && ((okmask >> src) & 1u);
            if (!ok) x[u] = 0.f;
            w[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok && vb < nvec) {
              w[u] = ld_row4(t.rows[owner] + slot * t.stride + vb * 4);
              if (t.lazy) {
                const int st = touch_state(t.state[owner], slot, FPS_ROW_PULLED);
                if (st == FPS_ROW_PULLED && !(t.init_lo == 0.f && t.init_hi == 0.f)) {
                  const float4 i4 = init_block(t, fid, vb);
                  w[u].x += i4.x; w[u].y += i4.y; w[u].z += i4.z; w[u].w += i4.w;
                }
              }
            }
          }
#pragma unroll
          for (int u = 0; u < PULL_U; ++u) {
            acc.x = fmaf(x[u], w[u].x, acc.x); acc.y = fmaf(x[u], w[u].y, acc.y);
            acc.z = fmaf(x[u], w[u].z, acc.z); acc.w = fmaf(x[u], w[u].w, acc.w);
          }
        }
      }
      if (vb < nvec) *reinterpret_cast<float4*>(s_part + warp * t.stride + vb * 4) = acc;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    if (lane == 0) s_n2[warp] = n2;
    __syncthreads();
    for (int j = threadIdx.x; j < t.stride; j += PM_THREADS) {
      float d = 0.f;
#pragma unroll
      for (int w = 0; w < WPB; ++w) d += s_part[w * t.stride + j];
      s_d[j] = d;
    }
    __syncthreads();
    float norm2 = 0.f;
#pragma unroll
    for (int w = 0; w < WPB; ++w) norm2 += s_n2[w];
    const int y = p.predict_only ? 0 : labels[e];
    // ---- argmax family (thread 0; L is ~100: cheaper than a tree for this size)
    if (threadIdx.x == 0) {
      int am = 0;
      float best = s_d[0];
      for (int j = 1; j < L; ++j)
        if (s_d[j] > best) { best = s_d[j]; am = j; }
      if (pred != nullptr) pred[e] = am;
      int q = am;
      float tau = 0.f;
      if (!p.predict_only && p.multiclass >= 2) {
        if (p.multiclass == 3) {   // ML: argmax_i d_i - d_y + sqrt(cost(y, i))
          float bestv = -INFINITY;
          for (int j = 0; j < L; ++j) {
            const float cj = cost ? sqrtf(cost[y * L + j]) : (j == y ? 0.f : 1.f);
            const float v = s_d[j] - s_d[y] + cj;
            if (v > bestv) { bestv = v; q = j; }
          }
        }
        if (q != y) {
          const float cq = cost ? sqrtf(cost[y * L + q]) : 1.f;
          const float loss = s_d[q] - s_d[y] + cq;
          tau = norm2 > 0.f ? loss / (2.f * norm2) : 0.f;
          c.loss += fmaxf(loss, 0.f);
        }
      }
      s_q = q;
      s_tau = tau;
      ++c.records;
    }
    __syncthreads();
    if (!p.predict_only) {
      bool any = false;
      if (p.multiclass == 1) {
        // one-vs-all: per-class multiplier tau_c * y_c, written over s_d
        float lsum = 0.f;
        for (int j = threadIdx.x; j < t.stride; j += PM_THREADS) {
          float mult = 0.f;
          if (j < L) {
            const float yc = (j == y) ? 1.f : -1.f;
            const float loss = fmaxf(0.f, 1.f - s_d[j] * yc);
            mult = tau_of(p.variant, p.C, loss, norm2) * yc;
            lsum += loss;
          }
          s_d[j] = mult;
        }
        c.loss += lsum;
        __syncthreads();
        any = true;   // rows of all-zero multipliers are rare (every class already separated)
      }
      if (p.multiclass == 1 && any) {
        // same dealing as the pull: lane u resolves the warp's u-th feature, the warp walks them with shuffles
        const float4 m0 = lane < nvec ? *reinterpret_cast<const float4*>(s_d + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        for (long long base = b; base < end; base += 32ll * WPB) {
          const long long mk = base + warp + (long long)lane * WPB;
          const bool mhave = mk < end;
          const long long mid = mhave ? (long long)indices[mk] : 0;
          const float mx = mhave ? values[mk] : 0.f;
          const RowRef mr = locate(t, mid);
          const bool mok = mhave && in_table(t, mid, mr);
          if (mok) {
            if (t.lazy) touch_state(t.state[mr.owner], mr.slot, FPS_ROW_PUSHED);
            ++c.pushes;
          }
          const unsigned okmask = __ballot_sync(0xffffffffu, mok);
          const int rem = (int)min(end - base, 32ll * WPB) - warp;
          const int nfeat = rem > 0 ? (rem + WPB - 1) / WPB : 0;
          for (int u = 0; u < nfeat; ++u) {
            const int owner = __shfl_sync(0xffffffffu, mr.owner, u);
            const long long row = __shfl_sync(0xffffffffu, mr.slot, u) * t.stride;
            const float x = __shfl_sync(0xffffffffu, mx, u);
            if (!((okmask >> u) & 1u)) continue;
            if (lane < nvec) push_add4(t, owner, row + lane * 4, make_float4(x * m0.x, x * m0.y, x * m0.z, x * m0.w));
            for (int vb = lane + 32; vb < nvec; vb += 32) {
              const float4 m4 = *reinterpret_cast<const float4*>(s_d + vb * 4);
              push_add4(t, owner, row + vb * 4, make_float4(x * m4.x, x * m4.y, x * m4.z, x * m4.w));
            }
          }
        }
      } else if (p.multiclass >= 2 && s_q != y) {
        const float tau = s_tau;
        const int q = s_q;
        for (long long k = b + threadIdx.x; k < end; k += PM_THREADS) {
          const long long id = indices[k];
          const float x = values[k];
          const RowRef r = locate(t, id);
          if (!in_table(t, id, r)) continue;
          const long long row = r.slot * t.stride;
          if (t.lazy) touch_state(t.state[r.owner], r.slot, FPS_ROW_PUSHED);
          push_add1(t, r.owner, row + y, tau * x);
          push_add1(t, r.owner, row + q, -tau * x);
          ++c.pushes;
        }
      }
    }
    __syncthreads();
  }
  if (p.pull_limit > 0) flow_release(flow, credits);
  pa_flush<PM_THREADS>(c, flow);
}

}  // namespace

extern "C" int fps_pa_step(const FpsTableView* weights, const int64_t* d_indptr, const int32_t* d_indices,
                           const float* d_values, const int32_t* d_labels, int64_t n_examples,
                           const FpsPaParams* params, int32_t* d_pred, FpsFlow* d_flow, void* stream) {
  return fps_pa_step_cost(weights, d_indptr, d_indices, d_values, d_labels, n_examples, params, nullptr,
                          d_pred, d_flow, stream);
}

extern "C" int fps_pa_step_cost(const FpsTableView* weights, const int64_t* d_indptr, const int32_t* d_indices,
                                const float* d_values, const int32_t* d_labels, int64_t n_examples,
                                const FpsPaParams* params, const float* d_cost, int32_t* d_pred,
                                FpsFlow* d_flow, void* stream) {
  FPS_REQUIRE(weights && params && d_flow, FPS_E_INVALID, "fps_pa_step: null argument");
  FPS_REQUIRE(weights->n_shards >= 1 && weights->n_shards <= FPS_MAX_SHARDS, FPS_E_INVALID, "fps_pa_step: n_shards");
  FPS_REQUIRE(params->predict_only || d_labels != nullptr, FPS_E_INVALID, "fps_pa_step: labels required for training");
  if (n_examples <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int sms = fps_num_sms();
  if (params->multiclass == 0) {
    FPS_REQUIRE(weights->dim == 1, FPS_E_INVALID, "fps_pa_step: binary PA needs a dim-1 table");
    int warps_active = PA_THREADS / 32, lanes_active = 32;
    long long blocks = (n_examples + 7) / 8;
    if (blocks > (long long)sms * 8) blocks = (long long)sms * 8;
    if (params->pull_limit > 0) {
      const long long limit = params->pull_limit;
      if (limit < 32) { lanes_active = (int)limit; warps_active = 1; }
      else if (limit < 32 * (PA_THREADS / 32)) warps_active = (int)(limit / 32);
      const long long fit = limit / (warps_active * lanes_active);
      if (fit < (long long)sms * 8 && blocks > fit) blocks = fit > 0 ? fit : 1;   // else: only resident CTAs hold credits
    }
    k_pa_binary<<<(int)blocks, PA_THREADS, 0, st>>>(*weights, (const long long*)d_indptr, d_indices, d_values,
                                                    d_labels, n_examples, *params, d_pred, warps_active,
                                                    lanes_active, d_flow);
  } else {
    FPS_REQUIRE(weights->stride % 4 == 0 && weights->dim == params->n_classes, FPS_E_INVALID,
                "fps_pa_step: multiclass table must have dim == n_classes (%d vs %d)", weights->dim, params->n_classes);
    constexpr int PMT = PM_THREADS;
    const size_t smem = (size_t)(PMT / 32 + 1) * weights->stride * sizeof(float);
    FPS_REQUIRE(smem <= 200 * 1024, FPS_E_UNSUPPORTED, "fps_pa_step: too many classes (%d)", params->n_classes);
    if (smem > 48 * 1024) FPS_CUDA(cudaFuncSetAttribute(k_pa_multi, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long blocks = n_examples;
    if (blocks > (long long)sms * (2048 / PMT)) blocks = (long long)sms * (2048 / PMT);
    int warps_pulling = PMT / 32, pull_u = 4;        // = PULL_U of the kernel
    if (params->pull_limit > 0) {
      const long long limit = params->pull_limit;
      if (limit < 4) { pull_u = (int)limit; warps_pulling = 1; }
      else if (limit < 4 * (PMT / 32)) warps_pulling = (int)(limit / 4);
      const long long fit = limit / (warps_pulling * pull_u);
      // CTAs beyond the resident ones (64 registers: 1024 threads per SM) hold no credits until they start
      if (fit < (long long)sms * (1024 / PMT) && blocks > fit) blocks = fit > 0 ? fit : 1;
    }
    k_pa_multi<<<(int)blocks, PMT, smem, st>>>(*weights, (const long long*)d_indptr, d_indices, d_values,
                                                      d_labels, n_examples, *params, d_cost, d_pred, warps_pulling,
                                                      pull_u, d_flow);
  }
  FPS_LAUNCH_CHECK();
  return 0;
}
