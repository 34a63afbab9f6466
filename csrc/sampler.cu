// Negative sampling on the device (K4): the per-user recent-items memory and the rejection sampler
// of the MF workers, as two kernels that turn a batch of positive ratings into the expanded batch
// (all negatives with rating 0 first, draw by draw, then the positives) the fused SGD kernels consume.
//
// Reference behaviour (PSOnlineMatrixFactorizationWorker.scala:59-89,
// PSOfflineMatrixFactorizationWorker.scala:64-95): per rating, (1) the item enters the user's
// recent-items FIFO of `userMemory` entries and the worker's list of items seen so far, (2)
// min(#items seen - #items in the user's memory, negativeSampleRate) negatives are drawn uniformly from
// the items seen so far, re-drawing while the candidate is in the user's memory.
//
// Here the worker's state lives in device memory (caller-allocated, described by FpsNegSampler):
//   seen[slot, userMemory]   ring of the user's last items (slot = the user's row in the worker-resident
//                            user table), seen_count[slot] = items appended so far
//   item_bitmap / item_pool  set + list of distinct items this worker has seen, pool_count = its length
// Draws are Philox4x32-10 keyed by (seed, step tag, record, draw, retry): no RNG state; a pure function
// of those and of the pool (whose order is the order in which the atomics of k_neg_register landed).
// "Seen so far" has batch granularity: every record of a batch draws from the pool as it stands after
// the whole batch was registered.  A draw gives up after 32 rejected candidates (the reference's
// learner+generator worker does the same, ...AndTopKGeneratorWorker.scala:145-149; its plain online
// worker would loop forever); such a slot, like the slots beyond min(...), is marked FPS_SKIP_ID and
// the SGD kernels pass over it without counting it as a bad id.
#include "common.cuh"

namespace {

constexpr int NEG_MAX_RETRIES = 32;

// index of the user's ring: its row in the worker-resident user table (local_shard >= 0: the one local
// shard; -1: every shard is local, rings are laid out shard after shard), -1 for a user that is not ours
__device__ __forceinline__ long long ring_of(const FpsTableView& users, const FpsNegSampler& s, int u) {
  if (s.local_shard == -2) return (u >= 0 && u < s.n_rings) ? u : -1;   // one ring per user id (replicated memory)
  const RowRef ur = locate(users, u);
  if (!in_table(users, u, ur)) return -1;
  if (s.local_shard >= 0) return ur.owner == s.local_shard ? ur.slot : -1;
  return (long long)ur.owner * users.rows_per_shard + ur.slot;
}

__global__ void __launch_bounds__(256)
k_neg_register(const __grid_constant__ FpsTableView users, const FpsNegSampler s, const int* __restrict__ uid,
               const int* __restrict__ iid, long long n) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    const int u = uid[k], i = iid[k];
    if (u == FPS_SKIP_ID || i < 0 || i >= s.n_items) continue;
    const unsigned int bit = 1u << (i & 31);
    const unsigned int old = atomicOr(&s.item_bitmap[i >> 5], bit);
    if (!(old & bit)) s.item_pool[atomicAdd(s.pool_count, 1)] = i;
    const long long ring = ring_of(users, s, u);
    if (ring < 0) continue;
    const int pos = atomicAdd(&s.seen_count[ring], 1);
    s.seen[ring * s.user_memory + pos % s.user_memory] = i;
  }
}

// one thread per (record, draw)
__global__ void __launch_bounds__(256)
k_neg_draw(const __grid_constant__ FpsTableView users, const FpsNegSampler s, const int* __restrict__ uid,
           const int* __restrict__ iid, const float* __restrict__ rating, long long n, unsigned long long tag,
           const unsigned long long* __restrict__ d_tag, int* __restrict__ out_u, int* __restrict__ out_i,
           float* __restrict__ out_r) {
  if (d_tag != nullptr) tag = *d_tag;          // the step tag lives on the device (launches replayed from a CUDA graph)
  const int per = s.rate + 1;
  const long long total = n * per;
  const int pool_n = *s.pool_count;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long k = t / per;
    const int d = (int)(t - k * per);
    const int u = uid[k];
    // Layout: draw-major - all first negatives, all second negatives, ..., then all positives.  Record-major
    // blocks would put the rate + 1 records of one user next to each other, i.e. into sub-warps that run
    // at the same time and overwrite each other's update of that user's row (measured: the online job
    // stopped learning).  The order "negatives before their positive" of the reference is kept.
    const long long o = (long long)d * n + k;
    if (d == s.rate) {
      out_u[o] = u;
      out_i[o] = iid[k];
      out_r[o] = rating ? rating[k] : 1.f;
      continue;
    }
    int pick = FPS_SKIP_ID;
    const long long ring_ix = ring_of(users, s, u);
    if (ring_ix >= 0 && pool_n > 0) {
      const int have = min(s.seen_count[ring_ix], s.user_memory);
      const int* ring = s.seen + ring_ix * s.user_memory;
      if (d < pool_n - have) {              // min(itemIds.length - seenSet.size, negativeSampleRate) draws
        for (int retry = 0; retry < NEG_MAX_RETRIES && pick == FPS_SKIP_ID; retry += 4) {
          const uint4 w = philox4x32_10(make_uint4((uint32_t)k, (uint32_t)((unsigned long long)k >> 32),
                                                   (uint32_t)(d * NEG_MAX_RETRIES + retry), (uint32_t)tag),
                                        make_uint2((uint32_t)s.seed, (uint32_t)(s.seed >> 32) ^ (uint32_t)(tag >> 32)));
          const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (pick != FPS_SKIP_ID) break;
            const int cand = s.item_pool[ws[q] % (uint32_t)pool_n];
            bool seen = false;
            for (int m = 0; m < have; ++m) seen |= ring[m] == cand;
            if (!seen) pick = cand;
          }
        }
      }
    }
    out_u[o] = pick == FPS_SKIP_ID ? FPS_SKIP_ID : u;
    out_i[o] = pick == FPS_SKIP_ID ? 0 : pick;
    out_r[o] = 0.f;
  }
}

// the merger's / worker's per-user memory when every rank sees every event: ring = user id
__global__ void __launch_bounds__(256)
k_ring_append(int* __restrict__ seen, int* __restrict__ seen_count, int user_memory, long long n_rings,
              const int* __restrict__ uid, const int* __restrict__ iid, long long n) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    const int u = uid[k];
    if (u < 0 || u >= n_rings) continue;
    const int pos = atomicAdd(&seen_count[u], 1);
    seen[(long long)u * user_memory + pos % user_memory] = iid[k];
  }
}

// an item becomes a top-K candidate with its first rating on the worker that owns it
__global__ void __launch_bounds__(256)
k_mark_live(const __grid_constant__ FpsTableView items, int local_shard, const int* __restrict__ uid,
            const int* __restrict__ iid, long long n, int* __restrict__ live_ids) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
    if (uid[k] == FPS_SKIP_ID) continue;
    const int i = iid[k];
    const RowRef r = locate(items, i);
    if (!in_table(items, i, r)) continue;
    if (local_shard >= 0) {
      if (r.owner == local_shard) live_ids[r.slot] = i;
    } else {
      live_ids[(long long)r.owner * items.rows_per_shard + r.slot] = i;
    }
  }
}

}  // namespace

extern "C" {

int fps_seen_ring_append(int32_t* d_seen, int32_t* d_seen_count, int32_t user_memory, int64_t n_rings,
                         const int32_t* d_user_ids, const int32_t* d_item_ids, int64_t n, void* stream) {
  FPS_REQUIRE(d_seen && d_seen_count && user_memory > 0 && n_rings > 0, FPS_E_INVALID, "fps_seen_ring_append: bad arguments");
  if (n <= 0) return 0;
  long long b = (n + 255) / 256;
  if (b > fps_num_sms() * 16) b = fps_num_sms() * 16;
  k_ring_append<<<(int)b, 256, 0, (cudaStream_t)stream>>>(d_seen, d_seen_count, user_memory, n_rings, d_user_ids, d_item_ids, n);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_mark_live_items(const FpsTableView* items, int32_t local_shard, const int32_t* d_user_ids,
                        const int32_t* d_item_ids, int64_t n, int32_t* d_live_ids, void* stream) {
  if (int e = fps_check_view(items, "fps_mark_live_items")) return e;
  FPS_REQUIRE(d_live_ids != nullptr, FPS_E_INVALID, "fps_mark_live_items: live_ids is required");
  if (n <= 0) return 0;
  long long b = (n + 255) / 256;
  if (b > fps_num_sms() * 16) b = fps_num_sms() * 16;
  k_mark_live<<<(int)b, 256, 0, (cudaStream_t)stream>>>(*items, local_shard, d_user_ids, d_item_ids, n, d_live_ids);
  FPS_LAUNCH_CHECK();
  return 0;
}

static int sample_negatives(const FpsTableView* users, const FpsNegSampler* sampler, const int32_t* d_user_ids,
                            const int32_t* d_item_ids, const float* d_ratings, int64_t n, uint64_t step_tag,
                            const uint64_t* d_step_tag, int32_t* d_out_users, int32_t* d_out_items,
                            float* d_out_ratings, void* stream) {
  if (int e = fps_check_view(users, "fps_mf_sample_negatives")) return e;
  FPS_REQUIRE(sampler != nullptr && sampler->seen && sampler->seen_count && sampler->item_bitmap &&
                  sampler->item_pool && sampler->pool_count,
              FPS_E_INVALID, "fps_mf_sample_negatives: sampler state pointers are required");
  FPS_REQUIRE(sampler->rate >= 0 && sampler->rate <= 1024 && sampler->user_memory > 0 && sampler->n_items > 0,
              FPS_E_INVALID, "fps_mf_sample_negatives: rate in [0, 1024], user_memory > 0, n_items > 0");
  FPS_REQUIRE(d_out_users && d_out_items && d_out_ratings, FPS_E_INVALID, "fps_mf_sample_negatives: outputs are required");
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int sms = fps_num_sms();
  long long b1 = (n + 255) / 256, b2 = (n * (sampler->rate + 1) + 255) / 256;
  if (b1 > sms * 16) b1 = sms * 16;
  if (b2 > sms * 32) b2 = sms * 32;
  k_neg_register<<<(int)b1, 256, 0, st>>>(*users, *sampler, d_user_ids, d_item_ids, n);
  k_neg_draw<<<(int)b2, 256, 0, st>>>(*users, *sampler, d_user_ids, d_item_ids, d_ratings, n, step_tag,
                                       (const unsigned long long*)d_step_tag, d_out_users, d_out_items, d_out_ratings);
  FPS_LAUNCH_CHECK();
  return 0;
}

int fps_mf_sample_negatives(const FpsTableView* users, const FpsNegSampler* sampler, const int32_t* d_user_ids,
                            const int32_t* d_item_ids, const float* d_ratings, int64_t n, uint64_t step_tag,
                            int32_t* d_out_users, int32_t* d_out_items, float* d_out_ratings, void* stream) {
  return sample_negatives(users, sampler, d_user_ids, d_item_ids, d_ratings, n, step_tag, nullptr, d_out_users,
                          d_out_items, d_out_ratings, stream);
}

int fps_mf_sample_negatives_dtag(const FpsTableView* users, const FpsNegSampler* sampler, const int32_t* d_user_ids,
                                 const int32_t* d_item_ids, const float* d_ratings, int64_t n,
                                 const uint64_t* d_step_tag, int32_t* d_out_users, int32_t* d_out_items,
                                 float* d_out_ratings, void* stream) {
  FPS_REQUIRE(d_step_tag != nullptr, FPS_E_INVALID, "fps_mf_sample_negatives_dtag: the device step tag is required");
  return sample_negatives(users, sampler, d_user_ids, d_item_ids, d_ratings, n, 0, d_step_tag, d_out_users,
                          d_out_items, d_out_ratings, stream);
}

}  // extern "C"
