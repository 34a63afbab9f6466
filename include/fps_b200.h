/* fps_b200 - C ABI of the B200-native parameter-server kernels (sm_100a).
 *
 * Every entry point returns 0 on success or a non-zero code (a cudaError_t value, or one of the
 * FPS_E_* codes below); fps_last_error() returns a human-readable description of the last
 * failure on the calling thread.  All pointers named d_* are device pointers.  `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  Nothing here synchronises the
 * stream unless stated.
 *
 * "Reference" citations (file:line) point at the JVM code whose behaviour the entry point
 * provides; M/ = src/main/scala/hu/sztaki/ilab/ps/ of flink-parameter-server.
 */
#ifndef FPS_B200_H
#define FPS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FPS_MAX_SHARDS 16

#define FPS_E_INVALID   10001   /* bad argument                                   */
#define FPS_E_TOO_LARGE 10002   /* size beyond a documented fixed maximum          */
#define FPS_E_UNSUPPORTED 10003 /* shape not supported by this build               */

/* How parameter ids map to (owner shard, slot).  Same formulas as the reference:
 *   HASH : owner = id % n_shards, slot = id / n_shards            (M/FlinkParameterServer.scala:126-134)
 *   RANGE: owner = id / div,      slot = id - owner * div         (M/passive/aggressive/PassiveAggressiveParameterServer.scala:372-375,
 *                                                                  M/server/RangePSLogicWithClose.scala:53-61)            */
enum { FPS_PART_HASH = 0, FPS_PART_RANGE = 1 };

/* Row life-cycle in lazy mode (state byte per row; SURVEY 2.9 "Parameter life-cycle"):
 *   0 untouched; 1 first touch was a pull -> value = init(id) + stored; 2 first touch was a push
 *   (or a model load) -> value = stored.  In eager mode `state` pointers are NULL and every row
 *   already holds init(id).                                                                      */
enum { FPS_ROW_UNTOUCHED = 0, FPS_ROW_PULLED = 1, FPS_ROW_PUSHED = 2 };

/* A sharded fp32 table as seen from one GPU: rows[s] is shard s' base pointer mapped into this
 * process (its own allocation for the local shard; a peer's shard through a VMM symmetric-memory mapping,
 * or a cudaIpc mapping with peer_mapping="ipc").            */
typedef struct FpsTableView {
  float*    rows[FPS_MAX_SHARDS];
  uint8_t*  state[FPS_MAX_SHARDS];   /* NULL everywhere = eager table                            */
  int32_t   n_shards;
  int32_t   dim;                     /* logical row length                                       */
  int32_t   stride;                  /* allocated row length in floats, multiple of 4            */
  int32_t   partition;               /* FPS_PART_*                                               */
  int64_t   div;                     /* RANGE divisor = ceil(n_ids / n_shards)                   */
  int64_t   rows_per_shard;          /* capacity of every shard (bounds checks)                  */
  float     init_lo, init_hi;        /* lazy init: value = lo + (hi-lo)*U(seed, stream, id, j)   */
  uint64_t  seed;
  int32_t   stream_tag;              /* Philox stream: 0 = PS-side table, 1 = worker-side table  */
  int32_t   lazy;                    /* 1 = use the state bytes                                  */
  /* Replicated table (small hot tables, e.g. 47K binary-PA weights): every GPU holds the whole
   * table (n_shards = 1, rows[0] = the local replica, pulls never leave the GPU) and a push is
   * applied to all replicas - with ONE multimem.red on `mc_rows`, the NVLS multicast alias of the
   * replicas (NVSwitch does the fan-out), or, without multicast support, one red.add per entry of
   * replica_rows.  With `delta_rows` (a local zeroed array of the table's shape) pushes are deferred:
   * a push adds to the local replica and to delta_rows, and fps_table_flush_delta sends the
   * accumulated deltas to the other replicas after the step - thousands of 4-byte pushes to one hot
   * weight cross NVLink as one add (measured: single-word multimem.red / red.add top out near 5 G
   * packets/s per GPU whatever the payload).  n_replicas = 0 for an ordinary sharded table.  No
   * counterpart in the reference (its shards are always partitions).                           */
  float*    mc_rows;
  float*    delta_rows;
  float*    replica_rows[FPS_MAX_SHARDS];
  int32_t   n_replicas;
  int32_t   replica_rank;            /* index of the local replica in replica_rows                */
  int64_t   n_ids;                   /* ids are valid in [0, n_ids); 0 = only the shard capacity bounds them */
} FpsTableView;

/* Device-side flow control + observability block (one per worker; 256 B, zero it before use).
 * The pull limiter of M/WorkerLogic.scala:176-225 and the idle timeout of
 * M/FlinkParameterServer.scala:49-52 live here, in device memory, not on the host.              */
typedef struct FpsFlow {
  int32_t   in_flight;               /* unanswered pulls right now                               */
  int32_t   max_in_flight;           /* high-water mark (must stay <= pull_limit)                */
  uint64_t  limiter_stalls;          /* times a CTA had to wait for credits                      */
  uint64_t  pulls;                   /* rows pulled                                              */
  uint64_t  pushes;                  /* rows pushed                                              */
  uint64_t  remote_pulls;            /* of which from another shard than `local_shard`           */
  uint64_t  remote_pushes;
  uint64_t  nan_count;               /* non-finite factors seen (FactorIsNotANumberException)    */
  uint64_t  bad_ids;                 /* ids outside the table                                    */
  double    loss_sum;                /* sum of squared errors / hinge losses                     */
  uint64_t  records;                 /* records processed                                        */
  uint64_t  last_activity_ns;        /* globaltimer stamp of the last processed chunk            */
  uint64_t  ring_head;               /* streaming kernel: next record index to claim             */
  uint64_t  ring_done;               /* streaming kernel: records fully retired                  */
  int32_t   exit_reason;             /* 1 idle timeout, 2 stop flag, 3 lifetime cap              */
  int32_t   combined_groups;         /* hot-key groups pre-combined in shared memory             */
  uint64_t  reserved[18];            /* [0] = item rows actually pulled by the combine variant   */
} FpsFlow;

const char* fps_last_error(void);
int fps_version(void);

/* ---- runtime: memory, peer mapping, probes --------------------------------------------------- */
int fps_malloc(void** d_ptr, size_t bytes);                  /* cudaMalloc (IPC-exportable)      */
int fps_free(void* d_ptr);
int fps_memset(void* d_ptr, int value, size_t bytes, void* stream);
int fps_ipc_get_handle(void* d_ptr, unsigned char handle[64]);
int fps_ipc_open_handle(const unsigned char handle[64], void** d_ptr);
int fps_ipc_close_handle(void* d_ptr);
int fps_enable_peer_access(int peer_device);                 /* single-process multi-GPU         */
int fps_host_alloc_mapped(void** h_ptr, void** d_ptr, size_t bytes);  /* pinned + device alias   */
int fps_host_free(void* h_ptr);
/* P2P probe: moves `bytes` between local and (possibly peer-mapped) buffers with the three access
 * types the PS uses; mode 0 = 128-bit loads from src, 1 = 128-bit stores to dst, 2 = red.add.v4.f32
 * into dst.  Returns elapsed ms of `iters` launches via *ms.                                       */
int fps_p2p_probe(float* d_dst, const float* d_src, size_t bytes, int mode, int iters, float* ms,
                  void* stream);

/* Random-row probes (the PS access pattern, unlike the streaming probe above): n_rows rows of row_floats
 * floats at random offsets within span_rows rows of d_table (local or peer-mapped); mode as above;
 * `unroll` rows in flight per warp, blocks_per_sm CTAs per SM.  The mixed probe interleaves peer loads,
 * peer red.adds, local loads and local stores as selected by the bits of `flags` (1, 2, 4, 8; 128 = all
 * four on one row index; 256 = the peer push as one 512-byte cp.reduce.async.bulk per row instead of
 * 32 red.add.v4) - the mix the fused MF kernel issues.  Elapsed ms of `iters` launches in *ms.          */
int fps_p2p_probe_random(float* d_table, int64_t span_rows, int row_floats, int64_t n_rows, int mode,
                         int unroll, int blocks_per_sm, int iters, float* ms, void* stream);
int fps_p2p_probe_mixed(float* d_remote, float* d_local, int64_t span_rows, int64_t n_rows, int flags,
                        int iters, float* ms, void* stream);

/* ---- table: init (N9), load (N10), dump (N11) ------------------------------------------------ */
/* Fill local shard `shard` with init(id) for every slot (eager tables) - RangedRandomFactorInitializer
 * (M/matrix/factorization/factors/RangedRandomFactorInitializer.scala:7-9), made counter-based.  */
int fps_table_init_shard(const FpsTableView* view, int shard, int64_t n_rows, void* stream);
/* Gather rows by id (pull without compute): out[i, :dim] = value(ids[i]).  ids may be remote.    */
int fps_table_gather(const FpsTableView* view, const int64_t* d_ids, int64_t n, float* d_out,
                     int out_stride, FpsFlow* d_flow, void* stream);
/* Locked read-modify-write (N13; M/server/LockPSLogicA.scala:13-46 on the device): row <- alpha * row +
 * delta with the row's lock word held - an int32 in auxiliary column `lock_col` of the row itself
 * (dim <= lock_col < stride; auxiliary columns start as zero = unlocked), taken with a system-scope CAS
 * by one lane of the warp that runs the cycle, so it also excludes warps on other GPUs.  Sharded tables
 * only; lazy tables need the zero initialiser.  use_lock = 0 runs the same cycle unprotected (for tests).
 * d_flow (optional): pulls / pushes += cycles, limiter_stalls += lock retries.                        */
int fps_table_locked_axpby(const FpsTableView* view, const int64_t* d_ids, int64_t n, float alpha,
                           const float* d_deltas, int32_t delta_stride, int32_t lock_col, int32_t use_lock,
                           FpsFlow* d_flow, void* stream);
/* Push rows: mode 0 = add (vectorSum), 1 = overwrite, 2 = add then store the new L2 norm in
 * column `dim` (LengthAndVector tables, stride > dim).  value semantics of SimplePSLogic
 * (M/server/SimplePSLogic.scala:16-25): an untouched row becomes exactly the pushed value.       */
int fps_table_scatter(const FpsTableView* view, const int64_t* d_ids, int64_t n, const float* d_vals,
                      int val_stride, int mode, FpsFlow* d_flow, void* stream);
/* Dump the live rows of local shard `shard` (close() of the ...WithClose logics,
 * M/server/SimplePSLogicWithClose.scala:29-31): compacts (id, row) pairs into d_out_ids /
 * d_out_rows, count in *d_count (device int64, zero it first).  Eager tables dump every row.     */
int fps_table_dump_shard(const FpsTableView* view, int shard, int64_t n_rows, int64_t n_ids_total,
                         int64_t* d_out_ids, float* d_out_rows, int64_t capacity, int64_t* d_count,
                         void* stream);

/* The same for the slots [row_begin, row_end) of the shard only: a checkpoint streams a table through a
 * fixed-size buffer chunk by chunk instead of materialising it (M/FlinkParameterServer.scala:377-566 is
 * the matching streamed load).                                                                        */
int fps_table_dump_range(const FpsTableView* view, int shard, int64_t row_begin, int64_t row_end,
                         int64_t n_ids_total, int64_t* d_out_ids, float* d_out_rows, int64_t capacity,
                         int64_t* d_count, void* stream);

/* ---- matrix factorisation (N1, N2) ----------------------------------------------------------- */
typedef struct FpsMfParams {
  float   lr;
  float   lambda;          /* L2 regularisation, reference = 0                                    */
  int32_t pull_limit;      /* 0 = unlimited; else max unanswered pulls per worker                 */
  int32_t local_shard;     /* this worker's index (its user rows, and "remote" accounting)        */
  int32_t variant;         /* fps_mf_sgd_step: 0 = direct (sub-warp per rating, rows in registers)
                              2 = combine (round-1 path kept for comparison): cub radix sort by item,
                                  one pull + one combined push per run inside a 32-record slice
                              (the tile path, fps_mf_bucket + fps_mf_sgd_tiles, ignores this field)   */
  int32_t unroll;          /* ratings a sub-warp keeps in flight: 0 = default (1), 1, 2 or 4       */
  int32_t blocks_per_sm;   /* cap of the grid = SMs * this; 0 = default (64, i.e. not persistent)  */
  int32_t reserved;        /* tuning: 6 = the unroll-1 build without the 32-register launch bound  */
} FpsMfParams;

/* One fused pull-compute-push pass over n ratings: for each (user, item, r): v = pull(item) from
 * its owner shard (P2P load if remote), u = local user row; e = r - u.v; u += lr*(e*v - lambda*u);
 * push(item, lr*(e*u - lambda*v)) as a vector red.add on the owner.  SGDUpdater.scala:5-11 +
 * PSOfflineMatrixFactorizationWorker.scala:131-146 + SimplePSLogic.scala:16-25 in one kernel.    */
int fps_mf_sgd_step(const FpsTableView* items, const FpsTableView* users,
                    const int32_t* d_user_ids, const int32_t* d_item_ids, const float* d_ratings,
                    int64_t n, const FpsMfParams* params, FpsFlow* d_flow, void* stream);

/* Streaming worker (PSOnlineMatrixFactorizationWorker.scala:22-90 on device): a persistent kernel
 * that consumes a ring of ratings written by the host, exits by itself after the ring has been
 * idle for iteration_wait_ms (0 = until *stop != 0), never lives longer than max_lifetime_ms.
 * ring_* are device-visible (mapped pinned or device memory); `tail` is advanced by the producer.*/
int fps_mf_online_stream(const FpsTableView* items, const FpsTableView* users,
                         const int32_t* ring_users, const int32_t* ring_items, const float* ring_ratings,
                         uint64_t ring_capacity, const volatile uint64_t* tail,
                         const volatile int32_t* stop, const FpsMfParams* params,
                         uint32_t iteration_wait_ms, uint32_t max_lifetime_ms, FpsFlow* d_flow,
                         void* stream);

/* Variant "tiles" (csrc/mf_tiles.cu): hot-key pre-combining in shared memory with TMA-staged rows,
 * in two calls so that the first can overlap the previous step.
 *   fps_mf_bucket     groups the n ratings by (item mod NB) into packed records in caller-owned
 *                     scratch (fps_mf_tiles_scratch_bytes(n) bytes) with three small kernels of our
 *                     own (count / scan / scatter); depends on the inputs only.
 *   fps_mf_sgd_tiles  one CTA per tile of bucketed records: shared-memory sort by item, then every
 *                     warp walks a sorted segment carrying (item, row, combined delta); rows of up to
 *                     512 bytes arrive through per-warp cp.async.bulk rings.  Same update rule as
 *                     fps_mf_sgd_step; one pull per run of equal items, one push per <= 32 records of it.
 *                     Eager, sharded tables of rank <= 512 (else FPS_E_UNSUPPORTED).  params->reserved =
 *                     tile size (512 / 1024 / 2048, 0 = 1024); params->unroll = 1 forces the register
 *                     path.  Pull limiter: every CTA takes its allowance from FpsFlow.in_flight once.
 * Replaces X1 / X2 of M/FlinkParameterServer.scala:265-317 and SimplePSLogic.scala:16-25 like
 * fps_mf_sgd_step, with far fewer messages for popular items.                                      */
size_t fps_mf_tiles_scratch_bytes(int64_t n);
int fps_mf_bucket(const int32_t* d_user_ids, const int32_t* d_item_ids, const float* d_ratings, int64_t n,
                  void* d_scratch, size_t scratch_bytes, void* stream);
int fps_mf_sgd_tiles(const FpsTableView* items, const FpsTableView* users, void* d_scratch, int64_t n,
                     const FpsMfParams* params, FpsFlow* d_flow, void* stream);

/* Appends (user, item) to per-user-id rings of user_memory entries (the seen filter of
 * CollectTopKFromEachWorker.scala:50-55 kept on the device; same layout as FpsNegSampler.seen).         */
int fps_seen_ring_append(int32_t* d_seen, int32_t* d_seen_count, int32_t user_memory, int64_t n_rings,
                         const int32_t* d_user_ids, const int32_t* d_item_ids, int64_t n, void* stream);
/* live_ids[slot of item in the local shard(s)] = item id for every record whose item this worker owns
 * (local_shard >= 0: that shard; -1: all shards, laid out shard after shard) and whose user id is not
 * FPS_SKIP_ID: an item becomes a top-K candidate with its first rating
 * (PSOnlineMatrixFactorizationAndTopKGeneratorWorker.scala:169-187).                                   */
int fps_mark_live_items(const FpsTableView* items, int32_t local_shard, const int32_t* d_user_ids,
                        const int32_t* d_item_ids, int64_t n, int32_t* d_live_ids, void* stream);

/* ---- negative sampling on the device (K4) ------------------------------------------------------ */
/* A record whose user id is FPS_SKIP_ID is passed over by every MF kernel without being counted.   */
#define FPS_SKIP_ID (-2147483647 - 1)
/* Device-resident state of one worker's sampler; the caller allocates and zeroes the arrays (seen may
 * hold anything: seen_count says how much of a ring is valid).
 *   seen        [n_rings * user_memory] int32 - ring of the last user_memory items of every local user
 *   seen_count  [n_rings] int32              - items appended to the ring so far
 *   item_bitmap [ceil(n_items / 32)] uint32   - items this worker has seen
 *   item_pool   [n_items] int32, pool_count [1] int32 - the same as a list, in order of first sight
 * n_rings = rows_per_shard of the user table (local_shard >= 0) or n_shards * rows_per_shard (-1).    */
typedef struct FpsNegSampler {
  int32_t*  seen;
  int32_t*  seen_count;
  uint32_t* item_bitmap;
  int32_t*  item_pool;
  int32_t*  pool_count;
  int64_t   n_items;
  uint64_t  seed;
  int32_t   user_memory;     /* userMemory of the reference (default 128)                           */
  int32_t   rate;            /* negativeSampleRate                                                  */
  int32_t   local_shard;     /* shard of the user table that lives here, -1 = all of them, -2 = rings
                                are indexed by the user id itself (n_rings of them; a memory kept on
                                every rank, as in the learner+generator job where all ranks see all events) */
  int32_t   reserved;
  int64_t   n_rings;         /* local_shard == -2 only                                              */
} FpsNegSampler;
/* Registers the n positive ratings (item -> the user's recent-items ring and the worker's item pool),
 * then writes the expanded batch of n * (rate + 1) records: slot d * n + k holds negative d of rating k
 * (rating 0, drawn from the items seen so far, rejected while in the user's ring, at most 32 candidates
 * each), slots rate * n + k the ratings themselves; unused negative slots get user id FPS_SKIP_ID.  Draws are keyed by
 * (seed, step_tag, record, draw).  PSOnlineMatrixFactorizationWorker.scala:59-89,
 * PSOfflineMatrixFactorizationWorker.scala:64-95.                                                     */
int fps_mf_sample_negatives(const FpsTableView* users, const FpsNegSampler* sampler,
                            const int32_t* d_user_ids, const int32_t* d_item_ids, const float* d_ratings,
                            int64_t n, uint64_t step_tag, int32_t* d_out_users, int32_t* d_out_items,
                            float* d_out_ratings, void* stream);
/* The same with the step tag read from device memory at launch time: for step sequences replayed from a
 * CUDA graph, where a by-value tag would be frozen at capture (the caller bumps *d_step_tag on the stream). */
int fps_mf_sample_negatives_dtag(const FpsTableView* users, const FpsNegSampler* sampler,
                                 const int32_t* d_user_ids, const int32_t* d_item_ids, const float* d_ratings,
                                 int64_t n, const uint64_t* d_step_tag, int32_t* d_out_users,
                                 int32_t* d_out_items, float* d_out_ratings, void* stream);

/* Replicated table with deferred pushes: add the locally accumulated deltas to every other replica
 * (red.add.v4 per peer, all-zero 16-byte chunks are skipped) and clear them.  Run it on the stream of
 * the step kernels, after them.                                                                    */
int fps_table_flush_delta(const FpsTableView* t, void* stream);

/* Predict r_hat = u.v for n (user, item) pairs (RMSE evaluation).                                */
int fps_mf_predict(const FpsTableView* items, const FpsTableView* users, const int32_t* d_user_ids,
                   const int32_t* d_item_ids, int64_t n, float* d_out, void* stream);

/* ---- passive-aggressive (N3, N4, N5) --------------------------------------------------------- */
typedef struct FpsPaParams {
  int32_t variant;        /* 0 = PA, 1 = PA-I, 2 = PA-II  (PassiveAggressiveBinaryAlgorithm.scala:95-112) */
  float   C;              /* aggressiveness                                                       */
  int32_t multiclass;     /* 0 binary, 1 one-vs-all, 2 cost-based PB, 3 cost-based ML             */
  int32_t n_classes;
  int32_t pull_limit;     /* counted in features (rows), like the reference limiter               */
  int32_t local_shard;
  int32_t predict_only;   /* 1 = unlabelled: write predictions, push nothing                      */
  int32_t reserved;
} FpsPaParams;

/* CSR mini-batch: example e has features d_indices[d_indptr[e] .. d_indptr[e+1]) with values
 * d_values; label d_labels[e] (+1/-1 binary, class index multiclass).  Fused gather (pull of every
 * active feature), margin / tau, and scatter red.add (push of tau*y*x_i).  d_pred (may be NULL)
 * receives the prediction made with the pulled (pre-update) weights.
 * PassiveAggressiveParameterServer.scala:283-338 + algorithm/*.scala.                            */
int fps_pa_step(const FpsTableView* weights, const int64_t* d_indptr, const int32_t* d_indices,
                const float* d_values, const int32_t* d_labels, int64_t n_examples,
                const FpsPaParams* params, int32_t* d_pred, FpsFlow* d_flow, void* stream);

/* Same, with an optional class-cost matrix d_cost[n_classes * n_classes] (row = true label) for
 * the cost-based variants (PassiveAggressiveCostBased.scala:54-64); NULL = 0/1 cost.             */
int fps_pa_step_cost(const FpsTableView* weights, const int64_t* d_indptr, const int32_t* d_indices,
                     const float* d_values, const int32_t* d_labels, int64_t n_examples,
                     const FpsPaParams* params, const float* d_cost, int32_t* d_pred, FpsFlow* d_flow,
                     void* stream);

/* ---- LEMP top-K (N6, N7) ---------------------------------------------------------------------- */
/* Scores n_q queries against n_items local item rows on the tensor cores (tcgen05.mma kind::tf32,
 * 3xTF32 split for fp32 fidelity, accumulators in TMEM), prunes tiles by the length bound
 * ||q||*max||p|| <= theta (LEMP LENGTH test, PSTopKGeneratorWorker.scala:72-76) and keeps a
 * per-query top-k.  Items must be sorted by length descending (d_item_len) for the bound to bite.
 * Partial top-k lists per (query, item-split) go to caller-provided scratch, then are merged.
 * Outputs: d_out_scores/d_out_ids [n_q, k] sorted descending, ids are d_item_ids[slot] or the
 * slot itself when d_item_ids is NULL; unused entries have id -1.
 * scratch_bytes must be >= fps_topk_scratch_bytes(n_q, k, n_items).
 * use_pruning: bit 0 = the LENGTH bound; bit 1 = items whose d_item_ids entry is negative are skipped
 * (scoring a worker-resident table in place, where not every slot is a live item).                 */
size_t fps_topk_scratch_bytes(int64_t n_q, int32_t k, int64_t n_items);
/* With a buffer of this size (>= the one above) fps_lemp_topk / fps_topk_simt cut the items into up to 64
 * splits per query tile instead of 16, so that a small query batch still occupies the SMs; with anything in
 * between they take as many splits as fit.                                                              */
size_t fps_topk_scratch_bytes_full(int64_t n_q, int32_t k, int64_t n_items);
int fps_lemp_topk(const float* d_queries, int32_t q_stride, int64_t n_q,
                  const float* d_items, int32_t item_stride, const float* d_item_len,
                  const int32_t* d_item_ids, int64_t n_items, int32_t dim, int32_t k,
                  int32_t use_pruning, void* d_scratch, size_t scratch_bytes,
                  float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats, void* stream);
/* Second-generation path: pack the (length-sorted) items once into per-tile tf32 hi/lo images in
 * the swizzled layout the tensor cores read (2x the fp32 bytes; fps_topk_packed_bytes), then score
 * query batches with fps_lemp_topk_packed, which streams the tiles with TMA bulk copies into a
 * warp-specialised tcgen05 pipeline (four TMEM accumulators, 1 / 2 / 4 groups of selection warps;
 * env FPS_TOPK_GROUPS).  Same results as fps_lemp_topk.  Scratch: fps_topk_packed_scratch_bytes
 * lets every selection group keep its own candidate lists; with less (down to
 * fps_topk_scratch_bytes) the kernel runs with fewer groups.                                      */
size_t fps_topk_packed_bytes(int64_t n_items, int32_t dim);
size_t fps_topk_packed_scratch_bytes(int64_t n_q, int32_t k, int64_t n_items);
int fps_topk_pack_items(const float* d_items, int32_t item_stride, int64_t n_items, int32_t dim,
                        void* d_packed, void* stream);
int fps_lemp_topk_packed(const float* d_queries, int32_t q_stride, int64_t n_q, const void* d_packed,
                         const float* d_item_len, const int32_t* d_item_ids, int64_t n_items,
                         int32_t dim, int32_t k, int32_t use_pruning, void* d_scratch,
                         size_t scratch_bytes, float* d_out_scores, int32_t* d_out_ids,
                         uint64_t* d_stats, void* stream);
/* Wide variant of the packed path: 128-item tiles (one tcgen05.mma covers N = 128, twice the math per
 * instruction), shared-memory stages of one K block, two TMEM accumulators.  Own packed layout
 * (fps_topk_pack_items_wide), own scratch size; same results and arguments as fps_lemp_topk_packed.   */
size_t fps_topk_wide_packed_bytes(int64_t n_items, int32_t dim);
size_t fps_topk_wide_scratch_bytes(int64_t n_q, int32_t k, int64_t n_items);
int fps_topk_pack_items_wide(const float* d_items, int32_t item_stride, int64_t n_items, int32_t dim,
                             void* d_packed, void* stream);
/* The same for a table scored in place: d_live_ids[slot] < 0 marks a slot that is not a candidate.  The packed
 * image gets one more column (rank + 1 <= 128: size it with fps_topk_wide_packed_bytes(n_items, dim + 1)) that
 * holds 0 for a live slot and -1e30 for a dead one; run fps_lemp_topk_wide with dim + 1 and queries whose column
 * `dim` is 1: live slots score exactly their inner product, dead ones about -1e30 (ids from d_live_ids: < 0).  */
int fps_topk_pack_items_wide_live(const float* d_items, int32_t item_stride, int64_t n_items, int32_t dim,
                                  const int32_t* d_live_ids, void* d_packed, void* stream);
int fps_lemp_topk_wide(const float* d_queries, int32_t q_stride, int64_t n_q, const void* d_packed,
                       const float* d_item_len, const int32_t* d_item_ids, int64_t n_items,
                       int32_t dim, int32_t k, int32_t use_pruning, void* d_scratch,
                       size_t scratch_bytes, float* d_out_scores, int32_t* d_out_ids,
                       uint64_t* d_stats, void* stream);
/* fps_lemp_topk for callers that cannot allocate device memory - no counterpart in the reference.
 * Meant for request paths where a cudaMalloc may fail (all device memory is already spoken for by
 * a training job or another allocator) or would stall (it synchronises the device).  The scratch
 * lives in an array built into the library (FPS_TOPK_STATIC_SCRATCH_BYTES, held by every GPU that
 * loads the library); a problem needing more than that (fps_topk_scratch_bytes; about 256 queries
 * per call) is refused with FPS_E_TOO_LARGE.  Calls share the array: issue them on one stream, never
 * overlapping.  The first call on each device does a one-time address look-up and should be made
 * outside stream capture.  Results are identical to fps_lemp_topk.                                */
#define FPS_TOPK_STATIC_SCRATCH_BYTES (16u << 20)
int fps_lemp_topk_static(const float* d_queries, int32_t q_stride, int64_t n_q,
                         const float* d_items, int32_t item_stride, const float* d_item_len,
                         const int32_t* d_item_ids, int64_t n_items, int32_t dim, int32_t k,
                         int32_t use_pruning, float* d_out_scores, int32_t* d_out_ids,
                         uint64_t* d_stats, void* stream);
/* CUDA-core fp32 reference scorer with the same interface (baseline + oracle on device).          */
int fps_topk_simt(const float* d_queries, int32_t q_stride, int64_t n_q,
                  const float* d_items, int32_t item_stride, const int32_t* d_item_ids,
                  int64_t n_items, int32_t dim, int32_t k, void* d_scratch, size_t scratch_bytes,
                  float* d_out_scores, int32_t* d_out_ids, void* stream);
/* LEMP candidate filters executed on the device (E11 / E12; csrc/lemp_filters.cu).  strategy: 0 LENGTH,
 * 1 COORD, 2 INCR(n_focus), 3 LC(switch_threshold), 4 LI(n_focus, switch_threshold)
 * (pruning/LEMPPruningStrategy.scala:6-50); n_focus <= 16.  Items sorted by length descending, cut into
 * buckets of bucket_size; a bucket's predicate is chosen and parameterised as in
 * PSTopKGeneratorWorker.scala:82-96.
 * fps_lemp_keep_mask: d_keep[q, i] = 1 if item i passes its bucket's predicate for query q at d_theta[q].
 * fps_lemp_topk_filtered: the reference's whole retrieval loop, one warp per query, scalar fp32 dot
 *   products for the survivors - the parity path of the tensor-core kernels above, which apply only the
 *   bucket-level LENGTH bound.  d_part_scores / d_part_ids: [n_q, k] scratch; d_stats (optional, 4 x
 *   uint64, accumulated): candidates seen, candidates pruned, dot products, buckets skipped.          */
int fps_lemp_keep_mask(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                       int32_t item_stride, const float* d_item_len, int64_t n_items, int32_t dim,
                       const float* d_theta, int32_t bucket_size, int32_t strategy, int32_t n_focus,
                       float switch_threshold, uint8_t* d_keep, void* stream);
int fps_lemp_topk_filtered(const float* d_queries, int32_t q_stride, int64_t n_q, const float* d_items,
                           int32_t item_stride, const float* d_item_len, const int32_t* d_item_ids,
                           int64_t n_items, int32_t dim, int32_t k, int32_t bucket_size, int32_t strategy,
                           int32_t n_focus, float switch_threshold, float* d_part_scores,
                           int32_t* d_part_ids, float* d_out_scores, int32_t* d_out_ids, uint64_t* d_stats,
                           void* stream);
/* Microbenchmark: cycles per tcgen05.mma kind::tf32 instruction (M128, K8) for a long chain issued by one
 * thread per CTA on n_ctas CTAs.  mode 0 / 1: A from TMEM, one / two accumulators (N 128); 2 / 3: A from
 * shared memory, one / two accumulators; 4: A from TMEM, N 256; 5: A from TMEM, four accumulators of N 64.
 * d_cycles_per_mma: n_ctas floats.  (scripts/umma_probe.py, profiles/r2_topk/README.md)                 */
int fps_umma_probe(int32_t mode, int32_t n_mma, int32_t n_ctas, float* d_cycles_per_mma, void* stream);
/* Merge n_parts partial top-k lists per query (X8: CollectTopKFromEachWorker.scala:40-48) with an
 * optional per-query seen-set filter (d_seen: [n_q, seen_stride] item ids, -1 padded).            */
int fps_topk_merge(const float* d_part_scores, const int32_t* d_part_ids, int32_t n_parts,
                   int64_t n_q, int32_t k_in, int32_t k_out, const int32_t* d_seen,
                   int32_t seen_stride, float* d_out_scores, int32_t* d_out_ids, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FPS_B200_H */
