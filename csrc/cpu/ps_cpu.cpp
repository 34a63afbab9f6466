// BL2 - multi-threaded C++ restatement of the reference's parameter-server loop (CPU baseline
// and second oracle).  "CPU restatement, not Flink": the Flink job cannot be built or run in this
// environment (no JVM / sbt / jars), so this stands in for it - same roles and message flow as
// M/FlinkParameterServer.scala:195-336, without JVM, serialisation or network costs (i.e. an
// optimistic bound for the reference):
//   * W worker threads and S server threads; one FIFO ring per (worker, shard) and per (shard, worker);
//   * a worker sends Pull(item), the owning shard answers with the row (lazy init on first pull),
//     the worker runs the SGD step (SGDUpdater.scala:5-11), updates its resident user row and
//     sends Push(item, delta), the shard adds it (SimplePSLogic.scala:16-25);
//   * unanswered pulls per worker never exceed pull_limit (WorkerLogic.scala:176-225);
//   * fp64 like the reference.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct alignas(64) Ring {
  std::atomic<uint64_t> head{0};
  char pad0[56];
  std::atomic<uint64_t> tail{0};
  char pad1[56];
  uint64_t cap = 0, slot = 0;
  char* buf = nullptr;
  void init(uint64_t capacity_pow2, uint64_t slot_bytes) {
    cap = capacity_pow2;
    slot = slot_bytes;
    buf = (char*)aligned_alloc(64, cap * slot);
  }
  ~Ring() { free(buf); }
  char* try_reserve() {   // producer
    const uint64_t t = tail.load(std::memory_order_relaxed);
    if (t - head.load(std::memory_order_acquire) >= cap) return nullptr;
    return buf + (t & (cap - 1)) * slot;
  }
  void commit() { tail.store(tail.load(std::memory_order_relaxed) + 1, std::memory_order_release); }
  char* try_front() {     // consumer
    const uint64_t h = head.load(std::memory_order_relaxed);
    if (h == tail.load(std::memory_order_acquire)) return nullptr;
    return buf + (h & (cap - 1)) * slot;
  }
  void pop() { head.store(head.load(std::memory_order_relaxed) + 1, std::memory_order_release); }
};

struct Msg {
  int32_t type;   // 0 pull, 1 push / answer
  int32_t id;
  double row[1];
};

inline uint64_t splitmix(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

uint64_t pow2_at_least(uint64_t x) {
  uint64_t p = 1;
  while (p < x) p <<= 1;
  return p;
}

}  // namespace

extern "C" {

// Runs W workers over n ratings each (arrays are [W][n] flattened; users are worker-local slots,
// items are global ids) against S shards.  Returns elapsed seconds; *out_sse receives the summed
// squared error.  user_table [W][n_users_per_worker][dim], item_table [S][items_per_shard][dim]
// are caller-allocated fp64 (so tests can inspect the result); item rows are owned by shard
// (item % S) at slot (item / S).
double fps_cpu_mf_run(int W, int S, int dim, int64_t n_per_worker, const int32_t* users, const int32_t* items,
                      const float* ratings, double* user_table, int64_t users_per_worker, double* item_table,
                      int64_t items_per_shard, double lr, int pull_limit, double* out_sse) {
  const uint64_t slot = (sizeof(Msg) + (size_t)(dim - 1) * sizeof(double) + 63) & ~(size_t)63;
  const uint64_t cap = pow2_at_least((uint64_t)2 * pull_limit + 4);
  std::vector<Ring> req((size_t)W * S), ans((size_t)S * W);
  for (auto& r : req) r.init(cap, slot);
  for (auto& r : ans) r.init(cap, slot);
  std::atomic<int> workers_done{0};
  std::vector<double> sse(W, 0.0);
  auto t0 = std::chrono::steady_clock::now();

  auto shard_fn = [&](int s) {
    double* table = item_table + (size_t)s * items_per_shard * dim;
    while (true) {
      bool any = false;
      for (int w = 0; w < W; ++w) {
        Ring& in = req[(size_t)w * S + s];
        for (int burst = 0; burst < 64; ++burst) {
          char* p = in.try_front();
          if (!p) break;
          Msg* m = (Msg*)p;
          double* row = table + (size_t)(m->id / S) * dim;
          if (m->type == 0) {
            Ring& out = ans[(size_t)s * W + w];
            char* q;
            while (!(q = out.try_reserve())) std::this_thread::yield();
            Msg* a = (Msg*)q;
            a->type = 1;
            a->id = m->id;
            memcpy(a->row, row, sizeof(double) * dim);
            out.commit();
          } else {
            for (int j = 0; j < dim; ++j) row[j] += m->row[j];
          }
          in.pop();
          any = true;
        }
      }
      if (!any) {
        if (workers_done.load(std::memory_order_acquire) == W) {
          bool empty = true;
          for (int w = 0; w < W; ++w) empty = empty && req[(size_t)w * S + s].try_front() == nullptr;
          if (empty) break;
        }
        std::this_thread::yield();
      }
    }
  };

  auto worker_fn = [&](int w) {
    const int32_t* u = users + (size_t)w * n_per_worker;
    const int32_t* it = items + (size_t)w * n_per_worker;
    const float* rt = ratings + (size_t)w * n_per_worker;
    double* utab = user_table + (size_t)w * users_per_worker * dim;
    // per-shard FIFO of rating indices whose pull is outstanding (per-channel order == answer order)
    std::vector<std::vector<int64_t>> pend(S);
    std::vector<size_t> pend_head(S, 0);
    int64_t issued = 0, in_flight = 0;
    double local_sse = 0.0;
    std::vector<double> delta(dim);
    while (issued < n_per_worker || in_flight > 0) {
      for (int s = 0; s < S; ++s) {
        Ring& in = ans[(size_t)s * W + w];
        char* p;
        while ((p = in.try_front())) {
          Msg* a = (Msg*)p;
          const int64_t r = pend[s][pend_head[s]++];
          double* urow = utab + (size_t)u[r] * dim;
          double dot = 0.0;
          for (int j = 0; j < dim; ++j) dot += urow[j] * a->row[j];
          const double e = (double)rt[r] - dot;
          local_sse += e * e;
          for (int j = 0; j < dim; ++j) {
            delta[j] = lr * e * urow[j];
            urow[j] += lr * e * a->row[j];
          }
          const int id = a->id;
          in.pop();
          Ring& out = req[(size_t)w * S + s];
          char* q;
          while (!(q = out.try_reserve())) std::this_thread::yield();
          Msg* m = (Msg*)q;
          m->type = 1;
          m->id = id;
          memcpy(m->row, delta.data(), sizeof(double) * dim);
          out.commit();
          --in_flight;
        }
      }
      while (issued < n_per_worker && in_flight < pull_limit) {
        const int s = it[issued] % S;
        Ring& out = req[(size_t)w * S + s];
        char* q = out.try_reserve();
        if (!q) break;
        Msg* m = (Msg*)q;
        m->type = 0;
        m->id = it[issued];
        out.commit();
        pend[s].push_back(issued);
        ++issued;
        ++in_flight;
      }
    }
    sse[w] = local_sse;
    workers_done.fetch_add(1, std::memory_order_release);
  };

  std::vector<std::thread> th;
  for (int s = 0; s < S; ++s) th.emplace_back(shard_fn, s);
  for (int w = 0; w < W; ++w) th.emplace_back(worker_fn, w);
  for (auto& t : th) t.join();
  auto t1 = std::chrono::steady_clock::now();
  double tot = 0.0;
  for (double x : sse) tot += x;
  if (out_sse) *out_sse = tot;
  return std::chrono::duration<double>(t1 - t0).count();
}

// Binary passive-aggressive (PassiveAggressiveParameterServer.scala:283-338 with RangePSLogicWithClose):
// per example one Pull per active feature, the shard answers with the weight, the worker assembles the
// nnz answers (per-channel FIFO order), computes margin / tau (PA, PA-I, PA-II) and sends one Push per
// feature.  The limiter counts features, like the reference.  Arrays are [W] blocks: indptr
// [W][n_per_worker + 1] (offsets local to the worker's block of indices / values, which hold
// nnz_per_worker entries each), labels [W][n_per_worker] (+1 / -1).  weights [S][div] fp64, range
// partitioned: owner = id / div.  Returns elapsed seconds; *out_loss = summed hinge loss,
// *out_mistakes = examples whose prediction (with the pulled weights) had the wrong sign.
double fps_cpu_pa_binary_run(int W, int S, int64_t n_per_worker, int64_t nnz_per_worker, const int64_t* indptr,
                             const int32_t* indices, const float* values, const int32_t* labels, double* weights,
                             int64_t div, int variant, double C, int pull_limit, double* out_loss,
                             int64_t* out_mistakes) {
  const uint64_t slot = 64;
  const uint64_t cap = pow2_at_least((uint64_t)2 * pull_limit + 4);
  std::vector<Ring> req((size_t)W * S), ans((size_t)S * W);
  for (auto& r : req) r.init(cap, slot);
  for (auto& r : ans) r.init(cap, slot);
  std::atomic<int> workers_done{0};
  std::vector<double> loss(W, 0.0);
  std::vector<int64_t> wrong(W, 0);
  auto t0 = std::chrono::steady_clock::now();

  auto shard_fn = [&](int s) {
    double* table = weights + (size_t)s * div;
    while (true) {
      bool any = false;
      for (int w = 0; w < W; ++w) {
        Ring& in = req[(size_t)w * S + s];
        for (int burst = 0; burst < 64; ++burst) {
          char* p = in.try_front();
          if (!p) break;
          Msg* m = (Msg*)p;
          double* cell = table + ((int64_t)m->id - (int64_t)s * div);
          if (m->type == 0) {
            Ring& out = ans[(size_t)s * W + w];
            char* q;
            while (!(q = out.try_reserve())) std::this_thread::yield();
            Msg* a = (Msg*)q;
            a->type = 1;
            a->id = m->id;
            a->row[0] = *cell;
            out.commit();
          } else {
            *cell += m->row[0];
          }
          in.pop();
          any = true;
        }
      }
      if (!any) {
        if (workers_done.load(std::memory_order_acquire) == W) {
          bool empty = true;
          for (int w = 0; w < W; ++w) empty = empty && req[(size_t)w * S + s].try_front() == nullptr;
          if (empty) break;
        }
        std::this_thread::yield();
      }
    }
  };

  auto worker_fn = [&](int w) {
    const int64_t* ip = indptr + (size_t)w * (n_per_worker + 1);
    const int32_t* idx = indices + (size_t)w * nnz_per_worker;
    const float* val = values + (size_t)w * nnz_per_worker;
    const int32_t* lab = labels + (size_t)w * n_per_worker;
    std::vector<double> got((size_t)nnz_per_worker, 0.0);       // pulled weight per (example, feature) position
    std::vector<int32_t> missing((size_t)n_per_worker, 0);      // answers an example still waits for
    // per-shard FIFO of nnz positions whose pull is outstanding (channel order == answer order)
    std::vector<std::vector<int64_t>> pend(S);
    std::vector<size_t> pend_head(S, 0);
    std::vector<int64_t> owner_example((size_t)nnz_per_worker);
    int64_t next_ex = 0, next_pos = 0, in_flight = 0, done_ex = 0;
    double local_loss = 0.0;
    int64_t local_wrong = 0;
    while (done_ex < n_per_worker) {
      // ---- answers: assemble; a complete example computes and pushes
      for (int s = 0; s < S; ++s) {
        Ring& in = ans[(size_t)s * W + w];
        char* p;
        while ((p = in.try_front())) {
          Msg* a = (Msg*)p;
          const int64_t pos = pend[s][pend_head[s]++];
          got[pos] = a->row[0];
          in.pop();
          --in_flight;
          const int64_t e = owner_example[pos];
          if (--missing[e] == 0) {
            double m = 0.0, n2 = 0.0;
            for (int64_t k = ip[e]; k < ip[e + 1]; ++k) {
              m += (double)val[k] * got[k];
              n2 += (double)val[k] * (double)val[k];
            }
            const double y = lab[e] > 0 ? 1.0 : -1.0;
            if ((m > 0.0) != (y > 0.0)) ++local_wrong;
            const double l = std::max(0.0, 1.0 - y * m);
            local_loss += l;
            double tau = 0.0;
            if (n2 > 0.0) tau = variant == 0 ? l / n2 : variant == 1 ? std::min(C, l / n2) : l / (n2 + 1.0 / (2.0 * C));
            if (tau != 0.0) {
              for (int64_t k = ip[e]; k < ip[e + 1]; ++k) {
                const int so = (int)((int64_t)idx[k] / div);
                Ring& out = req[(size_t)w * S + so];
                char* q;
                while (!(q = out.try_reserve())) std::this_thread::yield();
                Msg* msg = (Msg*)q;
                msg->type = 1;
                msg->id = idx[k];
                msg->row[0] = tau * y * (double)val[k];
                out.commit();
              }
            }
            ++done_ex;
          }
        }
      }
      // ---- pulls, feature by feature, while credits last
      while (next_ex < n_per_worker && in_flight < pull_limit) {
        if (next_pos == ip[next_ex]) {
          missing[next_ex] = (int32_t)(ip[next_ex + 1] - ip[next_ex]);
          if (missing[next_ex] == 0) { ++done_ex; ++next_ex; continue; }
        }
        const int so = (int)((int64_t)idx[next_pos] / div);
        Ring& out = req[(size_t)w * S + so];
        char* q = out.try_reserve();
        if (!q) break;
        Msg* msg = (Msg*)q;
        msg->type = 0;
        msg->id = idx[next_pos];
        out.commit();
        pend[so].push_back(next_pos);
        owner_example[next_pos] = next_ex;
        ++in_flight;
        if (++next_pos == ip[next_ex + 1]) ++next_ex;
      }
    }
    loss[w] = local_loss;
    wrong[w] = local_wrong;
    workers_done.fetch_add(1, std::memory_order_release);
  };

  std::vector<std::thread> th;
  for (int s = 0; s < S; ++s) th.emplace_back(shard_fn, s);
  for (int w = 0; w < W; ++w) th.emplace_back(worker_fn, w);
  for (auto& t : th) t.join();
  auto t1 = std::chrono::steady_clock::now();
  double tl = 0.0;
  int64_t tw = 0;
  for (int w = 0; w < W; ++w) { tl += loss[w]; tw += wrong[w]; }
  if (out_loss) *out_loss = tl;
  if (out_mistakes) *out_mistakes = tw;
  return std::chrono::duration<double>(t1 - t0).count();
}

// Multiclass one-vs-all passive-aggressive (PassiveAggressiveOneVersusAll.scala:52-69 through the same
// worker / shard protocol): a parameter is the row of L class weights of one feature; Pull / PullAnswer /
// Push messages carry L doubles.  d = W^T x over the pulled rows, y_c = +1 for the label else -1,
// loss_c = max(0, 1 - d_c y_c), tau_c per variant, push of x_i * (tau o y) per active feature.
// labels are class indices.  weights [S][div][L].  *out_mistakes = examples whose argmax was wrong.
double fps_cpu_pa_ova_run(int W, int S, int L, int64_t n_per_worker, int64_t nnz_per_worker, const int64_t* indptr,
                          const int32_t* indices, const float* values, const int32_t* labels, double* weights,
                          int64_t div, int variant, double C, int pull_limit, double* out_loss,
                          int64_t* out_mistakes) {
  const uint64_t slot = (sizeof(Msg) + (size_t)(L - 1) * sizeof(double) + 63) & ~(size_t)63;
  const uint64_t cap = pow2_at_least((uint64_t)2 * pull_limit + 4);
  std::vector<Ring> req((size_t)W * S), ans((size_t)S * W);
  for (auto& r : req) r.init(cap, slot);
  for (auto& r : ans) r.init(cap, slot);
  std::atomic<int> workers_done{0};
  std::vector<double> loss(W, 0.0);
  std::vector<int64_t> wrong(W, 0);
  auto t0 = std::chrono::steady_clock::now();

  auto shard_fn = [&](int s) {
    double* table = weights + (size_t)s * div * L;
    while (true) {
      bool any = false;
      for (int w = 0; w < W; ++w) {
        Ring& in = req[(size_t)w * S + s];
        for (int burst = 0; burst < 64; ++burst) {
          char* p = in.try_front();
          if (!p) break;
          Msg* m = (Msg*)p;
          double* row = table + ((int64_t)m->id - (int64_t)s * div) * L;
          if (m->type == 0) {
            Ring& out = ans[(size_t)s * W + w];
            char* q;
            while (!(q = out.try_reserve())) std::this_thread::yield();
            Msg* a = (Msg*)q;
            a->type = 1;
            a->id = m->id;
            memcpy(a->row, row, (size_t)L * sizeof(double));
            out.commit();
          } else {
            for (int c = 0; c < L; ++c) row[c] += m->row[c];
          }
          in.pop();
          any = true;
        }
      }
      if (!any) {
        if (workers_done.load(std::memory_order_acquire) == W) {
          bool empty = true;
          for (int w = 0; w < W; ++w) empty = empty && req[(size_t)w * S + s].try_front() == nullptr;
          if (empty) break;
        }
        std::this_thread::yield();
      }
    }
  };

  auto worker_fn = [&](int w) {
    const int64_t* ip = indptr + (size_t)w * (n_per_worker + 1);
    const int32_t* idx = indices + (size_t)w * nnz_per_worker;
    const float* val = values + (size_t)w * nnz_per_worker;
    const int32_t* lab = labels + (size_t)w * n_per_worker;
    std::vector<double> dvec((size_t)n_per_worker * L, 0.0);   // decision vector, accumulated as the answers arrive
    std::vector<double> n2((size_t)n_per_worker, 0.0);
    std::vector<int32_t> missing((size_t)n_per_worker, 0);
    std::vector<std::vector<int64_t>> pend(S);
    std::vector<size_t> pend_head(S, 0);
    std::vector<int64_t> owner_example((size_t)nnz_per_worker);
    std::vector<double> mult(L);
    int64_t next_ex = 0, next_pos = 0, in_flight = 0, done_ex = 0;
    double local_loss = 0.0;
    int64_t local_wrong = 0;
    while (done_ex < n_per_worker) {
      for (int s = 0; s < S; ++s) {
        Ring& in = ans[(size_t)s * W + w];
        char* p;
        while ((p = in.try_front())) {
          Msg* a = (Msg*)p;
          const int64_t pos = pend[s][pend_head[s]++];
          const int64_t e = owner_example[pos];
          double* d = &dvec[(size_t)e * L];
          const double x = (double)val[pos];
          for (int c = 0; c < L; ++c) d[c] += x * a->row[c];
          n2[e] += x * x;
          in.pop();
          --in_flight;
          if (--missing[e] == 0) {
            int am = 0;
            for (int c = 1; c < L; ++c) if (d[c] > d[am]) am = c;
            if (am != lab[e]) ++local_wrong;
            bool any = false;
            for (int c = 0; c < L; ++c) {
              const double y = c == lab[e] ? 1.0 : -1.0;
              const double l = std::max(0.0, 1.0 - d[c] * y);
              local_loss += l;
              double tau = 0.0;
              if (n2[e] > 0.0) tau = variant == 0 ? l / n2[e] : variant == 1 ? std::min(C, l / n2[e]) : l / (n2[e] + 1.0 / (2.0 * C));
              mult[c] = tau * y;
              any = any || mult[c] != 0.0;
            }
            if (any) {
              for (int64_t k = ip[e]; k < ip[e + 1]; ++k) {
                const int so = (int)((int64_t)idx[k] / div);
                Ring& out = req[(size_t)w * S + so];
                char* q;
                while (!(q = out.try_reserve())) std::this_thread::yield();
                Msg* msg = (Msg*)q;
                msg->type = 1;
                msg->id = idx[k];
                for (int c = 0; c < L; ++c) msg->row[c] = mult[c] * (double)val[k];
                out.commit();
              }
            }
            ++done_ex;
          }
        }
      }
      while (next_ex < n_per_worker && in_flight < pull_limit) {
        if (next_pos == ip[next_ex]) {
          missing[next_ex] = (int32_t)(ip[next_ex + 1] - ip[next_ex]);
          if (missing[next_ex] == 0) { ++done_ex; ++next_ex; continue; }
        }
        const int so = (int)((int64_t)idx[next_pos] / div);
        Ring& out = req[(size_t)w * S + so];
        char* q = out.try_reserve();
        if (!q) break;
        Msg* msg = (Msg*)q;
        msg->type = 0;
        msg->id = idx[next_pos];
        out.commit();
        pend[so].push_back(next_pos);
        owner_example[next_pos] = next_ex;
        ++in_flight;
        if (++next_pos == ip[next_ex + 1]) ++next_ex;
      }
    }
    loss[w] = local_loss;
    wrong[w] = local_wrong;
    workers_done.fetch_add(1, std::memory_order_release);
  };

  std::vector<std::thread> th;
  for (int s = 0; s < S; ++s) th.emplace_back(shard_fn, s);
  for (int w = 0; w < W; ++w) th.emplace_back(worker_fn, w);
  for (auto& t : th) t.join();
  auto t1 = std::chrono::steady_clock::now();
  double tl = 0.0;
  int64_t tw = 0;
  for (int w = 0; w < W; ++w) { tl += loss[w]; tw += wrong[w]; }
  if (out_loss) *out_loss = tl;
  if (out_mistakes) *out_mistakes = tw;
  return std::chrono::duration<double>(t1 - t0).count();
}

// deterministic table fill in [lo, hi)
void fps_cpu_fill_uniform(double* p, int64_t n, double lo, double hi, uint64_t seed) {
  uint64_t s = seed;
  for (int64_t i = 0; i < n; ++i) p[i] = lo + (hi - lo) * ((splitmix(s) >> 11) * (1.0 / 9007199254740992.0));
}

}  // extern "C"
