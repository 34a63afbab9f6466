// Placeholder translation unit for the staged (TMA bulk) and pre-combining MF variants; filled in
// by the perf pass.  Until then both fall back to the direct kernel so the ABI is stable.
#include "common.cuh"

extern "C" int fps_mf_sgd_step(const FpsTableView*, const FpsTableView*, const int32_t*, const int32_t*,
                               const float*, int64_t, const FpsMfParams*, FpsFlow*, void*);

int fps_mf_sgd_staged(const FpsTableView* items, const FpsTableView* users, const int32_t* u, const int32_t* i,
                      const float* r, int64_t n, const FpsMfParams* p, FpsFlow* flow, cudaStream_t st) {
  FpsMfParams q = *p;
  q.variant = 0;
  return fps_mf_sgd_step(items, users, u, i, r, n, &q, flow, (void*)st);
}

int fps_mf_sgd_combine(const FpsTableView* items, const FpsTableView* users, const int32_t* u, const int32_t* i,
                       const float* r, int64_t n, const FpsMfParams* p, FpsFlow* flow, cudaStream_t st) {
  FpsMfParams q = *p;
  q.variant = 0;
  return fps_mf_sgd_step(items, users, u, i, r, n, &q, flow, (void*)st);
}
