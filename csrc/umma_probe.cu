// Microbenchmark of tcgen05.mma kind::tf32 issue shapes (why the top-K kernels sit at ~125 cycles per
// M128 x N128 x K8 instruction where the nominal rate is 64): one elected thread per CTA issues a long chain
// of MMAs on operands that stay in place, commits once, and the CTA reports cycles per instruction.
//   mode 0: A from TMEM (TS), every MMA accumulates into the SAME accumulator
//   mode 1: TS, MMAs alternate between two accumulators
//   mode 2: A from shared memory (SS), one accumulator      mode 3: SS, two accumulators
//   mode 4: TS, N = 256, one accumulator                    mode 5: TS, four accumulators of N = 64
//   mode 6: TS, N = 128, TWO issuing warps, each with its own accumulator and half of the instructions
#include "topk_core.cuh"

namespace {

__global__ void __launch_bounds__(128, 1)
k_umma_probe(int mode, int n_mma, float* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  float* sA = (float*)base;                         // [128 x 128 B] one K block, swizzled K-major
  float* sB = sA + 128 * 32;                        // [256 x 128 B]
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ __align__(8) uint64_t s_bar2;
  __shared__ long long s_cyc2;
  __shared__ uint32_t s_tmem;
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < (128 + 256) * 32; i += 128) sA[i] = to_tf32(1.0f + 0.001f * (float)((i * 2654435761u) >> 20));
  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (t == 0) {
    mbar_init(&s_bar, 1);
    mbar_init(&s_bar2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = s_tmem;
  const uint32_t a_tmem = tb + 384;                 // 32 columns of "A" (whatever TMEM holds)
  long long cycles = 0;
  if (mode == 6 && (t == 0 || t == 32)) {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t db = make_desc(smem_u32(sB));
    const uint32_t acc = tb + (t == 0 ? 0u : 128u);
    const long long t0 = clock64();
    for (int i = 0; i < n_mma / 2; ++i) {
      const uint32_t ks = (uint32_t)(i & 3);
      const uint32_t accumulate = i > 0 ? 1u : 0u;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
          ::"r"(acc), "r"(a_tmem + ks * 8), "l"(db + (uint64_t)(ks * 2)), "r"(idesc), "r"(accumulate) : "memory");
    }
    umma_commit(t == 0 ? &s_bar : &s_bar2);
    mbar_wait(t == 0 ? &s_bar : &s_bar2, 0);
    cycles = clock64() - t0;
    if (t == 32) s_cyc2 = cycles;
  } else if (mode != 6 && t == 0) {
    const int N = mode == 4 ? 256 : mode == 5 ? 64 : 128;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
    const bool ts = mode == 0 || mode == 1 || mode == 4 || mode == 5;
    const int n_acc = (mode == 1 || mode == 3) ? 2 : mode == 5 ? 4 : 1;
    const long long t0 = clock64();
    for (int i = 0; i < n_mma; ++i) {
      const uint32_t ks = (uint32_t)(i & 3);
      const uint32_t acc = tb + (uint32_t)(i % n_acc) * (uint32_t)N;
      const uint32_t accumulate = i >= n_acc ? 1u : 0u;
      if (ts) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
            ::"r"(acc), "r"(a_tmem + ks * 8), "l"(db + (uint64_t)(ks * 2)), "r"(idesc), "r"(accumulate) : "memory");
      } else {
        umma_tf32(acc, da + (uint64_t)(ks * 2), db + (uint64_t)(ks * 2), idesc, accumulate);
      }
    }
    umma_commit(&s_bar);
    mbar_wait(&s_bar, 0);
    cycles = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (t == 0) out[blockIdx.x] = (float)(mode == 6 ? max(cycles, s_cyc2) : cycles) / (float)n_mma;
  if (warp == 0) tmem_dealloc(tb, 512);
}

}  // namespace

extern "C" int fps_umma_probe(int32_t mode, int32_t n_mma, int32_t n_ctas, float* d_cycles_per_mma, void* stream) {
  FPS_REQUIRE(mode >= 0 && mode <= 6 && n_mma > 0 && n_ctas > 0 && d_cycles_per_mma, FPS_E_INVALID, "fps_umma_probe: bad arguments");
  const size_t smem = 1024 + (128 + 256) * 128;
  FPS_CUDA(cudaFuncSetAttribute(k_umma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_umma_probe<<<n_ctas, 128, smem, (cudaStream_t)stream>>>(mode, n_mma, d_cycles_per_mma);
  FPS_LAUNCH_CHECK();
  return 0;
}
