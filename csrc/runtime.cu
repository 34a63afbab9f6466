// Runtime plumbing: error strings, IPC-exportable allocations, peer mapping, mapped host memory,
// and the P2P bandwidth probe.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void fps_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fps_num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

extern "C" {

const char* fps_last_error(void) { return g_err; }
int fps_version(void) { return 100; }

int fps_malloc(void** d_ptr, size_t bytes) {
  FPS_REQUIRE(d_ptr != nullptr, FPS_E_INVALID, "fps_malloc: null out pointer");
  FPS_CUDA(cudaMalloc(d_ptr, bytes));
  return 0;
}

int fps_free(void* d_ptr) {
  FPS_CUDA(cudaFree(d_ptr));
  return 0;
}

int fps_memset(void* d_ptr, int value, size_t bytes, void* stream) {
  FPS_CUDA(cudaMemsetAsync(d_ptr, value, bytes, (cudaStream_t)stream));
  return 0;
}

int fps_ipc_get_handle(void* d_ptr, unsigned char handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  cudaIpcMemHandle_t h;
  FPS_CUDA(cudaIpcGetMemHandle(&h, d_ptr));
  memcpy(handle, &h, 64);
  return 0;
}

int fps_ipc_open_handle(const unsigned char handle[64], void** d_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  FPS_CUDA(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int fps_ipc_close_handle(void* d_ptr) {
  FPS_CUDA(cudaIpcCloseMemHandle(d_ptr));
  return 0;
}

int fps_enable_peer_access(int peer_device) {
  int dev = 0;
  FPS_CUDA(cudaGetDevice(&dev));
  if (dev == peer_device) return 0;
  int can = 0;
  FPS_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  FPS_REQUIRE(can, FPS_E_UNSUPPORTED, "device %d cannot access peer %d", dev, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  FPS_CUDA(e);
  return 0;
}

int fps_host_alloc_mapped(void** h_ptr, void** d_ptr, size_t bytes) {
  FPS_CUDA(cudaHostAlloc(h_ptr, bytes, cudaHostAllocMapped | cudaHostAllocPortable));
  FPS_CUDA(cudaHostGetDevicePointer(d_ptr, *h_ptr, 0));
  return 0;
}

int fps_host_free(void* h_ptr) {
  FPS_CUDA(cudaFreeHost(h_ptr));
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// P2P probe kernels: the three access types of the PS data path.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) probe_load(const float4* __restrict__ src, float4* sink, size_t n4) {
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const size_t step = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * step < n4; i += 4 * step) {
    const float4 a = ld_row4((const float*)(src + i));
    const float4 b = ld_row4((const float*)(src + i + step));
    const float4 c = ld_row4((const float*)(src + i + 2 * step));
    const float4 d = ld_row4((const float*)(src + i + 3 * step));
    acc.x += a.x + b.x + c.x + d.x;
    acc.y += a.y + b.y + c.y + d.y;
    acc.z += a.z + b.z + c.z + d.z;
    acc.w += a.w + b.w + c.w + d.w;
  }
  for (; i < n4; i += step) {
    const float4 a = ld_row4((const float*)(src + i));
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
  }
  if (acc.x + acc.y + acc.z + acc.w == 123.456f) sink[0] = acc;  // keep the loads alive
}

__global__ void __launch_bounds__(256) probe_store(float4* dst, size_t n4) {
  const size_t step = (size_t)gridDim.x * blockDim.x;
  const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += step)
    st_row4((float*)(dst + i), v);
}

__global__ void __launch_bounds__(256) probe_red(float4* dst, size_t n4) {
  const size_t step = (size_t)gridDim.x * blockDim.x;
  const float4 v = make_float4(1.f, 1.f, 1.f, 1.f);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += step)
    red_add4((float*)(dst + i), v);
}

extern "C" int fps_p2p_probe(float* d_dst, const float* d_src, size_t bytes, int mode, int iters,
                             float* ms, void* stream) {
  FPS_REQUIRE(bytes % 16 == 0 && iters > 0 && ms != nullptr, FPS_E_INVALID, "fps_p2p_probe: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n4 = bytes / 16;
  const int grid = fps_num_sms() * 8;
  cudaEvent_t e0, e1;
  FPS_CUDA(cudaEventCreate(&e0));
  FPS_CUDA(cudaEventCreate(&e1));
  for (int it = -1; it < iters; ++it) {   // one untimed warm-up
    if (it == 0) FPS_CUDA(cudaEventRecord(e0, st));
    if (mode == 0) probe_load<<<grid, 256, 0, st>>>((const float4*)d_src, (float4*)d_dst, n4);
    else if (mode == 1) probe_store<<<grid, 256, 0, st>>>((float4*)d_dst, n4);
    else probe_red<<<grid, 256, 0, st>>>((float4*)d_dst, n4);
  }
  FPS_LAUNCH_CHECK();
  FPS_CUDA(cudaEventRecord(e1, st));
  FPS_CUDA(cudaEventSynchronize(e1));
  FPS_CUDA(cudaEventElapsedTime(ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Random-row probe: the PS access pattern (one warp per 512-byte-ish row at a random slot inside
// `span_rows`), to separate link bandwidth from address-translation / request-size effects.
//   mode 0 = pull only (row load), 1 = push only (red.add.v4), 2 = pull then push on the same row
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

template <int UNROLL>
__global__ void __launch_bounds__(256)
probe_random_rows(float* table, long long span_rows, int row_floats, long long n_rows, int mode,
                  unsigned long long seed, float* sink) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int nvec = row_floats >> 2;
  float acc = 0.f;
  for (long long r0 = warp * UNROLL; r0 < n_rows; r0 += n_warps * UNROLL) {
    float4 v[UNROLL];
    float* row[UNROLL];
#pragma unroll
    for (int q = 0; q < UNROLL; ++q) {
      const long long slot = (long long)(mix64(seed + (unsigned long long)(r0 + q)) % (unsigned long long)span_rows);
      row[q] = table + slot * row_floats;
      v[q] = make_float4(1.f, 1.f, 1.f, 1.f);
      if (mode != 1 && lane < nvec) v[q] = ld_row4(row[q] + lane * 4);
    }
#pragma unroll
    for (int q = 0; q < UNROLL; ++q) {
      acc += v[q].x;
      if (mode != 0 && lane < nvec) red_add4(row[q] + lane * 4, make_float4(1e-9f * v[q].x, 0.f, 0.f, 0.f));
    }
  }
  if (acc == 123.456f) sink[0] = acc;
}

extern "C" int fps_p2p_probe_random(float* d_table, int64_t span_rows, int row_floats, int64_t n_rows,
                                    int mode, int unroll, int blocks_per_sm, int iters, float* ms, void* stream) {
  FPS_REQUIRE(row_floats % 4 == 0 && row_floats <= 128 && span_rows > 0 && iters > 0, FPS_E_INVALID,
              "fps_p2p_probe_random: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = fps_num_sms() * (blocks_per_sm > 0 ? blocks_per_sm : 8);
  float* sink = nullptr;
  FPS_CUDA(cudaMalloc(&sink, 256));
  cudaEvent_t e0, e1;
  FPS_CUDA(cudaEventCreate(&e0));
  FPS_CUDA(cudaEventCreate(&e1));
  for (int it = -1; it < iters; ++it) {
    if (it == 0) FPS_CUDA(cudaEventRecord(e0, st));
    const unsigned long long seed = 0x1234567ull * (unsigned long long)(it + 2);
    if (unroll >= 4) probe_random_rows<4><<<grid, 256, 0, st>>>(d_table, span_rows, row_floats, n_rows, mode, seed, sink);
    else if (unroll >= 2) probe_random_rows<2><<<grid, 256, 0, st>>>(d_table, span_rows, row_floats, n_rows, mode, seed, sink);
    else probe_random_rows<1><<<grid, 256, 0, st>>>(d_table, span_rows, row_floats, n_rows, mode, seed, sink);
  }
  FPS_LAUNCH_CHECK();
  FPS_CUDA(cudaEventRecord(e1, st));
  FPS_CUDA(cudaEventSynchronize(e1));
  FPS_CUDA(cudaEventElapsedTime(ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(sink);
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Mixed probe: per row, any combination of {remote load, remote red, local load, local store,
// local red} selected by `flags`, to find which mix of apertures inside one warp is slow.
//   bit0 remote load | bit1 remote red.add | bit2 local load | bit3 local store | bit4 local red
//   bit5: use weak (non-.cg) local stores / loads    bit6: remote red issued BEFORE the local store
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
probe_mixed_rows(float* remote, float* local, long long span_rows, long long n_rows, int flags,
                 unsigned long long seed, float* sink) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  float acc = 0.f;
  for (long long r = warp; r < n_rows; r += n_warps) {
    const unsigned long long h = mix64(seed + (unsigned long long)r);
    float* rr = remote + (long long)(h % (unsigned long long)span_rows) * 128 + lane * 4;
    float* lr = local + (long long)((h >> 20) % (unsigned long long)span_rows) * 128 + lane * 4;
    float4 a = make_float4(1.f, 1.f, 1.f, 1.f), b = a;
    if (flags & 1) a = ld_row4(rr);
    if (flags & 4) b = (flags & 32) ? *reinterpret_cast<const float4*>(lr) : ld_row4(lr);
    const float4 d = make_float4(1e-6f * (a.x + b.x), 1e-6f, 0.f, 0.f);
    acc += a.x + b.x;
    if (flags & 128) acc += a.y + a.z + a.w + b.y + b.z + b.w;   // keeps the loads 128-bit wide
    if ((flags & 2) && (flags & 64)) red_add4(rr, d);
    if (flags & 8) {
      if (flags & 32) *reinterpret_cast<float4*>(lr) = d; else st_row4(lr, d);
    }
    if (flags & 16) red_add4(lr, d);
    if ((flags & 2) && !(flags & 64) && !(flags & 256)) red_add4(rr, d);
    if ((flags & 2) && (flags & 256)) {
      // push as ONE bulk reduction of the whole 512-byte row from shared memory (TMA reduce) instead of
      // 32 x 16-byte red.add: does the wire carry it better?
      __shared__ __align__(128) float s_row[8][2][128];
      const int w = threadIdx.x >> 5;
      const int slot = (int)((r / n_warps) & 1);
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the copy that used this slot two rows ago
      __syncwarp();
      *reinterpret_cast<float4*>(&s_row[w][slot][lane * 4]) = d;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 512;"
                     ::"l"(rr - lane * 4), "r"((uint32_t)__cvta_generic_to_shared(&s_row[w][slot][0])) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }
  if (flags & 256) {
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
  }
  if (acc == 123.456f) sink[0] = acc;
}

extern "C" int fps_p2p_probe_mixed(float* d_remote, float* d_local, int64_t span_rows, int64_t n_rows,
                                   int flags, int iters, float* ms, void* stream) {
  FPS_REQUIRE(span_rows > 0 && iters > 0 && ms != nullptr, FPS_E_INVALID, "fps_p2p_probe_mixed: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = fps_num_sms() * 8;
  float* sink = nullptr;
  FPS_CUDA(cudaMalloc(&sink, 256));
  cudaEvent_t e0, e1;
  FPS_CUDA(cudaEventCreate(&e0));
  FPS_CUDA(cudaEventCreate(&e1));
  for (int it = -1; it < iters; ++it) {
    if (it == 0) FPS_CUDA(cudaEventRecord(e0, st));
    probe_mixed_rows<<<grid, 256, 0, st>>>(d_remote, d_local, span_rows, n_rows, flags,
                                           0x9876543ull * (unsigned long long)(it + 2), sink);
  }
  FPS_LAUNCH_CHECK();
  FPS_CUDA(cudaEventRecord(e1, st));
  FPS_CUDA(cudaEventSynchronize(e1));
  FPS_CUDA(cudaEventElapsedTime(ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(sink);
  return 0;
}
