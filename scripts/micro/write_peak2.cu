// Micro-benchmark 2: store strategies for a write-only stream on B200 (sm_100a).
//   A  st.global.v8.f32 (256-bit per thread), warp-contiguous 1 KB per instruction
//   B  TMA 1-D bulk store (cp.async.bulk.global.shared::cta) of tiles staged in shared memory, per warp
//   C  persistent grid-stride variant of the 128-bit store
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/wp2 scripts/micro/write_peak2.cu && /tmp/wp2
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void st_v8(float* p, float4 a, float4 b) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z),
               "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
               : "memory");
}

__global__ void __launch_bounds__(256) fill_v8(float* __restrict__ out, int64_t n8, int per_warp8) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * per_warp8;
  const float4 v = make_float4(1.0f, 0.0f, (float)(lane & 1), 0.0f);
  for (int i = lane; i < per_warp8; i += 32) {
    const int64_t k = base + i;
    if (k < n8) st_v8(out + k * 8, v, v);
  }
}

__global__ void __launch_bounds__(256) fill_persist(float4* __restrict__ out, int64_t n4) {
  const float4 v = make_float4(1.0f, 0.0f, (float)(threadIdx.x & 1), 0.0f);
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += (int64_t)gridDim.x * blockDim.x)
    __stcs(out + k, v);
}

// per-warp TMA bulk stores: each warp owns a contiguous run of `per_warp_bytes`; tiles of TILE bytes are
// written into a double-buffered shared-memory stage with STS.128 and pushed with one bulk copy each
template <int TILE, bool REFILL>
__global__ void __launch_bounds__(256) fill_bulk(char* __restrict__ out, int64_t bytes, int64_t per_warp_bytes) {
  extern __shared__ __align__(128) char smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  char* stage = smem + wib * 2 * TILE;
  const int64_t base = warp * per_warp_bytes;
  if (base >= bytes) return;
  int64_t len = per_warp_bytes;
  if (base + len > bytes) len = bytes - base;
  const float4 v = make_float4(1.0f, 0.0f, (float)(lane & 1), 0.0f);
  int buf = 0;
  if (!REFILL) {
    for (int o = lane * 16; o < 2 * TILE; o += 512) *reinterpret_cast<float4*>(stage + o) = v;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
  }
  for (int64_t o = 0; o < len; o += TILE) {
    const int sz = (int)((len - o) < TILE ? (len - o) : TILE);
    char* s = stage + buf * TILE;
    if (REFILL) {
      // the buffer used two tiles ago must have been read by the TMA engine
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncwarp();
      for (int q = lane * 16; q < sz; q += 512) *reinterpret_cast<float4*>(s + q) = v;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }
    if (lane == 0) {
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(s);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + base + o), "r"(sa),
                   "r"(sz)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    buf ^= 1;
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
float time_it(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e9f;
  for (int it = 0; it < 12; ++it) {
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (it >= 2 && ms < best) best = ms;
  }
  return best;
}

int main() {
  const int64_t bytes = 2000000000LL / 4096 * 4096;
  char* buf;
  cudaMalloc(&buf, bytes);
  const double gb = (double)bytes / 1e6;
  for (int pw8 : {906, 3626, 32768}) {
    const int64_t n8 = bytes / 32;
    const int64_t warps = (n8 + pw8 - 1) / pw8;
    const unsigned grid = (unsigned)((warps * 32 + 255) / 256);
    const float ms = time_it([&] { fill_v8<<<grid, 256>>>((float*)buf, n8, pw8); });
    printf("A v8 per_warp %6d x32B grid %7u: %.3f ms %.1f GB/s\n", pw8, grid, ms, gb / ms);
  }
  for (int mult : {4, 8, 16, 32}) {
    const unsigned grid = 148 * mult;
    const float ms = time_it([&] { fill_persist<<<grid, 256>>>((float4*)buf, bytes / 16); });
    printf("C persistent grid %5u: %.3f ms %.1f GB/s\n", grid, ms, gb / ms);
  }
  for (int64_t pwb : {29008LL, 116032LL, 1048576LL}) {
    const int64_t warps = (bytes + pwb - 1) / pwb;
    const unsigned grid = (unsigned)((warps + 7) / 8);
#define RUN_BULK(TILE, REFILL)                                                                              \
  {                                                                                                         \
    cudaFuncSetAttribute(fill_bulk<TILE, REFILL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * TILE); \
    const float ms = time_it([&] { fill_bulk<TILE, REFILL><<<grid, 256, 8 * 2 * TILE>>>(buf, bytes, pwb); }); \
    printf("B bulk tile %5d refill %d per_warp %8lld grid %7u: %.3f ms %.1f GB/s\n", TILE, (int)REFILL,       \
           (long long)pwb, grid, ms, gb / ms);                                                               \
  }
    RUN_BULK(2048, false)
    RUN_BULK(4096, false)
    RUN_BULK(8192, false)
    RUN_BULK(2048, true)
    RUN_BULK(4096, true)
    RUN_BULK(8192, true)
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
