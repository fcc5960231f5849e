// Micro-benchmark: what does a pure streaming-store kernel reach on this GPU?  (the ceiling the perspective
// writer is measured against, next to MEASURED_PEAKS.json's copy bandwidth).  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/write_peak scripts/micro/write_peak.cu && gpurun_out/write_peak
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>   // 0: st.global.cs.v4 (evict-first)  1: plain st.global.v4   2: st.global.wt? (write-through)
__global__ void __launch_bounds__(256) fill_kernel(float4* __restrict__ out, int64_t n4, int per_warp) {
  // every warp owns a contiguous run of per_warp float4 (like one env's perspective run) and writes it in
  // 512-byte warp-wide store instructions
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * per_warp;
  const float4 v = make_float4(1.0f, 0.0f, (float)(lane & 1), 0.0f);
  for (int i = lane; i < per_warp; i += 32) {
    const int64_t k = base + i;
    if (k < n4) {
      if (MODE == 0) __stcs(out + k, v);
      else if (MODE == 1) out[k] = v;
      else __stwt(out + k, v);
    }
  }
}

int main() {
  const int64_t bytes = 2000000000LL;            // ~ one step of the headline workload
  const int64_t n4 = bytes / 16;
  float4* buf;
  cudaMalloc(&buf, n4 * 16);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int per_warps[] = {1813, 7252 / 4 * 4, 65536};   // 74 perspectives x 98 floats / 4 ; large runs
  for (int mode = 0; mode < 3; ++mode) {
    for (int pw : per_warps) {
      const int64_t warps = (n4 + pw - 1) / pw;
      const unsigned grid = (unsigned)((warps * 32 + 255) / 256);
      float best = 1e9f;
      for (int it = 0; it < 12; ++it) {
        cudaEventRecord(e0);
        if (mode == 0) fill_kernel<0><<<grid, 256>>>(buf, n4, pw);
        else if (mode == 1) fill_kernel<1><<<grid, 256>>>(buf, n4, pw);
        else fill_kernel<2><<<grid, 256>>>(buf, n4, pw);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (it >= 2 && ms < best) best = ms;
      }
      printf("mode %d per_warp_float4 %6d grid %8u : %.3f ms  %.1f GB/s\n", mode, pw, grid, best,
             (double)n4 * 16 / best / 1e6);
    }
  }
  {
    float best = 1e9f;
    for (int it = 0; it < 8; ++it) {
      cudaEventRecord(e0);
      cudaMemsetAsync(buf, 0, n4 * 16);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (it >= 2 && ms < best) best = ms;
    }
    printf("cudaMemset : %.3f ms  %.1f GB/s\n", best, (double)n4 * 16 / best / 1e6);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
