// Microbenchmark: tcgen05.ld / tcgen05.st bandwidth per SM (TMEM <-> registers), 32x32b shapes.
// One CTA per SM; W warps (4, 8, 16) each read its lane quarter's 128 columns per iteration.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../video_style_transfer_b200/csrc/common.cuh"

template <int MODE>   // 0: ld x32 (4 per iter, one wait), 1: ld x32 with a wait after each, 2: st x32 (4 per iter)
__global__ void k(int iters, long long* out, uint32_t* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = slot + (uint32_t((warp & 3) * 32) << 16) + ((warp >> 2) & 3) * 128;
  uint32_t r[32], acc = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = lane + i;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 2) {
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_st32(base + c * 32, r);
      tmem_st_wait();
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(base + c * 32, v);
        if (MODE == 1) tmem_ld_wait();
        if (MODE == 0 && c == 3) tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= v[i];
      }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + r[0];
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

void vst_set_error(const char*, ...) {}
int main() {
  long long* d; uint32_t* sink;
  cudaMalloc(&d, 8); cudaMalloc(&sink, 148 * 1024 * 4);
  const int iters = 2000;
  for (int mode = 0; mode < 3; ++mode)
    for (int w = 4; w <= 16; w *= 2) {
      if (mode == 0) k<0><<<148, 32 * w>>>(iters, d, sink);
      if (mode == 1) k<1><<<148, 32 * w>>>(iters, d, sink);
      if (mode == 2) k<2><<<148, 32 * w>>>(iters, d, sink);
      cudaError_t e = cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      const double bytes = double(iters) * w * 4 * 32 * 32 * 4;
      printf("mode %d (%s) warps %2d: %.1f B/clk/SM  (%.0f clk per 128x128 fp32 tile)  %s\n", mode,
             mode == 0 ? "ld x32 x4, one wait" : mode == 1 ? "ld x32, wait each" : "st x32 x4", w, bytes / h,
             65536.0 * h / bytes, cudaGetErrorString(e));
    }
  return 0;
}
