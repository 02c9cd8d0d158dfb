// Microbenchmark: MUFU.EX2 throughput per SM sub-partition, alone and inside the softmax instruction mix.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }

__device__ __forceinline__ uint32_t ex2_h2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
template <int MODE>
__global__ void k(int iters, float seed, long long* out, float* sink) {
  float x[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) x[i] = seed * (i + threadIdx.x);
  float2 l2 = make_float2(0.f, 0.f);
  uint32_t acc = 0;
  const float2 sc2 = make_float2(seed, seed), nm2 = make_float2(-seed, -seed);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 3) {
      // f16x2 softmax mix per pair: FFMA2, cvt.f16x2, MUFU.f16x2, HADD2 (+ FMNMX3)
      float mx = -1e30f;
      uint32_t hs = 0;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        mx = fmaxf(mx, fmaxf(x[2 * i], x[2 * i + 1]));
        const float2 sv = __ffma2_rn(make_float2(x[2 * i], x[2 * i + 1]), sc2, nm2);
        const uint32_t p = ex2_h2(pack_h2(sv.x, sv.y));
        asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(hs) : "r"(p));
        acc ^= p;
        x[2 * i] = __uint_as_float(p) * 1e-30f; x[2 * i + 1] = sv.y;
      }
      l2.x += mx + __uint_as_float(hs);
    } else if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 32; ++i) x[i] = ex2(x[i]);
    } else {
      // softmax mix per pair: FFMA2, 2 MUFU, FADD2, F2FP (+ FMNMX3 in mode 2)
      float mx = -1e30f;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        if (MODE == 2) mx = fmaxf(mx, fmaxf(x[2 * i], x[2 * i + 1]));
        const float2 sv = __ffma2_rn(make_float2(x[2 * i], x[2 * i + 1]), sc2, nm2);
        const float2 pv = make_float2(ex2(sv.x), ex2(sv.y));
        l2 = __fadd2_rn(l2, pv);
        acc ^= pack_bf16(pv.x, pv.y);
        x[2 * i] = pv.x; x[2 * i + 1] = pv.y;
      }
      if (MODE == 2) l2.x += mx;
    }
  }
  long long t1 = clock64();
  float s = l2.x + l2.y + __uint_as_float(acc);
#pragma unroll
  for (int i = 0; i < 32; ++i) s += x[i];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
}
int main() {
  long long* d; float* sink; cudaMalloc(&d, 8); cudaMalloc(&sink, 148 * 1024 * 4);
  const int iters = 4000;
  for (int mode = 0; mode < 4; ++mode)
    for (int w = 4; w <= 32; w *= 2) {
      if (mode == 0) k<0><<<148, 32 * w>>>(iters, 0.001f, d, sink);
      if (mode == 1) k<1><<<148, 32 * w>>>(iters, 0.001f, d, sink);
      if (mode == 2) k<2><<<148, 32 * w>>>(iters, 0.001f, d, sink);
      if (mode == 3) k<3><<<148, 32 * w>>>(iters, 0.001f, d, sink);
      cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      const double mufu_per_smsp = double(iters) * (mode == 3 ? 16 : 32) * (w / 4);   // warp-level MUFU instr per SMSP
      printf("mode %d warps/SMSP %d: %.2f clk per warp-MUFU per SMSP  (%.2f ex2/clk/SM)\n", mode, w / 4, h / mufu_per_smsp,
             32.0 * 4 * mufu_per_smsp / h);
    }
  return 0;
}
