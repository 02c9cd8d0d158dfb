// Microbenchmark: issue rate of tcgen05.mma shapes used by the attention kernels (sm_100a).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I. scripts/ubench/mma_rate.cu -o gpurun_out/mma_rate -lcuda
#include "../../video_style_transfer_b200/csrc/common.cuh"
#include <stdarg.h>
void vst_set_error(const char* fmt, ...) {}

template <bool kUniform>
__global__ void __launch_bounds__(128) mma_rate(int mode, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_proxy_async_smem();
  if (threadIdx.x < 32) tmem_alloc(&slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (kUniform ? (threadIdx.x < 32) : (threadIdx.x == 0)) {
    const bool issuer = kUniform ? elect_one() : true;
    const uint64_t da = umma_desc_sw128(smem_u32(smem), 1024);
    const uint64_t db = umma_desc_sw128(smem_u32(smem + 16384), 1024);
    const uint64_t dbm = umma_desc_sw128(smem_u32(smem + 32768), 1024, 8192);
    constexpr uint32_t i128 = umma_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t i64m = umma_idesc_bf16(128, 64, 0, 1);
    constexpr uint32_t i256 = umma_idesc_bf16(128, 256, 0, 0);
    constexpr uint32_t i64k = umma_idesc_bf16(128, 64, 0, 0);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (mode == 0) {        // S-like: SS 128x128x16 x4
        for (int k = 0; k < 4; ++k) if (issuer) umma_bf16_ss(tm, da + 2 * k, db + 2 * k, i128, k != 0);
      } else if (mode == 1) { // PV-like: TS 128x64x16 x8, B MN-major
        for (int k = 0; k < 8; ++k) if (issuer) umma_bf16_ts(tm + 128, tm + 192 + 8 * k, dbm + uint64_t((k * 2048) >> 4), i64m, 1);
      } else if (mode == 2) { // SS 128x64x16 x8, B MN-major (P from smem)
        for (int k = 0; k < 8; ++k) if (issuer) umma_bf16_ss(tm + 128, da + 2 * (k & 3), dbm + uint64_t((k * 2048) >> 4), i64m, 1);
      } else if (mode == 3) { // SS 128x256x16 x4 (N=256: B 256 rows)
        for (int k = 0; k < 4; ++k) if (issuer) umma_bf16_ss(tm, da + 2 * k, db + 2 * k, i256, k != 0);
      } else if (mode == 4) { // TS 128x128x16 x4, B K-major (Q in TMEM)
        for (int k = 0; k < 4; ++k) if (issuer) umma_bf16_ts(tm, tm + 192 + 8 * k, db + 2 * k, i128, k != 0);
      } else if (mode == 5) { // kernel mix
        for (int k = 0; k < 4; ++k) if (issuer) umma_bf16_ss(tm, da + 2 * k, db + 2 * k, i128, k != 0);
        for (int k = 0; k < 8; ++k) if (issuer) umma_bf16_ts(tm + 128, tm + 192 + 8 * k, dbm + uint64_t((k * 2048) >> 4), i64m, 1);
      } else if (mode == 6) { // TS 128x64x16 x8, B K-major
        for (int k = 0; k < 8; ++k) if (issuer) umma_bf16_ts(tm + 128, tm + 192 + 8 * k, db + 2 * (k & 3), i64k, 1);
      }
    }
    if (issuer) umma_commit(&bar);
    mbar_wait(&bar, 0, 1);
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 256); }
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(mma_rate<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66 * 1024);
  cudaFuncSetAttribute(mma_rate<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66 * 1024);
  const int iters = 2000;
  const char* names[] = {"SS 128x128x16 x4 (S)", "TS 128x64x16 x8 B MN-major (PV)", "SS 128x64x16 x8 B MN-major", "SS 128x256x16 x4",
                         "TS 128x128x16 x4", "mix S+PV (12 instr)", "TS 128x64x16 x8 B K-major"};
  const int ninstr[] = {4, 8, 8, 4, 4, 12, 8};
  for (int uni = 0; uni < 2; ++uni)
  for (int ncta = 1; ncta <= 2; ++ncta)
    for (int mode = 0; mode < 7; ++mode) {
      if (uni) mma_rate<true><<<148 * ncta, 128, 66 * 1024>>>(mode, iters, d);
      else mma_rate<false><<<148 * ncta, 128, 66 * 1024>>>(mode, iters, d);
      cudaError_t e = cudaDeviceSynchronize();
      long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      printf("uniform %d ctas/SM %d  %-36s  %8.1f clk/group  %7.1f clk/instr  (%s)\n", uni, ncta, names[mode], double(h) / iters,
             double(h) / iters / ninstr[mode], cudaGetErrorString(e));
    }
  return 0;
}
