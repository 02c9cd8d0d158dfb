// How long does one mbarrier.try_wait poll suspend when the phase never completes?  (bounds the
// poll-count watchdog of csrc/common.cuh: VST_WATCHDOG_SPINS polls x this time)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o trywait_limit.bin trywait_limit.cu && ./trywait_limit.bin
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k(unsigned hint_ns, int polls, long long* out) {
  __shared__ uint64_t bar;
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(&bar);
  if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(addr));
  __syncthreads();
  uint64_t t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  const long long c0 = clock64();
  uint32_t done = 0, seen = 0;
  for (int i = 0; i < polls; ++i) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(0u), "r"(hint_ns)
        : "memory");
    seen += done;
  }
  const long long c1 = clock64();
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  if (threadIdx.x == 0) {
    out[0] = c1 - c0;
    out[1] = (long long)(t1 - t0);
    out[2] = seen;
  }
}

int main() {
  long long* d;
  cudaMalloc(&d, 24);
  for (unsigned hint : {0u, 1000u, 1000000u, 100000000u}) {
    const int polls = 64;
    k<<<1, 32>>>(hint, polls, d);
    long long h[3];
    if (cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("launch failed\n"); return 1; }
    printf("hint %9u ns: %8.1f clk, %10.1f ns per unsuccessful poll (%lld completed)\n", hint,
           (double)h[0] / polls, (double)h[1] / polls, h[2]);
  }
  return 0;
}
