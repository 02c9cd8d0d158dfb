// How long does one mbarrier.try_wait poll suspend when the phase never completes?  (bounds the
// poll-count watchdog of csrc/common.cuh: VST_WATCHDOG_SPINS polls x this time)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o trywait_limit.bin trywait_limit.cu && ./trywait_limit.bin
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_nohint(int polls, long long* out) {
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
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(0u)
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

// wake-up latency: lane 0 of warp 1 arrives `delay` clk after warp 0 started waiting; warp 0 reports how long
// after the arrive instruction it left the wait
template <bool HINT>
__global__ void k_wake(int delay, long long* out) {
  __shared__ uint64_t bar;
  __shared__ long long t_arrive;
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(&bar);
  if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(addr));
  __syncthreads();
  if (threadIdx.x == 32) {
    const long long c0 = clock64();
    while (clock64() - c0 < delay) {}
    t_arrive = clock64();
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
  } else if (threadIdx.x == 0) {
    uint32_t done = 0;
    while (!done) {
      if (HINT)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(0u), "r"(1000000u) : "memory");
      else
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(0u) : "memory");
    }
    const long long c1 = clock64();
    out[0] = c1 - *(volatile long long*)&t_arrive;
  }
}

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
  {
    const int polls = 64;
    k_nohint<<<1, 32>>>(polls, d);
    long long h[3];
    if (cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("launch failed\n"); return 1; }
    printf("no hint operand:   %8.1f clk, %10.1f ns per unsuccessful poll (%lld completed)\n",
           (double)h[0] / polls, (double)h[1] / polls, h[2]);
  }
  for (int delay : {200, 1000, 3000, 10000, 30000}) {
    long long a, b;
    k_wake<true><<<1, 64>>>(delay, d);
    cudaMemcpy(&a, d, 8, cudaMemcpyDeviceToHost);
    k_wake<false><<<1, 64>>>(delay, d);
    cudaMemcpy(&b, d, 8, cudaMemcpyDeviceToHost);
    printf("arrive %6d clk after the wait began: waiter leaves %6lld clk after the arrive with the hint, %6lld without\n",
           delay, a, b);
  }
  return 0;
}
