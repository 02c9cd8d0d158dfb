// §8e: sequence-parallel plumbing for ONE video split over the GPUs of an NVSwitch box.
// Spatial work is sharded over frames, temporal work over spatial positions; the two all-to-alls of
// every motion module are not separate collectives: they are the motion module's own layout kernels
// (layout.cu: GroupNorm-apply + rearrange on the way in, inverse rearrange + residual on the way
// out) storing to / loading from PEER memory over NVLink, tile by tile, so the transpose and the
// exchange are one pass over the data.  This file holds what those kernels need around them:
//   * CUDA-IPC export / import of each rank's symmetric buffer (one process per GPU),
//   * the communicator (peer pointers, per-peer flags, epoch),
//   * a flag barrier over NVLink (st.release.sys / ld.acquire.sys on peer-mapped words).
// The reference has no collective at all: this path exists only here.
// NCCL (through torch.distributed) is used by the host side only to exchange the 64-byte IPC
// handles and the 256-byte GroupNorm partials.
#include <string.h>

#include "common.cuh"

namespace {

struct BarrierArgs {
  uint32_t* flags[VST_MAX_RANKS];
  int rank, world;
  uint32_t epoch;
};

// One thread per peer: publish "rank reached epoch" in the peer's flag word for this rank, then
// wait for the peer's word in our own flags.  Every rank runs this kernel on its own GPU.
__global__ void comm_barrier_kernel(BarrierArgs a) {
  const int r = threadIdx.x;
  if (r >= a.world) return;
  __threadfence_system();   // everything this GPU wrote before the barrier is visible system-wide
  uint32_t* theirs = a.flags[r] + a.rank;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(a.epoch) : "memory");
  const uint32_t* mine = a.flags[a.rank] + r;
  const uint64_t t0 = globaltimer_ns();
  uint32_t v;
  do {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    if ((int32_t)(v - a.epoch) >= 0) break;
    if (globaltimer_ns() - t0 > 20000000000ull) {   // 20 s: a peer died; fail loudly, do not hang the box
      printf("vst comm barrier: rank %d timed out waiting for rank %d (epoch %u, saw %u)\n", a.rank, r, a.epoch, v);
      __trap();
    }
  } while (true);
}

}  // namespace

// The symmetric buffer is its own cudaMalloc allocation (an IPC handle maps a whole allocation, so
// it must not be a slice of a framework's cached block); zero-filled, flags included.
extern "C" int vst_comm_alloc(int64_t bytes, void** out) {
  VST_REQUIRE(bytes > 0 && out, VST_E_BADARG, "comm_alloc: bad argument");
  VST_CUDA_OK(cudaMalloc(out, (size_t)bytes));
  VST_CUDA_OK(cudaMemset(*out, 0, (size_t)bytes));
  VST_CUDA_OK(cudaDeviceSynchronize());
  return VST_OK;
}

extern "C" int vst_comm_free(void* ptr) {
  if (ptr) VST_CUDA_OK(cudaFree(ptr));
  return VST_OK;
}

extern "C" int vst_comm_ipc_get(const void* dev_ptr, void* handle64) {
  VST_REQUIRE(dev_ptr && handle64, VST_E_BADARG, "comm_ipc_get: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  VST_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr)));
  memcpy(handle64, &h, 64);
  return VST_OK;
}

extern "C" int vst_comm_ipc_open(const void* handle64, void** out) {
  VST_REQUIRE(handle64 && out, VST_E_BADARG, "comm_ipc_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  VST_CUDA_OK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return VST_OK;
}

extern "C" int vst_comm_ipc_close(void* ptr) {
  if (ptr) VST_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return VST_OK;
}

// bufs[r] / flags[r]: device addresses, valid on THIS GPU, of rank r's data buffer and of its
// `world` 32-bit flag words (zero-initialised by the owner before the handles are exchanged).
extern "C" int vst_comm_init(int rank, int world, void* const* bufs, void* const* flags, void** comm_out) {
  VST_REQUIRE(world >= 1 && world <= VST_MAX_RANKS && rank >= 0 && rank < world && bufs && flags && comm_out,
              VST_E_BADARG, "comm_init: bad argument (world <= %d)", VST_MAX_RANKS);
  VstComm* c = new VstComm();
  c->rank = rank;
  c->world = world;
  c->epoch = 0;
  for (int r = 0; r < world; ++r) {
    VST_REQUIRE(bufs[r] && flags[r], VST_E_BADARG, "comm_init: null peer pointer for rank %d", r);
    c->buf[r] = bufs[r];
    c->flags[r] = reinterpret_cast<uint32_t*>(flags[r]);
  }
  *comm_out = c;
  return VST_OK;
}

extern "C" int vst_comm_fini(void* comm) {
  delete reinterpret_cast<VstComm*>(comm);
  return VST_OK;
}

extern "C" int vst_comm_barrier(void* comm, void* stream) {
  VST_REQUIRE(comm, VST_E_BADARG, "comm_barrier: null communicator");
  VstComm* c = reinterpret_cast<VstComm*>(comm);
  BarrierArgs a;
  for (int r = 0; r < VST_MAX_RANKS; ++r) a.flags[r] = r < c->world ? c->flags[r] : nullptr;
  a.rank = c->rank;
  a.world = c->world;
  a.epoch = ++c->epoch;
  comm_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
  VST_LAUNCH_OK();
  return VST_OK;
}
