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
  uint32_t* epoch_ctr;     // this rank's barrier count, in its own flag block (device memory, so a
                           // captured CUDA graph advances it on every replay)
  int rank, world;
};

// One thread per peer: publish "rank reached epoch" in the peer's flag word for this rank, then
// wait for the peer's word in our own flags.  Every rank runs this kernel on its own GPU.
__global__ void comm_barrier_kernel(BarrierArgs a) {
  __shared__ uint32_t s_epoch;
  const int r = threadIdx.x;
  if (r == 0) s_epoch = *a.epoch_ctr + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (r < a.world) {
    __threadfence_system();   // everything this GPU wrote before the barrier is visible system-wide
    uint32_t* theirs = a.flags[r] + a.rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(epoch) : "memory");
    const uint32_t* mine = a.flags[a.rank] + r;
    const uint64_t t0 = globaltimer_ns();
    uint32_t v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
      if ((int32_t)(v - epoch) >= 0) break;
      if (globaltimer_ns() - t0 > 20000000000ull) {   // 20 s: a peer died; fail loudly, do not hang the box
        printf("vst comm barrier: rank %d timed out waiting for rank %d (epoch %u, saw %u)\n", a.rank, r, epoch, v);
        __trap();
      }
    } while (true);
  }
  __syncthreads();
  if (r == 0) *a.epoch_ctr = epoch;
}

struct GatherArgs {
  void* buf[VST_MAX_RANKS];
  int rank, world;
};

// GroupNorm partials of the frames-sharded layout (the one reduction the math
// forces): rank p's (rows, nparts, 3) fp32 block lands at columns [p*nparts, (p+1)*nparts) of a
// (rows, world*nparts, 3) array in EVERY rank's symmetric buffer, i.e. exactly the partial order of
// the unsharded statistics kernel, so vst_groupnorm5d_finalize gives bit-identical mean / rstd.
// grid = world blocks (block r stores into rank r's buffer over NVLink).
__global__ void comm_allgather_partials_kernel(GatherArgs a, const float* __restrict__ src, int rows,
                                               int nparts, int64_t dst_offset_bytes) {
  float* dst = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(a.buf[blockIdx.x]) + dst_offset_bytes);
  const int per_row = nparts * 3;
  const int64_t total = (int64_t)rows * per_row;
  for (int64_t i = threadIdx.x; i < total; i += blockDim.x) {
    const int64_t row = i / per_row;
    const int e = (int)(i - row * per_row);
    dst[(row * a.world + a.rank) * per_row + e] = src[i];
  }
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
  a.epoch_ctr = c->flags[c->rank] + VST_COMM_EPOCH_WORD;
  comm_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
  VST_LAUNCH_OK();
  return VST_OK;
}

extern "C" int vst_comm_allgather_partials(void* comm, const float* partials, int rows, int nparts,
                                           int64_t dst_offset_bytes, void* stream) {
  VST_REQUIRE(comm && partials && rows > 0 && nparts > 0 && dst_offset_bytes >= 0 &&
                  dst_offset_bytes % 16 == 0,
              VST_E_BADARG, "comm_allgather_partials: bad argument");
  VstComm* c = reinterpret_cast<VstComm*>(comm);
  GatherArgs a;
  for (int r = 0; r < VST_MAX_RANKS; ++r) a.buf[r] = r < c->world ? c->buf[r] : nullptr;
  a.rank = c->rank;
  a.world = c->world;
  comm_allgather_partials_kernel<<<c->world, 256, 0, (cudaStream_t)stream>>>(a, partials, rows, nparts,
                                                                             dst_offset_bytes);
  VST_LAUNCH_OK();
  return VST_OK;
}
