// Parameter-side gradient kernels of the LoRA projections: the adjoint of vst_lora_pack_up and the
// bias gradient.  Reference: what autograd derives for unziplora_linear_layer.py:236-247 (hard
// mask, merger, up-projection), lora_linear.py:81 (scale), temporal_lora.py:29-32.
// Memory-bound, tiny next to the token GEMMs: one pass over dy for the column sum, parameter-sized
// work for the pack-up adjoint.
#include "common.cuh"

namespace {

// B'[n, br*r + j] = scale * m_br[n] * mask_br[n] * B_br[n, j]   (vst_lora_pack_up)
//   dB_br[n, j] = scale * m_br[n] * mask_br[n] * dB'[n, br*rp + j]
//   dm_br[n]    = scale * mask_br[n] * sum_j dB'[n, br*rp + j] * B_br[n, j]
// One warp per (row n, branch).  rp = padded rank (row stride of dB' per branch).
template <typename T>
__global__ void pack_up_bwd_kernel(const float* __restrict__ dBp, int64_t ld, int rp,
                                   const T* __restrict__ Bc, const T* __restrict__ mc,
                                   const uint8_t* __restrict__ maskc, const T* __restrict__ Bs,
                                   const T* __restrict__ ms, const uint8_t* __restrict__ masks,
                                   float scale, float* __restrict__ dBc, float* __restrict__ dmc,
                                   float* __restrict__ dBs, float* __restrict__ dms, int N, int r,
                                   int nbr) {
  const int warp = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (warp >= N * nbr) return;
  const int n = warp / nbr, br = warp % nbr;
  const bool second = (Bc != nullptr) && br == 1;
  const bool style = Bc == nullptr || second;
  const T* B = style ? Bs : Bc;
  const T* m = style ? ms : mc;
  const uint8_t* mk = style ? masks : maskc;
  float* dB = style ? dBs : dBc;
  float* dm = style ? dms : dmc;
  const float keep = (mk != nullptr && mk[n] == 0) ? 0.f : 1.f;
  const float mv = m != nullptr ? (float)m[n] : 1.f;
  const float* g = dBp + (int64_t)n * ld + br * rp;
  float dot = 0.f;
  for (int j = lane; j < r; j += 32) {
    const float gv = g[j];
    dot += gv * (float)B[(int64_t)n * r + j];
    if (dB != nullptr) dB[(int64_t)n * r + j] = scale * mv * keep * gv;
  }
  dot = warp_sum(dot);
  if (lane == 0 && dm != nullptr) dm[n] = scale * keep * dot;
}

// Column sums of a bf16 matrix in two deterministic stages: partial[c][n] over row chunk c, then a
// fixed-order sum over chunks.  Block = 64 columns x one row chunk; 256 threads = 8 row lanes x 32
// column pairs, 4-byte loads (a warp reads one full 128-byte line of a row).
constexpr int kColsPerBlock = 64;
__global__ void colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int64_t M,
                                      int N, int64_t rows_per_chunk, float* __restrict__ partial) {
  const int c0 = blockIdx.x * kColsPerBlock + (threadIdx.x & 31) * 2;
  const int rl = threadIdx.x >> 5;   // 0..7
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t r1 = min(M, r0 + rows_per_chunk);
  float2 s = make_float2(0.f, 0.f);
  if (c0 < N) {
    for (int64_t r = r0 + rl; r < r1; r += 8) {
      const uint32_t v = *reinterpret_cast<const uint32_t*>(x + r * ld + c0);
      const float2 e = unpack_bf16(v);
      s.x += e.x;
      s.y += e.y;
    }
  }
  __shared__ float2 sm[8][32];
  sm[rl][threadIdx.x & 31] = s;
  __syncthreads();
  if (rl == 0 && c0 < N) {
    float2 t = sm[0][threadIdx.x];
#pragma unroll
    for (int i = 1; i < 8; ++i) {
      t.x += sm[i][threadIdx.x].x;
      t.y += sm[i][threadIdx.x].y;
    }
    *reinterpret_cast<float2*>(partial + (int64_t)blockIdx.y * N + c0) = t;
  }
}
// Same sums with 16-byte loads (N, ld multiples of 8 elements, 16-byte aligned base): block = 128 columns x one row
// chunk, 256 threads = 16 row lanes x 16 column octets, four rows in flight per thread.  The 4-byte form above
// measured 0.34 of HBM over the 60 bias-gradient launches of a config-4 step (r02, profiles/cfg4_kernels_r02.tsv).
constexpr int kVecCols = 128;
__global__ void __launch_bounds__(256)
colsum_partial_vec_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int64_t M, int N,
                          int64_t rows_per_chunk, float* __restrict__ partial) {
  const int cg = threadIdx.x & 15, rl = threadIdx.x >> 4;   // column octet, row lane 0..15
  const int c0 = blockIdx.x * kVecCols + cg * 8;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t r1 = min(M, r0 + rows_per_chunk);
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  auto add = [&](const uint4& u) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 e = unpack_bf16(w[q]);
      s[2 * q] += e.x;
      s[2 * q + 1] += e.y;
    }
  };
  if (c0 < N) {
    const __nv_bfloat16* p = x + c0;
    int64_t r = r0 + rl;
    for (; r + 48 < r1; r += 64) {
      const uint4 a = ld_stream16(p + r * ld), b = ld_stream16(p + (r + 16) * ld);
      const uint4 c = ld_stream16(p + (r + 32) * ld), d = ld_stream16(p + (r + 48) * ld);
      add(a); add(b); add(c); add(d);
    }
    for (; r < r1; r += 16) add(ld_stream16(p + r * ld));
  }
  __shared__ float sm[16][kVecCols + 4];
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[rl][cg * 8 + j] = s[j];
  __syncthreads();
  if (threadIdx.x < kVecCols && blockIdx.x * kVecCols + threadIdx.x < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) t += sm[i][threadIdx.x];
    partial[(int64_t)blockIdx.y * N + blockIdx.x * kVecCols + threadIdx.x] = t;
  }
}
__global__ void colsum_final_kernel(const float* __restrict__ partial, int chunks, int N,
                                    float alpha, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int c = 0; c < chunks; ++c) s += partial[(int64_t)c * N + n];
  out[n] = alpha * s;
}

void colsum_launch_partial(const void* x, int64_t ld, int64_t M, int N, int chunks, float* partial,
                           cudaStream_t st) {
  const int64_t rows_per_chunk = (M + chunks - 1) / chunks;
  if (N % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    dim3 grid((N + kVecCols - 1) / kVecCols, chunks);
    colsum_partial_vec_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, ld, M, N, rows_per_chunk, partial);
  } else {
    dim3 grid((N + kColsPerBlock - 1) / kColsPerBlock, chunks);
    colsum_partial_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, ld, M, N, rows_per_chunk, partial);
  }
}

int colsum_chunks(int64_t M, int N) {
  // enough (column block, row chunk) blocks for ~8 resident blocks of 256 threads per SM
  const int col_blocks = (N + kVecCols - 1) / kVecCols;
  int64_t chunks = (8 * vst_sm_count() + col_blocks - 1) / col_blocks;
  const int64_t max_chunks = (M + 63) / 64;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  return (int)chunks;
}

}  // namespace

extern "C" int vst_lora_pack_up_bwd(const float* dBp, int64_t ld, int rp, const void* Bc,
                                    const void* mc, const uint8_t* maskc, const void* Bs,
                                    const void* ms, const uint8_t* masks, float scale, float* dBc,
                                    float* dmc, float* dBs, float* dms, int N, int r,
                                    int param_dtype, void* stream) {
  VST_REQUIRE(N > 0 && r > 0 && rp >= r && dBp && (Bc || Bs), VST_E_BADARG, "lora_pack_up_bwd: bad argument");
  const int nbr = (Bc ? 1 : 0) + (Bs ? 1 : 0);
  const int64_t threads = (int64_t)N * nbr * 32;
  const unsigned grid = (unsigned)((threads + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (param_dtype == VST_BF16)
    pack_up_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        dBp, ld, rp, (const __nv_bfloat16*)Bc, (const __nv_bfloat16*)mc, maskc,
        (const __nv_bfloat16*)Bs, (const __nv_bfloat16*)ms, masks, scale, dBc, dmc, dBs, dms, N, r, nbr);
  else if (param_dtype == VST_F32)
    pack_up_bwd_kernel<float><<<grid, 256, 0, st>>>(dBp, ld, rp, (const float*)Bc, (const float*)mc,
                                                    maskc, (const float*)Bs, (const float*)ms, masks,
                                                    scale, dBc, dmc, dBs, dms, N, r, nbr);
  else {
    vst_set_error("lora_pack_up_bwd: bad param dtype %d", param_dtype);
    return VST_E_BADARG;
  }
  VST_LAUNCH_OK();
  return VST_OK;
}

extern "C" int64_t vst_colsum_workspace_bytes(int64_t M, int N) {
  if (M <= 0 || N <= 0) return 0;
  return (int64_t)colsum_chunks(M, N) * N * 4;
}

extern "C" int vst_colsum(const void* x, int64_t ld, int64_t M, int N, float alpha, float* out,
                          void* workspace, int64_t workspace_bytes, void* stream) {
  VST_REQUIRE(M > 0 && N > 0 && x && out, VST_E_BADARG, "colsum: bad argument");
  VST_REQUIRE(N % 2 == 0 && ld % 2 == 0, VST_E_UNSUPPORTED, "colsum: N and ld must be even");
  const int chunks = colsum_chunks(M, N);
  VST_REQUIRE(workspace && workspace_bytes >= (int64_t)chunks * N * 4, VST_E_BADARG,
              "colsum: workspace of vst_colsum_workspace_bytes(M, N) bytes required");
  cudaStream_t st = (cudaStream_t)stream;
  colsum_launch_partial(x, ld, M, N, chunks, (float*)workspace, st);
  VST_LAUNCH_OK();
  colsum_final_kernel<<<(N + 255) / 256, 256, 0, st>>>((const float*)workspace, chunks, N, alpha, out);
  VST_LAUNCH_OK();
  return VST_OK;
}


// ---- vst_colsum_noalloc: vst_colsum for callers that cannot allocate ---------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Same two
// kernels as vst_colsum; the fp32 partials live in a file-scope device array instead of a caller workspace.
//   * VST_COLSUM_NOALLOC_SCRATCH_BYTES (1 MiB) is the hard maximum: a shape whose partials
//     (vst_colsum_workspace_bytes(M, N)) exceed it is refused with VST_E_UNSUPPORTED.  Every GPU that loads the
//     library holds the array until the process exits.
//   * The array is shared by every call and keeps the last call's contents.  Nothing has to start at zero: the first
//     kernel writes every (chunk, column) slot the second one reads.  Calls that OVERLAP on the GPU (two streams,
//     two host threads) race on it: serialise them.
//   * In host code the array's NAME is a host-side stand-in, not a device address; the address is taken once per
//     device, with that device current, by cudaGetSymbolAddress (no allocation), and passed to the kernels.
static __device__ __align__(256) unsigned char g_colsum_noalloc_scratch[VST_COLSUM_NOALLOC_SCRATCH_BYTES];

static int colsum_noalloc_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static float* ptrs[kMaxDevices] = {nullptr};
  int dev = -1;
  VST_CUDA_OK(cudaGetDevice(&dev));
  VST_REQUIRE(dev >= 0 && dev < kMaxDevices, VST_E_UNSUPPORTED, "colsum_noalloc: device ordinal %d out of range", dev);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    VST_CUDA_OK(cudaGetSymbolAddress(&p, g_colsum_noalloc_scratch));
    VST_REQUIRE(p != nullptr, VST_E_CUDA, "colsum_noalloc: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = reinterpret_cast<float*>(p);
  }
  *out = ptrs[dev];
  return VST_OK;
}

extern "C" int vst_colsum_noalloc(const void* x, int64_t ld, int64_t M, int N, float alpha, float* out,
                                  void* stream) {
  VST_REQUIRE(M > 0 && N > 0 && x && out, VST_E_BADARG, "colsum_noalloc: bad argument");
  VST_REQUIRE(N % 2 == 0 && ld % 2 == 0, VST_E_UNSUPPORTED, "colsum_noalloc: N and ld must be even");
  const int chunks = colsum_chunks(M, N);
  const int64_t need = (int64_t)chunks * N * 4;
  VST_REQUIRE(need <= (int64_t)VST_COLSUM_NOALLOC_SCRATCH_BYTES, VST_E_UNSUPPORTED,
              "colsum_noalloc: M=%lld N=%d needs %lld bytes of scratch, the built-in array holds %lld "
              "(use vst_colsum with a caller workspace)",
              (long long)M, N, (long long)need, (long long)VST_COLSUM_NOALLOC_SCRATCH_BYTES);
  float* scratch = nullptr;
  int rc = colsum_noalloc_scratch(&scratch);
  if (rc != VST_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  colsum_launch_partial(x, ld, M, N, chunks, scratch, st);      // the same kernels as vst_colsum
  VST_LAUNCH_OK();
  colsum_final_kernel<<<(N + 255) / 256, 256, 0, st>>>(scratch, chunks, N, alpha, out);
  VST_LAUNCH_OK();
  return VST_OK;
}
