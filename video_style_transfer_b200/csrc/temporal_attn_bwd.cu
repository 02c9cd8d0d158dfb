// a8 backward: gradients of the frame-axis attention softmax(q k^T / sqrt(d)) v, F <= 32 frames per
// (spatial position, head).  Reference: autograd through animatediff/temporal_transformer.py:45-50
// and the diffusers motion-module attn1 / attn2 (train_animatediff.py trains the temporal LoRA
// through them).
//
// Per (sequence, head): 10*F^2*d FLOP against 14*F*d bytes (q, k, v, dO in; dq, dk, dv out) =
// 0.7*F FLOP/B (11 at F = 16): HBM-bound, like the forward.  One warp owns one (sequence, head):
// 16-byte loads of the four [F x d] slices into shared memory, the F x F probabilities recomputed
// in fp32 (no saved statistics), dS = P (.) (dP - rowsum(P (.) dP)) * scale, then the three
// [F x d] products with lanes along d so the 4-byte stores of a warp are contiguous.
#include "common.cuh"

namespace {

__device__ __forceinline__ float2 ldpair(const __nv_bfloat16* p) {
  return unpack_bf16(*reinterpret_cast<const uint32_t*>(p));
}

__global__ void temporal_attn_bwd_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                                         const __nv_bfloat16* __restrict__ v, const __nv_bfloat16* __restrict__ d_o,
                                         __nv_bfloat16* __restrict__ dq, __nv_bfloat16* __restrict__ dk,
                                         __nv_bfloat16* __restrict__ dv, int64_t pairs, int F, int heads, int d,
                                         int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq,
                                         int64_t lddk, int64_t lddv, float scale, int warps, int warp_bytes) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t pair = (int64_t)blockIdx.x * warps + warp;
  if (pair >= pairs) return;
  const int64_t seq = pair / heads;
  const int head = (int)(pair % heads);
  const int LD = d + 8;                  // bf16 row stride: 16-byte multiple, spreads banks
  const int LF = F + 1;                  // fp32 row stride of the F x F matrices
  __nv_bfloat16* sq = reinterpret_cast<__nv_bfloat16*>(smem_raw + (size_t)warp * warp_bytes);
  __nv_bfloat16* sk = sq + F * LD;
  __nv_bfloat16* sv = sk + F * LD;
  __nv_bfloat16* sdo = sv + F * LD;
  float* sP = reinterpret_cast<float*>(sdo + F * LD);
  float* sdS = sP + F * LF;

  const int64_t col0 = (int64_t)head * d;
  const int CH = d / 8;
  for (int idx = lane; idx < F * CH; idx += 32) {
    const int row = idx / CH, ch = idx % CH;
    const int64_t r = seq * F + row;
    *reinterpret_cast<uint4*>(sq + row * LD + ch * 8) = ld_stream16(q + r * ldq + col0 + ch * 8);
    *reinterpret_cast<uint4*>(sk + row * LD + ch * 8) = ld_stream16(k + r * ldk + col0 + ch * 8);
    *reinterpret_cast<uint4*>(sv + row * LD + ch * 8) = ld_stream16(v + r * ldv + col0 + ch * 8);
    *reinterpret_cast<uint4*>(sdo + row * LD + ch * 8) = ld_stream16(d_o + r * lddo + col0 + ch * 8);
  }
  __syncwarp();

  // S = scale * q k^T and dP = dO v^T
  for (int o = lane; o < F * F; o += 32) {
    const int i = o / F, j = o % F;
    float s = 0.f, dp = 0.f;
    for (int c = 0; c < d; c += 2) {
      const float2 a = ldpair(sq + i * LD + c), b = ldpair(sk + j * LD + c);
      const float2 g = ldpair(sdo + i * LD + c), w = ldpair(sv + j * LD + c);
      s = fmaf(a.x, b.x, fmaf(a.y, b.y, s));
      dp = fmaf(g.x, w.x, fmaf(g.y, w.y, dp));
    }
    sP[i * LF + j] = s * scale;
    sdS[i * LF + j] = dp;
  }
  __syncwarp();
  // row softmax and dS = P (.) (dP - sum_j P dP) * scale
  if (lane < F) {
    float* prow = sP + lane * LF;
    float* drow = sdS + lane * LF;
    float mx = -INFINITY;
    for (int j = 0; j < F; ++j) mx = fmaxf(mx, prow[j]);
    float sum = 0.f;
    for (int j = 0; j < F; ++j) {
      const float e = __expf(prow[j] - mx);
      prow[j] = e;
      sum += e;
    }
    const float inv = 1.f / sum;
    float dot = 0.f;
    for (int j = 0; j < F; ++j) {
      const float p = prow[j] * inv;
      prow[j] = p;
      dot = fmaf(p, drow[j], dot);
    }
    for (int j = 0; j < F; ++j) drow[j] = prow[j] * (drow[j] - dot) * scale;
  }
  __syncwarp();
  // dQ = dS k, dK = dS^T q, dV = P^T dO : lanes along d
  const int half = d / 2;
  for (int idx = lane; idx < F * half; idx += 32) {
    const int r = idx / half, c = (idx % half) * 2;
    float2 aq = make_float2(0.f, 0.f), ak = aq, av = aq;
    for (int t = 0; t < F; ++t) {
      const float ds_rt = sdS[r * LF + t], ds_tr = sdS[t * LF + r], p_tr = sP[t * LF + r];
      const float2 kk = ldpair(sk + t * LD + c), qq = ldpair(sq + t * LD + c), gg = ldpair(sdo + t * LD + c);
      aq.x = fmaf(ds_rt, kk.x, aq.x); aq.y = fmaf(ds_rt, kk.y, aq.y);
      ak.x = fmaf(ds_tr, qq.x, ak.x); ak.y = fmaf(ds_tr, qq.y, ak.y);
      av.x = fmaf(p_tr, gg.x, av.x);  av.y = fmaf(p_tr, gg.y, av.y);
    }
    const int64_t row = seq * F + r;
    *reinterpret_cast<uint32_t*>(dq + row * lddq + col0 + c) = pack_bf16(aq.x, aq.y);
    *reinterpret_cast<uint32_t*>(dk + row * lddk + col0 + c) = pack_bf16(ak.x, ak.y);
    *reinterpret_cast<uint32_t*>(dv + row * lddv + col0 + c) = pack_bf16(av.x, av.y);
  }
}

}  // namespace

extern "C" int vst_temporal_attn_bwd(const void* q, const void* k, const void* v, const void* d_o,
                                     void* dq, void* dk, void* dv, int64_t N, int F, int heads,
                                     int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo,
                                     int64_t lddq, int64_t lddk, int64_t lddv, float scale,
                                     void* stream) {
  VST_REQUIRE(N >= 0 && F > 0 && heads > 0 && d > 0, VST_E_BADARG, "temporal attention bwd: bad shape");
  if (N == 0) return VST_OK;
  VST_REQUIRE(q && k && v && d_o && dq && dk && dv, VST_E_BADARG, "temporal attention bwd: null pointer");
  VST_REQUIRE(F <= 32, VST_E_UNSUPPORTED, "temporal attention bwd: F=%d > 32", F);
  VST_REQUIRE(d % 8 == 0 && d <= 160, VST_E_UNSUPPORTED,
              "temporal attention bwd: head dim %d must be a multiple of 8, <= 160", d);
  VST_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && lddo % 8 == 0 && lddq % 2 == 0 &&
                  lddk % 2 == 0 && lddv % 2 == 0,
              VST_E_UNSUPPORTED, "temporal attention bwd: row strides must keep 16-byte (inputs) / 4-byte (outputs) alignment");
  VST_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) |
                reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(d_o)) & 15) == 0 &&
                  ((reinterpret_cast<uintptr_t>(dq) | reinterpret_cast<uintptr_t>(dk) |
                    reinterpret_cast<uintptr_t>(dv)) & 3) == 0,
              VST_E_UNSUPPORTED, "temporal attention bwd: pointer alignment");
  int warp_bytes = 4 * F * (d + 8) * 2 + 2 * F * (F + 1) * 4;
  warp_bytes = (warp_bytes + 15) / 16 * 16;
  int warps = (200 * 1024) / warp_bytes;
  if (warps > 8) warps = 8;
  VST_REQUIRE(warps >= 1, VST_E_UNSUPPORTED, "temporal attention bwd: tile does not fit shared memory");
  const int smem = warps * warp_bytes;
  static int configured = 0;
  if (smem > configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(temporal_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = smem;
  }
  const int64_t pairs = N * heads;
  const int64_t grid = (pairs + warps - 1) / warps;
  VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "temporal attention bwd: grid too large");
  temporal_attn_bwd_kernel<<<(unsigned)grid, warps * 32, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v,
      (const __nv_bfloat16*)d_o, (__nv_bfloat16*)dq, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, pairs, F,
      heads, d, ldq, ldk, ldv, lddo, lddq, lddk, lddv, scale, warps, warp_bytes);
  VST_LAUNCH_OK();
  return VST_OK;
}
