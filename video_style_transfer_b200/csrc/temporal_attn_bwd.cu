// a8 backward: gradients of the frame-axis attention softmax(q k^T / sqrt(d)) v, F <= 32 frames per
// (spatial position, head).  Reference: autograd through animatediff/temporal_transformer.py:45-50
// and the diffusers motion-module attn1 / attn2 (train_animatediff.py trains the temporal LoRA
// through them).
//
// Per (sequence, head): 10*F^2*d FLOP against 14*F*d bytes (q, k, v, dO in; dq, dk, dv out) =
// 0.7*F FLOP/B (11 at F = 16): HBM-bound, like the forward.  One warp owns one (sequence, head):
// 16-byte loads of the four [F x d] slices into shared memory, then five tiny contractions on
// warp-level tensor-core MMAs (m16n8k16, as the forward; a 128-row tcgen05 tile would be 7/8
// padding):
//   S = Q K^T, dP = dO V^T          -> P = softmax(S * scale), dS = P (.) (dP - rowsum(P (.) dP)) * scale
//   dV = P^T dO, dQ = dS K, dK = dS^T Q   (P / dS staged in smem as bf16; the transposed operands
//                                          come from ldmatrix.trans)
// The probabilities are recomputed (no saved statistics).  Outputs are staged in the input tiles
// that are dead by then and stored with 16-byte coalesced writes.  The first version of this
// kernel did the contractions with scalar fp32 loops and ran at 0.11 of HBM.
#include "common.cuh"

namespace {

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t* r) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t* r) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// out[FP x DPAD] = A[FP x FP] * B[FP x DPAD] with A, B bf16 in smem and the result written (bf16)
// into `dst` (row stride LD).  TRANS_A: A is stored transposed (memory row = contraction index).
// B is stored [contraction index][d] (row-major), so its fragments come from ldmatrix.trans.
template <int DPAD, int FP, bool TRANS_A>
__device__ __forceinline__ void small_mm(uint32_t a_addr, int lda, uint32_t b_addr, int ldb,
                                         __nv_bfloat16* dst, int ldd, int lane) {
  constexpr int MT = FP / 16, KT = FP / 16, DN = DPAD / 8;
  uint32_t a[MT][KT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      if (TRANS_A) {
        // memory holds A^T: rows = k, cols = m.  lane group q = lane / 8: tile (m-block q & 1,
        // k-block q >> 1); each lane passes one k-row of its tile
        const int q = lane >> 3;
        ldsm_x4_t(a_addr + ((kt * 16 + (q >> 1) * 8 + (lane & 7)) * lda + mt * 16 + (q & 1) * 8) * 2, a[mt][kt]);
      } else {
        ldsm_x4(a_addr + ((mt * 16 + (lane & 15)) * lda + kt * 16 + (lane >> 4) * 8) * 2, a[mt][kt]);
      }
    }
  const int g = lane >> 2, t4 = lane & 3;
#pragma unroll
  for (int np = 0; np < DN / 2; ++np) {
    float acc[MT][2][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[mt][j][i] = 0.f;
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      uint32_t b[4];
      ldsm_x4_t(b_addr + ((kt * 16 + (lane & 15)) * ldb + np * 16 + (lane >> 4) * 8) * 2, b);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_bf16(acc[mt][0], a[mt][kt], b[0], b[1]);
        mma_bf16(acc[mt][1], a[mt][kt], b[2], b[3]);
      }
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int col = np * 16 + j * 8 + 2 * t4;
        *reinterpret_cast<uint32_t*>(dst + (mt * 16 + g) * ldd + col) = pack_bf16(acc[mt][j][0], acc[mt][j][1]);
        *reinterpret_cast<uint32_t*>(dst + (mt * 16 + g + 8) * ldd + col) = pack_bf16(acc[mt][j][2], acc[mt][j][3]);
      }
  }
}

// DPAD: head dim rounded up to a multiple of 16; FP: frames rounded up to 16 or 32.
template <int DPAD, int FP, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
temporal_attn_bwd_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                         const __nv_bfloat16* __restrict__ v, const __nv_bfloat16* __restrict__ d_o,
                         __nv_bfloat16* __restrict__ dq, __nv_bfloat16* __restrict__ dk,
                         __nv_bfloat16* __restrict__ dv, int64_t pairs, int F, int heads, int d,
                         int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq,
                         int64_t lddk, int64_t lddv, float scale, float scale_log2) {
  constexpr int LD = DPAD + 8;           // smem row stride (elements) of the [F x d] tiles
  constexpr int LP = FP + 8;             // row stride of the [F x F] P / dS tiles
  constexpr int MT = FP / 16, NT = FP / 8, DK = DPAD / 16, CH = DPAD / 8;
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t pair = (int64_t)blockIdx.x * WARPS + warp;
  if (pair >= pairs) return;
  const int64_t seq = pair / heads;
  const int head = (int)(pair % heads);
  __nv_bfloat16* sq = reinterpret_cast<__nv_bfloat16*>(smem_raw) + (size_t)warp * (4 * FP * LD + 2 * FP * LP);
  __nv_bfloat16* sk = sq + FP * LD;
  __nv_bfloat16* sv = sk + FP * LD;
  __nv_bfloat16* sdo = sv + FP * LD;
  __nv_bfloat16* sP = sdo + FP * LD;
  __nv_bfloat16* sdS = sP + FP * LP;

  const int64_t col0 = (int64_t)head * d;
  const __nv_bfloat16* gq = q + seq * F * ldq + col0;
  const __nv_bfloat16* gk = k + seq * F * ldk + col0;
  const __nv_bfloat16* gv = v + seq * F * ldv + col0;
  const __nv_bfloat16* gdo = d_o + seq * F * lddo + col0;
  for (int idx = lane; idx < FP * CH; idx += 32) {
    const int row = idx / CH, ch = idx % CH;
    uint4 a = make_uint4(0, 0, 0, 0), b = a, c = a, e = a;
    if (row < F && ch * 8 < d) {
      a = ld_stream16(gq + row * ldq + ch * 8);
      b = ld_stream16(gk + row * ldk + ch * 8);
      c = ld_stream16(gv + row * ldv + ch * 8);
      e = ld_stream16(gdo + row * lddo + ch * 8);
    }
    *reinterpret_cast<uint4*>(sq + row * LD + ch * 8) = a;
    *reinterpret_cast<uint4*>(sk + row * LD + ch * 8) = b;
    *reinterpret_cast<uint4*>(sv + row * LD + ch * 8) = c;
    *reinterpret_cast<uint4*>(sdo + row * LD + ch * 8) = e;
  }
  __syncwarp();

  // ---- S = Q K^T and dP = dO V^T (C fragments: rows g / g+8 of each m-tile, 2 keys per lane) ----
  float s[MT][NT][4], dp[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) { s[mt][nt][i] = 0.f; dp[mt][nt][i] = 0.f; }
  const uint32_t sq_a = smem_u32(sq), sk_a = smem_u32(sk), sv_a = smem_u32(sv), sdo_a = smem_u32(sdo);
#pragma unroll
  for (int kt = 0; kt < DK; ++kt) {
    uint32_t aq[MT][4], ado[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const int off = ((mt * 16 + (lane & 15)) * LD + kt * 16 + (lane >> 4) * 8) * 2;
      ldsm_x4(sq_a + off, aq[mt]);
      ldsm_x4(sdo_a + off, ado[mt]);
    }
#pragma unroll
    for (int np = 0; np < NT / 2; ++np) {
      const int off = ((np * 16 + (lane & 7) + ((lane >> 4) << 3)) * LD + kt * 16 + ((lane >> 3) & 1) * 8) * 2;
      uint32_t bk[4], bv[4];
      ldsm_x4(sk_a + off, bk);
      ldsm_x4(sv_a + off, bv);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_bf16(s[mt][2 * np], aq[mt], bk[0], bk[1]);
        mma_bf16(s[mt][2 * np + 1], aq[mt], bk[2], bk[3]);
        mma_bf16(dp[mt][2 * np], ado[mt], bv[0], bv[1]);
        mma_bf16(dp[mt][2 * np + 1], ado[mt], bv[2], bv[3]);
      }
    }
  }

  // ---- P = softmax(S * scale); dS = P (.) (dP - sum_j P dP) * scale; both to smem as bf16 ----
  const int g = lane >> 2, t4 = lane & 3;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float mx = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int key = nt * 8 + 2 * t4 + i;
          const float val = key < F ? s[mt][nt][2 * h + i] * scale_log2 : -INFINITY;
          s[mt][nt][2 * h + i] = val;
          mx = fmaxf(mx, val);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float p = exp2f(s[mt][nt][2 * h + i] - mx);
          s[mt][nt][2 * h + i] = p;
          sum += p;
        }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      const float inv = 1.f / sum;
      float dot = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float p = s[mt][nt][2 * h + i] * inv;
          s[mt][nt][2 * h + i] = p;
          dot = fmaf(p, dp[mt][nt][2 * h + i], dot);
        }
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      dot += __shfl_xor_sync(0xffffffffu, dot, 2);
      const int row = mt * 16 + g + 8 * h;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const float p0 = s[mt][nt][2 * h], p1 = s[mt][nt][2 * h + 1];
        const float d0 = p0 * (dp[mt][nt][2 * h] - dot) * scale;
        const float d1 = p1 * (dp[mt][nt][2 * h + 1] - dot) * scale;
        const int col = nt * 8 + 2 * t4;
        // rows >= F hold q = dO = 0: P is a uniform row there, dS = 0; they only meet zero rows
        *reinterpret_cast<uint32_t*>(sP + row * LP + col) = row < F ? pack_bf16(p0, p1) : 0u;
        *reinterpret_cast<uint32_t*>(sdS + row * LP + col) = row < F ? pack_bf16(d0, d1) : 0u;
      }
    }
  }
  __syncwarp();

  const uint32_t sP_a = smem_u32(sP), sdS_a = smem_u32(sdS);
  // dV = P^T dO   -> sv (V is dead)
  small_mm<DPAD, FP, true>(sP_a, LP, sdo_a, LD, sv, LD, lane);
  __syncwarp();
  // dQ = dS K     -> sdo (dO is dead)
  small_mm<DPAD, FP, false>(sdS_a, LP, sk_a, LD, sdo, LD, lane);
  __syncwarp();
  // dK = dS^T Q   -> sk (K is dead)
  small_mm<DPAD, FP, true>(sdS_a, LP, sq_a, LD, sk, LD, lane);
  __syncwarp();

  __nv_bfloat16* oq = dq + seq * F * lddq + col0;
  __nv_bfloat16* ok = dk + seq * F * lddk + col0;
  __nv_bfloat16* ov = dv + seq * F * lddv + col0;
  for (int idx = lane; idx < FP * CH; idx += 32) {
    const int row = idx / CH, ch = idx % CH;
    if (row < F && ch * 8 < d) {
      st_stream16(oq + row * lddq + ch * 8, *reinterpret_cast<const uint4*>(sdo + row * LD + ch * 8));
      st_stream16(ok + row * lddk + ch * 8, *reinterpret_cast<const uint4*>(sk + row * LD + ch * 8));
      st_stream16(ov + row * lddv + ch * 8, *reinterpret_cast<const uint4*>(sv + row * LD + ch * 8));
    }
  }
}

template <int DPAD, int FP>
int launch(const void* q, const void* k, const void* v, const void* d_o, void* dq, void* dk, void* dv,
           int64_t N, int F, int heads, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo,
           int64_t lddq, int64_t lddk, int64_t lddv, float scale, cudaStream_t st) {
  constexpr int per_warp = (4 * FP * (DPAD + 8) + 2 * FP * (FP + 8)) * 2;
  constexpr int WARPS = per_warp * 8 <= 100 * 1024 ? 8 : 4;
  constexpr int smem = WARPS * per_warp;
  auto kern = temporal_attn_bwd_kernel<DPAD, FP, WARPS>;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  const int64_t pairs = N * heads;
  const int64_t grid = (pairs + WARPS - 1) / WARPS;
  VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "temporal attention bwd: grid too large");
  kern<<<(unsigned)grid, WARPS * 32, smem, st>>>(
      (const __nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v,
      (const __nv_bfloat16*)d_o, (__nv_bfloat16*)dq, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, pairs, F,
      heads, d, ldq, ldk, ldv, lddo, lddq, lddk, lddv, scale, scale * 1.4426950408889634f);
  VST_LAUNCH_OK();
  return VST_OK;
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
  VST_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && lddo % 8 == 0 && lddq % 8 == 0 &&
                  lddk % 8 == 0 && lddv % 8 == 0,
              VST_E_UNSUPPORTED, "temporal attention bwd: row strides must be multiples of 8 elements");
  VST_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) |
                reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(d_o) |
                reinterpret_cast<uintptr_t>(dq) | reinterpret_cast<uintptr_t>(dk) |
                reinterpret_cast<uintptr_t>(dv)) & 15) == 0,
              VST_E_UNSUPPORTED, "temporal attention bwd: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
#define TA(DP, FPP)                                                                                 \
  return launch<DP, FPP>(q, k, v, d_o, dq, dk, dv, N, F, heads, d, ldq, ldk, ldv, lddo, lddq, lddk, \
                         lddv, scale, st)
  if (F <= 16) {
    if (d <= 32) TA(32, 16);
    if (d <= 48) TA(48, 16);
    if (d <= 80) TA(80, 16);
    if (d <= 128) TA(128, 16);
    TA(160, 16);
  } else {
    if (d <= 32) TA(32, 32);
    if (d <= 48) TA(48, 32);
    if (d <= 80) TA(80, 32);
    if (d <= 128) TA(128, 32);
    TA(160, 32);
  }
#undef TA
}
