// a2: spatial self-attention / text cross-attention forward, head_dim 64, on tcgen05 + TMEM + TMA.
// Reference: animatediff/attention_processor.py:74-83 (head split, F.scaled_dot_product_attention,
// head merge) -- the head split/merge transposes disappear into the TMA tensor maps, and the
// text repeat_interleave of :62-66 into the kv batch index (b / kv_group).
//
// One CTA = 128 query rows of one (batch, head); two CTAs are resident per SM (256 TMEM columns
// each: S [0,128) fp32, O [128,192) fp32, P [192,256) bf16 pairs).  192 threads:
//   warps 0..3  softmax: thread t owns query row t (TMEM lane t).  Per 128-key block ONE pass:
//               S row -> registers (S is released at once, so the next S = Q K^T overlaps this
//               block's softmax), row max, P = exp2(s*scale - m) with packed f32x2 math, bf16 P
//               written back to TMEM.  O accumulates in TMEM across blocks; it is rescaled only
//               when the running max moves by more than 2^8 (lazy rescale), by the same warps.
//   warp 4      TMA producer: Q once, K / V tiles [128 x 64] (2 stages each).
//   warp 5      MMA issuer: S = Q K^T (4 x tcgen05.mma 128x128x16, both operands K-major smem)
//               and O += P V (8 x 128x64x16, A = P from TMEM, B = V MN-major straight from its
//               row-major tile), so P never touches shared memory.
// Roofline: tensor pipe for self-attention (S/2 FLOP/B); FLOPs = 4*Bt*h*Sq*Sk*d.
#include <type_traits>

#include "common.cuh"

namespace {

constexpr int BQ = 128;   // query rows per CTA
constexpr int BKV = 128;  // keys per block
constexpr int HD = 64;    // head dim
constexpr int kThreads = 192;
constexpr int KST = 2;    // K and V stages
constexpr int kTile = 128 * HD * 2;  // 16 KB
constexpr int kSmemBytes = kTile * (1 + 2 * KST) + 1024 + 128;
constexpr float kRescaleThreshold = 8.f;  // log2 units: P stays below 2^8, safe in bf16 / fp32

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct AttnArgs {
  float* lse;
  int Sq, Sk, heads, kv_group;
  float scale_log2;
};

__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kTile;
  uint8_t* sV = sK + KST * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + KST * kTile);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;          // [KST]
  uint64_t* k_empty = k_full + KST;     // [KST]
  uint64_t* v_full = k_empty + KST;     // [KST]
  uint64_t* v_empty = v_full + KST;     // [KST]
  uint64_t* s_full = v_empty + KST;     // S_j complete in TMEM
  uint64_t* s_free = s_full + 1;        // S_j copied to registers by all 128 rows
  uint64_t* p_full = s_free + 1;        // P_j in TMEM (and O rescaled if needed)
  uint64_t* o_full = p_full + 1;        // O += P_j V_j complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, head = blockIdx.y, b = blockIdx.z;
  const int nblk = (a.Sk + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < KST; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 128);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base;        // columns [0,128): S (fp32)
  const uint32_t tO = tmem_base + 128;  // columns [128,192): O accumulator (fp32)
  const uint32_t tP = tmem_base + 192;  // columns [192,256): P (bf16 pairs, 2 keys per column)

  if (warp == 4) {
    // ================= TMA producer =================
    if (lane == 0) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      mbar_arrive_expect_tx(q_full, kTile);
      tma_load_3d(sQ, &tmQ, q_full, head * HD, q0, b);
      const int bk = b / a.kv_group;
      for (int j = 0; j < nblk; ++j) {
        const int st = j % KST;
        const uint32_t ph = ((j / KST) & 1) ^ 1;
        mbar_wait(&k_empty[st], ph, 10);
        mbar_arrive_expect_tx(&k_full[st], kTile);
        tma_load_3d(sK + st * kTile, &tmK, &k_full[st], head * HD, j * BKV, bk);
        mbar_wait(&v_empty[st], ph, 11);
        mbar_arrive_expect_tx(&v_full[st], kTile);
        tma_load_3d(sV + st * kTile, &tmV, &v_full[st], head * HD, j * BKV, bk);
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, 0, 0);
      constexpr uint32_t idesc_o = umma_idesc_bf16(BQ, HD, 0, 1);  // B (= V) is MN-major
      const uint64_t dq = umma_desc_sw128(smem_u32(sQ), 1024);
      mbar_wait(q_full, 0, 20);
      auto issue_s = [&](int j) {
        const int st = j % KST;
        mbar_wait(&k_full[st], (j / KST) & 1, 21);
        tc_fence_after();
        const uint64_t dk = umma_desc_sw128(smem_u32(sK + st * kTile), 1024);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16_ss(tS, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
        umma_commit(&k_empty[st]);
        umma_commit(s_full);
      };
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) {
          mbar_wait(s_free, j & 1, 24);  // S_j is in registers: S_{j+1} overlaps softmax j
          issue_s(j + 1);
        }
        const int st = j % KST;
        mbar_wait(p_full, j & 1, 22);
        mbar_wait(&v_full[st], (j / KST) & 1, 23);
        tc_fence_after();
        const uint64_t dv = umma_desc_sw128(smem_u32(sV + st * kTile), 1024, 8192);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          // A: 16 keys of P = 8 TMEM columns; B: 16 key rows of V = 16 x 128 B further on
          umma_bf16_ts(tO, tP + 8 * k, dv + uint64_t((k * 2048) >> 4), idesc_o, (j | k) != 0);
        }
        umma_commit(&v_empty[st]);
        umma_commit(o_full);
      }
    }
  } else {
    // ================= softmax + output (warps 0..3) =================
    // Instruction budget per score (the SM issues 128 lanes/clk; the MMAs need ~512 clk per
    // 128x128 block): 0.5 FMNMX3 + 0.5 FFMA2 + 1 MUFU.EX2 + 0.5 FADD2 + 0.5 F2FP.
    const int row = warp * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_off = uint32_t(warp * 32) << 16;
    float m = -INFINITY;               // running (possibly stale) row max, log2 domain
    float2 l2 = make_float2(0.f, 0.f);
    const float2 sc2 = make_float2(a.scale_log2, a.scale_log2);

    for (int j = 0; j < nblk; ++j) {
      const int kv0 = j * BKV;
      const bool tail = kv0 + BKV > a.Sk;
      mbar_wait(s_full, j & 1, 30);
      tc_fence_after();
      uint32_t r[BKV];
#pragma unroll
      for (int c = 0; c < BKV / 32; ++c) tmem_ld32(tS + lane_off + c * 32, r + c * 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_free);             // the tensor pipe may overwrite S now
      if (tail) {
#pragma unroll
        for (int i = 0; i < BKV; ++i)
          if (kv0 + i >= a.Sk) r[i] = 0xff800000u;  // -inf: excluded from max, exp2 -> 0
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < BKV; i += 2)
        mx = fmaxf(mx, fmaxf(__uint_as_float(r[i]), __uint_as_float(r[i + 1])));
      // lazy rescale: keep the stale max unless the new one is > 2^8 larger
      const float m_cand = fmaxf(m, mx * a.scale_log2);
      const bool grow = m_cand - m > kRescaleThreshold;   // true on the first block (m = -inf)
      const float m_use = grow ? m_cand : m;
      const float alpha = grow ? ex2(m - m_use) : 1.f;
      m = m_use;
      l2 = __fmul2_rn(l2, make_float2(alpha, alpha));
      const float2 nm2 = make_float2(-m_use, -m_use);
#pragma unroll
      for (int i = 0; i < BKV / 2; ++i) {
        const float2 sv = __ffma2_rn(
            make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), sc2, nm2);
        const float2 pv = make_float2(ex2(sv.x), ex2(sv.y));
        l2 = __fadd2_rn(l2, pv);
        r[i] = pack_bf16(pv.x, pv.y);   // in place: word i <- keys 2i, 2i+1 (already consumed)
      }
      if (j > 0) {
        // P V of the previous block must have finished reading P (and writing O)
        mbar_wait(o_full, (j - 1) & 1, 31);
        tc_fence_after();
        if (__any_sync(0xffffffffu, grow)) {
          // O *= alpha for this warp's 32 rows (alpha = 1 for rows that did not grow)
          const float2 al2 = make_float2(alpha, alpha);
#pragma unroll
          for (int c = 0; c < HD / 32; ++c) {
            uint32_t o[32];
            tmem_ld32(tO + lane_off + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float2 t = __fmul2_rn(
                  make_float2(__uint_as_float(o[2 * i]), __uint_as_float(o[2 * i + 1])), al2);
              o[2 * i] = __float_as_uint(t.x);
              o[2 * i + 1] = __float_as_uint(t.y);
            }
            tmem_st32(tO + lane_off + c * 32, o);
          }
        }
      }
      tmem_st32(tP + lane_off, r);
      tmem_st32(tP + lane_off + 32, r + 32);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    // ---- normalise, stage O through smem (the Q tile is dead), TMA store ----
    mbar_wait(o_full, (nblk - 1) & 1, 32);
    tc_fence_after();
    const float l = l2.x + l2.y;
    const float inv = 1.f / l;
    uint8_t* orow = sQ + row * 128;
    const int rx = row & 7;
#pragma unroll
    for (int c = 0; c < HD / 32; ++c) {
      uint32_t o[32];
      tmem_ld32(tO + lane_off + c * 32, o);
      tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 w;
        w.x = pack_bf16(__uint_as_float(o[8 * q]) * inv, __uint_as_float(o[8 * q + 1]) * inv);
        w.y = pack_bf16(__uint_as_float(o[8 * q + 2]) * inv, __uint_as_float(o[8 * q + 3]) * inv);
        w.z = pack_bf16(__uint_as_float(o[8 * q + 4]) * inv, __uint_as_float(o[8 * q + 5]) * inv);
        w.w = pack_bf16(__uint_as_float(o[8 * q + 6]) * inv, __uint_as_float(o[8 * q + 7]) * inv);
        *reinterpret_cast<uint4*>(orow + (((c * 4 + q) ^ rx) * 16)) = w;
      }
    }
    if (a.lse != nullptr && q0 + row < a.Sq)
      a.lse[((int64_t)b * a.heads + head) * a.Sq + q0 + row] =
          (m + log2f(l)) * 0.6931471805599453f;
    fence_proxy_async_smem();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (threadIdx.x == 0) {
      tma_store_4d(&tmO, sQ, head * HD, q0, b, 0);
      tma_store_commit();
      tma_store_wait<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

int make_qkv_map(CUtensorMap* m, const void* p, int64_t B, int64_t S, int width, int64_t ld,
                 int rank4) {
  const uint64_t dims[4] = {(uint64_t)width, (uint64_t)S, (uint64_t)B, 1};
  const uint64_t strides[4] = {2, (uint64_t)ld * 2, (uint64_t)S * ld * 2,
                               (uint64_t)B * S * ld * 2};
  const uint32_t box[4] = {HD, 128, 1, 1};
  return vst_make_tmap(m, p, rank4 ? 4 : 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace

extern "C" int vst_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse,
                            int Bt, int heads, int Sq, int Sk, int d, int64_t ldq, int64_t ldk,
                            int64_t ldv, int64_t ldo, int kv_group, float scale, void* stream) {
  VST_REQUIRE(Bt >= 0 && heads > 0 && Sq >= 0 && Sk > 0, VST_E_BADARG, "attn_fwd: bad shape");
  if (Bt == 0 || Sq == 0) return VST_OK;
  VST_REQUIRE(q && k && v && o, VST_E_BADARG, "attn_fwd: null pointer");
  VST_REQUIRE(d == HD, VST_E_UNSUPPORTED, "attn_fwd: head dim %d unsupported (SDXL spatial d=64)", d);
  VST_REQUIRE(kv_group >= 1 && Bt % kv_group == 0, VST_E_BADARG, "attn_fwd: Bt %% kv_group != 0");
  VST_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0, VST_E_UNSUPPORTED,
              "attn_fwd: row strides must be multiples of 8 elements");
  VST_REQUIRE(heads <= 65535 && Bt <= 65535, VST_E_UNSUPPORTED, "attn_fwd: grid too large");
  CUtensorMap tQ, tK, tV, tO;
  int rc;
  const int width = heads * HD;
  if ((rc = make_qkv_map(&tQ, q, Bt, Sq, width, ldq, 0))) return rc;
  if ((rc = make_qkv_map(&tK, k, Bt / kv_group, Sk, width, ldk, 0))) return rc;
  if ((rc = make_qkv_map(&tV, v, Bt / kv_group, Sk, width, ldv, 0))) return rc;
  if ((rc = make_qkv_map(&tO, o, Bt, Sq, width, ldo, 1))) return rc;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kSmemBytes));
    configured = true;
  }
  AttnArgs a{lse, Sq, Sk, heads, kv_group, scale * 1.4426950408889634f};
  dim3 grid((Sq + BQ - 1) / BQ, heads, Bt);
  attn_fwd_kernel<<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(tQ, tK, tV, tO, a);
  VST_LAUNCH_OK();
  return VST_OK;
}
