// a2: spatial self-attention / text cross-attention forward, head_dim 64, on tcgen05 + TMEM + TMA.
// Reference: animatediff/attention_processor.py:74-83 (head split, F.scaled_dot_product_attention,
// head merge) -- the head split/merge transposes disappear into the TMA tensor maps, and the
// text repeat_interleave of :62-66 into the kv batch index (b / kv_group).
//
// One CTA = 128 query rows of one (batch, head); two CTAs are resident per SM so one CTA's
// softmax overlaps the other's MMAs.  192 threads:
//   warps 0..3  softmax: thread t owns query row t (TMEM lane t).  Per 128-key block: two passes
//               over S in TMEM (row max; exp2 + row sum + bf16 P -> 128B-swizzled smem), then
//               O_reg = O_reg * alpha + (P V) read back from TMEM.
//   warp 4      TMA producer: Q once, K (2 stages) and V (1 stage) tiles [128 x 64] per block.
//   warp 5      MMA issuer: S = Q K^T (4 x tcgen05.mma 128x128x16, K-major operands) and
//               O_blk = P V (8 x 128x64x16, V consumed MN-major straight from its row-major tile).
// Roofline: tensor pipe for self-attention (S/2 FLOP/B); FLOPs = 4*Bt*h*Sq*Sk*d.
#include "common.cuh"

namespace {

constexpr int BQ = 128;   // query rows per CTA
constexpr int BKV = 128;  // keys per block
constexpr int HD = 64;    // head dim
constexpr int kThreads = 192;
constexpr int KST = 2;    // K stages
constexpr int kTile = 128 * HD * 2;  // 16 KB
constexpr int kSmemBytes = kTile * (1 + KST + 1) + BQ * BKV * 2 + 1024 + 128;

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
  uint8_t* sP = sV + kTile;  // two 64-key chunks of [128 rows x 128 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + BQ * BKV * 2);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;          // [KST]
  uint64_t* k_empty = k_full + KST;     // [KST]
  uint64_t* v_full = k_empty + KST;
  uint64_t* v_empty = v_full + 1;
  uint64_t* s_full = v_empty + 1;
  uint64_t* p_full = s_full + 1;
  uint64_t* o_full = p_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, head = blockIdx.y, b = blockIdx.z;
  const int nblk = (a.Sk + BKV - 1) / BKV;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < KST; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    mbar_init(v_full, 1);
    mbar_init(v_empty, 1);
    mbar_init(s_full, 1);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base;        // columns [0,128): S
  const uint32_t tO = tmem_base + 128;  // columns [128,192): O_blk

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
        const int ks = j % KST;
        mbar_wait(&k_empty[ks], ((j / KST) & 1) ^ 1, 10);
        mbar_arrive_expect_tx(&k_full[ks], kTile);
        tma_load_3d(sK + ks * kTile, &tmK, &k_full[ks], head * HD, j * BKV, bk);
        mbar_wait(v_empty, (j & 1) ^ 1, 11);
        mbar_arrive_expect_tx(v_full, kTile);
        tma_load_3d(sV, &tmV, v_full, head * HD, j * BKV, bk);
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, 0, 0);
      constexpr uint32_t idesc_o = umma_idesc_bf16(BQ, HD, 0, 1);  // B (= V) is MN-major
      const uint64_t dq = umma_desc_sw128(smem_u32(sQ), 1024);
      const uint64_t dp0 = umma_desc_sw128(smem_u32(sP), 1024);
      const uint64_t dp1 = umma_desc_sw128(smem_u32(sP) + BQ * 128, 1024);
      const uint64_t dv = umma_desc_sw128(smem_u32(sV), 1024, 8192);
      mbar_wait(q_full, 0, 20);
      auto issue_s = [&](int j) {
        const int ks = j % KST;
        mbar_wait(&k_full[ks], (j / KST) & 1, 21);
        tc_fence_after();
        const uint64_t dk = umma_desc_sw128(smem_u32(sK + ks * kTile), 1024);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16_ss(tS, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
        umma_commit(&k_empty[ks]);
        umma_commit(s_full);
      };
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        mbar_wait(p_full, j & 1, 22);  // P_j in smem; S and O_blk in TMEM are free again
        mbar_wait(v_full, j & 1, 23);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          const uint64_t dp = (k < 4 ? dp0 : dp1) + 2 * (k & 3);
          // V tile rows are keys: each 16-key k-step is 16 rows x 128 B = 2048 B further on
          umma_bf16_ss(tO, dp, dv + uint64_t((k * 2048) >> 4), idesc_o, k != 0);
        }
        umma_commit(v_empty);
        umma_commit(o_full);
        if (j + 1 < nblk) issue_s(j + 1);
      }
    }
  } else {
    // ================= softmax + output (warps 0..3) =================
    const int row = warp * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_off = uint32_t(warp * 32) << 16;
    float o[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f, alpha = 1.f;
    uint8_t* prow = sP + row * 128;
    const int rx = row & 7;
    for (int j = 0; j < nblk; ++j) {
      mbar_wait(s_full, j & 1, 30);
      tc_fence_after();
      const int kv0 = j * BKV;
      const bool tail = kv0 + BKV > a.Sk;
      // ---- pass 1: row max ----
      float mx = m;
#pragma unroll 1
      for (int c = 0; c < BKV / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(tS + lane_off + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float sv = __uint_as_float(r[i]) * a.scale_log2;
          if (tail && kv0 + c * 32 + i >= a.Sk) sv = -INFINITY;
          mx = fmaxf(mx, sv);
        }
      }
      const float alpha_new = ex2(m - mx);  // m = -inf on the first block -> 0
      // ---- fold in the previous block's P V (computed against the previous max) ----
      if (j > 0) {
        mbar_wait(o_full, (j - 1) & 1, 31);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(tO + lane_off + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[c * 32 + i] = o[c * 32 + i] * alpha + __uint_as_float(r[i]);
        }
      }
      alpha = alpha_new;
      l *= alpha_new;
      m = mx;
      // ---- pass 2: P = exp2(s - m), row sum, bf16 P -> swizzled smem ----
#pragma unroll 1
      for (int c = 0; c < BKV / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(tS + lane_off + c * 32, r);
        tmem_ld_wait();
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float s0 = __uint_as_float(r[2 * i]) * a.scale_log2 - m;
          float s1 = __uint_as_float(r[2 * i + 1]) * a.scale_log2 - m;
          float p0 = ex2(s0), p1 = ex2(s1);
          if (tail) {
            if (kv0 + c * 32 + 2 * i >= a.Sk) p0 = 0.f;
            if (kv0 + c * 32 + 2 * i + 1 >= a.Sk) p1 = 0.f;
          }
          l += p0 + p1;
          w[i] = pack_bf16(p0, p1);
        }
        // 32 keys = 4 x 16-byte chunks; chunk index within the 64-key half: (c & 1) * 4 + q
        uint8_t* half = prow + (c >> 1) * (BQ * 128);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int ch = ((c & 1) * 4 + q) ^ rx;
          *reinterpret_cast<uint4*>(half + ch * 16) =
              make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(p_full);
    }
    // ---- last block's P V, normalise, stage O through smem, TMA store ----
    mbar_wait(o_full, (nblk - 1) & 1, 32);
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < HD / 32; ++c) {
      uint32_t r[32];
      tmem_ld32(tO + lane_off + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[c * 32 + i] = o[c * 32 + i] * alpha + __uint_as_float(r[i]);
    }
    const float inv = 1.f / l;
    uint8_t* orow = sQ + row * 128;  // every S MMA has completed: the Q tile is dead
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      uint4 w;
      w.x = pack_bf16(o[8 * q] * inv, o[8 * q + 1] * inv);
      w.y = pack_bf16(o[8 * q + 2] * inv, o[8 * q + 3] * inv);
      w.z = pack_bf16(o[8 * q + 4] * inv, o[8 * q + 5] * inv);
      w.w = pack_bf16(o[8 * q + 6] * inv, o[8 * q + 7] * inv);
      *reinterpret_cast<uint4*>(orow + ((q ^ rx) * 16)) = w;
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
