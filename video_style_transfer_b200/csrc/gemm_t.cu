// Backward-pass GEMMs of the LoRA projections (a3/a4/a5/a7 gradients) on tcgen05: a GEMM whose
// operands may be given transposed, with deterministic split-K.
//
//   D[M,N] (+)= alpha * ( opA(A1) opB(B1)^T + opA(A2) opB(B2)^T )
//     opA(A) is [M,K]:  transA = 0 -> A stored [M,K] (row stride lda);  1 -> stored [K,M]
//     opB(B) is [N,K]:  transB = 0 -> B stored [N,K] (row stride ldb);  1 -> stored [K,N]
//
// What the reference leaves to autograd (unziplora_unet/unziplora_linear_layer.py:221-263,
// animatediff/temporal_lora.py:29-32, lora_linear.py:74-81) becomes
//   dx  = dy W + du [A_c; A_s]            (transB = 1, two operand pairs in one accumulator)
//   du  = dy B'                           (transB = 1)
//   dB' = dy^T u,  dA = du^T x,  dW = dy^T x   (transA = transB = 1: the contraction runs over the
//                                          token axis, i.e. over the rows of both operands)
// A transposed operand is an MN-major tcgen05 operand: TMA brings [64 k-rows x 64 mn-columns]
// boxes (128-byte swizzle) and the shared-memory descriptor walks them with LBO = one box; no
// transposed copy of an activation is ever made.
// The token-axis contractions have tiny outputs (N x 2r) and K = 16k..262k tokens: the k-blocks of
// a tile are split over CTAs, fp32 partials go to a caller-provided workspace and are summed in a
// fixed order by gemm_reduce_kernel (deterministic, no atomics).
//
// Kernel shape: persistent, one CTA per SM, 192 threads: warp 0 TMA producer, warp 1 MMA issuer
// (whole warp in the loop, one elected lane issues), warps 2..5 epilogue (TMEM -> registers ->
// global).  Roofline: tensor pipe for dx / dW; HBM (one pass over dy and u / x) for the rank-sized
// products.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 192;
constexpr int kSmemBudget = 200 * 1024;

template <int BN>
struct TCfg {
  static constexpr int kStageBytes = (BM + BN) * BK * 2;
  static constexpr int kStages = (kSmemBudget / kStageBytes) > 8 ? 8 : (kSmemBudget / kStageBytes);
  static constexpr int kAccStride = BN <= 64 ? 64 : (BN <= 128 ? 128 : 256);
  static constexpr int kTmemCols = 2 * kAccStride;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

struct TArgs {
  void* D;            // final output (splits == 1) ...
  float* partial;     // ... or fp32 partials [splits][M][N]
  int64_t ldd, M;
  int N;
  int bn;             // pair kernel: tile width actually used (<= 256, multiple of 32): N = 320 runs as 2 x 160
  int nk1, nk2;       // k-blocks of the two operand pairs
  int m_tiles, n_tiles, splits, kb_per_split;
  int transA, transB;
  float alpha;
  int accumulate, out_f32;
};

// One accumulator row chunk (32 fp32 columns starting at `col`) to its destination: the fp32 split-K partials,
// or D in fp32 / bf16 with alpha and the optional accumulate.
__device__ __forceinline__ void store_row_chunk(const TArgs& a, const uint32_t* r, int64_t row, int col, int split) {
  const bool empty_split = split * a.kb_per_split >= a.nk1 + a.nk2;   // cannot happen by construction
  if (a.splits > 1) {
    float* dst = a.partial + ((int64_t)split * a.M + row) * a.N + col;
#pragma unroll
    for (int v = 0; v < 8; ++v) {
      if (col + 4 * v >= a.N) break;
      *reinterpret_cast<float4*>(dst + 4 * v) =
          empty_split ? make_float4(0.f, 0.f, 0.f, 0.f)
                      : make_float4(__uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]),
                                    __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3]));
    }
  } else if (a.out_f32) {
    float* dst = reinterpret_cast<float*>(a.D) + row * a.ldd + col;
#pragma unroll
    for (int v = 0; v < 8; ++v) {
      if (col + 4 * v >= a.N) break;
      float4 o = make_float4(__uint_as_float(r[4 * v]) * a.alpha, __uint_as_float(r[4 * v + 1]) * a.alpha,
                             __uint_as_float(r[4 * v + 2]) * a.alpha, __uint_as_float(r[4 * v + 3]) * a.alpha);
      if (a.accumulate) {
        const float4 p = *reinterpret_cast<const float4*>(dst + 4 * v);
        o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
      }
      *reinterpret_cast<float4*>(dst + 4 * v) = o;
    }
  } else {
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(a.D) + row * a.ldd + col;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      if (col + 8 * v >= a.N) break;
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = __uint_as_float(r[8 * v + j]) * a.alpha;
      if (a.accumulate) {
        const uint4 p = *reinterpret_cast<const uint4*>(dst + 8 * v);
        const uint32_t pw[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 e = unpack_bf16(pw[j]);
          o[2 * j] += e.x;
          o[2 * j + 1] += e.y;
        }
      }
      uint4 wv;
      wv.x = pack_bf16(o[0], o[1]); wv.y = pack_bf16(o[2], o[3]);
      wv.z = pack_bf16(o[4], o[5]); wv.w = pack_bf16(o[6], o[7]);
      *reinterpret_cast<uint4*>(dst + 8 * v) = wv;
    }
  }
}

template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_t_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
              const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
              TArgs a) {
  using C = TCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* empty = full + C::kStages;
  uint64_t* acc_full = empty + C::kStages;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = a.nk1 + a.nk2;
  const int total = a.m_tiles * a.n_tiles * a.splits;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, C::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const int tile = w / a.splits, split = w % a.splits;
        const int m0 = (tile / a.n_tiles) * BM, n0 = (tile % a.n_tiles) * BN;
        const int kb0 = split * a.kb_per_split;
        const int kb1 = min(nk, kb0 + a.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + BM * BK * 2;
          mbar_arrive_expect_tx(&full[stage], C::kStageBytes);
          const bool second = kb >= a.nk1;
          const CUtensorMap* mA = second ? &tmA2 : &tmA1;
          const CUtensorMap* mB = second ? &tmB2 : &tmB1;
          const int k0 = (second ? kb - a.nk1 : kb) * BK;
          if (a.transA) {
#pragma unroll
            for (int blk = 0; blk < BM / 64; ++blk)
              tma_load_2d(sa + blk * 8192, mA, &full[stage], m0 + 64 * blk, k0);
          } else {
            tma_load_2d(sa, mA, &full[stage], k0, m0);
          }
          if (a.transB) {
#pragma unroll
            for (int blk = 0; blk < BN / 64; ++blk)
              tma_load_2d(sb + blk * 8192, mB, &full[stage], n0 + 64 * blk, k0);
          } else {
            tma_load_2d(sb, mB, &full[stage], k0, n0);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (warp-uniform loop, elected lane issues) =================
    const bool issuer = elect_one();
    const uint32_t idesc = umma_idesc_bf16(BM, BN, a.transA, a.transB);
    // K-major: 16 k = 32 B inside the swizzle row; MN-major: 16 k-rows of 128 B
    const uint64_t stepA = a.transA ? (2048 >> 4) : 2, stepB = a.transB ? (2048 >> 4) : 2;
    const uint32_t lboA = a.transA ? 8192 : 0, lboB = a.transB ? 8192 : 0;
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const int split = w % a.splits;
      const int kb0 = split * a.kb_per_split;
      const int kb1 = min(nk, kb0 + a.kb_per_split);
      mbar_wait(&acc_empty[acc], acc_phase ^ 1, 2);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * C::kAccStride;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full[stage], phase, 3);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
        const uint32_t sb = sa + BM * BK * 2;
        const uint64_t da = umma_desc_sw128(sa, 1024, lboA);
        const uint64_t db = umma_desc_sw128(sb, 1024, lboB);
        if (issuer) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16_ss(d_tmem, da + stepA * k, db + stepB * k, idesc, (kb > kb0) || (k != 0));
          umma_commit(&empty[stage]);
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      if (issuer) umma_commit(&acc_full[acc]);
      __syncwarp();
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ================= epilogue (warps 2..5): thread = accumulator row =================
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const int tile = w / a.splits, split = w % a.splits;
      const int m0 = (tile / a.n_tiles) * BM, n0 = (tile % a.n_tiles) * BN;
      const int64_t row = (int64_t)m0 + quarter * 32 + lane;
      const bool row_ok = row < a.M;
      mbar_wait(&acc_full[acc], acc_phase, 4);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * C::kAccStride + (uint32_t(quarter * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + c0, r);
        tmem_ld_wait();
        if (!row_ok || n0 + c0 >= a.N) continue;
        store_row_chunk(a, r, row, n0 + c0, split);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// CTA-pair form (cta_group::2) of the kernel above for N > 128 and M > 128: a 256 x 256 tile over two
// SMs, each CTA bringing its own 128 rows of A and 128 of the 256 columns of B, the leader issuing one
// M = 256 tcgen05.mma.  The single-CTA 128 x 256 tile is bound by the shared-memory feed (49 KB per
// k-block for 4.2 MFLOP; narrower tiles measured no faster, scripts/gpu_r02r.sh), which is what held dx /
// dW 20-35 % behind the library; the pair moves 64 KB per 8.4 MFLOP.  Barrier protocol as
// lora_gemm2_kernel (gemm.cu): full[s] lives in the leader and counts both CTAs' TMA bytes, empty[s] and
// acc_full[a] are released in both CTAs by multicast commits, acc_empty[a] lives in the leader and takes
// the 4 + 4 epilogue-warp arrivals.  Operand transposes, split-K and the epilogue are the single-CTA ones
// (a CTA drains its own 128 TMEM lanes).
// ------------------------------------------------------------------------------------------------
struct TCfg2 {
  static constexpr int BN = 256;
  static constexpr int kBHalf = BN / 2;
  static constexpr int kStageBytes = (BM + kBHalf) * BK * 2;
  static constexpr int kStages = (kSmemBudget / kStageBytes) > 8 ? 8 : (kSmemBudget / kStageBytes);
  static constexpr int kAccStride = 256;
  static constexpr int kTmemCols = 2 * kAccStride;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_t2_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
               const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
               TArgs a) {
  using C = TCfg2;
  constexpr int BN = C::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* empty = full + C::kStages;
  uint64_t* acc_full = empty + C::kStages;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int nk = a.nk1 + a.nk2;
  const int total = a.m_tiles * a.n_tiles * a.splits;      // m_tiles counts 256-row tiles
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 8);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_2sm(tmem_slot, C::kTmemCols);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer (both CTAs) =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = cluster_id; w < total; w += n_clusters) {
        const int tile = w / a.splits, split = w % a.splits;
        const int m0 = (tile / a.n_tiles) * (2 * BM) + (int)rank * BM;
        const int half = a.bn >> 1;                                         // this CTA's half of the B tile
        const int nb = (tile % a.n_tiles) * a.bn + (int)rank * half;
        // an MN-major half arrives as whole 64-column boxes (the last may reach past it: never read)
        const int b_boxes = (half + 63) >> 6;
        const uint32_t cta_tx = BM * BK * 2 + (a.transB ? b_boxes * 8192 : half * BK * 2);
        const int kb0 = split * a.kb_per_split;
        const int kb1 = min(nk, kb0 + a.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 11);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + BM * BK * 2;
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * cta_tx);
          const bool second = kb >= a.nk1;
          const CUtensorMap* mA = second ? &tmA2 : &tmA1;
          const CUtensorMap* mB = second ? &tmB2 : &tmB1;
          const int k0 = (second ? kb - a.nk1 : kb) * BK;
          if (a.transA) {
#pragma unroll
            for (int blk = 0; blk < BM / 64; ++blk)
              tma_load_2d_2sm(sa + blk * 8192, mA, &full[stage], m0 + 64 * blk, k0);
          } else {
            tma_load_2d_2sm(sa, mA, &full[stage], k0, m0);
          }
          if (a.transB) {
#pragma unroll
            for (int blk = 0; blk < C::kBHalf / 64; ++blk)
              if (blk < b_boxes) tma_load_2d_2sm(sb + blk * 8192, mB, &full[stage], nb + 64 * blk, k0);
          } else {
            tma_load_2d_2sm(sb, mB, &full[stage], k0, nb);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader CTA, one lane) =================
    if (leader && lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(2 * BM, a.bn, a.transA, a.transB);
      const uint64_t stepA = a.transA ? (2048 >> 4) : 2, stepB = a.transB ? (2048 >> 4) : 2;
      const uint32_t lboA = a.transA ? 8192 : 0, lboB = a.transB ? 8192 : 0;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int w = cluster_id; w < total; w += n_clusters) {
        const int split = w % a.splits;
        const int kb0 = split * a.kb_per_split;
        const int kb1 = min(nk, kb0 + a.kb_per_split);
        mbar_wait(&acc_empty[acc], acc_phase ^ 1, 12);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * C::kAccStride;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase, 13);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t sb = sa + BM * BK * 2;
          const uint64_t da = umma_desc_sw128(sa, 1024, lboA);
          const uint64_t db = umma_desc_sw128(sb, 1024, lboB);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16_ss_2sm(d_tmem, da + stepA * k, db + stepB * k, idesc, (kb > kb0) || (k != 0));
          umma_commit_2sm(&empty[stage]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&acc_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ================= epilogue (warps 2..5 of both CTAs; own 128 rows): thread = accumulator row =================
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int w = cluster_id; w < total; w += n_clusters) {
      const int tile = w / a.splits, split = w % a.splits;
      const int m0 = (tile / a.n_tiles) * (2 * BM) + (int)rank * BM, n0 = (tile % a.n_tiles) * a.bn;
      const int64_t row = (int64_t)m0 + quarter * 32 + lane;
      const bool row_ok = row < a.M;
      mbar_wait(&acc_full[acc], acc_phase, 14);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * C::kAccStride + (uint32_t(quarter * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < a.bn; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + c0, r);
        tmem_ld_wait();
        if (!row_ok || n0 + c0 >= a.N) continue;
        store_row_chunk(a, r, row, n0 + c0, split);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&acc_empty[acc], 0);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, C::kTmemCols);
  }
}

// D = alpha * sum_s partial[s] (+ D), fixed summation order.
__global__ void gemm_reduce_kernel(const float* __restrict__ partial, void* __restrict__ D,
                                   int64_t ldd, int64_t M, int N, int splits, float alpha,
                                   int accumulate, int out_f32) {
  const int64_t idx = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (idx >= M * N) return;
  const int64_t row = idx / N;
  const int col = (int)(idx % N);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int k = 0; k < splits; ++k) {
    const float4 p = *reinterpret_cast<const float4*>(partial + (int64_t)k * M * N + idx);
    s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w;
  }
  s.x *= alpha; s.y *= alpha; s.z *= alpha; s.w *= alpha;
  if (out_f32) {
    float* dst = reinterpret_cast<float*>(D) + row * ldd + col;
    if (accumulate) {
      const float4 p = *reinterpret_cast<const float4*>(dst);
      s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w;
    }
    *reinterpret_cast<float4*>(dst) = s;
  } else {
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(D) + row * ldd + col;
    if (accumulate) {
      const uint2 p = *reinterpret_cast<const uint2*>(dst);
      const float2 e0 = unpack_bf16(p.x), e1 = unpack_bf16(p.y);
      s.x += e0.x; s.y += e0.y; s.z += e1.x; s.w += e1.y;
    }
    *reinterpret_cast<uint2*>(dst) = make_uint2(pack_bf16(s.x, s.y), pack_bf16(s.z, s.w));
  }
}

int make_map(CUtensorMap* m, const void* p, int64_t rows, int64_t K, int64_t ld, int trans,
             int box_rows) {
  if (trans) {   // stored [K][rows]: innermost = the MN axis
    const uint64_t dims[2] = {(uint64_t)rows, (uint64_t)K};
    const uint64_t strides[2] = {2, (uint64_t)ld * 2};
    const uint32_t box[2] = {64, BK};
    return vst_make_tmap(m, p, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
  }
  const uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  const uint64_t strides[2] = {2, (uint64_t)ld * 2};
  const uint32_t box[2] = {BK, (uint32_t)box_rows};
  return vst_make_tmap(m, p, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int pick_bn(int N) { return N <= 64 ? 64 : (N <= 128 ? 128 : 256); }

// VST_GEMM_T_PAIR=0: single-CTA kernels only (A/B switch)
int pair_mode() {
  static const int mode = [] {
    const char* e = getenv("VST_GEMM_T_PAIR");
    return e ? atoi(e) : 1;
  }();
  return mode;
}

// Tiling and k-split of one product: CTA pairs (256-row tiles) when the output is wider than 128 columns and
// taller than one CTA's 128 rows; k is split only when the tiles alone leave more than half of the scheduling
// units (SMs, or SM pairs) idle.
struct Plan {
  int pair, bn, m_tiles, n_tiles, splits, kb_per;
};
Plan pick_plan(int64_t M, int N, int nk) {
  Plan p;
  int bn = pick_bn(N);
  p.pair = pair_mode() && bn == 256 && M > BM;
  if (p.pair) {
    // the narrowest multiple of 32 that covers N with the same number of column tiles (N = 320: 2 x 160 instead
    // of 256 + 64, N = 640: 3 x 224): the pair kernel is MMA-bound, columns past N would be paid in full
    const int tiles = (N + 255) / 256;
    const int w = ((N + tiles - 1) / tiles + 31) / 32 * 32;
    bn = w < 256 ? w : 256;
  }
  p.bn = bn;
  const int bm = p.pair ? 2 * BM : BM;
  p.m_tiles = (int)((M + bm - 1) / bm);
  p.n_tiles = (N + bn - 1) / bn;
  const int64_t tiles = (int64_t)p.m_tiles * p.n_tiles;
  const int units = p.pair ? vst_sm_count() / 2 : vst_sm_count();
  int s = 1;
  if (tiles * 2 <= units && nk >= 8) {
    s = (int)(units / tiles);
    if (s > nk / 4) s = nk / 4;
    if (s < 1) s = 1;
  }
  p.kb_per = (nk + s - 1) / s;
  p.splits = (nk + p.kb_per - 1) / p.kb_per;
  return p;
}

template <int BN>
int launch_t(const CUtensorMap& tA1, const CUtensorMap& tB1, const CUtensorMap& tA2,
             const CUtensorMap& tB2, const TArgs& a, cudaStream_t st) {
  using C = TCfg<BN>;
  auto kern = gemm_t_kernel<BN>;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    configured = true;
  }
  const int64_t total = (int64_t)a.m_tiles * a.n_tiles * a.splits;
  const int grid = (int)(total < vst_sm_count() ? total : vst_sm_count());
  kern<<<grid, kThreads, C::kSmemBytes, st>>>(tA1, tB1, tA2, tB2, a);
  VST_LAUNCH_OK();
  return VST_OK;
}

int launch_t2(const CUtensorMap& tA1, const CUtensorMap& tB1, const CUtensorMap& tA2,
              const CUtensorMap& tB2, const TArgs& a, cudaStream_t st) {
  using C = TCfg2;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(gemm_t2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    configured = true;
  }
  const int64_t total = (int64_t)a.m_tiles * a.n_tiles * a.splits;
  const int pairs = vst_sm_count() / 2;
  const int grid = 2 * (int)(total < pairs ? total : pairs);
  gemm_t2_kernel<<<grid, kThreads, C::kSmemBytes, st>>>(tA1, tB1, tA2, tB2, a);
  VST_LAUNCH_OK();
  return VST_OK;
}

}  // namespace

extern "C" int64_t vst_gemm_workspace_bytes(int64_t M, int N, int64_t K1, int64_t K2) {
  if (M <= 0 || N <= 0 || K1 <= 0) return 0;
  const int nk = (int)((K1 + BK - 1) / BK + (K2 + BK - 1) / BK);
  const Plan p = pick_plan(M, N, nk);
  return p.splits > 1 ? (int64_t)p.splits * M * N * 4 : 0;
}

extern "C" int vst_gemm(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int64_t K1,
                        const void* A2, int64_t lda2, const void* B2, int64_t ldb2, int64_t K2,
                        int transA, int transB, void* D, int64_t ldd, int64_t M, int N, float alpha,
                        int accumulate, int out_dtype, void* workspace, int64_t workspace_bytes,
                        void* stream) {
  VST_REQUIRE(M >= 0 && N > 0 && K1 > 0 && K2 >= 0, VST_E_BADARG, "gemm: bad shape");
  if (M == 0) return VST_OK;
  VST_REQUIRE(A1 && B1 && D && (K2 == 0 || (A2 && B2)), VST_E_BADARG, "gemm: null pointer");
  VST_REQUIRE(N % 8 == 0, VST_E_UNSUPPORTED, "gemm: N=%d must be a multiple of 8", N);
  VST_REQUIRE(lda1 % 8 == 0 && ldb1 % 8 == 0 && (K2 == 0 || (lda2 % 8 == 0 && ldb2 % 8 == 0)),
              VST_E_UNSUPPORTED, "gemm: operand row strides must be multiples of 8 elements");
  VST_REQUIRE(out_dtype == VST_F32 || out_dtype == VST_BF16, VST_E_BADARG, "gemm: out dtype");
  VST_REQUIRE(ldd % (out_dtype == VST_F32 ? 4 : 8) == 0 && (reinterpret_cast<uintptr_t>(D) & 15) == 0,
              VST_E_UNSUPPORTED, "gemm: D must keep 16-byte row alignment");
  VST_REQUIRE(transA == 0 || transA == 1, VST_E_BADARG, "gemm: transA");
  VST_REQUIRE(transB == 0 || transB == 1, VST_E_BADARG, "gemm: transB");
  // a K-major operand needs K % 8 (16-byte rows); an MN-major one needs its MN extent % 8
  VST_REQUIRE(transA ? (M % 8 == 0) : (K1 % 8 == 0 && K2 % 8 == 0), VST_E_UNSUPPORTED,
              "gemm: A extent along its contiguous axis must be a multiple of 8");
  VST_REQUIRE(transB ? true : (K1 % 8 == 0 && K2 % 8 == 0), VST_E_UNSUPPORTED,
              "gemm: B extent along its contiguous axis must be a multiple of 8");
  const int bn = pick_bn(N);
  const int nk_all = (int)((K1 + BK - 1) / BK + (K2 + BK - 1) / BK);
  Plan plan = pick_plan(M, N, nk_all);
  const int b_box = plan.pair ? plan.bn / 2 : bn;        // rows of a K-major B box: a pair CTA loads half the tile
  CUtensorMap tA1, tB1, tA2, tB2;
  int rc;
  if ((rc = make_map(&tA1, A1, M, K1, lda1, transA, BM))) return rc;
  if ((rc = make_map(&tB1, B1, N, K1, ldb1, transB, b_box))) return rc;
  if (K2 > 0) {
    if ((rc = make_map(&tA2, A2, M, K2, lda2, transA, BM))) return rc;
    if ((rc = make_map(&tB2, B2, N, K2, ldb2, transB, b_box))) return rc;
  } else {
    tA2 = tA1;
    tB2 = tB1;
  }
  TArgs a;
  memset(&a, 0, sizeof(a));
  a.D = D;
  a.ldd = ldd;
  a.M = M;
  a.N = N;
  a.nk1 = (int)((K1 + BK - 1) / BK);
  a.nk2 = (int)((K2 + BK - 1) / BK);
  a.bn = plan.bn;
  a.m_tiles = plan.m_tiles;
  a.n_tiles = plan.n_tiles;
  a.splits = plan.splits;
  a.kb_per_split = plan.kb_per;
  if (a.splits > 1 && (workspace == nullptr || workspace_bytes < (int64_t)a.splits * M * N * 4)) {
    a.splits = 1;                      // no (or too small a) workspace: one CTA walks all of K
    a.kb_per_split = a.nk1 + a.nk2;
  }
  a.partial = reinterpret_cast<float*>(workspace);
  a.transA = transA;
  a.transB = transB;
  a.alpha = alpha;
  a.accumulate = accumulate;
  a.out_f32 = out_dtype == VST_F32;
  cudaStream_t st = (cudaStream_t)stream;
  if (plan.pair) rc = launch_t2(tA1, tB1, tA2, tB2, a, st);
  else if (bn == 64) rc = launch_t<64>(tA1, tB1, tA2, tB2, a, st);
  else if (bn == 128) rc = launch_t<128>(tA1, tB1, tA2, tB2, a, st);
  else rc = launch_t<256>(tA1, tB1, tA2, tB2, a, st);
  if (rc) return rc;
  if (a.splits > 1) {
    const int64_t quads = (M * N + 3) / 4;
    gemm_reduce_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, st>>>(
        a.partial, D, ldd, M, N, a.splits, alpha, accumulate, a.out_f32);
    VST_LAUNCH_OK();
  }
  return VST_OK;
}
