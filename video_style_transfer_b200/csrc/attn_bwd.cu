// a2 backward: gradients of softmax(Q K^T * scale) V for head dim 64 on tcgen05 + TMEM + TMA.
// Reference: autograd through F.scaled_dot_product_attention, animatediff/attention_processor.py
// :78-80 (train_animatediff.py back-propagates through every spatial attention to reach the
// motion modules and the mergers).
//
// Standard flash backward with the probabilities recomputed from the saved log-sum-exp:
//   D_i = sum_d dO_id O_id                                   (attn_bwd_dot_kernel)
//   P = exp(S * scale - LSE),  dP = dO V^T,  dS = P (.) (dP - D) * scale
//   dV = P^T dO,  dK = dS^T Q,  dQ = dS K
// One CTA owns a 128-key tile of one (batch, head) and walks the query tiles.  Five tcgen05 GEMMs
// per (query tile, key tile) pair, all operands straight from their row-major TMA tiles:
//   S  = Q K^T        A = Q   [q x d]  K-major      B = K  [k x d]  K-major
//   dP = dO V^T       A = dO  [q x d]  K-major      B = V  [k x d]  K-major
//   dV += P^T dO      A = P   [q x k] as MN-major   B = dO [q x d]  as MN-major   (contraction: q)
//   dK += dS^T Q      A = dS  [q x k] as MN-major   B = Q  [q x d]  as MN-major
//   dQ  = dS K        A = dS  [q x k]  K-major      B = K  [k x d]  as MN-major   (contraction: k)
// P and dS are written once to shared memory (bf16, 128-byte swizzle, [q rows][64-key blocks]); the
// same bytes serve as the MN-major A of dV / dK and the K-major A of dQ.  dV / dK accumulate in
// TMEM over the whole walk; dQ of each pair is added to an fp32 accumulator in global memory
// (vector atomics, [Bt, heads, Sq, 64] so a warp adds whole 128-byte lines) and converted to bf16
// afterwards.  TMEM: S [0,128) dP [128,256) dV [256,320) dK [320,384) dQ [384,448).
// (kSoftmaxWarps + 6) warps: softmax warps (warp w: TMEM lane quarter w%4, key slice w/4), then one TMA warp,
// one MMA-issue warp (elected lane), four dQ drain warps.
// Roofline: tensor pipe; FLOPs = 2.5 x forward = 10 * Bt*h*Sq*Sk*d.
#include <type_traits>

#include "common.cuh"

// -DVST_BWD_ABLATE=<bits>: timing-only builds that drop one stage (WRONG results; never shipped):
// 1 = no dQ atomics, 2 = no exponential / dS arithmetic, 8 = no dV / dK / dQ MMAs, 16 = no P / dS stores.
#ifndef VST_BWD_ABLATE
#define VST_BWD_ABLATE 0
#endif

// VST_BWD_WARP_ARRIVE: one elected mbarrier arrival per warp (after __syncwarp) instead of one per thread
#ifndef VST_BWD_ORDER
#define VST_BWD_ORDER 0
#endif
#ifndef VST_BWD_ELECT_WAIT
#define VST_BWD_ELECT_WAIT 0
#endif
#ifndef VST_BWD_WARP_ARRIVE
#define VST_BWD_WARP_ARRIVE 1
#endif

// -DVST_BWD_TRACE: CTA (0,0,0) records clock64() stamps of its pipeline events (scripts/attn_bwd_trace.py)
#ifdef VST_BWD_TRACE
__device__ long long g_bwd_trace[16 * 64];   // [slot][pair index]
#define BTRACE(slot, idx) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (idx) < 64 && \
    (threadIdx.x & 31) == 0) g_bwd_trace[(slot) * 64 + (idx)] = clock64(); } while (0)
#else
#define BTRACE(slot, idx) do { } while (0)
#endif

namespace {

#if VST_BWD_WARP_ARRIVE
constexpr int kArrivePerWarp = 1;
__device__ __forceinline__ void warp_arrive(uint64_t* bar) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
}
#else
constexpr int kArrivePerWarp = 32;
__device__ __forceinline__ void warp_arrive(uint64_t* bar) { mbar_arrive(bar); }
#endif

constexpr int BQ = 128, BKV = 128, HD = 64;
constexpr int kTile = 128 * HD * 2;     // 16 KB: one [128 x 64] bf16 tile
constexpr int kPBytes = 128 * 128 * 2;  // 32 KB: P or dS, two 64-key blocks of [128 q x 128 B]
// Softmax warps: 4 TMEM lane quarters x (kSoftmaxWarps / 4) key slices.  Eight: each thread handles 64 keys in two
// 32-key rounds, and the second round's TMEM load waits out the previous pair's gradient MMAs (r02 trace).  Sixteen
// (32 keys per thread, one load round; 704 threads at 80 registers) was measured: every stage slows down with 22
// warps on the SM (pair period 3990 against 3560 clk, S = 1024 shape 494 against 449 us), also with the gradient
// MMAs held back until the loads are done (VST_BWD_ORDER=1: 5070 clk).
#ifndef VST_BWD_SOFTMAX_WARPS
#define VST_BWD_SOFTMAX_WARPS 8
#endif
constexpr int kSoftmaxWarps = VST_BWD_SOFTMAX_WARPS;
constexpr int kKeysPerThread = BKV / (kSoftmaxWarps / 4);       // 64 or 32
constexpr int kChunks = kKeysPerThread / 32;                    // TMEM load rounds per pair and thread
constexpr int kTmaWarp = kSoftmaxWarps, kMmaWarp = kSoftmaxWarps + 1, kDrainWarp0 = kSoftmaxWarps + 2;   // + 4 dQ drain warps
constexpr int kThreads = 32 * (kSoftmaxWarps + 6);
// Q / dO stages: with two, the load of pair i+1 can only start when the gradient MMAs of pair i-1 release their
// stage and its ~1300 clk of TMA latency sits in front of the S / dP issue (three: L1 self 2411 -> 2361 us).
#ifndef VST_BWD_QSTAGES
#define VST_BWD_QSTAGES 3
#endif
constexpr int kQS = VST_BWD_QSTAGES;
// VST_BWD_DQ_TMA: the dQ tile of a pair leaves through a swizzled fp32 staging buffer and ONE bulk tensor reduction
// (cp.reduce.async.bulk.tensor ... .add, done by the TMA unit in L2) per 32-column half instead of 2048 RED.128
// instructions per pair from the drain warps: the r02 trace showed the softmax warps' shared-memory / barrier
// instructions queueing behind those atomics in the SM's memory pipe for ~2000 clk per pair.
#ifndef VST_BWD_DQ_TMA
#define VST_BWD_DQ_TMA 1
#endif
constexpr int kDQStage = VST_BWD_DQ_TMA ? 128 * HD * 4 : 0;   // 32 KB: two [128 rows x 128 B] swizzled boxes
constexpr int kSmemBytes = 2 * kTile /*K,V*/ + 2 * kQS * kTile /*Q,dO*/ + 2 * kPBytes + kDQStage + 1024 + 256;

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// exp2 on the FMA pipe for a share of the pairs (same polynomial as csrc/attn.cu: Cody-Waite split + degree-3
// minimax, |rel err| < 7.6e-5, far below the bf16 rounding of P): the 16384 exponentials of a pair are 1024 clk of
// MUFU.EX2 at 16/clk/SM.  x <= ~0 here (P = exp(s - lse) <= 1); x is clamped to >= -126.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  const float2 magic = make_float2(12582912.f, 12582912.f);      // 1.5 * 2^23
  const float2 nmagic = make_float2(-12582912.f, -12582912.f);
  x.x = fmaxf(x.x, -126.f);
  x.y = fmaxf(x.y, -126.f);
  const float2 xf = __fadd2_rn(x, magic);
  const float2 xi = __fadd2_rn(xf, nmagic);
  const float2 f = __fadd2_rn(x, make_float2(-xi.x, -xi.y));
  const float2 c3 = make_float2(0.05520550534129143f, 0.05520550534129143f);
  const float2 c2 = make_float2(0.24261397123336792f, 0.24261397123336792f);
  const float2 c1 = make_float2(0.6932547688484192f, 0.6932547688484192f);
  const float2 c0 = make_float2(0.9999276995658875f, 0.9999276995658875f);
  float2 p = __ffma2_rn(c3, f, c2);
  p = __ffma2_rn(p, f, c1);
  p = __ffma2_rn(p, f, c0);
  p.x = __int_as_float(__float_as_int(p.x) + (__float_as_int(xf.x) << 23));
  p.y = __int_as_float(__float_as_int(p.y) + (__float_as_int(xf.y) << 23));
  return p;
}
#ifndef VST_BWD_POLY
#define VST_BWD_POLY 2          // of every 8 pairs
#endif

struct BwdArgs {
  const float* lse;   // [Bt, heads, Sq] natural log
  const float* dvec;  // [Bt, heads, Sq] D = rowsum(dO . O)
  float* dq_acc;      // [Bt, heads, Sq, 64] fp32, zeroed
  void* dk;           // bf16 [Bk, Sk, C] (kv_group == 1) or fp32 accumulator [Bk, Sk, C]
  void* dv;
  int64_t ldk, ldv;   // row strides (elements) of dk / dv
  int Sq, Sk, heads, kv_group;
  int qsplit;         // CTAs per key tile, each walking 1/qsplit of the query tiles (only with fp32 dK / dV accumulators)
  float scale, scale_log2;
  const float* mask;                  // optional additive mask of the forward (natural-log domain)
  int64_t mask_sb, mask_sh, mask_sq;  // element strides over (batch, head, query); 0 = broadcast
};

// ---- D = rowsum(dO . O): 8 lanes per (row, head), 16-byte loads ---------------------------------
__global__ void attn_bwd_dot_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                    float* __restrict__ dvec, int64_t rows, int Sq, int heads,
                                    int64_t ldo, int64_t lddo) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t item = gid >> 3;          // (row, head)
  const int sub = (int)(gid & 7);
  if (item >= rows * heads) return;
  const int64_t row = item / heads;
  const int h = (int)(item % heads);
  const uint4 a = *reinterpret_cast<const uint4*>(o + row * ldo + h * HD + sub * 8);
  const uint4 b = *reinterpret_cast<const uint4*>(d_o + row * lddo + h * HD + sub * 8);
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = unpack_bf16(aw[i]), y = unpack_bf16(bw[i]);
    s += x.x * y.x + x.y * y.y;
  }
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  if (sub == 0) {
    const int64_t b_ = row / Sq, q = row % Sq;
    dvec[(b_ * heads + h) * Sq + q] = s;
  }
}

// ---- fp32 [B, heads, S, 64] accumulator -> bf16 [B, S, heads*64] (row stride ld) -------------------
__global__ void attn_bwd_dq_convert_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out,
                                           int64_t rows, int S, int heads, int64_t ld) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // 8 elements each
  const int64_t item = gid >> 3;
  const int sub = (int)(gid & 7);
  if (item >= rows * heads) return;
  const int64_t row = item / heads;
  const int h = (int)(item % heads);
  const int64_t b_ = row / S, q = row % S;
  const float* src = acc + (((b_ * heads + h) * S + q) * HD) + sub * 8;
  const float4 x = *reinterpret_cast<const float4*>(src), y = *reinterpret_cast<const float4*>(src + 4);
  uint4 w;
  w.x = pack_bf16(x.x, x.y); w.y = pack_bf16(x.z, x.w);
  w.z = pack_bf16(y.x, y.y); w.w = pack_bf16(y.z, y.w);
  *reinterpret_cast<uint4*>(out + row * ld + h * HD + sub * 8) = w;
}
// fp32 [rows, C] (row stride C) -> bf16 [rows, C] (row stride ld)
__global__ void attn_bwd_rows_convert_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out,
                                             int64_t rows, int C, int64_t ld) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // 8 elements each
  const int per_row = C / 8;
  if (gid >= rows * per_row) return;
  const int64_t row = gid / per_row;
  const int c = (int)(gid % per_row) * 8;
  const float* src = acc + row * C + c;
  const float4 x = *reinterpret_cast<const float4*>(src), y = *reinterpret_cast<const float4*>(src + 4);
  uint4 w;
  w.x = pack_bf16(x.x, x.y); w.y = pack_bf16(x.z, x.w);
  w.z = pack_bf16(y.x, y.y); w.w = pack_bf16(y.z, y.w);
  *reinterpret_cast<uint4*>(out + row * ld + c) = w;
}

__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                const __grid_constant__ CUtensorMap tmdQ, BwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + kTile;
  uint8_t* sQ = sV + kTile;          // [kQS]
  uint8_t* sdO = sQ + kQS * kTile;   // [kQS]
  uint8_t* sP = sdO + kQS * kTile;     // [2 key blocks][128 q][128 B]
  uint8_t* sdS = sP + kPBytes;
  uint8_t* sDQ = sdS + kPBytes;      // [2 column halves][128 q][128 B] fp32 dQ staging (VST_BWD_DQ_TMA)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDQ + kDQStage);
  uint64_t* kv_full = bars;
  uint64_t* q_full = bars + 1;       // [kQS]
  uint64_t* q_empty = q_full + kQS;  // [kQS]
  uint64_t* sdp_full = q_empty + kQS;  // S_i, dP_i complete in TMEM
  uint64_t* sdp_free = sdp_full + 1; // ... and copied to registers by the 256 softmax threads
  uint64_t* pds_full = sdp_free + 1; // P_i, dS_i in shared memory
  uint64_t* pds_free = pds_full + 1; // dV / dK / dQ MMAs of pair i have read them
  uint64_t* dq_full = pds_free + 1;  // dQ_i complete in TMEM
  uint64_t* dq_free = dq_full + 1;   // ... and drained by the 128 dQ threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dq_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) BTRACE(12, 0);          // CTA lifetime stamps: entry, set-up done, K/V landed, dV/dK stored, exit
  const int kv0 = (blockIdx.x / a.qsplit) * BKV, head = blockIdx.y, b = blockIdx.z;
  const int bk = b / a.kv_group;
  // this CTA's query tiles: [qt0, qt0 + nq) -- all of them unless the launch splits the walk (a.qsplit > 1)
  const int nq_all = (a.Sq + BQ - 1) / BQ;
  const int q_per = (nq_all + a.qsplit - 1) / a.qsplit;
  const int qt0 = (blockIdx.x % a.qsplit) * q_per;
  const int nq = min(q_per, nq_all - qt0);      // >= 1 by the host's choice of qsplit

  // Set-up and the first loads by ONE thread (lane 0 of the TMA warp), before the CTA-wide barrier: the K / V tiles
  // take ~4300 clk to land (cold) and a CTA of the S = 1024 shape lives only ~41000 clk (r02 trace), so they must
  // not wait for the TMEM allocation and the __syncthreads as well.
  if (threadIdx.x == kTmaWarp * 32) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < kQS; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_free, kArrivePerWarp * kSoftmaxWarps);
    mbar_init(pds_full, kArrivePerWarp * kSoftmaxWarps);
    mbar_init(pds_free, 1);
    mbar_init(dq_full, 1);
    mbar_init(dq_free, kArrivePerWarp * 4);
    fence_mbar_init();
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmdO);
    mbar_arrive_expect_tx(kv_full, 2 * kTile);
    tma_load_3d(sK, &tmK, kv_full, head * HD, kv0, bk);
    tma_load_3d(sV, &tmV, kv_full, head * HD, kv0, bk);
    for (int i = 0; i < kQS && i < nq; ++i) {          // first fill of the Q / dO ring: every stage is free
      mbar_arrive_expect_tx(&q_full[i], 2 * kTile);
      tma_load_3d(sQ + i * kTile, &tmQ, &q_full[i], head * HD, (qt0 + i) * BQ, b);
      tma_load_3d(sdO + i * kTile, &tmdO, &q_full[i], head * HD, (qt0 + i) * BQ, b);
    }
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) BTRACE(12, 1);
  const uint32_t tS = tmem_base, tdP = tmem_base + 128, tdV = tmem_base + 256,
                 tdK = tmem_base + 320, tdQ = tmem_base + 384;

  if (warp == kTmaWarp) {
    // ================= TMA producer =================
    if (lane == 0) {
      for (int i = kQS; i < nq; ++i) {           // (K / V and the first kQS pairs were issued during set-up)
        const int st = i % kQS;
        mbar_wait(&q_empty[st], ((i / kQS) & 1) ^ 1, 10);
        BTRACE(9, i);
        mbar_arrive_expect_tx(&q_full[st], 2 * kTile);
        tma_load_3d(sQ + st * kTile, &tmQ, &q_full[st], head * HD, (qt0 + i) * BQ, b);
        tma_load_3d(sdO + st * kTile, &tmdO, &q_full[st], head * HD, (qt0 + i) * BQ, b);
      }
    }
  } else if (warp == kMmaWarp) {
    // ================= MMA issuer (warp-uniform loop, elected lane issues) =================
    const bool issuer = elect_one();
    constexpr uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, 0, 0);    // S, dP: both K-major
    constexpr uint32_t idesc_t = umma_idesc_bf16(BKV, HD, 1, 1);    // dV, dK: both MN-major
    constexpr uint32_t idesc_q = umma_idesc_bf16(BQ, HD, 0, 1);     // dQ: A K-major, B MN-major
    constexpr uint64_t kRow16 = 2048 >> 4;                          // 16 rows of 128 B (MN-major k-step)
    const uint64_t dK_k = umma_desc_sw128(smem_u32(sK), 1024);          // K as K-major B (S)
    const uint64_t dK_mn = umma_desc_sw128(smem_u32(sK), 1024, 8192);   // K as MN-major B (dQ)
    const uint64_t dV_k = umma_desc_sw128(smem_u32(sV), 1024);
    const uint64_t dP_mn = umma_desc_sw128(smem_u32(sP), 1024, 16384);  // P^T: MN blocks 16 KB apart
    const uint64_t dS_mn = umma_desc_sw128(smem_u32(sdS), 1024, 16384);
    mbar_wait(kv_full, 0, 20);
    BTRACE(12, 2);
    auto issue_sdp = [&](int i) {
      const int st = i % kQS;
      BTRACE(11, i);
      mbar_wait(&q_full[st], (i / kQS) & 1, 21);
      tc_fence_after();
      const uint64_t dQ_k = umma_desc_sw128(smem_u32(sQ + st * kTile), 1024);
      const uint64_t ddO_k = umma_desc_sw128(smem_u32(sdO + st * kTile), 1024);
      BTRACE(5, i);
      if (issuer) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16_ss(tS, dQ_k + 2 * k, dK_k + 2 * k, idesc_s, k != 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16_ss(tdP, ddO_k + 2 * k, dV_k + 2 * k, idesc_s, k != 0);
        umma_commit(sdp_full);
      }
      __syncwarp();
    };
    // gradient MMAs of pair i: dV, dK (accumulated over the walk) and dQ (fresh), then the three releases
    auto issue_grads = [&](int i) {
      const int st = i % kQS;
      mbar_wait(pds_full, i & 1, 23);
      BTRACE(10, i);
      if (i > 0) mbar_wait(dq_free, (i - 1) & 1, 24);
      tc_fence_after();
      const uint64_t ddO_mn = umma_desc_sw128(smem_u32(sdO + st * kTile), 1024, 8192);
      const uint64_t dQ_mn = umma_desc_sw128(smem_u32(sQ + st * kTile), 1024, 8192);
      BTRACE(6, i);
      if (issuer) {
#if VST_BWD_ABLATE & 8
        if (i < 0)
#endif
        {
#pragma unroll
        for (int k = 0; k < BQ / 16; ++k)     // dV += P^T dO : contraction over the 128 query rows
          umma_bf16_ss(tdV, dP_mn + kRow16 * k, ddO_mn + kRow16 * k, idesc_t, (i | k) != 0);
#pragma unroll
        for (int k = 0; k < BQ / 16; ++k)     // dK += dS^T Q
          umma_bf16_ss(tdK, dS_mn + kRow16 * k, dQ_mn + kRow16 * k, idesc_t, (i | k) != 0);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {  // dQ = dS K : contraction over the 128 keys
          const uint64_t dS_k = umma_desc_sw128(smem_u32(sdS + (k >> 2) * 16384), 1024);
          umma_bf16_ss(tdQ, dS_k + 2 * (k & 3), dK_mn + kRow16 * k, idesc_q, k != 0);
        }
        }
        umma_commit(&q_empty[st]);
        umma_commit(pds_free);
        umma_commit(dq_full);
      }
      __syncwarp();
    };
    issue_sdp(0);
#if VST_BWD_ORDER == 1
    // (measured, slower: 5080 against 3860 clk per pair) gradient MMAs of pair i-1 issued only once S / dP of pair i
    // are in registers, then S / dP of pair i+1 -- the P / dS buffers are then held through the next pair's loads.
    for (int i = 0; i <= nq; ++i) {
      if (i < nq) mbar_wait(sdp_free, i & 1, 22);
      if (i >= 1) issue_grads(i - 1);
      if (i + 1 < nq) issue_sdp(i + 1);
    }
#else
    for (int i = 0; i < nq; ++i) {
      if (i + 1 < nq) {
        mbar_wait(sdp_free, i & 1, 22);
        issue_sdp(i + 1);
      }
      issue_grads(i);
    }
#endif
  } else if (warp < kSoftmaxWarps) {
    // ================= softmax warps: P and dS of this warp's 32 rows x 64 keys =================
    const int quarter = warp & 3, ks = warp >> 2;       // key slice: columns [ks * kKeysPerThread, + kKeysPerThread)
    const int col0 = ks * kKeysPerThread;
    const int row = quarter * 32 + lane;
    const uint32_t lane_off = uint32_t(quarter * 32) << 16;
    const int64_t vec_base = ((int64_t)b * a.heads + head) * a.Sq;
    uint8_t* prow = sP + (col0 >> 6) * 16384 + row * 128;      // [64-key block][row][128 B]
    uint8_t* dsrow = sdS + (col0 >> 6) * 16384 + row * 128;
    const int cb = (col0 & 63) >> 3;                           // first 16-byte chunk of this thread in the row
    const int rx = row & 7;
    // The arithmetic is the longest stage of a pair (r02 clock64 trace, scripts/attn_bwd_trace.py: ~3000 of 5550 clk
    // with scalar math and per-element bounds / mask predicates), so the common case -- no additive mask, key tile
    // inside Sk (CTA-uniform) -- runs on packed f32x2 math without predicates, VST_BWD_POLY of every 8 exponential
    // pairs on the FMA pipe, and the per-row LSE / D of the NEXT pair are fetched under this pair's arithmetic (their
    // L2 latency was a 730 clk stall at the top of every pair).
    const bool fast = a.mask == nullptr;
    const bool tail = kv0 + BKV > a.Sk;       // CTA-uniform: keys past Sk get P = 0 (their K / V rows are zero-filled)
    auto row_stats = [&](int i, float& lse2, float& dv) {
      const int q = (qt0 + i) * BQ + row;
      // rows past Sq: LSE = +inf makes P = 0 (their Q / dO rows are zero-filled by TMA)
      lse2 = q < a.Sq ? a.lse[vec_base + q] * 1.4426950408889634f : INFINITY;
      dv = q < a.Sq ? a.dvec[vec_base + q] : 0.f;
    };
    float lse2, dv;
    row_stats(0, lse2, dv);
    for (int i = 0; i < nq; ++i) {
      const int q = (qt0 + i) * BQ + row;
      const float* mrow = a.mask == nullptr ? nullptr
                              : a.mask + (int64_t)b * a.mask_sb + (int64_t)head * a.mask_sh +
                                    (int64_t)min(q, a.Sq - 1) * a.mask_sq;
      uint32_t pw[16 * kChunks], dsw[16 * kChunks];
      mbar_wait(sdp_full, i & 1, 30);
      if (warp == 0) BTRACE(0, i);
      tc_fence_after();
      float lse_raw_n = 0.f, dv_n = 0.f;      // next pair's row statistics: RAW loads only, no arithmetic on them
      bool valid_n = false;                   // before the end of this pair (an FMUL here would stall on L2 latency)
      const float2 sc2 = make_float2(a.scale_log2, a.scale_log2), nl2 = make_float2(-lse2, -lse2);
      const float2 s2 = make_float2(a.scale, a.scale), nd2 = make_float2(-dv * a.scale, -dv * a.scale);
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        uint32_t s[32], dp[32];
        tmem_ld32(tS + lane_off + col0 + c * 32, s);
        tmem_ld32(tdP + lane_off + col0 + c * 32, dp);
        tmem_ld_wait();
        if (c == kChunks - 1 && i + 1 < nq) {
          // In flight under the chunk-1 arithmetic and the stores.  AFTER the pair's last tcgen05.wait::ld: with the
          // loads issued before it, that wait also sat out their L2 latency (r02 trace: 997 clk for the first
          // chunk's TMEM load against 99 for the second).
          const int qn = (qt0 + i + 1) * BQ + row;
          valid_n = qn < a.Sq;
          if (valid_n) {
            lse_raw_n = __ldg(a.lse + vec_base + qn);
            dv_n = __ldg(a.dvec + vec_base + qn);
          }
        }
        if (c == kChunks - 1) {
          tc_fence_before();
          warp_arrive(sdp_free);       // S / dP may be overwritten by the next pair
          if (warp == 0) BTRACE(1, i);
        }
        const int key0 = kv0 + col0 + c * 32;
        if (fast) {
          auto chunk = [&](auto tail_c) {
#pragma unroll
            for (int e = 0; e < 16; ++e) {
#if VST_BWD_ABLATE & 2
              pw[c * 16 + e] = s[2 * e] ^ __float_as_uint(nl2.x);
              dsw[c * 16 + e] = dp[2 * e + 1] ^ __float_as_uint(nd2.x);
              continue;
#endif
              const float2 x = __ffma2_rn(make_float2(__uint_as_float(s[2 * e]), __uint_as_float(s[2 * e + 1])), sc2, nl2);
              const bool poly = ((e * VST_BWD_POLY) & 7) + VST_BWD_POLY > 7 && VST_BWD_POLY > 0;
              float2 pv = poly ? ex2_poly2(x) : make_float2(ex2f(x.x), ex2f(x.y));
              if (decltype(tail_c)::value) {
                if (key0 + 2 * e >= a.Sk) pv.x = 0.f;
                if (key0 + 2 * e + 1 >= a.Sk) pv.y = 0.f;
              }
              // dS = P (dP - D) scale
              const float2 t = __ffma2_rn(make_float2(__uint_as_float(dp[2 * e]), __uint_as_float(dp[2 * e + 1])), s2, nd2);
              const float2 d = __fmul2_rn(pv, t);
              pw[c * 16 + e] = pack_bf16(pv.x, pv.y);
              dsw[c * 16 + e] = pack_bf16(d.x, d.y);
            }
          };
          if (tail) chunk(std::true_type{}); else chunk(std::false_type{});
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            float b0 = -lse2, b1 = -lse2;
            if (mrow != nullptr) {       // warp-uniform branch; P = exp(s * scale + mask - lse)
              if (key0 + 2 * e < a.Sk) b0 = fmaf(__ldg(mrow + key0 + 2 * e), 1.4426950408889634f, -lse2);
              if (key0 + 2 * e + 1 < a.Sk) b1 = fmaf(__ldg(mrow + key0 + 2 * e + 1), 1.4426950408889634f, -lse2);
            }
            float p0 = ex2f(fmaf(__uint_as_float(s[2 * e]), a.scale_log2, b0));
            float p1 = ex2f(fmaf(__uint_as_float(s[2 * e + 1]), a.scale_log2, b1));
            if (key0 + 2 * e >= a.Sk) p0 = 0.f;
            if (key0 + 2 * e + 1 >= a.Sk) p1 = 0.f;
            const float d0 = p0 * (__uint_as_float(dp[2 * e]) - dv) * a.scale;
            const float d1 = p1 * (__uint_as_float(dp[2 * e + 1]) - dv) * a.scale;
            pw[c * 16 + e] = pack_bf16(p0, p1);
            dsw[c * 16 + e] = pack_bf16(d0, d1);
          }
        }
      }
      lse2 = valid_n ? lse_raw_n * 1.4426950408889634f : INFINITY;
      dv = dv_n;
      if (warp == 0) BTRACE(2, i);
#if VST_BWD_ELECT_WAIT
      if (i > 0) {
        if (lane == 0) mbar_wait(pds_free, (i - 1) & 1, 31);
        __syncwarp();
      }
#else
      if (i > 0) mbar_wait(pds_free, (i - 1) & 1, 31);   // previous pair's MMAs are done with sP / sdS
#endif
      if (warp == 0) BTRACE(3, i);
#pragma unroll
      for (int cc = 0; cc < ((VST_BWD_ABLATE & 16) ? 0 : 4 * kChunks); ++cc) {
        const int off = ((cb + cc) ^ rx) * 16;
        *reinterpret_cast<uint4*>(prow + off) = make_uint4(pw[4 * cc], pw[4 * cc + 1], pw[4 * cc + 2], pw[4 * cc + 3]);
        *reinterpret_cast<uint4*>(dsrow + off) = make_uint4(dsw[4 * cc], dsw[4 * cc + 1], dsw[4 * cc + 2], dsw[4 * cc + 3]);
      }
      fence_proxy_async_smem();
      warp_arrive(pds_full);
      if (warp == 0) BTRACE(4, i);
    }
    // ---- dV (warps 0..3) / dK (warps 4..7): TMEM lane = key row ----
    if (warp < 8) {
    const int half = warp >> 2;
    mbar_wait(pds_free, (nq - 1) & 1, 32);
    tc_fence_after();
    const int key = kv0 + row;
    const uint32_t tsrc = (half == 0 ? tdV : tdK) + lane_off;
    void* dst_base = half == 0 ? a.dv : a.dk;
    const int64_t ld = half == 0 ? a.ldv : a.ldk;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t o[32];
      tmem_ld32(tsrc + c * 32, o);
      tmem_ld_wait();
      if (key < a.Sk) {
        if (a.kv_group == 1) {
          __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(dst_base) +
                               ((int64_t)bk * a.Sk + key) * ld + head * HD + c * 32;
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            uint4 w;
            w.x = pack_bf16(__uint_as_float(o[8 * v]), __uint_as_float(o[8 * v + 1]));
            w.y = pack_bf16(__uint_as_float(o[8 * v + 2]), __uint_as_float(o[8 * v + 3]));
            w.z = pack_bf16(__uint_as_float(o[8 * v + 4]), __uint_as_float(o[8 * v + 5]));
            w.w = pack_bf16(__uint_as_float(o[8 * v + 6]), __uint_as_float(o[8 * v + 7]));
            *reinterpret_cast<uint4*>(dst + 8 * v) = w;
          }
        } else {
          // text K / V are shared by kv_group frames: sum their gradients in fp32
          float* dst = reinterpret_cast<float*>(dst_base) +
                       ((int64_t)bk * a.Sk + key) * (a.heads * HD) + head * HD + c * 32;
#pragma unroll
          for (int v = 0; v < 8; ++v)
            atomicAdd(reinterpret_cast<float4*>(dst + 4 * v),
                      make_float4(__uint_as_float(o[4 * v]), __uint_as_float(o[4 * v + 1]),
                                  __uint_as_float(o[4 * v + 2]), __uint_as_float(o[4 * v + 3])));
        }
      }
    }
    }
  } else {
    // ================= dQ drain (warps 10..13): add each pair's dQ tile to the fp32 accumulator ===
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_off = uint32_t(quarter * 32) << 16;
#if VST_BWD_DQ_TMA
    const bool leader = warp == kDrainWarp0 && lane == 0;
    const int bh = b * a.heads + head;
    const int rx = row & 7;
    if (leader) tma_prefetch_desc(&tmdQ);
#endif
    for (int i = 0; i < nq; ++i) {
      mbar_wait(dq_full, i & 1, 40);
      if (warp == kDrainWarp0) BTRACE(7, i);
      tc_fence_after();
      uint32_t o[64];
      tmem_ld32(tdQ + lane_off, o);
      tmem_ld32(tdQ + lane_off + 32, o + 32);
      tmem_ld_wait();
      tc_fence_before();
      warp_arrive(dq_free);
#if VST_BWD_DQ_TMA
      if (i > 0) {
        if (leader) tma_store_wait_read<0>();          // the previous pair's reduction has read the staging tile
        asm volatile("bar.sync 2, 128;" ::: "memory");
      }
#if !(VST_BWD_ABLATE & 1)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint8_t* drow = sDQ + h * 16384 + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(drow + ((j ^ rx) * 16)) =
              make_uint4(o[h * 32 + 4 * j], o[h * 32 + 4 * j + 1], o[h * 32 + 4 * j + 2], o[h * 32 + 4 * j + 3]);
      }
#endif
      fence_proxy_async_smem();
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (leader) {
#if !(VST_BWD_ABLATE & 1)
        // rows past Sq are clipped by the tensor map (dims {64, Sq, Bt * heads})
        tma_reduce_add_3d(&tmdQ, sDQ, 0, (qt0 + i) * BQ, bh);
        tma_reduce_add_3d(&tmdQ, sDQ + 16384, 32, (qt0 + i) * BQ, bh);
#endif
        tma_store_commit();
      }
#else
      const int q = (qt0 + i) * BQ + row;
      if (q < a.Sq && !(VST_BWD_ABLATE & 1)) {
        float* dst = a.dq_acc + (((int64_t)b * a.heads + head) * a.Sq + q) * HD;
#pragma unroll
        for (int v = 0; v < 16; ++v)
          atomicAdd(reinterpret_cast<float4*>(dst + 4 * v),
                    make_float4(__uint_as_float(o[4 * v]), __uint_as_float(o[4 * v + 1]),
                                __uint_as_float(o[4 * v + 2]), __uint_as_float(o[4 * v + 3])));
      }
#endif
      if (warp == kDrainWarp0) BTRACE(8, i);
    }
#if VST_BWD_DQ_TMA
    if (leader) tma_store_wait_read<0>();      // shared memory read out; the reductions complete with the kernel
#endif
  }

  if (warp == 0) BTRACE(12, 3);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) BTRACE(12, 4);
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

int make_map3(CUtensorMap* m, const void* p, int64_t B, int64_t S, int width, int64_t ld) {
  const uint64_t dims[3] = {(uint64_t)width, (uint64_t)S, (uint64_t)B};
  const uint64_t strides[3] = {2, (uint64_t)ld * 2, (uint64_t)S * ld * 2};
  const uint32_t box[3] = {HD, 128, 1};
  return vst_make_tmap(m, p, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int64_t align256(int64_t x) { return (x + 255) / 256 * 256; }

}  // namespace

#ifdef VST_BWD_TRACE
extern "C" int vst_attn_bwd_debug_trace(long long* host_dst) {
  return (int)cudaMemcpyFromSymbol(host_dst, g_bwd_trace, sizeof(long long) * 16 * 64);
}
#endif

extern "C" int64_t vst_attn_bwd_workspace_bytes(int Bt, int heads, int Sq, int Sk, int kv_group) {
  if (Bt <= 0 || heads <= 0 || Sq <= 0 || Sk <= 0 || kv_group <= 0) return 0;
  int64_t n = align256((int64_t)Bt * heads * Sq * 4);             // D
  n += align256((int64_t)Bt * heads * Sq * HD * 4);               // dQ accumulator
  if (kv_group > 1) n += 2 * align256((int64_t)(Bt / kv_group) * Sk * heads * HD * 4);   // dK, dV accumulators
  return n;
}

extern "C" int vst_attn_bwd(const void* q, const void* k, const void* v, const void* o,
                            const void* d_o, const float* lse, void* dq, void* dk, void* dv, int Bt,
                            int heads, int Sq, int Sk, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                            int64_t ldo, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv,
                            int kv_group, float scale, const float* mask, int64_t mask_sb,
                            int64_t mask_sh, int64_t mask_sq, void* workspace, int64_t workspace_bytes,
                            void* stream) {
  VST_REQUIRE(Bt >= 0 && heads > 0 && Sq >= 0 && Sk > 0, VST_E_BADARG, "attn_bwd: bad shape");
  if (Bt == 0 || Sq == 0) return VST_OK;
  VST_REQUIRE(q && k && v && o && d_o && lse && dq && dk && dv, VST_E_BADARG, "attn_bwd: null pointer");
  VST_REQUIRE(d == HD, VST_E_UNSUPPORTED, "attn_bwd: head dim %d unsupported (SDXL spatial d=64)", d);
  VST_REQUIRE(kv_group >= 1 && Bt % kv_group == 0, VST_E_BADARG, "attn_bwd: Bt %% kv_group != 0");
  VST_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && lddo % 8 == 0 &&
                  lddq % 8 == 0 && lddk % 8 == 0 && lddv % 8 == 0,
              VST_E_UNSUPPORTED, "attn_bwd: row strides must be multiples of 8 elements");
  VST_REQUIRE(heads <= 65535 && Bt <= 65535, VST_E_UNSUPPORTED, "attn_bwd: grid too large");
  const int64_t need = vst_attn_bwd_workspace_bytes(Bt, heads, Sq, Sk, kv_group);
  VST_REQUIRE(workspace && workspace_bytes >= need && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
              VST_E_BADARG, "attn_bwd: a 256-byte aligned workspace of %lld bytes is required", (long long)need);
  cudaStream_t st = (cudaStream_t)stream;
  const int C = heads * HD, Bk = Bt / kv_group;
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  float* dvec = reinterpret_cast<float*>(ws);
  ws += align256((int64_t)Bt * heads * Sq * 4);
  float* dq_acc = reinterpret_cast<float*>(ws);
  const int64_t dq_bytes = (int64_t)Bt * heads * Sq * HD * 4;
  ws += align256(dq_bytes);
  float *dk_acc = nullptr, *dv_acc = nullptr;
  const int64_t kv_bytes = (int64_t)Bk * Sk * C * 4;
  if (kv_group > 1) {
    dk_acc = reinterpret_cast<float*>(ws);
    ws += align256(kv_bytes);
    dv_acc = reinterpret_cast<float*>(ws);
    VST_CUDA_OK(cudaMemsetAsync(dk_acc, 0, kv_bytes, st));
    VST_CUDA_OK(cudaMemsetAsync(dv_acc, 0, kv_bytes, st));
  }
  VST_CUDA_OK(cudaMemsetAsync(dq_acc, 0, dq_bytes, st));

  const int64_t rows = (int64_t)Bt * Sq;
  {
    const int64_t threads = rows * heads * 8;
    attn_bwd_dot_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
        (const __nv_bfloat16*)o, (const __nv_bfloat16*)d_o, dvec, rows, Sq, heads, ldo, lddo);
    VST_LAUNCH_OK();
  }
  CUtensorMap tQ, tK, tV, tdO;
  int rc;
  if ((rc = make_map3(&tQ, q, Bt, Sq, C, ldq))) return rc;
  if ((rc = make_map3(&tK, k, Bk, Sk, C, ldk))) return rc;
  if ((rc = make_map3(&tV, v, Bk, Sk, C, ldv))) return rc;
  if ((rc = make_map3(&tdO, d_o, Bt, Sq, C, lddo))) return rc;
  CUtensorMap tdQ;
  {
    // fp32 accumulator [Bt * heads][Sq][64]: boxes of 32 columns (128 B, 128-byte swizzle) x 128 rows
    const uint64_t dims[3] = {(uint64_t)HD, (uint64_t)Sq, (uint64_t)Bt * heads};
    const uint64_t strides[3] = {4, (uint64_t)HD * 4, (uint64_t)Sq * HD * 4};
    const uint32_t box[3] = {32, 128, 1};
    if ((rc = vst_make_tmap_dt(&tdQ, dq_acc, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_DATA_TYPE_FLOAT32)))
      return rc;
  }
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    configured = true;
  }
  BwdArgs a;
  a.lse = lse;
  a.dvec = dvec;
  a.dq_acc = dq_acc;
  a.dk = kv_group > 1 ? (void*)dk_acc : dk;
  a.dv = kv_group > 1 ? (void*)dv_acc : dv;
  a.ldk = lddk;
  a.ldv = lddv;
  a.Sq = Sq;
  a.Sk = Sk;
  a.heads = heads;
  a.kv_group = kv_group;
  a.scale = scale;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.mask = mask;
  a.mask_sb = mask_sb;
  a.mask_sh = mask_sh;
  a.mask_sq = mask_sq;
  // Text cross-attention has ONE key tile per (batch, head): 160-320 CTAs of 8-32 pairs each, under two waves of
  // 148.  Its dK / dV already accumulate in fp32 with atomics (kv_group > 1), so the query walk can be split over
  // several CTAs per key tile until the grid is a few waves deep.
  const int kt = (Sk + BKV - 1) / BKV, nq_all = (Sq + BQ - 1) / BQ;
  int qsplit = 1;
  if (kv_group > 1) {
    const int64_t ctas = (int64_t)kt * heads * Bt;
    while (qsplit < 8 && ctas * qsplit < 4 * vst_sm_count() && (nq_all + 2 * qsplit - 1) / (2 * qsplit) >= 2) qsplit *= 2;
    while (qsplit > 1 && ((nq_all + qsplit - 1) / qsplit) * (qsplit - 1) >= nq_all) --qsplit;   // no empty CTA
  }
  a.qsplit = qsplit;
  dim3 grid(kt * qsplit, heads, Bt);
  attn_bwd_kernel<<<grid, kThreads, kSmemBytes, st>>>(tQ, tK, tV, tdO, tdQ, a);
  VST_LAUNCH_OK();
  {
    const int64_t threads = rows * heads * 8;
    attn_bwd_dq_convert_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
        dq_acc, (__nv_bfloat16*)dq, rows, Sq, heads, lddq);
    VST_LAUNCH_OK();
  }
  if (kv_group > 1) {
    const int64_t krows = (int64_t)Bk * Sk;
    const int64_t threads = krows * (C / 8);
    attn_bwd_rows_convert_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
        dk_acc, (__nv_bfloat16*)dk, krows, C, lddk);
    VST_LAUNCH_OK();
    attn_bwd_rows_convert_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
        dv_acc, (__nv_bfloat16*)dv, krows, C, lddv);
    VST_LAUNCH_OK();
  }
  return VST_OK;
}
