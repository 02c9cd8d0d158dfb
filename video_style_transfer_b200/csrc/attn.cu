// a2: spatial self-attention / text cross-attention forward, head_dim 64, on tcgen05 + TMEM + TMA.
// Reference: animatediff/attention_processor.py:74-83 (head split, F.scaled_dot_product_attention,
// head merge) -- the head split/merge transposes disappear into the TMA tensor maps, and the
// text repeat_interleave of :62-66 into the kv batch index (b / kv_group).
//
// A persistent CTA (two per SM) walks a strided list of items = 128 query rows of one (batch, head).
// TMEM per CTA: 256 columns (S [0,128) fp32, O [128,192) fp32, P [192,256) bf16 pairs).  Warps:
//   0..3  softmax (warpgroup 0, 208 registers by setmaxnreg): thread t owns query row t (TMEM lane
//         t).  Per 128-key block ONE pass: S row -> registers (S is released at once, so the next
//         S = Q K^T overlaps this block's softmax), row max (4 independent chains),
//         P = exp2(s*scale - m) with packed f32x2 math, kPoly of every 8 pairs on the FMA pipe
//         (ex2_poly2), bf16 P written back to TMEM.  O accumulates in TMEM across blocks; it is
//         rescaled only when the running max moves by more than 2^8 (lazy rescale).
//   4     TMA producer: Q ring of 3 tiles, K / V tiles [128 x 64] (2 stages each).
//   5     MMA issuer (whole warp in the loop, one elected lane issues): S = Q K^T
//         (4 x tcgen05.mma 128x128x16, both operands K-major smem) and O += P V
//         (8 x 128x64x16, A = P from TMEM, B = V MN-major straight from its row-major tile),
//         so P never touches shared memory.
//   6     store warp: staged O tiles leave by TMA.   7: idle (warpgroup 1 runs on 48 registers).
// Text cross-attention (64 < Sk <= 80, one 80-key block per item) is its own instantiation: an HBM-bound stream of
// Q tiles in and O tiles out, shaped by scripts/attn_trace_cross.py -- four Q buffers (the 80-key K / V slots are
// 10 KB and O has a staging buffer of its own, so a Q buffer returns to the ring when its S MMA retires), only the
// last 16 keys are tested for masking, P's completion is signalled at once and the Q ring has its own producer warp,
// so the MMA warp runs an item ahead (r02: 46.1 / 29.5 -> 38.9 / 24.6 us at the L1 / L2 shapes).
// What bounds it (r01 measurements: scripts/ubench/*.cu, scripts/attn_trace.py,
// profiles/attn_r01_notes.md): at head dim 64 a 128x128 score block is ~530 clk of tensor work but
// 1024 clk of MUFU.EX2 (16/clk/SM; ex2.f16x2 is two MUFU ops, no packed rate) plus ~500 issue
// slots per softmax warp, and with two softmax warps per scheduler the stream is latency-bound at
// ~2400 clk per block pair (xu pipe 58 %, issue 40 %).  Measured and slower or equal: two query
// tiles per CTA sharing K / V, 8 half-row softmax warps, a per-SM CTA stagger, an exp-phase
// mutex, a one-shot (non-persistent) CTA per item (shipped until the item-boundary trace).
// Roofline: tensor pipe for self-attention (S/2 FLOP/B); FLOPs = 4*Bt*h*Sq*Sk*d.
#include <stdlib.h>
#include <type_traits>

// Watchdog placement (r02, kbench attn on one box): with the counting watchdog in every role's waits the L1 / L2
// self shapes ran at 855 / 718 TFLOP/s, with none at 888 / 744 (the 208-register softmax loop and the 48-register
// auxiliary warps have no spare register for the poll counter; the producer's waits are on the critical path of
// the one-block text cross-attention items, 48 -> 45 us).  Every role but one therefore waits with mbar_wait_plain;
// the store warp keeps mbar_wait -- every deadlock of this protocol ends with the store warp waiting for an
// o_staged that never comes, so the launch still fails loudly instead of hanging.
#define VST_MBAR_QUIET 1
#include "common.cuh"

namespace {

constexpr int BQ = 128;   // query rows per CTA
constexpr int BKV = 128;  // keys per block
constexpr int HD = 64;    // head dim
constexpr int kThreads = 256;          // warpgroup 0: softmax; warpgroup 1: TMA, MMA, store, idle
#ifndef VST_ATTN_REGS_SM
#define VST_ATTN_REGS_SM 208
#define VST_ATTN_REGS_AUX 48
#endif
constexpr int kRegsSoftmax = VST_ATTN_REGS_SM;      // setmaxnreg: 128 * 208 + 128 * 48 = half the register file,
constexpr int kRegsAux = VST_ATTN_REGS_AUX;           // so two CTAs stay resident and the softmax loop does not spill
constexpr int KST = 2;    // K and V stages
constexpr int kTile = 128 * HD * 2;  // 16 KB
constexpr int kDefaultPoly = 2;
constexpr float kRescaleThreshold = 8.f;  // log2 units: P stays below 2^8, safe in bf16 / fp32

template <int N> __device__ __forceinline__ void setmaxnreg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N> __device__ __forceinline__ void setmaxnreg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// exp2 on the FMA pipe (Cody-Waite split + degree-3 minimax polynomial, |rel err| < 7.6e-5, far
// below the bf16 rounding of P).  The MUFU unit retires 16 ex2/clk/SM: at head dim 64 that is
// 1024 clk per 128x128 score block against 512 clk of tensor work, so the kernel was XU-bound
// (ncu r01: xu pipe 72 %, tensor 36 %).  kPoly of every 8 packed pairs take this path instead.
// x is clamped to >= -126 so the exponent add cannot wrap; x <= 2^8 by the lazy-rescale bound.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  const float2 magic = make_float2(12582912.f, 12582912.f);      // 1.5 * 2^23
  const float2 nmagic = make_float2(-12582912.f, -12582912.f);
  x.x = fmaxf(x.x, -126.f);
  x.y = fmaxf(x.y, -126.f);
  const float2 xf = __fadd2_rn(x, magic);        // low mantissa bits = round(x)
  const float2 xi = __fadd2_rn(xf, nmagic);      // round(x) as a float
  const float2 f = __fadd2_rn(x, make_float2(-xi.x, -xi.y));   // in [-0.5, 0.5]
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

// -DVST_ATTN_TRACE: CTA 0 records clock64() stamps of its pipeline events (tuning builds only).
#ifdef VST_ATTN_TRACE
__device__ long long g_attn_trace[4 * 12 * 1024];   // [warp & 3][slot][index]
#define TRACE(slot, idx) do { if (blockIdx.x == 0 && (idx) < 1024 && (threadIdx.x & 31) == 0) \
    g_attn_trace[((threadIdx.x >> 5) & 3) * 12 * 1024 + (slot) * 1024 + (idx)] = clock64(); } while (0)
#else
#define TRACE(slot, idx) do { } while (0)
#endif

struct AttnArgs {
  float* lse;
  int Sq, Sk, heads, kv_group;
  float scale_log2;
  const float* mask;                 // optional additive mask (natural-log domain), fp32
  int64_t mask_sb, mask_sh, mask_sq; // element strides over (batch, head, query); 0 = broadcast
  int mask_vec4;                     // base pointer and strides allow 16-byte loads
};

// ---- persistent form ---------------------------------------------------------------------------
// One CTA walks a strided list of (batch, head, query tile) items, two CTAs per SM.  The stream of
// score blocks never drains between items (clock64 traces of the first form, scripts/attn_trace.py:
// an item boundary cost ~3000 clk against a ~2400 clk block period, 11 % at 8 blocks per item):
//   * TMEM allocation and barrier set-up are paid once per CTA;
//   * Q lives in a ring of three buffers, so the Q of the next two items is in flight;
//   * the MMA warp issues the first score block of the NEXT item as soon as the last S of this
//     item is in registers (S = Q K^T never waits for an epilogue);
//   * the epilogue of item i is deferred: the softmax warps first run the exponentials of block 0 of
//     item i+1 -- by then the last P V of item i has long retired -- then read O out of TMEM,
//     normalise and stage it in item i's dead Q buffer, and only then store P;
//   * a dedicated store warp (warp 6) sends the staged tile by TMA and recycles the buffer, so the
//     softmax warps never wait for a bulk store to read shared memory.
// All barrier phases run on global counters (g: K/V blocks loaded, blk: score blocks, i: items).
// Cross-attention (80-key block) instantiation; -D overrides for tuning builds.  Measured, L1 / L2 text cross in us
// (kbench attn, B = 16): 3 Q buffers doubling as O stages 46.1 / 29.5 (round-2 start); + masking test on the last 16
// keys only, 4 Q buffers 39.9 / 25.5; 5 Q buffers with single-staged K / V 50.1 / 29.7 (k_full wait 1400 clk on the
// MMA warp's critical path); + P completion signalled at once 38.9 / 24.6; + Q ring fed by its own warp 37.9 / 25.6;
// rows stored straight from registers 49.1 / 28.7; shipped: 3 Q buffers released by the S MMA's commit + one
// dedicated O staging buffer + the three changes above, 38.9 / 24.6.  In that form the MMA warp no longer waits
// (q_full 130, s_free 84, k_full 180 clk per item) and CTA 0's steady state is 2700 clk per item, i.e. two resident
// CTAs move 64 KB per 2700 clk and SM = the HBM rate: what keeps the launch at 0.52-0.66 of HBM is the ramp
// (9-17 items per CTA at these sizes), not the pipeline.
#ifndef VST_ATTN_X_EAGER_P
#define VST_ATTN_X_EAGER_P 1
#endif
#ifndef VST_ATTN_X_SPLITQ
#define VST_ATTN_X_SPLITQ 1
#endif
#ifndef VST_ATTN_X_OWNO
#define VST_ATTN_X_OWNO 1
#endif
#ifndef VST_ATTN_LAG_PV
#define VST_ATTN_LAG_PV 0        // measured slower (see the MMA warp): tuning builds only
#endif
#ifndef VST_ATTN_X_NQ
#define VST_ATTN_X_NQ 3
#define VST_ATTN_X_KST 2
#endif

template <int BKVT_>
struct AttnSmem {
  static constexpr int kNQ = BKVT_ == BKV ? 3 : VST_ATTN_X_NQ;                      // Q ring depth
  static constexpr int kKST = BKVT_ == BKV ? KST : VST_ATTN_X_KST;                  // K and V stages
  static constexpr int kKVSlot = BKVT_ == BKV ? kTile : (BKVT_ * HD * 2 + 1023) / 1024 * 1024;   // bytes per K / V stage
  static constexpr int kOStage = BKVT_ == BKV ? 0 : VST_ATTN_X_OWNO;                // dedicated O staging buffers
  static constexpr int kBytes = kTile * (kNQ + kOStage) + 2 * kKST * kKVSlot + 256;
};
constexpr int kSmemBytesP = AttnSmem<BKV>::kBytes;
constexpr int kSmemBytes80 = AttnSmem<80>::kBytes;
static_assert(kSmemBytes80 <= kSmemBytesP, "the cross-attention instantiation must keep two CTAs per SM");

// BKVT: keys per block.  128 everywhere except text cross-attention (Sk = 77 <= 80: ONE block per item), which runs
// with an 80-key block: 5 instead of 8 score / PV k-steps and 40 instead of 64 exponential pairs per row (r02: the
// 128-wide block spent 40 % of its softmax and MMA work on masked columns; 70 launches, 5.9 ms of the step).
template <int kPoly, bool kMask, int BKVT = BKV>
__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_pers_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                     AttnArgs a, int n_items, int ntile) {
  static_assert(BKVT == BKV || (!kMask && BKVT % 16 == 0 && BKVT > 64 && BKVT < BKV), "BKVT: 128, or a cross-attention block in (64, 128)");
  // Text cross-attention (BKVT = 80) is ONE block per item, i.e. an HBM-bound stream of Q tiles in and O tiles out:
  // the clock64 trace (scripts/attn_trace_cross.py) showed the MMA warp waiting ~2800 clk per item for the next Q
  // tile -- its ring buffer doubles as the O staging buffer of the item three back, so with three buffers a Q load
  // could only be issued ~1.3 item periods ahead of its use, less than the loaded DRAM latency.  The 80-key K / V
  // slots are 10 KB instead of 16 KB, which pays for a fourth 16 KB buffer inside the same 2-CTA-per-SM budget
  // (a dedicated O stage, see kOwnO).
  constexpr int NQT = AttnSmem<BKVT>::kNQ, KSTT = AttnSmem<BKVT>::kKST, kKVSlot = AttnSmem<BKVT>::kKVSlot;
  constexpr bool kSplitQ = BKVT < BKV && VST_ATTN_X_SPLITQ;      // Q loads from their own warp (see warp 7)
  // one-block items: O is staged in a buffer of its own and the Q buffer goes back to the ring as soon as its S MMA
  // has read it.  Staged in its dead Q buffer (as the other instantiations do), an item's O held that buffer until
  // the TMA store had READ it, behind the loads queued on the same unit: the MMA warp still waited ~1200 clk per
  // item for q_full with four buffers and a dedicated Q producer (scripts/attn_trace_cross.py).  Storing the rows
  // straight from registers instead measured slower (49 vs 38 us at L1: 8 x 32-line stores per warp and item, the
  // softmax warps drift 2000 clk apart).
  constexpr bool kOwnO = BKVT < BKV && VST_ATTN_X_OWNO;
  constexpr bool kLagPV = BKVT == BKV && VST_ATTN_LAG_PV;          // P V issued one block late (see the MMA warp)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;                 // 128B-swizzled tiles need 1024-byte alignment (checked)
  uint8_t* sQ = smem;                       // [NQT] query tiles (each reused to stage its item's O)
  uint8_t* sO = sQ + NQT * kTile;            // (kOwnO) the O staging buffer
  uint8_t* sK = sO + (kOwnO ? kTile : 0);   // [KSTT]
  uint8_t* sV = sK + KSTT * kKVSlot;         // [KSTT]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + KSTT * kKVSlot);
  uint64_t* q_full = bars;                  // [NQT]
  uint64_t* q_empty = q_full + NQT;          // [NQT]  the O store staged in this buffer has been read
  uint64_t* o_staged = q_empty + NQT;        // [NQT]  the item's O is normalised and staged (128 rows)
  uint64_t* k_full = o_staged + NQT;         // [KSTT]
  uint64_t* k_empty = k_full + KSTT;         // [KSTT]
  uint64_t* v_full = k_empty + KSTT;         // [KSTT]
  uint64_t* v_empty = v_full + KSTT;         // [KSTT]
  uint64_t* s_full = v_empty + KSTT;         // S_j complete in TMEM
  uint64_t* s_free = s_full + 1;            // S_j copied to registers by all 128 rows
  uint64_t* p_full = s_free + 1;            // P_j in TMEM (and O rescaled if needed)
  uint64_t* o_full = p_full + 1;            // O += P_j V_j complete
  uint64_t* o_free = o_full + 1;            // the item's O has been read out of TMEM
  uint64_t* so_empty = o_free + 1;          // (kOwnO) the TMA store has read the O staging buffer
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(so_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nblk = (a.Sk + BKVT - 1) / BKVT;
  constexpr uint32_t kKVBytes = BKVT * HD * 2;           // bytes one K / V block load brings in

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) {
      printf("vst attn_fwd: dynamic shared memory is not 1024-byte aligned\n");
      __trap();
    }
    for (int i = 0; i < NQT; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&o_staged[i], 128);
    }
    for (int i = 0; i < KSTT; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 128);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    mbar_init(o_free, 128);
    mbar_init(so_empty, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    // ================= TMA producer =================
    setmaxnreg_dec<kRegsAux>();
    if (lane == 0) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      uint32_t g = 0;
      int i = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++i) {
        const int qtile = it % ntile, head = (it / ntile) % a.heads, b = it / (ntile * a.heads);
        const int bk = b / a.kv_group, buf = i % NQT;
        if constexpr (!kSplitQ) {
          mbar_wait_plain(&q_empty[buf], ((i / NQT) & 1) ^ 1);
          mbar_arrive_expect_tx(&q_full[buf], kTile);
          tma_load_3d(sQ + buf * kTile, &tmQ, &q_full[buf], head * HD, qtile * BQ, b);
        }
        for (int j = 0; j < nblk; ++j, ++g) {
          const int st = g % KSTT;
          const uint32_t ph = ((g / KSTT) & 1) ^ 1;
          mbar_wait_plain(&k_empty[st], ph);
          mbar_arrive_expect_tx(&k_full[st], kKVBytes);
          tma_load_3d(sK + st * kKVSlot, &tmK, &k_full[st], head * HD, j * BKVT, bk);
          mbar_wait_plain(&v_empty[st], ph);
          mbar_arrive_expect_tx(&v_full[st], kKVBytes);
          tma_load_3d(sV + st * kKVSlot, &tmV, &v_full[st], head * HD, j * BKVT, bk);
        }
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer (whole warp in the loop, one elected lane issues) ==========
    setmaxnreg_dec<kRegsAux>();
    const bool issuer = elect_one();
    const uint32_t tS = tmem_base;             // S (fp32) columns [0,128)
    const uint32_t tO = tS + 128;              // O accumulator (fp32), 64 columns
    const uint32_t tP = tS + 192;              // P (bf16 pairs), 64 columns
    constexpr uint32_t idesc_s = umma_idesc_bf16(BQ, BKVT, 0, 0);
    constexpr uint32_t idesc_o = umma_idesc_bf16(BQ, HD, 0, 1);  // B (= V) is MN-major
    uint32_t g = 0, blk = 0;
    int i = 0;
    auto issue_s = [&](uint64_t dq, uint32_t gg, uint32_t tr) {
      const int st = gg % KSTT;
      mbar_wait_plain(&k_full[st], (gg / KSTT) & 1);
      tc_fence_after();
      const uint64_t dk = umma_desc_sw128(smem_u32(sK + st * kKVSlot), 1024);
      TRACE(5, tr);
      if (issuer) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16_ss(tS, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
        umma_commit(&k_empty[st]);
        if (kOwnO) umma_commit(&q_empty[gg % NQT]);      // (one block per item: block index == item index)
        umma_commit(s_full);
      }
      __syncwarp();
    };
    auto issue_pv = [&](int pj, uint32_t pg, uint32_t pblk, int pi) {
      const int st = pg % KSTT;
      mbar_wait_plain(p_full, pblk & 1);
      if (pj == 0) mbar_wait_plain(o_free, (pi & 1) ^ 1);    // the previous item's O has left TMEM
      mbar_wait_plain(&v_full[st], (pg / KSTT) & 1);
      tc_fence_after();
      const uint64_t dv = umma_desc_sw128(smem_u32(sV + st * kKVSlot), 1024, 8192);
      TRACE(6, pblk);
      if (issuer) {
#pragma unroll
        for (int k = 0; k < BKVT / 16; ++k)
          umma_bf16_ts(tO, tP + 8 * k, dv + uint64_t((k * 2048) >> 4), idesc_o, (pj | k) != 0);
        umma_commit(&v_empty[st]);
        umma_commit(o_full);
      }
      __syncwarp();
    };
    bool pv_pending = false;
    int pv_j = 0, pv_i = 0;
    uint32_t pv_g = 0, pv_blk = 0;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++i) {
      const int buf = i % NQT;
      const uint64_t dq = umma_desc_sw128(smem_u32(sQ + buf * kTile), 1024);
      if (i == 0) {
        mbar_wait_plain(&q_full[0], 0);
        issue_s(dq, 0, 0);
      }
      const bool has_next = it + (int)gridDim.x < n_items;
      for (int j = 0; j < nblk; ++j) {
        // The score block that follows block j in this CTA's stream -- the next key block of this
        // item, or the FIRST block of the next item (its Q sits in the next ring buffer) -- is
        // issued as soon as S_j is in registers, so it overlaps softmax j also across items.
        if (j + 1 < nblk) {
          mbar_wait_plain(s_free, (blk + j) & 1);
          issue_s(dq, g + j + 1, blk + j + 1);
        } else if (has_next) {
          const int nb = (i + 1) % NQT;
          TRACE(9, blk + nblk);
          mbar_wait_plain(&q_full[nb], ((i + 1) / NQT) & 1);
          TRACE(10, blk + nblk);
          mbar_wait_plain(s_free, (blk + j) & 1);
          TRACE(11, blk + nblk);
          issue_s(umma_desc_sw128(smem_u32(sQ + nb * kTile), 1024), g + nblk, blk + nblk);
        }
        // P V of block j -- or, in a kLagPV build, of the block before.  The trace (scripts/attn_trace.py, S = 1024)
        // shows S(j+1) leaving ~880 clk after S(j) is in warp 0's registers and reaching the softmax warps ~220 clk
        // late; p_full(j-1) is signalled ~140 clk before s_free(j), so P V(j-1) (8 MMAs) is issued just ahead of
        // S(j+1).  Issuing P V one block late (S first) did NOT move the S issue (still +780 clk: it follows the
        // slowest of the four softmax warps' s_free arrivals, not the P V) and made o_full late (wait 105 -> 455
        // clk): 743 -> 683 TFLOP/s at the L2 shape.  Kept off.
        if constexpr (kLagPV) {
          if (pv_pending) issue_pv(pv_j, pv_g, pv_blk, pv_i);
          pv_pending = true; pv_j = j; pv_g = g + j; pv_blk = blk + j; pv_i = i;
        } else {
          issue_pv(j, g + j, blk + j, i);
        }
      }
      g += nblk;
      blk += nblk;
    }
    if (pv_pending) issue_pv(pv_j, pv_g, pv_blk, pv_i);
  } else if (warp == 6) {
    // ================= store warp: staged O tiles leave by TMA, buffers go back to the ring ======
    setmaxnreg_dec<kRegsAux>();
    if (lane == 0) {
      int i = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++i) {
        const int qtile = it % ntile, head = (it / ntile) % a.heads, b = it / (ntile * a.heads);
        const int buf = i % NQT;
        mbar_wait(&o_staged[buf], (i / NQT) & 1, 41);
        tma_store_4d(&tmO, kOwnO ? sO : sQ + buf * kTile, head * HD, qtile * BQ, b, 0);
        tma_store_commit();
        tma_store_wait_read<0>();          // smem read out: the buffer may take the Q of item i + NQT (or the next O)
        mbar_arrive(kOwnO ? so_empty : &q_empty[buf]);
      }
      tma_store_wait<0>();
    }
  } else if (warp > 6) {
    setmaxnreg_dec<kRegsAux>();          // idle warp of the auxiliary warpgroup, except:
    if constexpr (kSplitQ) {
      // one-block items: the Q ring is fed by this warp.  One producer thread issuing Q, K, V in item order
      // stalls on the K / V stages of item i + 2 (free only once P V of item i has retired) before it can ask
      // for the Q of item i + 3, which held the Q lead at about one item period whatever the ring depth
      // (q_full wait of the MMA warp: 1380 clk per item, scripts/attn_trace_cross.py).
      if (lane == 0) {
        tma_prefetch_desc(&tmQ);
        int i = 0;
        for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++i) {
          const int qtile = it % ntile, head = (it / ntile) % a.heads, b = it / (ntile * a.heads);
          const int buf = i % NQT;
          mbar_wait_plain(&q_empty[buf], ((i / NQT) & 1) ^ 1);
          mbar_arrive_expect_tx(&q_full[buf], kTile);
          tma_load_3d(sQ + buf * kTile, &tmQ, &q_full[buf], head * HD, qtile * BQ, b);
        }
      }
    }
  } else {
    // ================= softmax + output (warps 0..3) =================
    setmaxnreg_inc<kRegsSoftmax>();
    const int row = warp * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_off = uint32_t(warp * 32) << 16;
    const uint32_t tS = tmem_base + lane_off;
    const uint32_t tO = tS + 128;
    const uint32_t tP = tS + 192;
    const float2 sc2 = kMask ? make_float2(1.f, 1.f) : make_float2(a.scale_log2, a.scale_log2);
    uint32_t blk = 0;
    int i = 0;
    // deferred epilogue of the previous item
    int pend_it = -1;
    float pend_inv = 0.f, pend_lse = 0.f;
    bool p_pending = false;              // a P store is in flight, p_full not yet signalled
    auto epilogue = [&](int pit, int pi, float inv, float lse, uint32_t last_blk) {
      // O of item pit (ring slot pi % NQT): TMEM -> normalise -> its dead Q buffer -> store warp
      mbar_wait_plain(o_full, last_blk & 1);
      TRACE(7, pi);
      tc_fence_after();
      uint8_t* orow = (kOwnO ? sO : sQ + (pi % NQT) * kTile) + row * 128;
      const int rx = row & 7;
      if constexpr (kOwnO) mbar_wait_plain(so_empty, (pi & 1) ^ 1);    // the previous item's store has read the buffer
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t o[32];
        tmem_ld32(tO + c * 32, o);
        tmem_ld_wait();
        if (c == HD / 32 - 1) {
          tc_fence_before();
          mbar_arrive(o_free);           // the next item's first P V may overwrite O
        }
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
      fence_proxy_async_smem();
      mbar_arrive(&o_staged[pi % NQT]);
      TRACE(8, pi);
      if (a.lse != nullptr) {
        const int qt = (pit % ntile) * BQ;
        if (qt + row < a.Sq) {
          const int head = (pit / ntile) % a.heads, b = pit / (ntile * a.heads);
          a.lse[((int64_t)b * a.heads + head) * a.Sq + qt + row] = lse;
        }
      }
    };
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++i) {
      float m = -INFINITY;               // running (possibly stale) row max, log2 domain
      float2 l2 = make_float2(0.f, 0.f);
      for (int j = 0; j < nblk; ++j) {
        const int kv0 = j * BKVT;
        const bool tail = kv0 + BKVT > a.Sk;
        mbar_wait_plain(s_full, (blk + j) & 1);
        TRACE(0, blk + j);
        tc_fence_after();
        uint32_t r[BKV];
#pragma unroll
        for (int c = 0; c < (BKVT + 31) / 32; ++c) tmem_ld32(tS + c * 32, r + c * 32);   // (80: 96 columns read)
        if (p_pending) {
          // the previous block's P store completes under this block's S load (traces: the
          // tcgen05.wait::st + fence + arrive chain cost ~200 clk per block when it sat behind
          // the store itself)
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(p_full);
          p_pending = false;
        }
        tmem_ld_wait();
        TRACE(1, blk + j);
        tc_fence_before();
        mbar_arrive(s_free);             // the tensor pipe may overwrite S now
        if (tail) {
          // (an 80-key block is only launched for 64 < Sk <= 80: its masked keys are among the last 16)
#pragma unroll
          for (int e = (BKVT < BKV ? 64 : 0); e < BKVT; ++e)
            if (kv0 + e >= a.Sk) r[e] = 0xff800000u;  // -inf: excluded from max, exp2 -> 0
        }
        if constexpr (kMask) {
          const int qtile = it % ntile, head = (it / ntile) % a.heads, b = it / (ntile * a.heads);
          const int qrow = min(qtile * BQ + row, a.Sq - 1);
          const float* mrow = a.mask + (int64_t)b * a.mask_sb + (int64_t)head * a.mask_sh +
                              (int64_t)qrow * a.mask_sq + kv0;
          constexpr float kLog2e = 1.4426950408889634f;
          if (!tail && a.mask_vec4) {
#pragma unroll
            for (int e = 0; e < BKV; e += 4) {
              const float4 mk = __ldg(reinterpret_cast<const float4*>(mrow + e));
              r[e] = __float_as_uint(fmaf(__uint_as_float(r[e]), a.scale_log2, mk.x * kLog2e));
              r[e + 1] = __float_as_uint(fmaf(__uint_as_float(r[e + 1]), a.scale_log2, mk.y * kLog2e));
              r[e + 2] = __float_as_uint(fmaf(__uint_as_float(r[e + 2]), a.scale_log2, mk.z * kLog2e));
              r[e + 3] = __float_as_uint(fmaf(__uint_as_float(r[e + 3]), a.scale_log2, mk.w * kLog2e));
            }
          } else {
#pragma unroll
            for (int e = 0; e < BKV; ++e) {
              const float mk = kv0 + e < a.Sk ? __ldg(mrow + e) : 0.f;
              r[e] = __float_as_uint(fmaf(__uint_as_float(r[e]), a.scale_log2, mk * kLog2e));
            }
          }
        }
        const float sl = kMask ? 1.f : a.scale_log2;
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int e = 0; e < BKVT; e += 8) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
            mx4[c] = fmaxf(mx4[c], fmaxf(__uint_as_float(r[e + 2 * c]), __uint_as_float(r[e + 2 * c + 1])));
        }
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_cand = fmaxf(m, mx * sl);
        const bool grow = m_cand - m > kRescaleThreshold;   // true on the first block (m = -inf)
        const float m_use = grow ? m_cand : m;
        const float alpha = grow ? ex2(m - m_use) : 1.f;
        m = m_use;
        l2 = __fmul2_rn(l2, make_float2(alpha, alpha));
        const float2 nm2 = make_float2(-m_use, -m_use);
#pragma unroll
        for (int e = 0; e < BKVT / 2; ++e) {
          const float2 sv = __ffma2_rn(
              make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1])), sc2, nm2);
          const bool poly = ((e * kPoly) & 7) + kPoly > 7 && kPoly > 0;
          const float2 pv = poly ? ex2_poly2(sv) : make_float2(ex2(sv.x), ex2(sv.y));
          l2 = __fadd2_rn(l2, pv);
          r[e] = pack_bf16(pv.x, pv.y);
        }
        TRACE(2, blk + j);
        if (j > 0) {
          mbar_wait_plain(o_full, (blk + j - 1) & 1);   // P V of the previous block is done with P and O
          TRACE(3, blk + j);
          tc_fence_after();
          if (__any_sync(0xffffffffu, grow)) {
            const float2 al2 = make_float2(alpha, alpha);
#pragma unroll
            for (int c = 0; c < HD / 32; ++c) {
              uint32_t o[32];
              tmem_ld32(tO + c * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int e = 0; e < 16; ++e) {
                const float2 tt = __fmul2_rn(
                    make_float2(__uint_as_float(o[2 * e]), __uint_as_float(o[2 * e + 1])), al2);
                o[2 * e] = __float_as_uint(tt.x);
                o[2 * e + 1] = __float_as_uint(tt.y);
              }
              tmem_st32(tO + c * 32, o);
            }
          }
        } else if (pend_it >= 0) {
          // the previous item's last P V retired while this block's exponentials ran: its O leaves
          // TMEM now, before this item's first P V (gated by o_free) overwrites it
          epilogue(pend_it, i - 1, pend_inv, pend_lse, blk - 1);
        }
        tmem_st32(tP, r);
        if (BKVT > 64) tmem_st32(tP + 32, r + 32);      // (80 keys = 40 packed words: words 40..63 are never read)
        if constexpr (BKVT < BKV && VST_ATTN_X_EAGER_P) {
          // one-block items (text cross-attention): the deferred arrive reaches the MMA warp only when the NEXT
          // item's S is already being loaded, and that warp issues in program order -- P V of this item, then
          // the S of the item after next -- so every item's S was issued ~1600 clk after the previous S left
          // TMEM and the softmax warps idled 600-1000 clk per item (scripts/attn_trace_cross.py).  Paying the
          // store's completion here (~200 clk) lets the MMA warp run a whole item ahead.
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(p_full);
        } else {
          p_pending = true;              // completion + p_full arrive: under the next S load
        }
        TRACE(4, blk + j);
      }
      const float l = l2.x + l2.y;
      pend_it = it;
      pend_inv = 1.f / l;
      pend_lse = (m + log2f(l)) * 0.6931471805599453f;
      blk += nblk;
    }
    if (p_pending) {
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    if (pend_it >= 0) epilogue(pend_it, i - 1, pend_inv, pend_lse, blk - 1);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

int make_qkv_map(CUtensorMap* m, const void* p, int64_t B, int64_t S, int width, int64_t ld,
                 int rank4, int box_rows = 128) {
  const uint64_t dims[4] = {(uint64_t)width, (uint64_t)S, (uint64_t)B, 1};
  const uint64_t strides[4] = {2, (uint64_t)ld * 2, (uint64_t)S * ld * 2,
                               (uint64_t)B * S * ld * 2};
  const uint32_t box[4] = {HD, (uint32_t)box_rows, 1, 1};
  return vst_make_tmap(m, p, rank4 ? 4 : 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace

#ifdef VST_ATTN_TRACE
extern "C" int vst_attn_debug_trace(long long* host_dst) {
  return (int)cudaMemcpyFromSymbol(host_dst, g_attn_trace, sizeof(long long) * 4 * 12 * 1024);
}
#endif

extern "C" int vst_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse,
                            int Bt, int heads, int Sq, int Sk, int d, int64_t ldq, int64_t ldk,
                            int64_t ldv, int64_t ldo, int kv_group, float scale, const float* mask,
                            int64_t mask_sb, int64_t mask_sh, int64_t mask_sq, void* stream) {
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
  // text cross-attention (Sk = 77): one 80-key block per item instead of a 128-key one (VST_ATTN_BKV80=0: off)
  static int allow80 = -1;
  if (allow80 < 0) { const char* e = getenv("VST_ATTN_BKV80"); allow80 = (e && e[0] == '0') ? 0 : 1; }
  const bool kv80 = allow80 && mask == nullptr && Sk > 64 && Sk <= 80;
  if ((rc = make_qkv_map(&tK, k, Bt / kv_group, Sk, width, ldk, 0, kv80 ? 80 : 128))) return rc;
  if ((rc = make_qkv_map(&tV, v, Bt / kv_group, Sk, width, ldv, 0, kv80 ? 80 : 128))) return rc;
  if ((rc = make_qkv_map(&tO, o, Bt, Sq, width, ldo, 1))) return rc;
  // kPoly: how many of every 8 exp2 pairs run on the FMA pipe.  VST_ATTN_POLY (0..3) overrides the
  // default, VST_ATTN_CTAS_PER_SM the persistent grid (tuning only).
  static int poly = -1, per_sm = 2;
  if (poly < 0) {
    const char* e = getenv("VST_ATTN_POLY");
    poly = e ? atoi(e) : kDefaultPoly;
    if (poly < 0 || poly > 3) poly = kDefaultPoly;
    if ((e = getenv("VST_ATTN_CTAS_PER_SM")) && atoi(e) > 0) per_sm = atoi(e);
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_pers_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesP));
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_pers_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesP));
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_pers_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesP));
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_pers_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesP));
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_pers_kernel<kDefaultPoly, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesP));
    VST_CUDA_OK(cudaFuncSetAttribute(attn_fwd_pers_kernel<kDefaultPoly, false, 80>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes80));
  }
  const int vec4 = mask && (reinterpret_cast<uintptr_t>(mask) & 15) == 0 && mask_sb % 4 == 0 &&
                   mask_sh % 4 == 0 && mask_sq % 4 == 0;
  AttnArgs a{lse, Sq, Sk, heads, kv_group, scale * 1.4426950408889634f, mask, mask_sb, mask_sh, mask_sq, vec4};
  cudaStream_t st = (cudaStream_t)stream;
  const int ntile = (Sq + BQ - 1) / BQ;
  const int64_t items = (int64_t)ntile * heads * Bt;
  VST_REQUIRE(items < (1ll << 31), VST_E_UNSUPPORTED, "attn_fwd: too many tiles");
  const int slots = per_sm * vst_sm_count();
  const int pgrid = (int)(items < slots ? items : slots);
  const int n = (int)items;
  if (mask) {
    attn_fwd_pers_kernel<kDefaultPoly, true><<<pgrid, kThreads, kSmemBytesP, st>>>(tQ, tK, tV, tO, a, n, ntile);
  } else if (kv80) {
    attn_fwd_pers_kernel<kDefaultPoly, false, 80><<<pgrid, kThreads, kSmemBytes80, st>>>(tQ, tK, tV, tO, a, n, ntile);
  } else {
    switch (poly) {
      case 0: attn_fwd_pers_kernel<0, false><<<pgrid, kThreads, kSmemBytesP, st>>>(tQ, tK, tV, tO, a, n, ntile); break;
      case 1: attn_fwd_pers_kernel<1, false><<<pgrid, kThreads, kSmemBytesP, st>>>(tQ, tK, tV, tO, a, n, ntile); break;
      case 3: attn_fwd_pers_kernel<3, false><<<pgrid, kThreads, kSmemBytesP, st>>>(tQ, tK, tV, tO, a, n, ntile); break;
      default: attn_fwd_pers_kernel<2, false><<<pgrid, kThreads, kSmemBytesP, st>>>(tQ, tK, tV, tO, a, n, ntile); break;
    }
  }
  VST_LAUNCH_OK();
  return VST_OK;
}
