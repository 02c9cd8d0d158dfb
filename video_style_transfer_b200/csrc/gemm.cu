// a3/a4/a5/a7: the fused LoRA GEMM on tcgen05.
//
//   D[M,N] = (A1[M,K1] B1[N,K1]^T + A2[M,K2] B2[N,K2]^T + bias[N] + residual[M,N]) * out_scale
//
// The low-rank update rides in the same TMEM accumulator as extra K-blocks (A2 = u = x A^T,
// B2 = scale * m (.) B packed by vst_lora_pack_up), so W x + m_c(.)(B_c A_c x) + m_s(.)(B_s A_s x)
// is ONE pass over x and one write of y.  Reference arithmetic: unziplora_unet/lora_linear.py
// :74-81, unziplora_linear_layer.py:298-346 (which does 3 dense GEMMs + 2 in x out
// materialisations per call), lora_unzip.py:66-75, animatediff/temporal_lora.py:29-32.
//
// Kernel shape (sm_100a): persistent, one CTA per SM, 192 threads:
//   warp 0      TMA producer  (A/B k-blocks -> 128B-swizzled smem ring, mbarrier full/empty)
//   warp 1      MMA issuer    (one elected lane: tcgen05.mma 128 x BN x 16, accumulators in TMEM,
//                              two accumulator stages so the epilogue overlaps the next tile)
//   warps 2..5  epilogue      (tcgen05.ld -> +bias, +residual, *scale -> bf16/fp32 -> global)
// Roofline: tensor pipe; algorithmic FLOPs = 2*M*N*(K1+K2).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

static bool vst_force_single_cta() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("VST_GEMM_SINGLE_CTA"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // one 128-byte swizzle atom of bf16
constexpr int kEpiWarps = 8;                 // 2 per TMEM lane quarter, interleaved over column chunks
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kStageOutBytes = 32 * 32 * 2;  // one 32-row x 32-col bf16 chunk per TMA store
// RES kernels (bf16 output + residual): the residual chunk is brought in by TMA into the chunk's own
// staging buffer kResAhead chunks ahead, added in place and sent back out by the TMA store.
#ifndef VST_RES_BUFS
#define VST_RES_BUFS 4
#define VST_RES_AHEAD 2
#endif
constexpr int kResBufs = VST_RES_BUFS, kResAhead = VST_RES_AHEAD;
constexpr int kSmemTotal = 227 * 1024;       // per-CTA limit; 1024 alignment slack + 512 B of barriers
constexpr int kBarBytes = 512;

// staging buffers per epilogue warp of the kernels without a residual (tuning: -DVST_OUT_BUFS=n; four instead of two
// measured within 2 % either way on the K = 320 shapes and 2-4 % slower on the others -- fewer operand stages --, so
// the store read-back is not what bounds the short-K epilogue)
#ifndef VST_OUT_BUFS
#define VST_OUT_BUFS 2
#endif
constexpr int kOutBufs = VST_OUT_BUFS;
constexpr int out_bytes(bool res) { return kEpiWarps * (res ? kResBufs : kOutBufs) * kStageOutBytes; }
constexpr int ring_budget(bool res) { return kSmemTotal - 1024 - kBarBytes - out_bytes(res); }

template <int BN, bool RES = false>
struct Cfg {
  static constexpr int kStageBytes = (BM + BN) * BK * 2;
  static constexpr int kOutBytes = out_bytes(RES);
  static constexpr int kStages = (ring_budget(RES) / kStageBytes) > 8 ? 8 : (ring_budget(RES) / kStageBytes);
  static constexpr int kAccStride = BN <= 64 ? 64 : (BN <= 128 ? 128 : 256);
  static constexpr int kTmemCols = 2 * kAccStride;  // power of two >= 32
  static constexpr int kSmemBytes = kStages * kStageBytes + kOutBytes + 1024 /*align*/ + kBarBytes;
};

struct EpiArgs {
  const void* bias;
  const void* residual;
  void* D;
  int64_t ldr, ldd, M;
  int N;
  float out_scale;
  int bias_bf16, out_f32, act;
  int debug;   // VST_GEMM_DEBUG (profiling only; results are wrong when set): 1 = no stores,
               // 2 = no epilogue work at all, 4 = no MMAs, 8 = no TMA loads
};

// Bias of a warp's 32-column chunk: lane l holds column l's value (one coalesced load per tile and
// chunk, issued before the accumulator wait); the row-owning lanes read column j by shuffle.
__device__ __forceinline__ float bias_lane(const EpiArgs& ep, int col, int limit) {
  if (ep.bias == nullptr || col >= limit || (ep.debug & 64)) return 0.f;
  return ep.bias_bf16 ? __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(ep.bias) + col))
                      : __ldg(reinterpret_cast<const float*>(ep.bias) + col);
}

// One 32-column chunk of one accumulator row: +bias, act, +residual, *scale, store.
__device__ __forceinline__ void epilogue_chunk(const EpiArgs& ep, const uint32_t* r, int64_t row,
                                               bool row_ok, int n0, int c0, float bl) {
#pragma unroll
  for (int v8 = 0; v8 < 4; ++v8) {
    const int col = n0 + c0 + v8 * 8;
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = __uint_as_float(r[v8 * 8 + j]);
    if (ep.bias != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += __shfl_sync(0xffffffffu, bl, v8 * 8 + j);
    }
    if (!row_ok || col >= ep.N) continue;
    if (ep.act == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        o[j] = 0.5f * o[j] * (1.f + erff(o[j] * 0.70710678118654752f));  // exact GELU
    }
    if (ep.out_f32) {
      float* dp = reinterpret_cast<float*>(ep.D) + row * ep.ldd + col;
      if (ep.residual != nullptr) {
        const float* rp = reinterpret_cast<const float*>(ep.residual) + row * ep.ldr + col;
        const float4 r0 = *reinterpret_cast<const float4*>(rp);
        const float4 r1 = *reinterpret_cast<const float4*>(rp + 4);
        o[0] += r0.x; o[1] += r0.y; o[2] += r0.z; o[3] += r0.w;
        o[4] += r1.x; o[5] += r1.y; o[6] += r1.z; o[7] += r1.w;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] *= ep.out_scale;
      *reinterpret_cast<float4*>(dp) = make_float4(o[0], o[1], o[2], o[3]);
      *reinterpret_cast<float4*>(dp + 4) = make_float4(o[4], o[5], o[6], o[7]);
    } else {
      __nv_bfloat16* dp = reinterpret_cast<__nv_bfloat16*>(ep.D) + row * ep.ldd + col;
      if (ep.residual != nullptr) {
        const uint4 rv = *reinterpret_cast<const uint4*>(
            reinterpret_cast<const __nv_bfloat16*>(ep.residual) + row * ep.ldr + col);
        const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 e = unpack_bf16(rw[j]);
          o[2 * j] += e.x;
          o[2 * j + 1] += e.y;
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] *= ep.out_scale;
      uint4 w;
      w.x = pack_bf16(o[0], o[1]); w.y = pack_bf16(o[2], o[3]);
      w.z = pack_bf16(o[4], o[5]); w.w = pack_bf16(o[6], o[7]);
      *reinterpret_cast<uint4*>(dp) = w;
    }
  }
}

// bias / activation / residual / scale of one 32-column chunk, packed to bf16 (TMA-store path).
// res: the 32 residual values of this row chunk (bf16 pairs, zero where there is none).
__device__ __forceinline__ void epilogue_pack(const EpiArgs& ep, const uint32_t* r, float bl,
                                              const uint4* res, uint32_t* w) {
#pragma unroll
  for (int v8 = 0; v8 < 4; ++v8) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = __uint_as_float(r[v8 * 8 + j]);
    if (ep.bias != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += __shfl_sync(0xffffffffu, bl, v8 * 8 + j);
    }
    if (ep.act == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = 0.5f * o[j] * (1.f + erff(o[j] * 0.70710678118654752f));
    }
    if (res != nullptr) {
      const uint32_t rw[4] = {res[v8].x, res[v8].y, res[v8].z, res[v8].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 e = unpack_bf16(rw[j]);
        o[2 * j] += e.x;
        o[2 * j + 1] += e.y;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      w[v8 * 4 + j] = pack_bf16(o[2 * j] * ep.out_scale, o[2 * j + 1] * ep.out_scale);
  }
}

// Epilogue of one 128-row x BN accumulator by the 8 epilogue warps of a CTA.  Warp e owns TMEM
// lane quarter e % 4 and every second 32-column chunk.  bf16 output without residual goes
// TMEM -> registers -> 64B-swizzled smem chunk -> TMA store (full-line, asynchronous writes; the
// store clips rows >= M and columns >= N); otherwise rows are written directly.
// Latency hiding (r01 VST_GEMM_DEBUG measurements, profiles/gemm_r01_notes.md): the per-chunk bias
// loads used to sit on the critical path (L0 temporal projection: 18 of 75 us), so the bias of
// the warp's chunks is fetched by `epilogue_bias` BEFORE the accumulator wait, and the TMEM load of
// chunk i+1 is in flight while chunk i is packed and stored.  `release` is called as soon as the
// last TMEM load has landed, so the tensor pipe gets the accumulator back before the stores.
template <int BN>
struct EpiBias {
  static constexpr int kMax = (BN + 63) / 64;
  float v[kMax];
};

template <int BN>
__device__ __forceinline__ void epilogue_bias(const EpiArgs& ep, int n0, int ewarp, int lane,
                                              EpiBias<BN>& b) {
  const int half = ewarp >> 2;
#pragma unroll
  for (int i = 0; i < EpiBias<BN>::kMax; ++i) b.v[i] = bias_lane(ep, n0 + half * 32 + i * 64 + lane, ep.N);
}

template <int BN, typename Release>
__device__ __forceinline__ void epilogue_tile(const EpiArgs& ep, const CUtensorMap* tmD,
                                              uint32_t tacc, int m0, int n0, int ewarp, int lane,
                                              uint8_t* stage_out, int& buf, const EpiBias<BN>& bias,
                                              Release release) {
  // a warp can only read the TMEM lane quarter given by its hardware warp id % 4 (warps 2..9)
  const int quarter = (ewarp + 2) & 3, half = ewarp >> 2;
  const int64_t row = (int64_t)m0 + quarter * 32 + lane;
  const bool row_ok = row < ep.M;
  const uint32_t taddr = tacc + (uint32_t(quarter * 32) << 16);
  // bf16 output always leaves through the staged TMA store (bf16 + residual: epilogue_res_loop);
  // fp32 output (parameter-gradient GEMMs, config-1 fp32 activations) is written row by row
  const bool via_tma = !ep.out_f32 && !(ep.debug & 16);   // 16: direct stores
  constexpr int kMax = EpiBias<BN>::kMax;
  const int nch = (BN - half * 32 + 63) / 64;          // chunks of this warp: c0 = half*32 + 64 i
  uint32_t r[2][32];
  tmem_ld32(taddr + half * 32, r[0]);
#pragma unroll
  for (int i = 0; i < kMax; ++i) {
    if (i >= nch) break;
    const int c0 = half * 32 + i * 64;
    tmem_ld_wait();
    if (i + 1 < nch) {
      tmem_ld32(taddr + c0 + 64, r[(i + 1) & 1]);
    } else {
      tc_fence_before();
      __syncwarp();
      release();
    }
    const uint32_t* rc = r[i & 1];
    if (n0 + c0 >= ep.N) continue;
    if (ep.debug & 2) continue;
    if (via_tma) {
      uint32_t w[16];
      epilogue_pack(ep, rc, bias.v[i], nullptr, w);
      uint8_t* sbuf = stage_out + buf * kStageOutBytes;
      if (lane == 0) tma_store_wait_read<kOutBufs - 1>();   // the store that last used this buffer is done
      __syncwarp();
      uint8_t* srow = sbuf + lane * 64;
      const int sx = (lane >> 1) & 3;                // SWIZZLE_64B: 16-B chunk ^= (row >> 1) & 3
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<uint4*>(srow + ((q ^ sx) * 16)) =
            make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
      if (!(ep.debug & 32)) fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0 && !(ep.debug & 1)) {
        tma_store_2d(tmD, sbuf, n0 + c0, m0 + quarter * 32);
        tma_store_commit();
      }
      buf = buf + 1 == kOutBufs ? 0 : buf + 1;
    } else {
      epilogue_chunk(ep, rc, row, row_ok, n0, c0, bias.v[i]);
    }
  }
}

// ---- RES epilogue: bf16 output + residual, the whole persistent loop of one epilogue warp -----------
// The residual tile never goes through the register-addressed row loads of the first version (each
// lane fetching its own 64-byte row segment: 8 warps x 2 KB in flight per SM, L0 to_out + residual
// 154.7 us against 74.8 us without residual, r01).  Each 32 x 32 chunk of the residual is fetched by
// ONE TMA load (64B swizzle, the layout of the output staging buffer) into the buffer the chunk
// will be stored from, kResAhead chunks ahead of its use and across tile boundaries, so while the
// warp waits for an accumulator the residual of the coming chunks is already in flight.  The
// epilogue then reads its row's 64 bytes from shared memory, adds, packs, writes the result IN
// PLACE and sends the buffer out by TMA store.  A buffer is refilled once the store issued from it
// kResBufs - kResAhead chunks ago has finished reading (cp.async.bulk.wait_group.read).
// coords(tile, m0, n0) maps a tile index of this CTA's walk to its origin; release(acc) hands the
// accumulator stage back to the MMA issuer.
template <int BN, typename Coords, typename Release>
__device__ __forceinline__ void epilogue_res_loop(const EpiArgs& ep, const CUtensorMap* tmD,
                                                  const CUtensorMap* tmR, uint32_t tmem_base, int acc_stride,
                                                  uint64_t* acc_full, uint8_t* my_out, uint64_t* my_bar,
                                                  int ewarp, int lane, int first, int stride, int total,
                                                  Coords coords, Release release) {
  constexpr int kMax = EpiBias<BN>::kMax;
  const int quarter = (ewarp + 2) & 3, half = ewarp >> 2;
  const int nch = (BN - half * 32 + 63) / 64;             // chunks of this warp per tile
  const int my_tiles = first < total ? (total - first + stride - 1) / stride : 0;
  const int nseq = my_tiles * nch;
  auto issue = [&](int seq) {                             // lane 0: residual chunk `seq` -> its buffer
    const int tk = seq / nch, i = seq - tk * nch;
    int m0, n0;
    coords(first + tk * stride, m0, n0);
    const int b = seq % kResBufs;
    mbar_arrive_expect_tx(&my_bar[b], kStageOutBytes);
    tma_load_2d(my_out + b * kStageOutBytes, tmR, &my_bar[b], n0 + half * 32 + i * 64, m0 + quarter * 32);
  };
  if (lane == 0) {
    tma_prefetch_desc(tmD);
    tma_prefetch_desc(tmR);
#pragma unroll
    for (int a = 0; a < kResAhead; ++a)
      if (a < nseq) issue(a);
  }
  int seq = 0, acc = 0;
  uint32_t acc_phase = 0;
  for (int tk = 0; tk < my_tiles; ++tk) {
    int m0, n0;
    coords(first + tk * stride, m0, n0);
    EpiBias<BN> bias;
    epilogue_bias<BN>(ep, n0, ewarp, lane, bias);
    mbar_wait(&acc_full[acc], acc_phase, 4);
    tc_fence_after();
    const uint32_t taddr = tmem_base + acc * acc_stride + (uint32_t(quarter * 32) << 16);
    uint32_t r[2][32];
    tmem_ld32(taddr + half * 32, r[0]);
#pragma unroll
    for (int i = 0; i < kMax; ++i) {
      if (i >= nch) break;
      const int c0 = half * 32 + i * 64;
      const int b = seq % kResBufs;
      uint8_t* sbuf = my_out + b * kStageOutBytes;
      tmem_ld_wait();
      if (i + 1 < nch) {
        tmem_ld32(taddr + c0 + 64, r[(i + 1) & 1]);
      } else {
        tc_fence_before();
        __syncwarp();
        release(acc);
      }
      mbar_wait(&my_bar[b], (seq / kResBufs) & 1, 5);     // the residual chunk has landed
      uint8_t* srow = sbuf + lane * 64;
      const int sx = (lane >> 1) & 3;                    // SWIZZLE_64B: 16-B chunk ^= (row >> 1) & 3
      uint4 res[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) res[q] = *reinterpret_cast<const uint4*>(srow + ((q ^ sx) * 16));
      uint32_t w[16];
      epilogue_pack(ep, r[i & 1], bias.v[i], res, w);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<uint4*>(srow + ((q ^ sx) * 16)) =
            make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (n0 + c0 < ep.N) tma_store_2d(tmD, sbuf, n0 + c0, m0 + quarter * 32);
        tma_store_commit();                               // (an empty group keeps the count uniform)
        if (seq + kResAhead < nseq) {
          tma_store_wait_read<kResBufs - kResAhead>();    // the store that last used that buffer is out
          issue(seq + kResAhead);
        }
      }
      ++seq;
    }
    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
  }
  if (lane == 0) tma_store_wait<0>();
}

// ---- VST_ACT_GEGLU: D[:, j] = (acc[:, j] + bias[j]) * gelu(acc[:, BN/2 + j] + bias[N + j]) ----------
// The tile's B rows are [W_hidden[n0 : n0+BN/2] ; W_gate[n0 : n0+BN/2]] (the leader CTA loads the
// hidden half, its peer the gate half), so value and gate of an output column sit BN/2 accumulator
// columns apart and the 2*inner-wide pre-activation never goes to HBM (diffusers GEGLU:
// hidden, gate = proj(x).chunk(2); hidden * gelu(gate)).
// erf by Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7, below fp32 accumulation noise; the products are
// rounded to bf16): 1 MUFU.RCP + 1 MUFU.EX2 + 9 FMA-pipe ops instead of erff's ~25.
__device__ __forceinline__ float gelu_as(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.f)));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * z * z));
  const float erfa = fmaf(-p * t, e, 1.f);            // erf(|x| / sqrt 2)
  return 0.5f * fmaf(fabsf(x), erfa, x);              // 0.5 x (1 + sign(x) erf) = 0.5 (x + |x| erf)
}

// The same formula on a PAIR with packed f32x2 math, returning 2 gelu(x) = x + |x| erf(|x| / sqrt 2) (the caller
// folds the 0.5 into its output scale).  r02 launch list: the L0 GEGLU GEMM (K = 320) ran at 715 TFLOP/s against
// 1559 at L2 -- ~8000 clk of epilogue per 128 x 128 output tile against ~2600 clk of MMA: ~22 scalar FMA-pipe
// instructions per element plus four bias shuffles per pair.  Packed: 10 FMA-pipe instructions per PAIR + 4 MUFU.
__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }
// -DVST_GELU_ONE_MUFU=1 selects the one-MUFU form below.  Measured SLOWER (r02, kbench geglu L0: 980 vs 930 us):
// the epilogue is bound by dependent FMA-pipe latency at two warps per scheduler, and this form is one ten-deep
// serial chain where rcp + ex2 gives two shorter independent ones.  Kept for the record, off by default.
#ifndef VST_GELU_ONE_MUFU
#define VST_GELU_ONE_MUFU 0
#endif
#if VST_GELU_ONE_MUFU
// Abramowitz-Stegun 7.1.28: erf(z) = 1 - (1 + a1 z + ... + a6 z^6)^-16, |err| < 3e-7.  ONE MUFU (rcp) per element
// instead of rcp + ex2: 6 ffma2 + 4 fmul2 (squarings) per pair.  z large: the power overflows to +inf, rcp -> 0, erf -> 1.
__device__ __forceinline__ float2 gelu2x_as(float2 x) {
  const float2 ax = make_float2(fabsf(x.x), fabsf(x.y));
  const float2 z = __fmul2_rn(ax, f2(0.70710678118654752f));
  float2 d = __ffma2_rn(f2(0.0000430638f), z, f2(0.0002765672f));
  d = __ffma2_rn(d, z, f2(0.0001520143f));
  d = __ffma2_rn(d, z, f2(0.0092705272f));
  d = __ffma2_rn(d, z, f2(0.0422820123f));
  d = __ffma2_rn(d, z, f2(0.0705230784f));
  d = __ffma2_rn(d, z, f2(1.f));
  d = __fmul2_rn(d, d);
  d = __fmul2_rn(d, d);
  d = __fmul2_rn(d, d);
  d = __fmul2_rn(d, d);
  float2 r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(d.y));
  // x + |x| erf = (x + |x|) - |x| r
  return __ffma2_rn(make_float2(-ax.x, -ax.y), r, __fadd2_rn(x, ax));
}
#else
__device__ __forceinline__ float2 gelu2x_as(float2 x) {
  const float2 ax = make_float2(fabsf(x.x), fabsf(x.y));
  const float2 den = __ffma2_rn(f2(0.3275911f * 0.70710678118654752f), ax, f2(1.f));
  float2 t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(den.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(den.y));
  float2 p = __ffma2_rn(f2(1.061405429f), t, f2(-1.453152027f));
  p = __ffma2_rn(p, t, f2(1.421413741f));
  p = __ffma2_rn(p, t, f2(-0.284496736f));
  p = __ffma2_rn(p, t, f2(0.254829592f));
  // exp(-z^2), z = |x| / sqrt 2:  2^(-(|x| sqrt(log2 e / 2))^2)
  const float2 sx = __fmul2_rn(ax, f2(0.84932180028801904f));
  const float2 arg = __fmul2_rn(sx, make_float2(-sx.x, -sx.y));
  float2 e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(arg.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(arg.y));
  const float2 pt = __fmul2_rn(p, t);
  const float2 erfa = __ffma2_rn(make_float2(-pt.x, -pt.y), e, f2(1.f));
  return __ffma2_rn(ax, erfa, x);
}
#endif
// eight consecutive bias values, the same address in every lane (one broadcast load per 4 / 8 values)
__device__ __forceinline__ void bias8(const EpiArgs& ep, int col, float* b) {
  if (ep.bias_bf16) {
    const uint4 r = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(ep.bias) + col));
    float2 e;
    e = unpack_bf16(r.x); b[0] = e.x; b[1] = e.y;
    e = unpack_bf16(r.y); b[2] = e.x; b[3] = e.y;
    e = unpack_bf16(r.z); b[4] = e.x; b[5] = e.y;
    e = unpack_bf16(r.w); b[6] = e.x; b[7] = e.y;
  } else {
    const float4* q = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(ep.bias) + col);
    const float4 a = __ldg(q), c = __ldg(q + 1);
    b[0] = a.x; b[1] = a.y; b[2] = a.z; b[3] = a.w; b[4] = c.x; b[5] = c.y; b[6] = c.z; b[7] = c.w;
  }
}

template <int BN, typename Release>
__device__ __forceinline__ void epilogue_tile_geglu(const EpiArgs& ep, const CUtensorMap* tmD,
                                                    uint32_t tacc, int m0, int n0, int ewarp, int lane,
                                                    uint8_t* stage_out, int& buf, Release release) {
  constexpr int HALF = BN / 2;
  constexpr int kMax = (HALF + 63) / 64;
  const int quarter = (ewarp + 2) & 3, half = ewarp >> 2;
  const uint32_t taddr = tacc + (uint32_t(quarter * 32) << 16);
  const int nch = (HALF - half * 32 + 63) / 64;
#pragma unroll
  for (int i = 0; i < kMax; ++i) {
    if (i >= nch) break;
    const int c0 = half * 32 + i * 64;
    uint32_t rh[32], rg[32];
    tmem_ld32(taddr + c0, rh);
    tmem_ld32(taddr + HALF + c0, rg);
    tmem_ld_wait();
    if (i + 1 >= nch) {
      tc_fence_before();
      __syncwarp();
      release();
    }
    if (n0 + c0 >= ep.N) continue;
    uint32_t w[16];
    const float2 hs2 = f2(0.5f * ep.out_scale);        // gelu2x_as returns 2 gelu
    // N % 128 == 0 (checked by the entry point), so a chunk that starts below N lies inside [0, N): the broadcast
    // bias loads need no guard; bias (2N entries, 16-byte aligned) = [hidden | gate]
#pragma unroll
    for (int j8 = 0; j8 < 4; ++j8) {
      float bhv[8], bgv[8];
      if (ep.bias != nullptr) {
        bias8(ep, n0 + c0 + 8 * j8, bhv);
        bias8(ep, ep.N + n0 + c0 + 8 * j8, bgv);
      }
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int j = 4 * j8 + jj;
        float2 h = make_float2(__uint_as_float(rh[2 * j]), __uint_as_float(rh[2 * j + 1]));
        float2 g = make_float2(__uint_as_float(rg[2 * j]), __uint_as_float(rg[2 * j + 1]));
        if (ep.bias != nullptr) {
          h = __fadd2_rn(h, make_float2(bhv[2 * jj], bhv[2 * jj + 1]));
          g = __fadd2_rn(g, make_float2(bgv[2 * jj], bgv[2 * jj + 1]));
        }
        const float2 o = __fmul2_rn(__fmul2_rn(h, hs2), gelu2x_as(g));
        w[j] = pack_bf16(o.x, o.y);
      }
    }
    uint8_t* sbuf = stage_out + buf * kStageOutBytes;
    if (lane == 0) tma_store_wait_read<kOutBufs - 1>();
    __syncwarp();
    uint8_t* srow = sbuf + lane * 64;
    const int sx = (lane >> 1) & 3;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      *reinterpret_cast<uint4*>(srow + ((q ^ sx) * 16)) =
          make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(tmD, sbuf, n0 + c0, m0 + quarter * 32);
      tma_store_commit();
    }
    buf = buf + 1 == kOutBufs ? 0 : buf + 1;
  }
}

template <int BN, bool RES>
__global__ void __launch_bounds__(kThreads, 1)
lora_gemm_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                 const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                 const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmR,
                 int nk1, int nk2, int m_tiles, int n_tiles, int k2_group_n, int k2_cols,
                 EpiArgs ep) {
  using C = Cfg<BN, RES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* stage_out = smem + C::kStages * C::kStageBytes;   // 1024-aligned: stage sizes are k*1024
  uint64_t* full = reinterpret_cast<uint64_t*>(stage_out + C::kOutBytes);
  uint64_t* empty = full + C::kStages;
  uint64_t* acc_full = empty + C::kStages;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* res_bar = acc_empty + 2;                       // [kEpiWarps][kResBufs] (RES kernels)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + kEpiWarps * kResBufs);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = nk1 + nk2;
  const int total_tiles = m_tiles * n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    if (nk2 > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kEpiWarps);
    }
    if (RES)
      for (int i = 0; i < kEpiWarps * kResBufs; ++i) mbar_init(&res_bar[i], 1);
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
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + BM * BK * 2;
          mbar_arrive_expect_tx(&full[stage], C::kStageBytes);
          if (kb < nk1) {
            tma_load_2d(sa, &tmA1, &full[stage], kb * BK, m0);
            tma_load_2d(sb, &tmB1, &full[stage], kb * BK, n0);
          } else {
            // grouped low-rank operand: output columns of group g read A2 columns [g*K2, ...)
            const int a2_off = k2_group_n > 0 ? (n0 / k2_group_n) * k2_cols : 0;
            tma_load_2d(sa, &tmA2, &full[stage], a2_off + (kb - nk1) * BK, m0);
            tma_load_2d(sb, &tmB2, &full[stage], (kb - nk1) * BK, n0);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&acc_empty[acc], acc_phase ^ 1, 2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * C::kAccStride;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&full[stage], phase, 3);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t sb = sa + BM * BK * 2;
          const uint64_t da = umma_desc_sw128(sa, 1024);
          const uint64_t db = umma_desc_sw128(sb, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          umma_commit(&empty[stage]);  // frees the smem slot once these MMAs have read it
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&acc_full[acc]);   // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ================= epilogue (warps 2..9) =================
    const int ewarp = warp - 2;
    if constexpr (RES) {
      epilogue_res_loop<BN>(
          ep, &tmD, &tmR, tmem_base, C::kAccStride, acc_full, stage_out + ewarp * kResBufs * kStageOutBytes,
          res_bar + ewarp * kResBufs, ewarp, lane, blockIdx.x, gridDim.x, total_tiles,
          [&](int tile, int& m0, int& n0) { m0 = (tile / n_tiles) * BM; n0 = (tile % n_tiles) * BN; },
          [&](int acc) { if (lane == 0) mbar_arrive(&acc_empty[acc]); });
    } else {
      uint8_t* my_out = stage_out + ewarp * kOutBufs * kStageOutBytes;
      int buf = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      if (lane == 0) tma_prefetch_desc(&tmD);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
        EpiBias<BN> bias;
        epilogue_bias<BN>(ep, n0, ewarp, lane, bias);
        mbar_wait(&acc_full[acc], acc_phase, 4);
        tc_fence_after();
        uint64_t* done = &acc_empty[acc];
        epilogue_tile<BN>(ep, &tmD, tmem_base + acc * C::kAccStride, m0, n0, ewarp, lane, my_out, buf, bias,
                          [&]() { if (lane == 0) mbar_arrive(done); });
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (lane == 0) tma_store_wait<0>();
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
// CTA-pair variant (cta_group::2): a cluster of two CTAs (the two SMs of a TPC) owns a 256 x BN
// tile.  Each CTA stages its own 128 rows of A and HALF of the B tile (BN/2 rows), and the leader
// issues one tcgen05.mma M=256 that reads both halves: per SM the shared-memory traffic per MAC
// (TMA writes + MMA reads) drops by a third against the 128 x BN single-CTA tile, which is what
// bounds the single-CTA kernel at ~56 % of tensor peak.
//   full[s]      lives in the leader; both CTAs' TMA loads complete_tx on it
//   empty[s]     in both CTAs, released by a multicast tcgen05.commit
//   acc_full[a]  in both CTAs (multicast commit); acc_empty[a] in the leader, 8 warp arrivals
// ------------------------------------------------------------------------------------------------
template <int BN, bool RES = false>
struct Cfg2 {
  static constexpr int kBRows = BN / 2;
  static constexpr int kStageBytes = (BM + kBRows) * BK * 2;
  static constexpr int kOutBytes = out_bytes(RES);
  static constexpr int kStages = (ring_budget(RES) / kStageBytes) > 8 ? 8 : (ring_budget(RES) / kStageBytes);
  static constexpr int kAccStride = BN <= 64 ? 64 : (BN <= 128 ? 128 : 256);
  static constexpr int kTmemCols = 2 * kAccStride;
  static constexpr int kSmemBytes = kStages * kStageBytes + kOutBytes + 1024 + kBarBytes;
};

template <int BN, bool RES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
lora_gemm2_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                  const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                  const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmR,
                  int nk1, int nk2, int m_tiles, int n_tiles, int k2_group_n, int k2_cols,
                  EpiArgs ep) {
  using C = Cfg2<BN, RES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* stage_out = smem + C::kStages * C::kStageBytes;   // 1024-aligned: stage sizes are k*1024
  uint64_t* full = reinterpret_cast<uint64_t*>(stage_out + C::kOutBytes);
  uint64_t* empty = full + C::kStages;
  uint64_t* acc_full = empty + C::kStages;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* res_bar = acc_empty + 2;                       // [kEpiWarps][kResBufs] (RES kernels)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + kEpiWarps * kResBufs);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int nk = nk1 + nk2;
  const int total_tiles = m_tiles * n_tiles;          // m_tiles counts 256-row tiles
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    if (nk2 > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 2 * kEpiWarps);
    }
    if (RES)
      for (int i = 0; i < kEpiWarps * kResBufs; ++i) mbar_init(&res_bar[i], 1);
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
      for (int tile = cluster_id; tile < total_tiles; tile += n_clusters) {
        const int m0 = (tile / n_tiles) * (2 * BM) + rank * BM;
        const bool geglu = ep.act == 2;
        const int n0 = (tile % n_tiles) * (geglu ? BN / 2 : BN);
        // this CTA's half of the B tile (GEGLU: leader = hidden rows, peer = gate rows N + n0..)
        const int nb = geglu ? n0 + rank * ep.N : n0 + rank * C::kBRows;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + BM * BK * 2;
          if (ep.debug & 8) {
            if (leader) mbar_arrive(&full[stage]);
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
            continue;
          }
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * C::kStageBytes);
          if (kb < nk1) {
            tma_load_2d_2sm(sa, &tmA1, &full[stage], kb * BK, m0);
            tma_load_2d_2sm(sb, &tmB1, &full[stage], kb * BK, nb);
          } else {
            const int a2_off = k2_group_n > 0 ? (n0 / k2_group_n) * k2_cols : 0;
            tma_load_2d_2sm(sa, &tmA2, &full[stage], a2_off + (kb - nk1) * BK, m0);
            tma_load_2d_2sm(sb, &tmB2, &full[stage], (kb - nk1) * BK, nb);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader CTA only) =================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += n_clusters) {
        mbar_wait(&acc_empty[acc], acc_phase ^ 1, 2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * C::kAccStride;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&full[stage], phase, 3);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t sb = sa + BM * BK * 2;
          const uint64_t da = umma_desc_sw128(sa, 1024);
          const uint64_t db = umma_desc_sw128(sb, 1024);
          if (!(ep.debug & 4)) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_bf16_ss_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit_2sm(&empty[stage]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&acc_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ================= epilogue (warps 2..9 of both CTAs; own 128 rows) =================
    const int ewarp = warp - 2;
    if constexpr (RES) {
      epilogue_res_loop<BN>(
          ep, &tmD, &tmR, tmem_base, C::kAccStride, acc_full, stage_out + ewarp * kResBufs * kStageOutBytes,
          res_bar + ewarp * kResBufs, ewarp, lane, cluster_id, n_clusters, total_tiles,
          [&](int tile, int& m0, int& n0) { m0 = (tile / n_tiles) * (2 * BM) + rank * BM; n0 = (tile % n_tiles) * BN; },
          [&](int acc) { if (lane == 0) mbar_arrive_cluster(&acc_empty[acc], 0); });
    } else {
      uint8_t* my_out = stage_out + ewarp * kOutBufs * kStageOutBytes;
      int buf = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      if (lane == 0) tma_prefetch_desc(&tmD);
      for (int tile = cluster_id; tile < total_tiles; tile += n_clusters) {
        const int m0 = (tile / n_tiles) * (2 * BM) + rank * BM;
        uint64_t* done = &acc_empty[acc];
        auto release = [&]() { if (lane == 0) mbar_arrive_cluster(done, 0); };
        if (ep.act == 2) {
          const int n0 = (tile % n_tiles) * (BN / 2);
          mbar_wait(&acc_full[acc], acc_phase, 4);
          tc_fence_after();
          epilogue_tile_geglu<BN>(ep, &tmD, tmem_base + acc * C::kAccStride, m0, n0, ewarp, lane, my_out, buf,
                                  release);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
          continue;
        }
        const int n0 = (tile % n_tiles) * BN;
        EpiBias<BN> bias;
        epilogue_bias<BN>(ep, n0, ewarp, lane, bias);
        mbar_wait(&acc_full[acc], acc_phase, 4);
        tc_fence_after();
        epilogue_tile<BN>(ep, &tmD, tmem_base + acc * C::kAccStride, m0, n0, ewarp, lane, my_out, buf, bias, release);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (lane == 0) tma_store_wait<0>();
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, C::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// Single-launch fused LoRA GEMM (CTA pairs):  D = act(x W^T + (x A^T) B'^T + bias) * out_scale  with the
// down-projection u = x A^T computed INSIDE this kernel, once per (256-row tile, projection group):
//
//   unit (m, g)   U item      u[256 x K2] = x[256 x K1] A_g^T     (leading MMA, N = K2, into an accumulator stage)
//                             epilogue warps: TMEM -> bf16 -> THIS CTA's shared memory, laid out as the K-major,
//                             128B-swizzled A operand of K2/64 extra k-blocks (and, when the backward needs it,
//                             sent to u_out by TMA store from that very tile)
//                 TILE items  n-tiles of group g:  K1/64 k-blocks from the TMA ring, then the K2 k-blocks whose A
//                             operand is the resident u tile and whose B operand (B' rows of the tile) comes
//                             through the ring
//
// so the low-rank path costs its own FLOPs (2 M K1 K2 + 2 M K2 N) and nothing else: no second launch, no u
// round trip through HBM.  Reference arithmetic: unziplora_linear_layer.py:221-263,298-346 + lora_linear.py
// :74-81 (single stream), temporal_lora.py:29-32; grouped along N for q|k|v.  Tiles are dealt to the CTA pairs
// in CONTIGUOUS ranges of the (m, n) order (n fastest), so a pair recomputes u only when it enters a new (m, g)
// unit.  Barriers on top of lora_gemm2_kernel's: u_ready (leader CTA, 16 warp arrivals: both CTAs' u tiles
// are in shared memory and visible to the async proxy), u_free (both CTAs, multicast commit: the low-rank MMAs
// of the unit have read the tile).
// ------------------------------------------------------------------------------------------------
struct FusedSched {
  int t, t_end, n_tiles, tiles_per_group, prev_unit;
  __device__ FusedSched(int cluster, int n_clusters, int total, int n_tiles_, int tpg)
      : n_tiles(n_tiles_), tiles_per_group(tpg), prev_unit(-1) {
    t = (int)((int64_t)total * cluster / n_clusters);
    t_end = (int)((int64_t)total * (cluster + 1) / n_clusters);
  }
  __device__ bool done() const { return t >= t_end; }
  __device__ int unit_of(int tt) const { return (tt / n_tiles) * (n_tiles / tiles_per_group) + (tt % n_tiles) / tiles_per_group; }
  __device__ bool opens_unit() const { return unit_of(t) != prev_unit; }      // a U item precedes this tile
  __device__ bool closes_unit() const { return t + 1 >= t_end || unit_of(t + 1) != unit_of(t); }
  __device__ void advance() { prev_unit = unit_of(t); ++t; }
};

template <int BN>
struct CfgF {
  using C = Cfg2<BN, false>;
  static constexpr int kUBytes = 2 * BM * BK * 2;       // resident u tile: two k-blocks of [128 x 64] bf16
  static constexpr int kStages = (ring_budget(false) - kUBytes) / C::kStageBytes > 8
                                     ? 8 : (ring_budget(false) - kUBytes) / C::kStageBytes;
  static constexpr int kSmemBytes = kStages * C::kStageBytes + kUBytes + C::kOutBytes + 1024 + kBarBytes;
};

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
lora_gemm2_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                        const __grid_constant__ CUtensorMap tmAd, const __grid_constant__ CUtensorMap tmBp,
                        const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmU,
                        int nk1, int K2, int m_tiles, int n_tiles, int k2_group_n, int store_u, EpiArgs ep) {
  using C = Cfg2<BN, false>;
  constexpr int kUBytes = CfgF<BN>::kUBytes;
  constexpr int kStagesF = CfgF<BN>::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* u_smem = smem + kStagesF * C::kStageBytes;       // 1024-aligned (stage sizes are k * 1024)
  uint8_t* stage_out = u_smem + kUBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(stage_out + C::kOutBytes);
  uint64_t* empty = full + kStagesF;
  uint64_t* acc_full = empty + kStagesF;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* u_ready = acc_empty + 2;
  uint64_t* u_free = u_ready + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(u_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int total_tiles = m_tiles * n_tiles;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int tpg = k2_group_n > 0 ? k2_group_n / BN : n_tiles;          // n-tiles per projection group
  const int nk2 = K2 / BK;
  const int ubrows = K2 / 2;                                           // A_g rows each CTA stages for the U item

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmAd);
    tma_prefetch_desc(&tmBp);
    for (int i = 0; i < kStagesF; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 2 * kEpiWarps);
    }
    mbar_init(u_ready, 2 * kEpiWarps);
    mbar_init(u_free, 1);
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
      auto next = [&]() { if (++stage == kStagesF) { stage = 0; phase ^= 1; } };
      for (FusedSched sc(cluster_id, n_clusters, total_tiles, n_tiles, tpg); !sc.done(); sc.advance()) {
        const int m0 = (sc.t / n_tiles) * (2 * BM) + rank * BM;
        const int n0 = (sc.t % n_tiles) * BN;
        const int g = (sc.t % n_tiles) / tpg;
        if (sc.opens_unit()) {                           // U item: x k-blocks + this CTA's half of A_g
          const uint32_t bytes = BM * BK * 2 + ubrows * BK * 2;
          for (int kb = 0; kb < nk1; ++kb) {
            mbar_wait(&empty[stage], phase ^ 1, 1);
            uint8_t* sa = smem + stage * C::kStageBytes;
            if (leader) mbar_arrive_expect_tx(&full[stage], 2 * bytes);
            tma_load_2d_2sm(sa, &tmX, &full[stage], kb * BK, m0);
            tma_load_2d_2sm(sa + BM * BK * 2, &tmAd, &full[stage], kb * BK, g * K2 + rank * ubrows);
            next();
          }
        }
        const int nb = n0 + rank * C::kBRows;
        for (int kb = 0; kb < nk1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * C::kStageBytes);
          tma_load_2d_2sm(sa, &tmX, &full[stage], kb * BK, m0);
          tma_load_2d_2sm(sa + BM * BK * 2, &tmW, &full[stage], kb * BK, nb);
          next();
        }
        for (int kb = 0; kb < nk2; ++kb) {               // low-rank k-blocks: only B' travels, u is resident
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * C::kBRows * BK * 2);
          tma_load_2d_2sm(sa + BM * BK * 2, &tmBp, &full[stage], kb * BK, nb);
          next();
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader CTA only) =================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN);
      const uint32_t idesc_u = umma_idesc_bf16(2 * BM, K2);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0, unit_phase = 0;
      auto next = [&]() { if (++stage == kStagesF) { stage = 0; phase ^= 1; } };
      auto next_acc = [&]() { if (++acc == 2) { acc = 0; acc_phase ^= 1; } };
      for (FusedSched sc(cluster_id, n_clusters, total_tiles, n_tiles, tpg); !sc.done(); sc.advance()) {
        if (sc.opens_unit()) {
          mbar_wait(&acc_empty[acc], acc_phase ^ 1, 2);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * C::kAccStride;
          for (int kb = 0; kb < nk1; ++kb) {
            mbar_wait(&full[stage], phase, 3);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
            const uint64_t da = umma_desc_sw128(sa, 1024);
            const uint64_t db = umma_desc_sw128(sa + BM * BK * 2, 1024);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_bf16_ss_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc_u, (kb | k) != 0);
            umma_commit_2sm(&empty[stage]);
            next();
          }
          umma_commit_2sm(&acc_full[acc]);
          next_acc();
        }
        mbar_wait(&acc_empty[acc], acc_phase ^ 1, 2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * C::kAccStride;
        for (int kb = 0; kb < nk1; ++kb) {
          mbar_wait(&full[stage], phase, 3);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint64_t da = umma_desc_sw128(sa, 1024);
          const uint64_t db = umma_desc_sw128(sa + BM * BK * 2, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16_ss_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          umma_commit_2sm(&empty[stage]);
          next();
        }
        mbar_wait_cluster(u_ready, unit_phase, 6);       // both CTAs' u tiles of this unit are in place
        tc_fence_after();
        for (int kb = 0; kb < nk2; ++kb) {
          mbar_wait(&full[stage], phase, 3);
          tc_fence_after();
          const uint64_t da = umma_desc_sw128(smem_u32(u_smem + kb * BM * BK * 2), 1024);
          const uint64_t db = umma_desc_sw128(smem_u32(smem + stage * C::kStageBytes + BM * BK * 2), 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) umma_bf16_ss_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc, 1);
          umma_commit_2sm(&empty[stage]);
          next();
        }
        umma_commit_2sm(&acc_full[acc]);
        next_acc();
        if (sc.closes_unit()) {
          umma_commit_2sm(u_free);                       // the unit's low-rank MMAs have read the u tiles
          unit_phase ^= 1;
        }
      }
    }
  } else {
    // ================= epilogue (warps 2..9 of both CTAs; own 128 rows) =================
    const int ewarp = warp - 2;
    const int quarter = (ewarp + 2) & 3, half = ewarp >> 2;
    uint8_t* my_out = stage_out + ewarp * kOutBufs * kStageOutBytes;
    int buf = 0;
    int acc = 0;
    uint32_t acc_phase = 0, unit_phase = 0;
    if (lane == 0) tma_prefetch_desc(&tmD);
    auto next_acc = [&]() { if (++acc == 2) { acc = 0; acc_phase ^= 1; } };
    for (FusedSched sc(cluster_id, n_clusters, total_tiles, n_tiles, tpg); !sc.done(); sc.advance()) {
      const int m0 = (sc.t / n_tiles) * (2 * BM) + rank * BM;
      if (sc.opens_unit()) {
        // ---- U item: this CTA's 128 rows of u -> bf16 -> shared memory (K-major, 128B swizzle) ----
        const int g = (sc.t % n_tiles) / tpg;
        if (ewarp == 0 && lane == 0) tma_store_wait_read<0>();        // the previous unit's u_out store has left
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
        mbar_wait(u_free, unit_phase ^ 1, 7);                          // ... and its MMAs have read the tile
        mbar_wait(&acc_full[acc], acc_phase, 4);
        tc_fence_after();
        const uint32_t taddr = tmem_base + acc * C::kAccStride + (uint32_t(quarter * 32) << 16);
        const int row = quarter * 32 + lane;
        for (int c0 = half * 32; c0 < K2; c0 += 64) {
          uint32_t r[32];
          tmem_ld32(taddr + c0, r);
          tmem_ld_wait();
          uint8_t* dst = u_smem + (c0 >> 6) * (BM * BK * 2) + row * 128;
          const int cbase = (c0 & 32) ? 4 : 0;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 w;
            w.x = pack_bf16(__uint_as_float(r[8 * q]), __uint_as_float(r[8 * q + 1]));
            w.y = pack_bf16(__uint_as_float(r[8 * q + 2]), __uint_as_float(r[8 * q + 3]));
            w.z = pack_bf16(__uint_as_float(r[8 * q + 4]), __uint_as_float(r[8 * q + 5]));
            w.w = pack_bf16(__uint_as_float(r[8 * q + 6]), __uint_as_float(r[8 * q + 7]));
            *reinterpret_cast<uint4*>(dst + (((cbase + q) ^ (row & 7)) * 16)) = w;
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&acc_empty[acc], 0);
        next_acc();
        fence_proxy_async_smem();                                      // u tile -> visible to tcgen05 / TMA
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
        if (ewarp == 0 && lane == 0 && store_u) {
          for (int kb = 0; kb < nk2; ++kb)
            tma_store_2d(&tmU, u_smem + kb * BM * BK * 2, g * K2 + kb * BK, m0);
          tma_store_commit();
        }
        if (lane == 0) {                                                // release: the peer SM's tensor core reads it
          uint32_t remote;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(u_ready)), "r"(0));
          asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
        }
        unit_phase ^= 1;
      }
      const int n0 = (sc.t % n_tiles) * BN;
      EpiBias<BN> bias;
      epilogue_bias<BN>(ep, n0, ewarp, lane, bias);
      mbar_wait(&acc_full[acc], acc_phase, 4);
      tc_fence_after();
      uint64_t* done = &acc_empty[acc];
      epilogue_tile<BN>(ep, &tmD, tmem_base + acc * C::kAccStride, m0, n0, ewarp, lane, my_out, buf, bias,
                        [&]() { if (lane == 0) mbar_arrive_cluster(done, 0); });
      next_acc();
    }
    if (lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, C::kTmemCols);
  }
}

int make_operand_map(CUtensorMap* m, const void* p, int64_t rows, int K, int64_t ld, int box_rows) {
  const uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  const uint64_t strides[2] = {2, (uint64_t)ld * 2};
  const uint32_t box[2] = {BK, (uint32_t)box_rows};
  return vst_make_tmap(m, p, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_row_map(CUtensorMap* m, const void* p, int64_t ld, int64_t M, int N);

int make_out_map(CUtensorMap* m, const EpiArgs& ep) {
  if (ep.out_f32) {      // direct-store path: the map is never used
    memset(m, 0, sizeof(*m));
    return VST_OK;
  }
  return make_row_map(m, ep.D, ep.ldd, ep.M, ep.N);
}

int make_row_map(CUtensorMap* m, const void* p, int64_t ld, int64_t M, int N) {
  const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
  const uint64_t strides[2] = {2, (uint64_t)ld * 2};
  const uint32_t box[2] = {32, 32};
  return vst_make_tmap(m, p, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
}

template <int BN, bool RES>
int launch_t(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1, const void* A2,
             int64_t lda2, const void* B2, int64_t ldb2, int K2, int k2_group_n, const EpiArgs& ep,
             cudaStream_t st) {
  using C = Cfg<BN, RES>;
  CUtensorMap tA1, tB1, tA2, tB2;
  int rc;
  if ((rc = make_operand_map(&tA1, A1, ep.M, K1, lda1, BM))) return rc;
  if ((rc = make_operand_map(&tB1, B1, ep.N, K1, ldb1, BN))) return rc;
  if (K2 > 0) {
    const int groups = k2_group_n > 0 ? (ep.N + k2_group_n - 1) / k2_group_n : 1;
    if ((rc = make_operand_map(&tA2, A2, ep.M, K2 * groups, lda2, BM))) return rc;
    if ((rc = make_operand_map(&tB2, B2, ep.N, K2, ldb2, BN))) return rc;
  } else {
    tA2 = tA1;
    tB2 = tB1;
  }
  CUtensorMap tD, tR;
  if ((rc = make_out_map(&tD, ep))) return rc;
  tR = tD;
  if (RES && (rc = make_row_map(&tR, ep.residual, ep.ldr, ep.M, ep.N))) return rc;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(lora_gemm_kernel<BN, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     C::kSmemBytes));
    configured = true;
  }
  const int m_tiles = (int)((ep.M + BM - 1) / BM), n_tiles = (ep.N + BN - 1) / BN;
  const int64_t tiles = (int64_t)m_tiles * n_tiles;
  VST_REQUIRE(tiles < (1ll << 31), VST_E_UNSUPPORTED, "lora_gemm: too many tiles");
  const int grid = (int)(tiles < vst_sm_count() ? tiles : vst_sm_count());
  lora_gemm_kernel<BN, RES><<<grid, kThreads, C::kSmemBytes, st>>>(tA1, tB1, tA2, tB2, tD, tR, (K1 + BK - 1) / BK,
                                                                   (K2 + BK - 1) / BK, m_tiles, n_tiles,
                                                                   k2_group_n, K2, ep);
  VST_LAUNCH_OK();
  return VST_OK;
}

template <int BN>
int launch(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1, const void* A2,
           int64_t lda2, const void* B2, int64_t ldb2, int K2, int k2_group_n, const EpiArgs& ep,
           cudaStream_t st) {
  const bool res = ep.residual != nullptr && !ep.out_f32;    // bf16 + residual: TMA-fed residual epilogue
  return res ? launch_t<BN, true>(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, k2_group_n, ep, st)
             : launch_t<BN, false>(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, k2_group_n, ep, st);
}

template <int BN, bool RES>
int launch2_t(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1, const void* A2,
              int64_t lda2, const void* B2, int64_t ldb2, int K2, int k2_group_n, const EpiArgs& ep,
              cudaStream_t st) {
  using C = Cfg2<BN, RES>;
  CUtensorMap tA1, tB1, tA2, tB2;
  int rc;
  const bool geglu = ep.act == 2;     // B1 = [W_hidden ; W_gate]: 2N rows
  if ((rc = make_operand_map(&tA1, A1, ep.M, K1, lda1, BM))) return rc;
  if ((rc = make_operand_map(&tB1, B1, geglu ? 2 * ep.N : ep.N, K1, ldb1, C::kBRows))) return rc;
  if (K2 > 0) {
    const int groups = k2_group_n > 0 ? (ep.N + k2_group_n - 1) / k2_group_n : 1;
    if ((rc = make_operand_map(&tA2, A2, ep.M, K2 * groups, lda2, BM))) return rc;
    if ((rc = make_operand_map(&tB2, B2, ep.N, K2, ldb2, C::kBRows))) return rc;
  } else {
    tA2 = tA1;
    tB2 = tB1;
  }
  CUtensorMap tD, tR;
  if ((rc = make_out_map(&tD, ep))) return rc;
  tR = tD;
  if (RES && (rc = make_row_map(&tR, ep.residual, ep.ldr, ep.M, ep.N))) return rc;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(lora_gemm2_kernel<BN, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     C::kSmemBytes));
    configured = true;
  }
  // N need not be a multiple of BN: TMA zero-fills the B rows past N and clips the stores
  const int m_tiles = (int)((ep.M + 2 * BM - 1) / (2 * BM)), n_tiles = geglu ? ep.N / (BN / 2) : (ep.N + BN - 1) / BN;
  const int64_t tiles = (int64_t)m_tiles * n_tiles;
  VST_REQUIRE(tiles < (1ll << 31), VST_E_UNSUPPORTED, "lora_gemm: too many tiles");
  const int pairs = vst_sm_count() / 2;
  const int grid = 2 * (int)(tiles < pairs ? tiles : pairs);
  lora_gemm2_kernel<BN, RES><<<grid, kThreads, C::kSmemBytes, st>>>(tA1, tB1, tA2, tB2, tD, tR, (K1 + BK - 1) / BK,
                                                                    (K2 + BK - 1) / BK, m_tiles, n_tiles,
                                                                    k2_group_n, K2, ep);
  VST_LAUNCH_OK();
  return VST_OK;
}

template <int BN>
int launch2(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1, const void* A2,
            int64_t lda2, const void* B2, int64_t ldb2, int K2, int k2_group_n, const EpiArgs& ep,
            cudaStream_t st) {
  const bool res = ep.residual != nullptr && !ep.out_f32;
  return res ? launch2_t<BN, true>(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, k2_group_n, ep, st)
             : launch2_t<BN, false>(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, k2_group_n, ep, st);
}

template <int BN>
int launch2_fused(const void* X, int64_t ldx, const void* W, int64_t ldw, int K1, const void* Ad, int64_t lda,
                  const void* Bp, int64_t ldbp, int K2, int k2_group_n, void* U, int64_t ldu, const EpiArgs& ep,
                  cudaStream_t st) {
  using F = CfgF<BN>;
  using C = typename F::C;
  const int groups = k2_group_n > 0 ? ep.N / k2_group_n : 1;
  CUtensorMap tX, tW, tAd, tBp, tD, tU;
  int rc;
  if ((rc = make_operand_map(&tX, X, ep.M, K1, ldx, BM))) return rc;
  if ((rc = make_operand_map(&tW, W, ep.N, K1, ldw, C::kBRows))) return rc;
  if ((rc = make_operand_map(&tAd, Ad, (int64_t)groups * K2, K1, lda, K2 / 2))) return rc;
  if ((rc = make_operand_map(&tBp, Bp, ep.N, K2, ldbp, C::kBRows))) return rc;
  if ((rc = make_out_map(&tD, ep))) return rc;
  tU = tD;
  if (U != nullptr && (rc = make_operand_map(&tU, U, ep.M, groups * K2, ldu, BM))) return rc;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(lora_gemm2_fused_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     F::kSmemBytes));
    configured = true;
  }
  const int m_tiles = (int)((ep.M + 2 * BM - 1) / (2 * BM)), n_tiles = (ep.N + BN - 1) / BN;
  const int64_t tiles = (int64_t)m_tiles * n_tiles;
  VST_REQUIRE(tiles < (1ll << 31), VST_E_UNSUPPORTED, "lora_gemm_fused: too many tiles");
  const int pairs = vst_sm_count() / 2;
  const int grid = 2 * (int)(tiles < pairs ? tiles : pairs);
  lora_gemm2_fused_kernel<BN><<<grid, kThreads, F::kSmemBytes, st>>>(tX, tW, tAd, tBp, tD, tU, (K1 + BK - 1) / BK, K2,
                                                                     m_tiles, n_tiles, k2_group_n, U != nullptr, ep);
  VST_LAUNCH_OK();
  return VST_OK;
}

// ---- vst_lora_pack_up: B2'[n, br*r + j] = scale * m_br[n] * mask_br[n] * B_br[n, j] ------------
template <typename T>
__global__ void pack_up_kernel(const T* __restrict__ Bc, const T* __restrict__ mc,
                               const uint8_t* __restrict__ maskc, const T* __restrict__ Bs,
                               const T* __restrict__ ms, const uint8_t* __restrict__ masks,
                               float scale, __nv_bfloat16* __restrict__ out, int N, int r,
                               int nbr) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int width = nbr * r;
  if (idx >= (int64_t)N * width) return;
  const int n = (int)(idx / width), j = (int)(idx % width);
  const bool second = (Bc != nullptr) && j >= r;
  const T* Bm = (Bc == nullptr || second) ? Bs : Bc;
  const T* m = (Bc == nullptr || second) ? ms : mc;
  const uint8_t* mk = (Bc == nullptr || second) ? masks : maskc;
  const int jj = second ? j - r : j;
  float val = (float)Bm[(int64_t)n * r + jj] * scale;
  if (m != nullptr) val *= (float)m[n];
  if (mk != nullptr && mk[n] == 0) val = 0.f;
  out[idx] = __float2bfloat16_rn(val);
}

// ---- vst_lora_pack_down: A'[br * rp + j, :] = A_br[j, :] (bf16), rows j >= r of each branch zero ----------
template <typename T>
__global__ void pack_down_kernel(const T* __restrict__ Ac, const T* __restrict__ As, __nv_bfloat16* __restrict__ out,
                                 int r, int rp, int K, int nbr) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)nbr * rp * K) return;
  const int row = (int)(idx / K), k = (int)(idx % K);
  const int br = row / rp, j = row % rp;
  const T* src = (Ac != nullptr && br == 0) ? Ac : As;
  out[idx] = __float2bfloat16_rn(j < r ? (float)src[(int64_t)j * K + k] : 0.f);
}

}  // namespace

// Tile selection of vst_lora_gemm (also exported as vst_lora_gemm_plan so a caller -- bench.py's
// per-kernel roofline -- can name the kernel instantiation a shape runs on).  Returns BN and sets
// *pair = 1 for the CTA-pair kernel (lora_gemm2_kernel, 256-row tiles); 0 = unsupported grouping.
static int vst_lora_gemm_select(int64_t M, int N, int gn, int act, int* pair) {
  *pair = 0;
#define FITS(BNV) (N % BNV == 0 && (gn == 0 || gn % BNV == 0))
  if (act == VST_ACT_GEGLU) { *pair = 1; return 256; }
  // CTA pairs (256-row tiles) once there is at least one tile per SM pair to keep busy
  const bool big = M >= 2048 && !vst_force_single_cta();
  static int force_bn = -1;
  if (force_bn < 0) { const char* e = getenv("VST_GEMM_BN"); force_bn = e ? atoi(e) : 0; }   // tuning only
  const int half_sms = vst_sm_count() / 2;
  if (big && force_bn == 160 && FITS(160)) { *pair = 1; return 160; }
  if (big && force_bn == 128 && FITS(128)) { *pair = 1; return 128; }
  if (force_bn == 256 && FITS(256)) return 256;
  // 256-wide pair tiles move the fewest operand bytes per MAC (64 B/clk/SM against 83 for 160):
  // worth up to 20 % of padded columns (N = 640, 960, 1920 -> 768, 1024, 2048); r01 launch list:
  // the BN = 160 family ran at 760 TFLOP/s, the BN = 256 family at 1260
  if (big && (gn == 0 || gn % 256 == 0) && N > 256 && (int64_t)((N + 255) / 256) * 256 * 5 <= (int64_t)N * 6 &&
      (M / 256) * ((N + 255) / 256) >= half_sms) { *pair = 1; return 256; }
  if (big && FITS(256) && (M / 256) * (N / 256) >= half_sms) { *pair = 1; return 256; }
  if (big && FITS(160) && (M / 256) * (N / 160) >= half_sms) { *pair = 1; return 160; }
  if (big && FITS(128) && (M / 256) * (N / 128) >= half_sms) { *pair = 1; return 128; }
  // one row tile (the text K / V projections: M = 77 per prompt, K = 2048): the run time is the
  // weight read, and one SM pulls only ~0.15 TB/s -- narrow tiles spread it over 4x the SMs
  if (M <= 128 && (force_bn == 0 || force_bn == 64) && FITS(64) && N / 64 >= 16) return 64;
  if (FITS(256)) return 256;
  if (FITS(160)) return 160;
  if (FITS(192)) return 192;
  if (FITS(128)) return 128;
  if (FITS(64)) return 64;
#undef FITS
  if (gn != 0) return 0;
  if (N <= 64) return 64;
  if (N <= 128) return 128;
  if (N <= 160) return 160;
  return 256;
}

extern "C" int vst_lora_gemm_plan(int64_t M, int N, int K2, int k2_group_n, int act) {
  int pair = 0;
  const int gn = (K2 > 0 && k2_group_n > 0) ? k2_group_n : 0;
  const int bn = vst_lora_gemm_select(M, N, gn, act, &pair);
  return bn <= 0 ? VST_E_UNSUPPORTED : pair * 1000 + bn;
}

extern "C" int vst_lora_gemm(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1,
                             const void* A2, int64_t lda2, const void* B2, int64_t ldb2, int K2,
                             int k2_group_n, const void* bias, int bias_dtype, const void* residual, int64_t ldr,
                             float out_scale, int act, void* D, int64_t ldd, int64_t M, int N,
                             int out_dtype, void* stream) {
  VST_REQUIRE(M >= 0 && N > 0 && K1 > 0 && K2 >= 0, VST_E_BADARG, "lora_gemm: bad shape");
  if (M == 0) return VST_OK;
  VST_REQUIRE(A1 && B1 && D && (K2 == 0 || (A2 && B2)), VST_E_BADARG, "lora_gemm: null pointer");
  VST_REQUIRE(K1 % 8 == 0 && K2 % 8 == 0 && N % 8 == 0, VST_E_UNSUPPORTED,
              "lora_gemm: K1=%d, K2=%d, N=%d must be multiples of 8", K1, K2, N);
  VST_REQUIRE(lda1 % 8 == 0 && ldb1 % 8 == 0 && (K2 == 0 || (lda2 % 8 == 0 && ldb2 % 8 == 0)),
              VST_E_UNSUPPORTED, "lora_gemm: operand row strides must be multiples of 8 elements");
  VST_REQUIRE(out_dtype == VST_F32 || out_dtype == VST_BF16, VST_E_BADARG, "lora_gemm: out dtype");
  VST_REQUIRE(act == VST_ACT_NONE || act == VST_ACT_GELU || act == VST_ACT_GEGLU, VST_E_BADARG,
              "lora_gemm: bad act %d", act);
  if (act == VST_ACT_GEGLU)
    VST_REQUIRE(K2 == 0 && residual == nullptr && out_dtype == VST_BF16 && N % 128 == 0, VST_E_UNSUPPORTED,
                "lora_gemm: VST_ACT_GEGLU needs N %% 128 == 0, bf16 output, no low-rank pair, no residual");
  const int oa = out_dtype == VST_F32 ? 4 : 8;
  VST_REQUIRE(ldd % oa == 0 && (residual == nullptr || ldr % oa == 0), VST_E_UNSUPPORTED,
              "lora_gemm: output/residual row strides must keep 16-byte alignment");
  VST_REQUIRE((reinterpret_cast<uintptr_t>(D) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(residual) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(bias) & 15) == 0,
              VST_E_UNSUPPORTED, "lora_gemm: D/residual/bias must be 16-byte aligned");
  static int debug = -1;
  if (debug < 0) { const char* e = getenv("VST_GEMM_DEBUG"); debug = e ? atoi(e) : 0; }
  EpiArgs ep{bias, residual, D, ldr, ldd, M, N, out_scale, bias_dtype == VST_BF16,
             out_dtype == VST_F32, act, debug};
  cudaStream_t st = (cudaStream_t)stream;
  VST_REQUIRE(k2_group_n >= 0 && (k2_group_n == 0 || K2 > 0), VST_E_BADARG, "lora_gemm: bad k2_group_n");
  const int gn = (K2 > 0 && k2_group_n > 0) ? k2_group_n : 0;
  int pair = 0;
  const int bn = vst_lora_gemm_select(M, N, gn, act, &pair);
  VST_REQUIRE(bn > 0, VST_E_UNSUPPORTED,
              "lora_gemm: k2_group_n=%d must be a multiple of 64 that tiles N=%d", gn, N);
#define GO(BNV) return launch<BNV>(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, gn, ep, st)
#define GO2(BNV) return launch2<BNV>(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, gn, ep, st)
  if (pair) {
    switch (bn) {
      case 256: GO2(256);
      case 160: GO2(160);
      default: GO2(128);
    }
  }
  switch (bn) {
    case 256: GO(256);
    case 192: GO(192);
    case 160: GO(160);
    case 128: GO(128);
    default: GO(64);
  }
#undef GO2
#undef GO
}

extern "C" int vst_lora_pack_up(const void* Bc, const void* mc, const uint8_t* maskc,
                                const void* Bs, const void* ms, const uint8_t* masks, float scale,
                                void* out, int N, int r, int param_dtype, void* stream) {
  VST_REQUIRE(N > 0 && r > 0 && out && (Bc || Bs), VST_E_BADARG, "lora_pack_up: bad argument");
  const int nbr = (Bc ? 1 : 0) + (Bs ? 1 : 0);
  const int64_t total = (int64_t)N * nbr * r;
  const int threads = 256;
  const unsigned grid = (unsigned)((total + threads - 1) / threads);
  cudaStream_t st = (cudaStream_t)stream;
  if (param_dtype == VST_BF16)
    pack_up_kernel<__nv_bfloat16><<<grid, threads, 0, st>>>(
        (const __nv_bfloat16*)Bc, (const __nv_bfloat16*)mc, maskc, (const __nv_bfloat16*)Bs,
        (const __nv_bfloat16*)ms, masks, scale, (__nv_bfloat16*)out, N, r, nbr);
  else if (param_dtype == VST_F32)
    pack_up_kernel<float><<<grid, threads, 0, st>>>((const float*)Bc, (const float*)mc, maskc,
                                                    (const float*)Bs, (const float*)ms, masks,
                                                    scale, (__nv_bfloat16*)out, N, r, nbr);
  else {
    vst_set_error("lora_pack_up: bad param dtype %d", param_dtype);
    return VST_E_BADARG;
  }
  VST_LAUNCH_OK();
  return VST_OK;
}


extern "C" int vst_lora_pack_down(const void* Ac, const void* As, void* out, int r, int rp, int K, int param_dtype,
                                  void* stream) {
  VST_REQUIRE(out && (Ac || As) && r > 0 && rp >= r && K > 0, VST_E_BADARG, "lora_pack_down: bad argument");
  const int nbr = (Ac ? 1 : 0) + (As ? 1 : 0);
  const int64_t total = (int64_t)nbr * rp * K;
  const unsigned grid = (unsigned)((total + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (param_dtype == VST_BF16)
    pack_down_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)Ac, (const __nv_bfloat16*)As,
                                                          (__nv_bfloat16*)out, r, rp, K, nbr);
  else if (param_dtype == VST_F32)
    pack_down_kernel<float><<<grid, 256, 0, st>>>((const float*)Ac, (const float*)As, (__nv_bfloat16*)out, r, rp, K, nbr);
  else {
    vst_set_error("lora_pack_down: bad param dtype %d", param_dtype);
    return VST_E_BADARG;
  }
  VST_LAUNCH_OK();
  return VST_OK;
}

// != 0 (the tile width) when vst_lora_gemm_fused takes the shape: CTA-pair kernel, K2 in {64, 128}, tiles that do
// not straddle projection groups; else 0.
extern "C" int vst_lora_gemm_fused_supported(int64_t M, int N, int K1, int K2, int k2_group_n) {
  if (M < 2048 || N <= 0 || K1 <= 0 || K1 % 8 || N % 8) return 0;
  if (K2 != 64 && K2 != 128) return 0;
  const int gn = k2_group_n;
  if (gn < 0 || (gn > 0 && (N % gn != 0))) return 0;
  const int unit = gn > 0 ? gn : N;
  if (unit % 256 == 0 || (gn == 0 && N > 256)) return 256;
  if (unit % 160 == 0) return 160;
  if (unit % 128 == 0) return 128;
  return 0;
}

extern "C" int vst_lora_gemm_fused(const void* X, int64_t ldx, const void* W, int64_t ldw, int K1,
                                   const void* A_down, int64_t lda, const void* Bp, int64_t ldbp, int K2,
                                   int k2_group_n, const void* bias, int bias_dtype, float out_scale, int act,
                                   void* U_out, int64_t ldu, void* D, int64_t ldd, int64_t M, int N,
                                   void* stream) {
  VST_REQUIRE(X && W && A_down && Bp && D, VST_E_BADARG, "lora_gemm_fused: null pointer");
  const int bn = vst_lora_gemm_fused_supported(M, N, K1, K2, k2_group_n);
  VST_REQUIRE(bn > 0, VST_E_UNSUPPORTED,
              "lora_gemm_fused: M=%lld N=%d K1=%d K2=%d group=%d is not a fused shape (use vst_lora_gemm)",
              (long long)M, N, K1, K2, k2_group_n);
  VST_REQUIRE(act == VST_ACT_NONE || act == VST_ACT_GELU, VST_E_UNSUPPORTED, "lora_gemm_fused: act none / gelu");
  VST_REQUIRE(ldx % 8 == 0 && ldw % 8 == 0 && lda % 8 == 0 && ldbp % 8 == 0 && ldd % 8 == 0 &&
                  (U_out == nullptr || ldu % 8 == 0),
              VST_E_UNSUPPORTED, "lora_gemm_fused: row strides must be multiples of 8 elements");
  VST_REQUIRE((reinterpret_cast<uintptr_t>(D) & 15) == 0 && (reinterpret_cast<uintptr_t>(bias) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(U_out) & 15) == 0,
              VST_E_UNSUPPORTED, "lora_gemm_fused: D / bias / U_out must be 16-byte aligned");
  EpiArgs ep{bias, nullptr, D, 0, ldd, M, N, out_scale, bias_dtype == VST_BF16, 0, act, 0};
  cudaStream_t st = (cudaStream_t)stream;
  switch (bn) {
    case 256: return launch2_fused<256>(X, ldx, W, ldw, K1, A_down, lda, Bp, ldbp, K2, k2_group_n, U_out, ldu, ep, st);
    case 160: return launch2_fused<160>(X, ldx, W, ldw, K1, A_down, lda, Bp, ldbp, K2, k2_group_n, U_out, ldu, ep, st);
    default: return launch2_fused<128>(X, ldx, W, ldw, K1, A_down, lda, Bp, ldbp, K2, k2_group_n, U_out, ldu, ep, st);
  }
}
