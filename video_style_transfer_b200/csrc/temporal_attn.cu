// a8: frame-axis temporal self-attention, softmax(q k^T / sqrt(d)) v over F <= 32 frames for every
// (spatial position, head).  Reference: animatediff/temporal_transformer.py:45-50,69-71 and the
// diffusers motion-module attn1/attn2 (un-vendored).
//
// Roofline: 4*F^2*d FLOP against 8*F*d bytes per (sequence, head) = F/2 FLOP/B (8 at F=16), far
// below the B200 ridge (~210 FLOP/B) => HBM-bound as a standalone kernel.  One warp owns one
// (sequence, head): 16-byte loads of its Q/K/V slices into shared memory, the two tiny
// contractions on warp-level tensor-core MMAs (m16n8k16; a 128-row tcgen05 tile would be 7/8
// masked padding here and the op is bandwidth-bound either way), softmax in registers with
// quad shuffles, and 16-byte coalesced stores of O.
#include <stdlib.h>

#define VST_MBAR_QUIET 1   // watchdog traps without printf: see common.cuh (measured per kernel family)
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

// DPAD: head dim rounded up to a multiple of 16; FP: frames rounded up to 16 or 32.
template <int DPAD, int FP, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
temporal_attn_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                     const __nv_bfloat16* __restrict__ v, __nv_bfloat16* __restrict__ o,
                     int64_t pairs, int F, int heads, int d, int64_t ldq, int64_t ldk,
                     int64_t ldv, int64_t ldo, float scale_log2) {
  constexpr int LD = DPAD + 8;           // smem row stride (elements): 16-B multiple, odd in 16-B units
  constexpr int MT = FP / 16;            // query m-tiles
  constexpr int NT = FP / 8;             // key n-tiles of S
  constexpr int DK = DPAD / 16;          // k-steps of S = Q K^T
  constexpr int DN = DPAD / 8;           // n-tiles of O = P V
  constexpr int CH = DPAD / 8;           // 16-byte chunks per row
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t pair = (int64_t)blockIdx.x * WARPS + warp;
  if (pair >= pairs) return;
  const int64_t seq = pair / heads;
  const int head = (int)(pair % heads);
  __nv_bfloat16* sq = reinterpret_cast<__nv_bfloat16*>(smem_raw) + (size_t)warp * 3 * FP * LD;
  __nv_bfloat16* sk = sq + FP * LD;
  __nv_bfloat16* sv = sk + FP * LD;

  const __nv_bfloat16* gq = q + seq * F * ldq + (int64_t)head * d;
  const __nv_bfloat16* gk = k + seq * F * ldk + (int64_t)head * d;
  const __nv_bfloat16* gv = v + seq * F * ldv + (int64_t)head * d;
  for (int idx = lane; idx < FP * CH; idx += 32) {
    const int row = idx / CH, ch = idx % CH;
    uint4 a = make_uint4(0, 0, 0, 0), b = a, c = a;
    if (row < F && ch * 8 < d) {
      a = ld_stream16(gq + row * ldq + ch * 8);
      b = ld_stream16(gk + row * ldk + ch * 8);
      c = ld_stream16(gv + row * ldv + ch * 8);
    }
    *reinterpret_cast<uint4*>(sq + row * LD + ch * 8) = a;
    *reinterpret_cast<uint4*>(sk + row * LD + ch * 8) = b;
    *reinterpret_cast<uint4*>(sv + row * LD + ch * 8) = c;
  }
  __syncwarp();

  // ---- S = Q K^T -------------------------------------------------------------------------------
  float s[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) s[mt][nt][i] = 0.f;
  const uint32_t sq_a = smem_u32(sq), sk_a = smem_u32(sk), sv_a = smem_u32(sv);
#pragma unroll
  for (int kt = 0; kt < DK; ++kt) {
    uint32_t a[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
      ldsm_x4(sq_a + ((mt * 16 + (lane & 15)) * LD + kt * 16 + (lane >> 4) * 8) * 2, a[mt]);
#pragma unroll
    for (int np = 0; np < NT / 2; ++np) {
      // two key n-tiles per ldmatrix.x4: lanes 0-7 keys n0.., d k0 | 8-15 same keys, d k0+8 |
      // 16-23 keys n0+8.., d k0 | 24-31 keys n0+8.., d k0+8
      uint32_t b[4];
      ldsm_x4(sk_a + ((np * 16 + (lane & 7) + ((lane >> 4) << 3)) * LD + kt * 16 +
                      ((lane >> 3) & 1) * 8) * 2, b);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_bf16(s[mt][2 * np], a[mt], b[0], b[1]);
        mma_bf16(s[mt][2 * np + 1], a[mt], b[2], b[3]);
      }
    }
  }

  // ---- softmax over keys (rows g and g+8 of each m-tile; 4 lanes share a row) -------------------
  const int t4 = lane & 3;
  float inv[MT][2];
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
      inv[mt][h] = 1.f / sum;
    }
  }

  // ---- O = P V -----------------------------------------------------------------------------------
  uint32_t pa[MT][FP / 16][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int kt = 0; kt < FP / 16; ++kt) {
      pa[mt][kt][0] = pack_bf16(s[mt][2 * kt][0], s[mt][2 * kt][1]);
      pa[mt][kt][1] = pack_bf16(s[mt][2 * kt][2], s[mt][2 * kt][3]);
      pa[mt][kt][2] = pack_bf16(s[mt][2 * kt + 1][0], s[mt][2 * kt + 1][1]);
      pa[mt][kt][3] = pack_bf16(s[mt][2 * kt + 1][2], s[mt][2 * kt + 1][3]);
    }
  __syncwarp();  // everyone is done reading sq: it becomes the O staging buffer
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
    for (int kt = 0; kt < FP / 16; ++kt) {
      // lanes 0-7 keys k0.. , d n0 | 8-15 keys k0+8.., d n0 | 16-23 keys k0.., d n0+8 | 24-31 ...
      uint32_t b[4];
      ldsm_x4_t(sv_a + ((kt * 16 + (lane & 15)) * LD + np * 16 + (lane >> 4) * 8) * 2, b);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_bf16(acc[mt][0], pa[mt][kt], b[0], b[1]);
        mma_bf16(acc[mt][1], pa[mt][kt], b[2], b[3]);
      }
    }
    const int g = lane >> 2;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int col = np * 16 + j * 8 + 2 * t4;
        *reinterpret_cast<uint32_t*>(sq + (mt * 16 + g) * LD + col) =
            pack_bf16(acc[mt][j][0] * inv[mt][0], acc[mt][j][1] * inv[mt][0]);
        *reinterpret_cast<uint32_t*>(sq + (mt * 16 + g + 8) * LD + col) =
            pack_bf16(acc[mt][j][2] * inv[mt][1], acc[mt][j][3] * inv[mt][1]);
      }
  }
  __syncwarp();
  __nv_bfloat16* go = o + seq * F * ldo + (int64_t)head * d;
  for (int idx = lane; idx < FP * CH; idx += 32) {
    const int row = idx / CH, ch = idx % CH;
    if (row < F && ch * 8 < d)
      st_stream16(go + row * ldo + ch * 8, *reinterpret_cast<const uint4*>(sq + row * LD + ch * 8));
  }
}

// ---- TMA-fed persistent form (the one that ships) ----------------------------------------------
// Same math, but a warp never issues a global load or store itself: per (sequence, head) one lane
// posts three bulk tensor loads (box = [F frames x DPAD columns] of a 3-D (d, head, row) view of
// q / k / v; the columns past d are zero-filled by TMA, so head dims that are not multiples of 16
// need no padding pass) into a two-stage ring of its own, and the O tile leaves by a bulk tensor
// store.  The next pair's tiles are in flight while this one is computed, and the ~170 LDG / STS /
// index instructions per pair of the version above are gone (r01: 2.75 -> see profiles/ TB/s).
template <int DPAD, int FP, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
temporal_attn_tma_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                         const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                         int64_t pairs, int F, int heads, float scale_log2) {
  constexpr int LD = DPAD;               // dense TMA rows
  constexpr int MT = FP / 16, NT = FP / 8, DK = DPAD / 16, DN = DPAD / 8;
  constexpr int kTileElems = FP * LD;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full[WARPS][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __nv_bfloat16* base = reinterpret_cast<__nv_bfloat16*>(smem_raw) + (size_t)warp * 7 * kTileElems;
  __nv_bfloat16* so = base + 6 * kTileElems;        // O staging; stages: [q k v] x 2 before it

  // rows >= F of every tile stay zero for the whole kernel (TMA boxes have F rows)
  for (int i = lane; i < 7 * kTileElems / 8; i += 32) reinterpret_cast<uint4*>(base)[i] = make_uint4(0, 0, 0, 0);
  if (lane == 0) {
    mbar_init(&full[warp][0], 1);
    mbar_init(&full[warp][1], 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  __syncwarp();

  const int64_t stride = (int64_t)gridDim.x * WARPS;
  const int64_t first = (int64_t)blockIdx.x * WARPS + warp;
  const uint32_t tile_bytes = (uint32_t)F * LD * 2;
  auto issue = [&](int64_t pair, int st) {
    const int row0 = (int)(pair / heads) * F, head = (int)(pair % heads);
    __nv_bfloat16* sq = base + st * 3 * kTileElems;
    mbar_arrive_expect_tx(&full[warp][st], 3 * tile_bytes);
    tma_load_3d(sq, &tmQ, &full[warp][st], 0, head, row0);
    tma_load_3d(sq + kTileElems, &tmK, &full[warp][st], 0, head, row0);
    tma_load_3d(sq + 2 * kTileElems, &tmV, &full[warp][st], 0, head, row0);
  };
  if (first < pairs && lane == 0) issue(first, 0);

  int it = 0;
  for (int64_t pair = first; pair < pairs; pair += stride, ++it) {
    const int st = it & 1;
    if (pair + stride < pairs && lane == 0) issue(pair + stride, st ^ 1);   // its last readers are done
    mbar_wait(&full[warp][st], (it >> 1) & 1, 60);
    __nv_bfloat16* sq = base + st * 3 * kTileElems;
    const uint32_t sq_a = smem_u32(sq), sk_a = sq_a + kTileElems * 2, sv_a = sk_a + kTileElems * 2;

    float s[MT][NT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) s[mt][nt][i] = 0.f;
#pragma unroll
    for (int kt = 0; kt < DK; ++kt) {
      uint32_t a[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
        ldsm_x4(sq_a + ((mt * 16 + (lane & 15)) * LD + kt * 16 + (lane >> 4) * 8) * 2, a[mt]);
#pragma unroll
      for (int np = 0; np < NT / 2; ++np) {
        uint32_t b[4];
        ldsm_x4(sk_a + ((np * 16 + (lane & 7) + ((lane >> 4) << 3)) * LD + kt * 16 +
                        ((lane >> 3) & 1) * 8) * 2, b);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma_bf16(s[mt][2 * np], a[mt], b[0], b[1]);
          mma_bf16(s[mt][2 * np + 1], a[mt], b[2], b[3]);
        }
      }
    }
    const int t4 = lane & 3;
    float inv[MT][2];
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
        inv[mt][h] = 1.f / sum;
      }
    }
    uint32_t pa[MT][FP / 16][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int kt = 0; kt < FP / 16; ++kt) {
        pa[mt][kt][0] = pack_bf16(s[mt][2 * kt][0], s[mt][2 * kt][1]);
        pa[mt][kt][1] = pack_bf16(s[mt][2 * kt][2], s[mt][2 * kt][3]);
        pa[mt][kt][2] = pack_bf16(s[mt][2 * kt + 1][0], s[mt][2 * kt + 1][1]);
        pa[mt][kt][3] = pack_bf16(s[mt][2 * kt + 1][2], s[mt][2 * kt + 1][3]);
      }
    // the previous pair's O store must have read the staging tile before it is overwritten
    if (lane == 0) tma_store_wait_read<0>();
    __syncwarp();
    const int g = lane >> 2;
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
      for (int kt = 0; kt < FP / 16; ++kt) {
        uint32_t b[4];
        ldsm_x4_t(sv_a + ((kt * 16 + (lane & 15)) * LD + np * 16 + (lane >> 4) * 8) * 2, b);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma_bf16(acc[mt][0], pa[mt][kt], b[0], b[1]);
          mma_bf16(acc[mt][1], pa[mt][kt], b[2], b[3]);
        }
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int col = np * 16 + j * 8 + 2 * t4;
          *reinterpret_cast<uint32_t*>(so + (mt * 16 + g) * LD + col) =
              pack_bf16(acc[mt][j][0] * inv[mt][0], acc[mt][j][1] * inv[mt][0]);
          *reinterpret_cast<uint32_t*>(so + (mt * 16 + g + 8) * LD + col) =
              pack_bf16(acc[mt][j][2] * inv[mt][1], acc[mt][j][3] * inv[mt][1]);
        }
    }
    fence_proxy_async_smem();     // O tile (generic writes) -> TMA store; also orders this pair's
    __syncwarp();                 // ldmatrix reads before the loads that will refill the stage
    if (lane == 0) {
      tma_store_3d(&tmO, so, 0, (int)(pair % heads), (int)(pair / heads) * F);
      tma_store_commit();
    }
  }
  if (lane == 0) tma_store_wait<0>();
}

template <int DPAD, int FP>
int launch_tma(const void* q, const void* k, const void* v, void* o, int64_t N, int F, int heads, int d,
               int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, float scale, cudaStream_t st) {
  constexpr int per_warp = 7 * FP * DPAD * 2;
  constexpr int WARPS = per_warp * 8 <= 110 * 1024 ? 8 : (per_warp * 4 <= 110 * 1024 ? 4 : 2);
  constexpr int smem = WARPS * per_warp;
  auto kern = temporal_attn_tma_kernel<DPAD, FP, WARPS>;
  static int blocks_per_sm = 0;
  if (!blocks_per_sm) {
    VST_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    VST_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kern, WARPS * 32, smem));
    if (blocks_per_sm < 1) blocks_per_sm = 1;
  }
  CUtensorMap tQ, tK, tV, tO;
  const void* ptrs[4] = {q, k, v, o};
  const int64_t lds[4] = {ldq, ldk, ldv, ldo};
  CUtensorMap* maps[4] = {&tQ, &tK, &tV, &tO};
  for (int i = 0; i < 4; ++i) {
    const uint64_t dims[3] = {(uint64_t)d, (uint64_t)heads, (uint64_t)(N * F)};
    const uint64_t strides[3] = {2, (uint64_t)d * 2, (uint64_t)lds[i] * 2};
    const uint32_t box[3] = {(uint32_t)DPAD, 1, (uint32_t)F};
    int rc = vst_make_tmap(maps[i], ptrs[i], 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  const int64_t pairs = N * heads;
  int64_t grid = (pairs + WARPS - 1) / WARPS;
  const int64_t cap = (int64_t)vst_sm_count() * blocks_per_sm;
  if (grid > cap) grid = cap;
  kern<<<(unsigned)grid, WARPS * 32, smem, st>>>(tQ, tK, tV, tO, pairs, F, heads,
                                                 scale * 1.4426950408889634f);
  VST_LAUNCH_OK();
  return VST_OK;
}

template <int DPAD, int FP>
int launch(const void* q, const void* k, const void* v, void* o, int64_t N, int F, int heads,
           int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, float scale,
           cudaStream_t st) {
  constexpr int WARPS = (DPAD * FP <= 80 * 16) ? 8 : 4;
  constexpr int smem = WARPS * 3 * FP * (DPAD + 8) * 2;
  auto kern = temporal_attn_kernel<DPAD, FP, WARPS>;
  static bool configured = false;
  if (!configured) {
    VST_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  const int64_t pairs = N * heads;
  const int64_t grid = (pairs + WARPS - 1) / WARPS;
  VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "temporal attention: grid too large");
  kern<<<(unsigned)grid, WARPS * 32, smem, st>>>(
      (const __nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v,
      (__nv_bfloat16*)o, pairs, F, heads, d, ldq, ldk, ldv, ldo, scale * 1.4426950408889634f);
  VST_LAUNCH_OK();
  return VST_OK;
}

}  // namespace

extern "C" int vst_temporal_attn_fwd(const void* q, const void* k, const void* v, void* o,
                                     int64_t N, int F, int heads, int d, int64_t ldq, int64_t ldk,
                                     int64_t ldv, int64_t ldo, float scale, void* stream) {
  VST_REQUIRE(N >= 0 && F > 0 && heads > 0 && d > 0, VST_E_BADARG, "temporal attention: bad shape");
  if (N == 0) return VST_OK;
  VST_REQUIRE(q && k && v && o, VST_E_BADARG, "temporal attention: null pointer");
  VST_REQUIRE(F <= 32, VST_E_UNSUPPORTED, "temporal attention: F=%d > 32 (motion_max_seq_length)", F);
  VST_REQUIRE(d % 8 == 0 && d <= 160, VST_E_UNSUPPORTED,
              "temporal attention: head dim %d must be a multiple of 8, <= 160", d);
  VST_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0, VST_E_UNSUPPORTED,
              "temporal attention: row strides must be multiples of 8 elements");
  VST_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) |
                reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(o)) & 15) == 0,
              VST_E_UNSUPPORTED, "temporal attention: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  // TMA-fed form whenever the (d, head, row) view is expressible; VST_TATTN_TMA=0 forces the LDG form
  static int use_tma = -1;
  if (use_tma < 0) { const char* e = getenv("VST_TATTN_TMA"); use_tma = (e && e[0] == '0') ? 0 : 1; }
  const bool tma_ok = use_tma && N * F < (1ll << 31);
#define TA(DP, FPP)                                                                                    \
  do {                                                                                                 \
    if (tma_ok) return launch_tma<DP, FPP>(q, k, v, o, N, F, heads, d, ldq, ldk, ldv, ldo, scale, st); \
    return launch<DP, FPP>(q, k, v, o, N, F, heads, d, ldq, ldk, ldv, ldo, scale, st);                 \
  } while (0)
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
