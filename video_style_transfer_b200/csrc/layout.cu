// a9: (B,C,F,H,W) <-> (B*H*W, F, C) rearranges with the sinusoidal position add / residual add
// fused into the store.  Reference: animatediff/temporal_transformer.py:129-133,143-144 and
// PositionalEncoding.forward :27.  Pure data movement => bit-exact (the optional add is one
// fp32 add rounded once, exactly what torch does for same-dtype operands).
//
// HBM-bound: one read + one write per element.  Tile = 64 positions x 32 words staged in shared
// memory; 16-byte global loads along hw, 16-byte global stores along c; the shared tile is
// XOR-swizzled at 16-byte-chunk granularity so both phases are bank-conflict free for bf16.
#include <string.h>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kTileHW = 64;
constexpr int kTileWords = 32;  // 64 bf16 channels or 32 fp32 channels

__device__ __forceinline__ int swz(int row, int word) {
  return row * kTileWords + ((((word >> 2) ^ (row >> 3)) & 7) << 2) + (word & 3);
}

// a10 apply fused into the forward rearrange (MODE 2): y = x * a_c + b_c with
// a_c = rstd[b,g] * gamma[c], b_c = beta[c] - mean[b,g] * a_c   (fp32, rounded once).
struct NormArgs {
  const float* mean;
  const float* rstd;
  const void* gamma;
  const void* beta;
  int groups;
  int param_bf16;
};
__device__ __forceinline__ float2 norm_coeffs(const NormArgs& na, int b, int c, int C) {
  const int g = c / (C / na.groups);
  const float mu = __ldg(na.mean + b * na.groups + g), rs = __ldg(na.rstd + b * na.groups + g);
  float ga, be;
  if (na.param_bf16) {
    ga = __bfloat162float(((const __nv_bfloat16*)na.gamma)[c]);
    be = __bfloat162float(((const __nv_bfloat16*)na.beta)[c]);
  } else {
    ga = __ldg((const float*)na.gamma + c);
    be = __ldg((const float*)na.beta + c);
  }
  const float a = rs * ga;
  return make_float2(a, be - mu * a);
}
__device__ __forceinline__ uint32_t norm_word(uint32_t w, float2 ab) {
  const float2 e = unpack_bf16(w);
  return pack_bf16(fmaf(e.x, ab.x, ab.y), fmaf(e.y, ab.x, ab.y));
}

// Sequence-parallel addressing of the TOKEN side (§8e, comm.cu).  world == 0: the token tensor is
// local, (B*HW, F, C).  world > 0: this rank holds frames [f_offset, f_offset + F) of F_total; the
// token rows of position hw live on rank hw / hw_per_rank, in that rank's symmetric buffer laid out
// (B, hw_per_rank, F_total, C).  The forward modes then STORE to peer memory (push all-to-all fused
// into the GroupNorm-apply + rearrange), the backward mode LOADS from it (pull all-to-all fused
// into the inverse rearrange + residual add): the exchange costs no pass of its own.
struct ShardArgs {
  void* peer[VST_MAX_RANKS];
  int world, hw_per_rank, f_offset, f_total;
};
__device__ __forceinline__ int64_t tok_offset(const ShardArgs& sh, int b, int hw, int f, int F, int HW,
                                              int C, int* rank) {
  if (sh.world == 0) {
    *rank = -1;
    return (((int64_t)b * HW + hw) * F + f) * C;
  }
  const int r = hw / sh.hw_per_rank;
  *rank = r;
  return (((int64_t)b * sh.hw_per_rank + (hw - r * sh.hw_per_rank)) * sh.f_total + sh.f_offset + f) * C;
}

// ---- bf16: two channel rows per thread, 2x2 transposes with byte permutes -------------------
// MODE 0: forward (x[b,c,f,hw] -> y[(b,hw),f,c] + pe[f,c])
// MODE 1: backward (y[(b,hw),f,c] -> out[b,c,f,hw] + residual[b,c,f,hw])
// MODE 2: forward with the GroupNorm affine applied on the way in
// One CTA moves kNT 64 x 64 tiles that follow each other along hw, with the global loads of all of
// them issued before the first shared-memory pass: a one-tile CTA had 32 B per thread in flight and
// ran at 60-75 % of a plain copy (r01 kbench); kNT = 4 keeps 128 B (forward) / 256 B (backward +
// residual) per thread outstanding.
constexpr int kNT = 4;

template <int MODE>
__global__ void __launch_bounds__(kThreads)
rearrange_bf16_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                      const float* __restrict__ pe, const __nv_bfloat16* __restrict__ residual,
                      int C, int F, int HW, int64_t sB, int64_t sC, int64_t sF, int tilesHW,
                      int tilesC, NormArgs na, ShardArgs sh) {
  __shared__ __align__(16) uint32_t tiles[kNT][kTileHW * kTileWords];
  int t = blockIdx.x;
  const int thw = t % tilesHW;          // tilesHW counts GROUPS of kNT tiles
  t /= tilesHW;
  const int tc = t % tilesC;
  t /= tilesC;
  const int f = t % F;
  const int b = t / F;
  const int hwg = thw * (kNT * kTileHW), c0 = tc * 64;
  const int tid = threadIdx.x;

  // "planar" tensor P[b,c,f,hw] (strided) and "token" tensor T[(b,hw),f,c]
  const int64_t pbase = (int64_t)b * sB + (int64_t)f * sF;

  if (MODE == 0 || MODE == 2) {
    // load: thread = (channel pair cp, hw chunk hc); 16 B along hw from rows c, c+1
    const int hc = tid & 7, cp = tid >> 3;
    const int c = c0 + 2 * cp;
    uint4 r0[kNT], r1[kNT];
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
      const int hw = hwg + k * kTileHW + 8 * hc;
      r0[k] = make_uint4(0, 0, 0, 0);
      r1[k] = r0[k];
      if (hw < HW && c < C) {
        r0[k] = ld_stream16(src + pbase + (int64_t)c * sC + hw);
        if (c + 1 < C) r1[k] = ld_stream16(src + pbase + (int64_t)(c + 1) * sC + hw);
      }
    }
    float2 k0 = make_float2(1.f, 0.f), k1 = k0;
    if (MODE == 2 && c < C) {
      k0 = norm_coeffs(na, b, c, C);
      k1 = c + 1 < C ? norm_coeffs(na, b, c + 1, C) : k0;
    }
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
      const int hw = hwg + k * kTileHW + 8 * hc;
      uint32_t a[4] = {r0[k].x, r0[k].y, r0[k].z, r0[k].w}, bb[4] = {r1[k].x, r1[k].y, r1[k].z, r1[k].w};
      if (MODE == 2 && hw < HW && c < C) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          a[j] = norm_word(a[j], k0);
          bb[j] = norm_word(bb[j], k1);
        }
      }
      uint32_t* tile = tiles[k];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        tile[swz(8 * hc + 2 * j, cp)] = __byte_perm(a[j], bb[j], 0x5410);      // (c,hw) (c+1,hw)
        tile[swz(8 * hc + 2 * j + 1, cp)] = __byte_perm(a[j], bb[j], 0x7632);  // hw+1
      }
    }
    __syncthreads();
    // store: item = (hw row, 16-B chunk of 8 channels)
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
      const uint32_t* tile = tiles[k];
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        const int item = tid + it * kThreads;
        const int row = item >> 3, ch = item & 7;
        const int hwr = hwg + k * kTileHW + row, cc = c0 + 8 * ch;
        if (hwr < HW && cc < C) {
          uint4 v = *reinterpret_cast<const uint4*>(&tile[swz(row, 4 * ch)]);
          if (pe != nullptr) {
            const float4 p0 = __ldg(reinterpret_cast<const float4*>(pe + (int64_t)f * C + cc));
            const float4 p1 = __ldg(reinterpret_cast<const float4*>(pe + (int64_t)f * C + cc + 4));
            float2 e;
            e = unpack_bf16(v.x); v.x = pack_bf16(e.x + p0.x, e.y + p0.y);
            e = unpack_bf16(v.y); v.y = pack_bf16(e.x + p0.z, e.y + p0.w);
            e = unpack_bf16(v.z); v.z = pack_bf16(e.x + p1.x, e.y + p1.y);
            e = unpack_bf16(v.w); v.w = pack_bf16(e.x + p1.z, e.y + p1.w);
          }
          int rk;
          const int64_t toff = tok_offset(sh, b, hwr, f, F, HW, C, &rk);
          __nv_bfloat16* tdst = rk < 0 ? dst : reinterpret_cast<__nv_bfloat16*>(sh.peer[rk]);
          st_stream16(tdst + toff + cc, v);
        }
      }
    }
  } else {
    // load: item = (hw row, 16-B chunk of 8 channels) from the token tensor
    uint4 v[kNT][2];
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        const int item = tid + it * kThreads;
        const int row = item >> 3, ch = item & 7;
        const int hwr = hwg + k * kTileHW + row, cc = c0 + 8 * ch;
        v[k][it] = make_uint4(0, 0, 0, 0);
        if (hwr < HW && cc < C) {
          int rk;
          const int64_t toff = tok_offset(sh, b, hwr, f, F, HW, C, &rk);
          const __nv_bfloat16* tsrc = rk < 0 ? src : reinterpret_cast<const __nv_bfloat16*>(sh.peer[rk]);
          v[k][it] = ld_stream16(tsrc + toff + cc);
        }
      }
    }
    // the residual rows this thread will add: in flight under the shared-memory pass
    const int hc = tid & 7, cp = tid >> 3;
    const int c = c0 + 2 * cp;
    uint4 rr[kNT][2];
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
      const int hw = hwg + k * kTileHW + 8 * hc;
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        rr[k][r] = make_uint4(0, 0, 0, 0);
        if (residual != nullptr && hw < HW && c + r < C)
          rr[k][r] = ld_stream16(residual + pbase + (int64_t)(c + r) * sC + hw);
      }
    }
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        const int item = tid + it * kThreads;
        const int row = item >> 3, ch = item & 7;
        *reinterpret_cast<uint4*>(&tiles[k][swz(row, 4 * ch)]) = v[k][it];
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kNT; ++k) {
      const int hw = hwg + k * kTileHW + 8 * hc;
      if (hw < HW && c < C) {
        const uint32_t* tile = tiles[k];
        uint32_t lo[4], hi[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t w0 = tile[swz(8 * hc + 2 * j, cp)];      // (c,hw)   (c+1,hw)
          const uint32_t w1 = tile[swz(8 * hc + 2 * j + 1, cp)];  // (c,hw+1) (c+1,hw+1)
          lo[j] = __byte_perm(w0, w1, 0x5410);                    // row c  : hw, hw+1
          hi[j] = __byte_perm(w0, w1, 0x7632);                    // row c+1
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          if (c + r >= C) break;
          uint32_t* w = r == 0 ? lo : hi;
          const int64_t off = pbase + (int64_t)(c + r) * sC + hw;
          if (residual != nullptr) {
            const uint32_t rv[4] = {rr[k][r].x, rr[k][r].y, rr[k][r].z, rr[k][r].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 e = unpack_bf16(w[j]), q = unpack_bf16(rv[j]);
              w[j] = pack_bf16(e.x + q.x, e.y + q.y);
            }
          }
          st_stream16(dst + off, make_uint4(w[0], w[1], w[2], w[3]));
        }
      }
    }
  }
}

// ---- generic fallback (fp32, or shapes/pointers that are not 16-byte friendly) ---------------
template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f(float v);
template <>
__device__ __forceinline__ float from_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename T, int MODE>
__global__ void __launch_bounds__(kThreads)
rearrange_generic_kernel(const T* __restrict__ src, T* __restrict__ dst,
                         const float* __restrict__ pe, const T* __restrict__ residual, int C,
                         int F, int HW, int64_t sB, int64_t sC, int64_t sF, int tilesHW,
                         int tilesC, NormArgs na) {
  __shared__ T tile[32][33];
  int t = blockIdx.x;
  const int thw = t % tilesHW;
  t /= tilesHW;
  const int tc = t % tilesC;
  t /= tilesC;
  const int f = t % F;
  const int b = t / F;
  const int hw0 = thw * 32, c0 = tc * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int64_t pbase = (int64_t)b * sB + (int64_t)f * sF;
  const int64_t tbase = ((int64_t)b * HW * F + f) * C;
  if (MODE == 0 || MODE == 2) {
    for (int i = ty; i < 32; i += 8) {
      const int c = c0 + i, hw = hw0 + tx;
      if (c < C && hw < HW) {
        T v = src[pbase + (int64_t)c * sC + hw];
        if (MODE == 2) {
          const float2 k = norm_coeffs(na, b, c, C);
          v = from_f<T>(fmaf(to_f<T>(v), k.x, k.y));
        }
        tile[i][tx] = v;
      }
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
      const int hw = hw0 + i, c = c0 + tx;
      if (c < C && hw < HW) {
        T v = tile[tx][i];
        if (pe != nullptr) v = from_f<T>(to_f<T>(v) + pe[(int64_t)f * C + c]);
        dst[tbase + (int64_t)hw * F * C + c] = v;
      }
    }
  } else {
    for (int i = ty; i < 32; i += 8) {
      const int hw = hw0 + i, c = c0 + tx;
      if (c < C && hw < HW) tile[i][tx] = src[tbase + (int64_t)hw * F * C + c];
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
      const int c = c0 + i, hw = hw0 + tx;
      if (c < C && hw < HW) {
        const int64_t off = pbase + (int64_t)c * sC + hw;
        T v = tile[tx][i];
        if (residual != nullptr) v = from_f<T>(to_f<T>(v) + to_f<T>(residual[off]));
        dst[off] = v;
      }
    }
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int MODE>
int launch(const void* src, void* dst, const float* pe, const void* residual, int B, int C,
           int F, int HW, int64_t sB, int64_t sC, int64_t sF, int dtype, cudaStream_t st,
           NormArgs na = NormArgs{nullptr, nullptr, nullptr, nullptr, 1, 0},
           const ShardArgs* shard = nullptr) {
  if ((int64_t)B * C * F * HW == 0) return VST_OK;
  ShardArgs sh;
  memset(&sh, 0, sizeof(sh));
  if (shard) sh = *shard;
  const bool fast = dtype == VST_BF16 && C % 8 == 0 && HW % 8 == 0 && sB % 8 == 0 &&
                    sC % 8 == 0 && sF % 8 == 0 && aligned16(src) && aligned16(dst) &&
                    (residual == nullptr || aligned16(residual)) &&
                    (pe == nullptr || aligned16(pe));
  VST_REQUIRE(fast || shard == nullptr, VST_E_UNSUPPORTED,
              "sequence-parallel rearrange needs bf16 and 16-byte friendly shapes / pointers");
  if (fast) {
    const int tilesHW = (HW + kNT * kTileHW - 1) / (kNT * kTileHW), tilesC = (C + 63) / 64;   // groups of kNT tiles
    const int64_t grid = (int64_t)B * F * tilesHW * tilesC;
    VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "rearrange: grid too large");
    rearrange_bf16_kernel<MODE><<<(unsigned)grid, kThreads, 0, st>>>(
        (const __nv_bfloat16*)src, (__nv_bfloat16*)dst, pe, (const __nv_bfloat16*)residual, C, F,
        HW, sB, sC, sF, tilesHW, tilesC, na, sh);
  } else {
    const int tilesHW = (HW + 31) / 32, tilesC = (C + 31) / 32;
    const int64_t grid = (int64_t)B * F * tilesHW * tilesC;
    VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "rearrange: grid too large");
    if (dtype == VST_BF16)
      rearrange_generic_kernel<__nv_bfloat16, MODE><<<(unsigned)grid, kThreads, 0, st>>>(
          (const __nv_bfloat16*)src, (__nv_bfloat16*)dst, pe, (const __nv_bfloat16*)residual, C,
          F, HW, sB, sC, sF, tilesHW, tilesC, na);
    else
      rearrange_generic_kernel<float, MODE><<<(unsigned)grid, kThreads, 0, st>>>(
          (const float*)src, (float*)dst, pe, (const float*)residual, C, F, HW, sB, sC, sF,
          tilesHW, tilesC, na);
  }
  VST_LAUNCH_OK();
  return VST_OK;
}

int check(const void* a, const void* b, int B, int C, int F, int HW, int in_dtype,
          int out_dtype) {
  VST_REQUIRE(B >= 0 && C >= 0 && F >= 0 && HW >= 0, VST_E_BADARG, "rearrange: negative dim");
  VST_REQUIRE(in_dtype == out_dtype && (in_dtype == VST_F32 || in_dtype == VST_BF16),
              VST_E_UNSUPPORTED, "rearrange: in/out dtype must match and be f32 or bf16");
  VST_REQUIRE((a && b) || (int64_t)B * C * F * HW == 0, VST_E_BADARG, "rearrange: null pointer");
  return VST_OK;
}

}  // namespace

extern "C" int vst_rearrange_fwd(const void* x, void* y, const float* pe, int B, int C, int F,
                                 int HW, int64_t sB, int64_t sC, int64_t sF, int in_dtype,
                                 int out_dtype, void* stream) {
  int rc = check(x, y, B, C, F, HW, in_dtype, out_dtype);
  if (rc) return rc;
  return launch<0>(x, y, pe, nullptr, B, C, F, HW, sB, sC, sF, in_dtype, (cudaStream_t)stream);
}

extern "C" int vst_rearrange_bwd(const void* y, void* out, const void* residual, int B, int C,
                                 int F, int HW, int64_t sB, int64_t sC, int64_t sF, int in_dtype,
                                 int out_dtype, void* stream) {
  int rc = check(y, out, B, C, F, HW, in_dtype, out_dtype);
  if (rc) return rc;
  return launch<1>(y, out, nullptr, residual, B, C, F, HW, sB, sC, sF, in_dtype,
                   (cudaStream_t)stream);
}

extern "C" int vst_groupnorm5d_apply_rearrange(const void* x, void* y, const float* mean,
                                               const float* rstd, const void* gamma,
                                               const void* beta, int B, int C, int F, int HW,
                                               int64_t sB, int64_t sC, int64_t sF, int groups,
                                               int in_dtype, int out_dtype, int param_dtype,
                                               void* stream) {
  int rc = check(x, y, B, C, F, HW, in_dtype, out_dtype);
  if (rc) return rc;
  VST_REQUIRE(groups > 0 && C % groups == 0, VST_E_BADARG, "groupnorm: C %% groups != 0");
  VST_REQUIRE(mean && rstd && gamma && beta, VST_E_BADARG, "groupnorm: null stats/affine");
  VST_REQUIRE(param_dtype == VST_F32 || param_dtype == VST_BF16, VST_E_BADARG, "bad param dtype");
  NormArgs na{mean, rstd, gamma, beta, groups, param_dtype == VST_BF16};
  return launch<2>(x, y, nullptr, nullptr, B, C, F, HW, sB, sC, sF, in_dtype,
                   (cudaStream_t)stream, na);
}

// ---- §8e: the two all-to-alls of a sequence-parallel motion module ------------------------------
namespace {
int make_shard(ShardArgs* sh, const VstComm* c, int64_t offset_bytes, int F_local, int F_total, int HW) {
  VST_REQUIRE(c != nullptr, VST_E_BADARG, "a2a: null communicator");
  VST_REQUIRE(F_local > 0 && F_total == F_local * c->world, VST_E_BADARG,
              "a2a: F_total=%d must be world(%d) x F_local(%d)", F_total, c->world, F_local);
  VST_REQUIRE(HW % (c->world * kTileHW) == 0, VST_E_UNSUPPORTED,
              "a2a: HW=%d must be a multiple of world x %d", HW, kTileHW);
  VST_REQUIRE(offset_bytes >= 0 && offset_bytes % 16 == 0, VST_E_BADARG, "a2a: buffer offset must be a multiple of 16");
  memset(sh, 0, sizeof(*sh));
  for (int r = 0; r < c->world; ++r) sh->peer[r] = reinterpret_cast<uint8_t*>(c->buf[r]) + offset_bytes;
  sh->world = c->world;
  sh->hw_per_rank = HW / c->world;
  sh->f_offset = c->rank * F_local;
  sh->f_total = F_total;
  return VST_OK;
}
}  // namespace

extern "C" int vst_a2a_frames_to_hw(void* comm, const void* x, int64_t dst_offset_bytes,
                                    const float* mean, const float* rstd, const void* gamma,
                                    const void* beta, int B, int C, int F_local, int F_total, int HW,
                                    int64_t sB, int64_t sC, int64_t sF, int groups, int param_dtype,
                                    void* stream) {
  VST_REQUIRE(x != nullptr && B > 0 && C > 0, VST_E_BADARG, "a2a_frames_to_hw: bad argument");
  ShardArgs sh;
  int rc = make_shard(&sh, reinterpret_cast<const VstComm*>(comm), dst_offset_bytes, F_local, F_total, HW);
  if (rc) return rc;
  void* local = sh.peer[reinterpret_cast<const VstComm*>(comm)->rank];
  if (mean != nullptr) {
    VST_REQUIRE(rstd && gamma && beta && groups > 0 && C % groups == 0, VST_E_BADARG, "a2a_frames_to_hw: bad norm arguments");
    NormArgs na{mean, rstd, gamma, beta, groups, param_dtype == VST_BF16};
    return launch<2>(x, local, nullptr, nullptr, B, C, F_local, HW, sB, sC, sF, VST_BF16, (cudaStream_t)stream, na, &sh);
  }
  return launch<0>(x, local, nullptr, nullptr, B, C, F_local, HW, sB, sC, sF, VST_BF16, (cudaStream_t)stream,
                   NormArgs{nullptr, nullptr, nullptr, nullptr, 1, 0}, &sh);
}

extern "C" int vst_a2a_hw_to_frames(void* comm, int64_t src_offset_bytes, void* out, const void* residual,
                                    int B, int C, int F_local, int F_total, int HW, int64_t sB,
                                    int64_t sC, int64_t sF, void* stream) {
  VST_REQUIRE(out != nullptr && B > 0 && C > 0, VST_E_BADARG, "a2a_hw_to_frames: bad argument");
  ShardArgs sh;
  int rc = make_shard(&sh, reinterpret_cast<const VstComm*>(comm), src_offset_bytes, F_local, F_total, HW);
  if (rc) return rc;
  const void* local = sh.peer[reinterpret_cast<const VstComm*>(comm)->rank];
  return launch<1>(local, out, nullptr, residual, B, C, F_local, HW, sB, sC, sF, VST_BF16, (cudaStream_t)stream,
                   NormArgs{nullptr, nullptr, nullptr, nullptr, 1, 0}, &sh);
}


// ---- GroupNorm affine on the TOKEN layout ((B, rows_per_batch, C) bf16, in place) -----------------------
// Sequence parallelism (comm.cu): the frames -> positions push sends RAW activations together with the
// GroupNorm partials, so ONE flag barrier covers both; the owner of the positions then applies
// y = x * a_c + b_c here, with exactly the coefficients and the single fp32 fma + rounding of the
// fused apply+rearrange above (bit-identical tokens).
namespace {
__global__ void gn_tokens_apply_kernel(__nv_bfloat16* __restrict__ tok, int64_t rows_per_batch, int C,
                                       NormArgs na) {
  extern __shared__ float2 coef[];                 // [C] (a_c, b_c) of this block's batch
  const int b = blockIdx.y;
  for (int c = threadIdx.x; c < C; c += blockDim.x) coef[c] = norm_coeffs(na, b, c, C);
  __syncthreads();
  const int vec_per_row = C >> 3;
  const int64_t total = rows_per_batch * vec_per_row;
  uint4* base = reinterpret_cast<uint4*>(tok + (int64_t)b * rows_per_batch * C);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c0 = (int)(i % vec_per_row) * 8;
    uint4 w = base[i];
    uint32_t* ww = reinterpret_cast<uint32_t*>(&w);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 e = unpack_bf16(ww[j]);
      const float2 k0 = coef[c0 + 2 * j], k1 = coef[c0 + 2 * j + 1];
      ww[j] = pack_bf16(fmaf(e.x, k0.x, k0.y), fmaf(e.y, k1.x, k1.y));
    }
    base[i] = w;
  }
}
}  // namespace

extern "C" int vst_groupnorm_tokens_apply(void* tok, const float* mean, const float* rstd, const void* gamma,
                                          const void* beta, int B, int64_t rows_per_batch, int C, int groups,
                                          int param_dtype, void* stream) {
  VST_REQUIRE(tok && mean && rstd && gamma && beta && B > 0 && rows_per_batch >= 0 && C > 0, VST_E_BADARG,
              "groupnorm_tokens_apply: bad argument");
  VST_REQUIRE(groups > 0 && C % groups == 0 && C % 8 == 0, VST_E_UNSUPPORTED,
              "groupnorm_tokens_apply: C=%d must be a multiple of 8 and of groups=%d", C, groups);
  if (rows_per_batch == 0) return VST_OK;
  NormArgs na{mean, rstd, gamma, beta, groups, param_dtype == VST_BF16};
  const int64_t vecs = rows_per_batch * (C / 8);
  int gx = (int)((vecs + 256 * 8 - 1) / (256 * 8));
  const int cap = 8 * vst_sm_count();
  if (gx > cap) gx = cap;
  if (gx < 1) gx = 1;
  gn_tokens_apply_kernel<<<dim3(gx, B), 256, C * sizeof(float2), (cudaStream_t)stream>>>(
      reinterpret_cast<__nv_bfloat16*>(tok), rows_per_batch, C, na);
  VST_LAUNCH_OK();
  return VST_OK;
}
