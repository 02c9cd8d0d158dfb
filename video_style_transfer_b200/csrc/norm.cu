// a10 statistics (GroupNorm over the (B,C,F,H,W) view) and K9 LayerNorm (+ fused position add).
// Both are HBM-bound: one 16-byte-vectorised pass, warp-shuffle reductions, fp32 statistics.
//
// GroupNorm statistics are computed as (count, mean, M2) partials over slabs that a block holds
// entirely in registers (exact two-pass inside the slab) and merged with Chan's formula, so
// there is no E[x^2]-E[x]^2 cancellation and the result is deterministic (no atomics).  Under
// sequence parallelism the partials are what crosses the NVLink all-reduce.
#include "common.cuh"

namespace {

constexpr int kSlabThreads = 256;
constexpr int kVecPerThread = 4;                                      // 4 x 8 bf16 (8 per thread measured slower)
constexpr int kSlabElems = kSlabThreads * kVecPerThread * 8;          // 8192

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = l < nw ? red[l] : 0.f;
  return warp_sum(t);
}

template <typename T>
__device__ __forceinline__ void load8(const T* p, float* v);
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float* v) {
  const uint4 r = ld_stream16(p);
  float2 e;
  e = unpack_bf16(r.x); v[0] = e.x; v[1] = e.y;
  e = unpack_bf16(r.y); v[2] = e.x; v[3] = e.y;
  e = unpack_bf16(r.z); v[4] = e.x; v[5] = e.y;
  e = unpack_bf16(r.w); v[6] = e.x; v[7] = e.y;
}
template <>
__device__ __forceinline__ void load8<float>(const float* p, float* v) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// One work item (b,g,f,s) = elements [s*kSlabElems, ...) of the cg*HW chunk of one (group, frame)
// (32-bit index arithmetic: the host checks cg*HW < 2^31).  Blocks are persistent over a strided
// list of items and keep the NEXT item's 16-byte loads in flight (raw, in registers) while the
// current one goes through its two block reductions: one-shot blocks spent most of their life
// in the reductions with nothing outstanding (r01: 1.8 TB/s at L0).
template <typename T> struct Raw8;
template <> struct Raw8<__nv_bfloat16> {
  uint4 r;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { r = ld_stream16(p); }
  __device__ __forceinline__ void zero() { r = make_uint4(0, 0, 0, 0); }
  __device__ __forceinline__ void get(float* v) const {
    float2 e;
    e = unpack_bf16(r.x); v[0] = e.x; v[1] = e.y;
    e = unpack_bf16(r.y); v[2] = e.x; v[3] = e.y;
    e = unpack_bf16(r.z); v[4] = e.x; v[5] = e.y;
    e = unpack_bf16(r.w); v[6] = e.x; v[7] = e.y;
  }
};
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = *reinterpret_cast<const float4*>(p);
    b = *reinterpret_cast<const float4*>(p + 4);
  }
  __device__ __forceinline__ void zero() { a = b = make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ __forceinline__ void get(float* v) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};

// Vector path (HW, strides, base pointer all 16-byte friendly).  A block owns ONE (b, group) and walks its
// (frame, slab) items with a fixed stride, stepping (f, s) by add-and-carry: the first version decomposed a flat
// item index and every element index with runtime integer divisions -- ~45 SASS instructions between two
// 16-byte loads, ~500 per thread and slab, which held the kernel at 2.3 TB/s (0.35 of HBM) on issue slots and
// the I2F / MUFU.RCP latency chains, not on memory.  Here the address of a vector is `base + e` when the group's
// channels are contiguous (frames-major activations, sC == HW), a shift / mask when HW is a power of two, and a
// division only otherwise.  Arithmetic per slab and the partial order are unchanged (bit-identical partials).
template <typename T>
__global__ void __launch_bounds__(kSlabThreads, 4)
gn_stats_vec_kernel(const T* __restrict__ x, float* __restrict__ partials, int F, int HW, int hw_shift,
                    int contig, int64_t sB, int64_t sC, int64_t sF, int groups, int cg, int slabs,
                    int blocks_per_bg) {
  __shared__ float red[32];
  const int bg = blockIdx.x / blocks_per_bg, jb = blockIdx.x - bg * blocks_per_bg;
  const int b = bg / groups, g = bg - b * groups;
  const T* base_bg = x + (int64_t)b * sB + (int64_t)g * cg * sC;
  const int chunk = cg * HW;
  const int per_bg = F * slabs;
  const int step_f = blocks_per_bg / slabs, step_s = blocks_per_bg - step_f * slabs;
  const int tid8 = (int)threadIdx.x * 8;
  Raw8<T> nxt[kVecPerThread];
  auto fetch = [&](int f, int s) {
    const T* base = base_bg + (int64_t)f * sF;
    const int e0 = s * kSlabElems + tid8;
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i) {
      const int e = e0 + i * (kSlabThreads * 8);
      if (e < chunk) {
        if (contig) {
          nxt[i].load(base + e);
        } else {
          const int c = hw_shift >= 0 ? (e >> hw_shift) : (e / HW);
          nxt[i].load(base + (int64_t)c * sC + (e - c * HW));
        }
      } else {
        nxt[i].zero();
      }
    }
  };
  int j = jb, f = jb / slabs, s = jb - f * slabs;
  if (j < per_bg) fetch(f, s);
  while (j < per_bg) {
    float v[kVecPerThread][8];
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i) nxt[i].get(v[i]);
    const int nj = j + blocks_per_bg;
    int nf = f + step_f, ns = s + step_s;
    if (ns >= slabs) { ns -= slabs; ++nf; }
    if (nj < per_bg) fetch(nf, ns);                               // in flight under the reductions
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i)
#pragma unroll
      for (int k = 0; k < 8; ++k) sum += v[i][k];
    const int e0 = s * kSlabElems;
    const int rem = chunk - e0;
    const float n = (float)(rem < kSlabElems ? rem : kSlabElems);
    const float mean = block_sum(sum, red) / n;
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i) {
      // chunk is a multiple of 8 here: a vector is entirely inside the chunk or entirely outside
      if (e0 + tid8 + i * (kSlabThreads * 8) < chunk) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float d = v[i][k] - mean;
          m2 = fmaf(d, d, m2);
        }
      }
    }
    m2 = block_sum(m2, red);
    if (threadIdx.x == 0) {
      float* out = partials + ((int64_t)(bg * F + f) * slabs + s) * 3;      // item order, as the scalar kernel
      out[0] = n; out[1] = mean; out[2] = m2;
    }
    j = nj; f = nf; s = ns;
  }
}

// Scalar path (odd HW / strides / alignment): one work item (b,g,f,s) per loop trip of a persistent block.
template <typename T>
__global__ void __launch_bounds__(kSlabThreads, 4)
gn_stats_kernel(const T* __restrict__ x, float* __restrict__ partials, int C, int F, int HW,
                int64_t sB, int64_t sC, int64_t sF, int groups, int slabs, int total) {
  __shared__ float red[32];
  const int cg = C / groups;
  const int chunk = cg * HW;
  for (int t = blockIdx.x; t < total; t += gridDim.x) {
    int r = t;
    const int s = r % slabs; r /= slabs;
    const int f = r % F; r /= F;
    const int g = r % groups;
    const int b = r / groups;
    const T* base = x + (int64_t)b * sB + (int64_t)f * sF + (int64_t)g * cg * sC;
    const int e0 = s * kSlabElems;
    float v[kVecPerThread][8];
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i) {
      const int e = e0 + (i * kSlabThreads + (int)threadIdx.x) * 8;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int ek = e + k;
        if (ek < chunk) {
          const int c = ek / HW, hw = ek - c * HW;
          v[i][k] = (float)base[(int64_t)c * sC + hw];
        } else {
          v[i][k] = 0.f;
        }
      }
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i)
#pragma unroll
      for (int k = 0; k < 8; ++k) sum += v[i][k];
    const int rem = chunk - e0;
    const float n = (float)(rem < kSlabElems ? rem : kSlabElems);
    const float mean = block_sum(sum, red) / n;
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < kVecPerThread; ++i) {
      const int e = e0 + (i * kSlabThreads + (int)threadIdx.x) * 8;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (e + k < chunk) {
          const float d = v[i][k] - mean;
          m2 = fmaf(d, d, m2);
        }
    }
    m2 = block_sum(m2, red);
    if (threadIdx.x == 0) {
      // item t = ((b*groups + g)*F + f)*slabs + s: the partials are laid out in item order
      float* out = partials + (int64_t)t * 3;
      out[0] = n; out[1] = mean; out[2] = m2;
    }
  }
}

__device__ __forceinline__ void chan(float& n, float& mu, float& m2, float nb, float mub,
                                     float m2b) {
  if (nb == 0.f) return;
  const float nn = n + nb, d = mub - mu;
  mu = mu + d * (nb / nn);
  m2 = m2 + m2b + d * d * (n * nb / nn);
  n = nn;
}

// one warp per (b, group): fixed-order combine => deterministic
__global__ void gn_finalize_kernel(const float* __restrict__ partials, int nparts, float eps,
                                   float* __restrict__ mean, float* __restrict__ rstd) {
  const int bg = blockIdx.x, l = threadIdx.x;
  const float* p = partials + (int64_t)bg * nparts * 3;
  float n = 0.f, mu = 0.f, m2 = 0.f;
  for (int i = l; i < nparts; i += 32) chan(n, mu, m2, p[3 * i], p[3 * i + 1], p[3 * i + 2]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float nb = __shfl_xor_sync(0xffffffffu, n, o);
    const float mub = __shfl_xor_sync(0xffffffffu, mu, o);
    const float m2b = __shfl_xor_sync(0xffffffffu, m2, o);
    // combine in a lane-symmetric order so every lane ends with the same value
    float na = n, mua = mu, m2a = m2;
    if (l & o) { na = nb; mua = mub; m2a = m2b; chan(na, mua, m2a, n, mu, m2); }
    else chan(na, mua, m2a, nb, mub, m2b);
    n = na; mu = mua; m2 = m2a;
  }
  if (l == 0) {
    mean[bg] = mu;
    rstd[bg] = rsqrtf(m2 / n + eps);
  }
}

// ---- LayerNorm: one warp per ROWS rows, rows cached in registers --------------------------------
// Memory-bound (2 B read + 2 B written per element).  All of a warp's rows are loaded before the
// first reduction (ROWS x MAXV 16-byte loads in flight per lane), gamma / beta / the position row
// are read with 16-byte cached loads, and gamma / beta are shared by the warp's rows.  The first
// version issued 24 scalar parameter loads per 16 bytes of activation and ran at 0.25 of HBM.
template <typename TP>
__device__ __forceinline__ void loadp8(const TP* p, float* v);
template <>
__device__ __forceinline__ void loadp8<__nv_bfloat16>(const __nv_bfloat16* p, float* v) {
  const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
  float2 e;
  e = unpack_bf16(r.x); v[0] = e.x; v[1] = e.y;
  e = unpack_bf16(r.y); v[2] = e.x; v[3] = e.y;
  e = unpack_bf16(r.z); v[4] = e.x; v[5] = e.y;
  e = unpack_bf16(r.w); v[6] = e.x; v[7] = e.y;
}
template <>
__device__ __forceinline__ void loadp8<float>(const float* p, float* v) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

template <typename TI, typename TO, typename TP, int MAXV, int ROWS>
__global__ void __launch_bounds__(256)
layernorm_kernel(const TI* __restrict__ x, TO* __restrict__ y, const TP* __restrict__ gamma,
                 const TP* __restrict__ beta, const float* __restrict__ pe, int64_t rows, int C,
                 int F, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * ROWS;
  if (row0 >= rows) return;
  const int nvec = C >> 3;
  float v[ROWS][MAXV][8];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    const bool ok = row0 + r < rows;
    const TI* xr = x + (row0 + r) * C;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int vi = lane + 32 * i;
      if (ok && vi < nvec) {
        load8<TI>(xr + vi * 8, v[r][i]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[r][i][j] = 0.f;
      }
    }
  }
  float mean[ROWS], rstd[ROWS];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[r][i][j];
    mean[r] = warp_sum(sum) / (float)C;
  }
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      if (lane + 32 * i < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[r][i][j] - mean[r];
          m2 = fmaf(d, d, m2);
        }
      }
    }
    rstd[r] = rsqrtf(warp_sum(m2) / (float)C + eps);
  }
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int vi = lane + 32 * i;
    if (vi >= nvec) continue;
    float g[8], bt[8];
    if (gamma) loadp8<TP>(gamma + vi * 8, g);
    if (beta) loadp8<TP>(beta + vi * 8, bt);
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      if (row0 + r >= rows) continue;
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float t = (v[r][i][j] - mean[r]) * rstd[r];
        if (gamma) t = t * g[j];
        if (beta) t = t + bt[j];
        o[j] = t;
      }
      if (pe) {
        // LN output is first rounded to the activation dtype, then the position row is added
        // (diffusers: norm_hidden_states = pos_embed(norm1(x))) -- two roundings, as in torch.
        float pr[8];
        loadp8<float>(pe + (int64_t)((row0 + r) % F) * C + vi * 8, pr);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (float)(TO)o[j] + pr[j];
      }
      TO* yr = y + (row0 + r) * C;
      if (sizeof(TO) == 2) {
        uint4 w;
        w.x = pack_bf16(o[0], o[1]); w.y = pack_bf16(o[2], o[3]);
        w.z = pack_bf16(o[4], o[5]); w.w = pack_bf16(o[6], o[7]);
        st_stream16(yr + vi * 8, w);
      } else {
        float4* d = reinterpret_cast<float4*>(yr + vi * 8);
        d[0] = make_float4(o[0], o[1], o[2], o[3]);
        d[1] = make_float4(o[4], o[5], o[6], o[7]);
      }
    }
  }
}

// C in {320, 640, 1280} (the SDXL motion-module widths): 160 / NVEC rows per warp laid out as ONE
// flat span of 160 16-byte vectors = exactly 5 per lane, so no lane idles (the per-row mapping
// leaves 24 of 32 lanes idle on the second vector of a 40-vector row) and 5 loads are in flight
// per lane before the first reduction.
#ifndef LN_MIN_BLOCKS
#define LN_MIN_BLOCKS 4
#endif
template <typename TI, typename TO, typename TP, int NVEC>
__global__ void __launch_bounds__(256, LN_MIN_BLOCKS)
layernorm_flat_kernel(const TI* __restrict__ x, TO* __restrict__ y, const TP* __restrict__ gamma,
                      const TP* __restrict__ beta, const float* __restrict__ pe, int64_t rows, int F,
                      float eps) {
  constexpr int ROWS = 160 / NVEC, C = NVEC * 8;
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * ROWS;
  if (row0 >= rows) return;
  // frame index of the warp's first row: ONE 64-bit modulo per warp instead of one per vector
  const int f0 = pe ? (int)(row0 % F) : 0;
  float v[5][8];
  int rr[5], vv[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int f = i * 32 + lane;
    rr[i] = f / NVEC;
    vv[i] = f - rr[i] * NVEC;
    if (row0 + rr[i] < rows) {
      load8<TI>(x + (row0 + rr[i]) * C + vv[i] * 8, v[i]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
    }
  }
  float mean[ROWS], rstd[ROWS];
  {
    float part[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) part[r] = 0.f;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      float s8 = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s8 += v[i][j];
#pragma unroll
      for (int r = 0; r < ROWS; ++r) part[r] += rr[i] == r ? s8 : 0.f;
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) mean[r] = warp_sum(part[r]) * (1.f / (float)C);
  }
  float mu[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    mu[i] = mean[0];
#pragma unroll
    for (int r = 1; r < ROWS; ++r) mu[i] = rr[i] == r ? mean[r] : mu[i];
  }
  {
    float part[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) part[r] = 0.f;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      float s8 = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = v[i][j] - mu[i];
        s8 = fmaf(d, d, s8);
      }
#pragma unroll
      for (int r = 0; r < ROWS; ++r) part[r] += rr[i] == r ? s8 : 0.f;
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) rstd[r] = rsqrtf(warp_sum(part[r]) * (1.f / (float)C) + eps);
  }
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int64_t row = row0 + rr[i];
    if (row >= rows) continue;
    float rs = rstd[0];
#pragma unroll
    for (int r = 1; r < ROWS; ++r) rs = rr[i] == r ? rstd[r] : rs;
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mu[i]) * rs;
    if (gamma) {
      float g[8];
      loadp8<TP>(gamma + vv[i] * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] *= g[j];
    }
    if (beta) {
      float bt[8];
      loadp8<TP>(beta + vv[i] * 8, bt);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += bt[j];
    }
    if (pe) {
      float pr[8];
      int fr = f0 + rr[i];             // rr < 4 <= F in every caller; the loop covers tiny F too
      while (fr >= F) fr -= F;
      loadp8<float>(pe + fr * C + vv[i] * 8, pr);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (float)(TO)o[j] + pr[j];   // two roundings, as in torch
    }
    TO* yp = y + row * C + vv[i] * 8;
    if (sizeof(TO) == 2) {
      uint4 w;
      w.x = pack_bf16(o[0], o[1]); w.y = pack_bf16(o[2], o[3]);
      w.z = pack_bf16(o[4], o[5]); w.w = pack_bf16(o[6], o[7]);
      st_stream16(yp, w);
    } else {
      float4* d = reinterpret_cast<float4*>(yp);
      d[0] = make_float4(o[0], o[1], o[2], o[3]);
      d[1] = make_float4(o[4], o[5], o[6], o[7]);
    }
  }
}

// Persistent form of the flat kernel for 2-byte activations: a warp walks groups g, g + stride, ...
// with the NEXT group's five vectors already in flight while it reduces and writes the current
// one (the one-shot kernel serialises load -> two reductions -> parameter loads -> store per warp
// and measured 0.57-0.72 of HBM with parameters against 0.80 without them, scripts/ln_probe.py).
// The host sizes the grid so that stride * ROWS is a multiple of F: the frame of a (lane, i) slot
// is then the same in every group the warp visits, and its position-table offset is computed once.
// Measured (profiles/ln_probe_r02.log, gamma+beta+pe, L0/L1/L2): 178/78/43 us one-shot ->
// 162/74/41 us; holding gamma / beta packed in registers as well spills at 128 registers and was
// 40 % slower, three blocks per SM (80 registers) likewise.
#ifndef LN_PERS_BLOCKS
#define LN_PERS_BLOCKS 2
#endif
// empty asm: the compiler must treat the packed words as new values, so the fp32 copies of one
// pass are not kept alive (40 registers) for the next one but re-derived (one ALU op each)
__device__ __forceinline__ void opaque(uint4& r) {
  asm volatile("" : "+r"(r.x), "+r"(r.y), "+r"(r.z), "+r"(r.w));
}
__device__ __forceinline__ void unpack8(const uint4& r, float* v) {
  float2 e;
  e = unpack_bf16(r.x); v[0] = e.x; v[1] = e.y;
  e = unpack_bf16(r.y); v[2] = e.x; v[3] = e.y;
  e = unpack_bf16(r.z); v[4] = e.x; v[5] = e.y;
  e = unpack_bf16(r.w); v[6] = e.x; v[7] = e.y;
}

template <typename TP, int NVEC>
__global__ void __launch_bounds__(256, LN_PERS_BLOCKS)
layernorm_flat_pers_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                           const TP* __restrict__ gamma, const TP* __restrict__ beta,
                           const float* __restrict__ pe, int64_t rows, int F, float eps) {
  constexpr int ROWS = 160 / NVEC, C = NVEC * 8;
  const int lane = threadIdx.x & 31;
  const int64_t groups = (rows + ROWS - 1) / ROWS;
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  int64_t g = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (g >= groups) return;
  int rr[5], vv[5];
  int po[5];                      // position-table offset of slot i (elements; the table is F * C < 2^31)
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int f = i * 32 + lane;
    rr[i] = f / NVEC;
    vv[i] = f - rr[i] * NVEC;
    po[i] = pe ? (int)((g * ROWS + rr[i]) % F) * C + vv[i] * 8 : 0;
  }
  const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
  uint4 cur[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int64_t row = g * ROWS + rr[i];
    cur[i] = row < rows ? ld_stream16(x + row * C + vv[i] * 8) : zero4;
  }
  for (;;) {
    const int64_t gn = g + stride;
    const bool more = gn < groups;
    uint4 nx[5];
    if (more) {
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const int64_t row = gn * ROWS + rr[i];
        nx[i] = row < rows ? ld_stream16(x + row * C + vv[i] * 8) : zero4;
      }
    }
    float mean[ROWS], rstd[ROWS];
    {
      float part[ROWS];
#pragma unroll
      for (int r = 0; r < ROWS; ++r) part[r] = 0.f;
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        float s8 = 0.f, v[8];
        unpack8(cur[i], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) s8 += v[j];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) part[r] += rr[i] == r ? s8 : 0.f;
      }
#pragma unroll
      for (int r = 0; r < ROWS; ++r) mean[r] = warp_sum(part[r]) * (1.f / (float)C);
    }
    float mu[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      mu[i] = mean[0];
#pragma unroll
      for (int r = 1; r < ROWS; ++r) mu[i] = rr[i] == r ? mean[r] : mu[i];
    }
    {
      float part[ROWS];
#pragma unroll
      for (int r = 0; r < ROWS; ++r) part[r] = 0.f;
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        float s8 = 0.f, v[8];
        opaque(cur[i]);
        unpack8(cur[i], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[j] - mu[i];
          s8 = fmaf(d, d, s8);
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) part[r] += rr[i] == r ? s8 : 0.f;
      }
#pragma unroll
      for (int r = 0; r < ROWS; ++r) rstd[r] = rsqrtf(warp_sum(part[r]) * (1.f / (float)C) + eps);
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const int64_t row = g * ROWS + rr[i];
      if (row >= rows) continue;
      float rs = rstd[0];
#pragma unroll
      for (int r = 1; r < ROWS; ++r) rs = rr[i] == r ? rstd[r] : rs;
      float o[8];
      opaque(cur[i]);
      unpack8(cur[i], o);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (o[j] - mu[i]) * rs;
      if (gamma) {
        float gf[8];
        loadp8<TP>(gamma + vv[i] * 8, gf);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] *= gf[j];
      }
      if (beta) {
        float bf[8];
        loadp8<TP>(beta + vv[i] * 8, bf);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += bf[j];
      }
      if (pe) {
        float pr[8];
        loadp8<float>(pe + po[i], pr);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (float)(__nv_bfloat16)o[j] + pr[j];   // two roundings, as in torch
      }
      uint4 w;
      w.x = pack_bf16(o[0], o[1]); w.y = pack_bf16(o[2], o[3]);
      w.z = pack_bf16(o[4], o[5]); w.w = pack_bf16(o[6], o[7]);
      st_stream16(y + row * C + vv[i] * 8, w);
    }
    if (!more) break;
#pragma unroll
    for (int i = 0; i < 5; ++i) cur[i] = nx[i];
    g = gn;
  }
}

// ---- lane-owned LayerNorm (bf16 activations, C = 64 * KS) ----------------------------------------------------
// The flat kernels above are bound by instruction ISSUE, not by memory: per 16-byte vector they unpack the
// activations three times, fetch and unpack gamma / beta / the position row and run scalar fp32 math -- ~130
// thread instructions per vector, i.e. > 5 warp instructions per clock and SM at the HBM rate against 4 issue
// slots (measured: 0.58-0.66 of HBM with gamma + beta + positions, 0.81-0.86 with none, profiles/ln_probe_r02.log).
// Here a lane owns the SAME channel pairs {64 k + 2 lane, +1} of every row it touches and a warp only visits rows
// of ONE frame (row = position * F + frame), so gamma, beta and the position row are loaded once per kernel and
// stay in registers: nothing but activations moves in the loop.  A row is KS coalesced 128-byte warp loads of one
// packed bf16 pair per lane; the arithmetic runs on packed f32x2 instructions (FADD2 / FMUL2 / FFMA2).  R rows are
// processed per trip with the rows of the next TWO trips in flight (register ring, loop unrolled by three so the
// ring never moves).  Numerics as the flat kernels: fp32 two-pass statistics, ((x - mean) * rstd) * gamma + beta
// rounded to bf16, then + position row in fp32 and rounded again (the two roundings of the torch module).
__device__ __forceinline__ uint32_t ld_stream4(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream4(void* p, uint32_t v) {
  asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
template <typename TP>
__device__ __forceinline__ float2 ldparam2(const TP* p);
template <>
__device__ __forceinline__ float2 ldparam2<__nv_bfloat16>(const __nv_bfloat16* p) {
  return unpack_bf16(__ldg(reinterpret_cast<const uint32_t*>(p)));
}
template <>
__device__ __forceinline__ float2 ldparam2<float>(const float* p) {
  return __ldg(reinterpret_cast<const float2*>(p));
}

template <typename TP, int KS, int R>
__global__ void __launch_bounds__(256, 2)
layernorm_lane_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                      const TP* __restrict__ gamma, const TP* __restrict__ beta,
                      const float* __restrict__ pe, int rows, int Fq, float eps) {
  // 32-bit indices throughout (the host takes this kernel only when rows * C < 2^31)
  constexpr int C = 64 * KS;
  constexpr float kInvC = 1.f / (float)C;
  const int lane = threadIdx.x & 31;
  const int wg = blockIdx.x * 8 + (threadIdx.x >> 5), W = gridDim.x * 8;
  const int f = wg % Fq;                                  // the one frame this warp visits
  const int q = wg / Fq, Q = W / Fq;                      // position block of this warp, position blocks in flight
  const int npos = (rows + Fq - 1) / Fq;
  const int n_trips = (npos + Q * R - 1) / (Q * R);
  const int c0 = 2 * lane;

  float2 gm[KS], bt[KS], pp[KS];
#pragma unroll
  for (int k = 0; k < KS; ++k) {
    gm[k] = gamma ? ldparam2<TP>(gamma + 64 * k + c0) : make_float2(1.f, 1.f);
    bt[k] = beta ? ldparam2<TP>(beta + 64 * k + c0) : make_float2(0.f, 0.f);
    pp[k] = pe ? __ldg(reinterpret_cast<const float2*>(pe + (int64_t)f * C + 64 * k + c0)) : make_float2(0.f, 0.f);
  }

  auto row_of = [&](int t, int j) -> int {                // -1: no such row
    const int pos = (t * Q + q) * R + j;
    const int row = pos * Fq + f;
    return (pos < npos && row < rows) ? row : -1;
  };
  auto load = [&](uint32_t (&buf)[R][KS], int t) {
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int row = row_of(t, j);
#pragma unroll
      for (int k = 0; k < KS; ++k) buf[j][k] = row >= 0 ? ld_stream4(x + row * C + 64 * k + c0) : 0u;
    }
  };
  auto process = [&](uint32_t (&buf)[R][KS], int t) {
    float mean[R], rstd[R];
    {
      float part[R];
#pragma unroll
      for (int j = 0; j < R; ++j) {
        float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < KS; ++k) s2 = __fadd2_rn(s2, unpack_bf16(buf[j][k]));
        part[j] = s2.x + s2.y;
      }
#pragma unroll
      for (int j = 0; j < R; ++j) mean[j] = warp_sum(part[j]) * kInvC;
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const float2 nm = make_float2(-mean[j], -mean[j]);
        float2 m2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < KS; ++k) {
          const float2 d = __fadd2_rn(unpack_bf16(buf[j][k]), nm);
          m2 = __ffma2_rn(d, d, m2);
        }
        part[j] = m2.x + m2.y;
      }
#pragma unroll
      for (int j = 0; j < R; ++j) rstd[j] = rsqrtf(warp_sum(part[j]) * kInvC + eps);
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int row = row_of(t, j);
      if (row < 0) continue;
      const float2 nm = make_float2(-mean[j], -mean[j]), rs = make_float2(rstd[j], rstd[j]);
      __nv_bfloat16* yr = y + row * C + c0;
#pragma unroll
      for (int k = 0; k < KS; ++k) {
        const float2 d = __fmul2_rn(__fadd2_rn(unpack_bf16(buf[j][k]), nm), rs);
        const float2 o = __ffma2_rn(d, gm[k], bt[k]);
        uint32_t w = pack_bf16(o.x, o.y);
        if (pe) {
          const float2 e = __fadd2_rn(unpack_bf16(w), pp[k]);
          w = pack_bf16(e.x, e.y);
        }
        st_stream4(yr + 64 * k, w);
      }
    }
  };

  uint32_t b0[R][KS], b1[R][KS], b2[R][KS];
  if (n_trips > 0) load(b0, 0);
  if (n_trips > 1) load(b1, 1);
  for (int t = 0; t < n_trips; t += 3) {
    if (t + 2 < n_trips) load(b2, t + 2);
    process(b0, t);
    if (t + 1 >= n_trips) break;
    if (t + 3 < n_trips) load(b0, t + 3);
    process(b1, t + 1);
    if (t + 2 >= n_trips) break;
    if (t + 4 < n_trips) load(b1, t + 4);
    process(b2, t + 2);
  }
}

inline int ln_lane_mode() {
  static const int mode = [] {
    const char* e = getenv("VST_LN_LANE");
    return e ? atoi(e) : 1;
  }();
  return mode;
}

// The lane-owned kernel when the shape is one of its instantiations and there is work for two trips of every
// resident warp; otherwise the caller goes on to the flat kernels (kLnNotTaken).
constexpr int kLnNotTaken = 1;
template <typename TP, int KS, int R>
int try_launch_ln_lane(const void* x, void* y, const void* g, const void* b, const float* pe, int64_t rows,
                       int F, float eps, cudaStream_t st) {
  if (!ln_lane_mode()) return kLnNotTaken;
  const int Fq = pe ? F : 1;
  if (Fq <= 0) return kLnNotTaken;
  int a = Fq, c = 8;
  while (c) { const int t = a % c; a = c; c = t; }
  const int mult = Fq / a;                              // blocks * 8 warps must be a multiple of Fq
  int64_t blocks = (int64_t)vst_sm_count() * 2;
  blocks -= blocks % mult;
  if (blocks < mult || rows < blocks * 8 * R * 2 || (rows + 2 * Fq) * (64 * KS) >= (1ll << 31)) return kLnNotTaken;
  layernorm_lane_kernel<TP, KS, R><<<(unsigned)blocks, 256, 0, st>>>(
      (const __nv_bfloat16*)x, (__nv_bfloat16*)y, (const TP*)g, (const TP*)b, pe, (int)rows, Fq, eps);
  VST_LAUNCH_OK();
  return VST_OK;
}

inline int ln_pers_mode() {
  static const int mode = [] {
    const char* e = getenv("VST_LN_PERS");
    return e ? atoi(e) : 1;
  }();
  return mode;
}

template <typename TI, typename TO, typename TP, int NVEC>
int launch_ln_flat(const void* x, void* y, const void* g, const void* b, const float* pe, int64_t rows,
                   int F, float eps, cudaStream_t st) {
  constexpr int ROWS = 160 / NVEC;
  if constexpr (sizeof(TI) == 2 && sizeof(TO) == 2) {
    // VST_LN_LANE: 0 = flat kernels only; 1 (default) = 3 rows per trip at C = 320 (no spills at 128 registers;
    // L0 with gamma + beta + positions: 124 us = 0.83 of HBM against 162 us flat); 2 / 4 = 2 / 4 rows per trip
    // (tuning builds: 136 us, and 158 us with 92 bytes of spills)
    if constexpr (NVEC == 40) {
      const int mode = ln_lane_mode();
      const int rc = mode == 2 ? try_launch_ln_lane<TP, 5, 2>(x, y, g, b, pe, rows, F, eps, st)
                   : mode == 4 ? try_launch_ln_lane<TP, 5, 4>(x, y, g, b, pe, rows, F, eps, st)
                               : try_launch_ln_lane<TP, 5, 3>(x, y, g, b, pe, rows, F, eps, st);
      if (rc != kLnNotTaken) return rc;
    }
    if constexpr (NVEC == 80 && sizeof(TP) == 2) {
      // C = 640: the parameters fill 60 registers, one row per trip (two rows spill and measured 80 vs 67 us at
      // L1; gamma / beta packed and unpacked per use likewise); without a position table the flat kernel is the
      // faster one at this width (61 vs 65 us)
      const int rc = pe == nullptr ? kLnNotTaken : try_launch_ln_lane<TP, 10, 1>(x, y, g, b, pe, rows, F, eps, st);
      if (rc != kLnNotTaken) return rc;
    }
    const int64_t groups = (rows + ROWS - 1) / ROWS;
    // blocks of 8 warps; the warp count must be a multiple of F / gcd(F, ROWS) groups (see the kernel)
    int per = 1;
    if (pe) {
      int a = F, c = ROWS;
      while (c) { const int t = a % c; a = c; c = t; }
      per = F / a;
    }
    int a = per, c = 8;
    while (c) { const int t = a % c; a = c; c = t; }
    const int mult = per / a;                       // blocks must be a multiple of this
    int64_t want = (int64_t)vst_sm_count() * LN_PERS_BLOCKS;
    want -= want % mult;
    if (ln_pers_mode() && want >= mult && groups >= want * 8 * 2) {
      layernorm_flat_pers_kernel<TP, NVEC><<<(unsigned)want, 256, 0, st>>>(
          (const __nv_bfloat16*)x, (__nv_bfloat16*)y, (const TP*)g, (const TP*)b, pe, rows, F, eps);
      VST_LAUNCH_OK();
      return VST_OK;
    }
  }
  const int wpb = 8;
  const int64_t grid = (rows + (int64_t)wpb * ROWS - 1) / ((int64_t)wpb * ROWS);
  VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "layernorm: too many rows");
  layernorm_flat_kernel<TI, TO, TP, NVEC><<<(unsigned)grid, wpb * 32, 0, st>>>(
      (const TI*)x, (TO*)y, (const TP*)g, (const TP*)b, pe, rows, F, eps);
  VST_LAUNCH_OK();
  return VST_OK;
}

template <typename TI, typename TO, typename TP, int MAXV, int ROWS>
int launch_ln_cfg(const void* x, void* y, const void* g, const void* b, const float* pe, int64_t rows,
                  int C, int F, float eps, cudaStream_t st) {
  const int wpb = 8;
  const int64_t grid = (rows + (int64_t)wpb * ROWS - 1) / ((int64_t)wpb * ROWS);
  VST_REQUIRE(grid < (1ll << 31), VST_E_UNSUPPORTED, "layernorm: too many rows");
  layernorm_kernel<TI, TO, TP, MAXV, ROWS><<<(unsigned)grid, wpb * 32, 0, st>>>(
      (const TI*)x, (TO*)y, (const TP*)g, (const TP*)b, pe, rows, C, F, eps);
  VST_LAUNCH_OK();
  return VST_OK;
}

template <typename TI, typename TO, typename TP>
int launch_ln(const void* x, void* y, const void* g, const void* b, const float* pe, int64_t rows,
              int C, int F, float eps, cudaStream_t st) {
  const int nvec = C / 8;
  if (nvec == 40) return launch_ln_flat<TI, TO, TP, 40>(x, y, g, b, pe, rows, F, eps, st);
  if (nvec == 80) return launch_ln_flat<TI, TO, TP, 80>(x, y, g, b, pe, rows, F, eps, st);
  if (nvec == 160) return launch_ln_flat<TI, TO, TP, 160>(x, y, g, b, pe, rows, F, eps, st);
  if (nvec <= 64) return launch_ln_cfg<TI, TO, TP, 2, 4>(x, y, g, b, pe, rows, C, F, eps, st);
  if (nvec <= 96) return launch_ln_cfg<TI, TO, TP, 3, 2>(x, y, g, b, pe, rows, C, F, eps, st);
  if (nvec <= 160) return launch_ln_cfg<TI, TO, TP, 5, 1>(x, y, g, b, pe, rows, C, F, eps, st);
  return launch_ln_cfg<TI, TO, TP, 8, 1>(x, y, g, b, pe, rows, C, F, eps, st);
}

}  // namespace

extern "C" int vst_groupnorm5d_partials(int C, int F, int HW, int groups) {
  if (groups <= 0 || C % groups != 0) return VST_E_BADARG;
  const int64_t chunk = (int64_t)(C / groups) * HW;
  return (int)(F * ((chunk + kSlabElems - 1) / kSlabElems));
}

extern "C" int vst_groupnorm5d_stats(const void* x, float* partials, int B, int C, int F, int HW,
                                     int64_t sB, int64_t sC, int64_t sF, int groups, int dtype,
                                     void* stream) {
  VST_REQUIRE(groups > 0 && C % groups == 0, VST_E_BADARG, "groupnorm: C %% groups != 0");
  VST_REQUIRE(x && partials, VST_E_BADARG, "groupnorm: null pointer");
  VST_REQUIRE(dtype == VST_F32 || dtype == VST_BF16, VST_E_BADARG, "groupnorm: bad dtype");
  VST_REQUIRE((int64_t)(C / groups) * HW < (1ll << 31) - kSlabElems, VST_E_UNSUPPORTED,
              "groupnorm: a (group, frame) chunk must stay below 2^31 elements");
  const int slabs = vst_groupnorm5d_partials(C, F, HW, groups) / F;
  const int64_t grid = (int64_t)B * groups * F * slabs;
  VST_REQUIRE(grid > 0 && grid < (1ll << 31), VST_E_UNSUPPORTED, "groupnorm: bad grid");
  const bool vec = HW % 8 == 0 && sB % 8 == 0 && sC % 8 == 0 && sF % 8 == 0 &&
                   (reinterpret_cast<uintptr_t>(x) & 31) == 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t resident = (int64_t)vst_sm_count() * 4;     // 4 x 256 threads per SM (64 registers)
  if (vec) {
    // blocks of one (b, group) walk its F * slabs items; as many blocks per (b, group) as stay resident
    const int64_t bgs = (int64_t)B * groups, per_bg = (int64_t)F * slabs;
    int64_t bpb = resident / bgs;
    if (bpb < 1) bpb = 1;
    if (bpb > per_bg) bpb = per_bg;
    int hw_shift = -1;
    if ((HW & (HW - 1)) == 0) { hw_shift = 0; while ((1 << hw_shift) < HW) ++hw_shift; }
    const int cg = C / groups;
    const int contig = sC == HW;
    const unsigned vgrid = (unsigned)(bgs * bpb);
    if (dtype == VST_BF16)
      gn_stats_vec_kernel<__nv_bfloat16><<<vgrid, kSlabThreads, 0, st>>>(
          (const __nv_bfloat16*)x, partials, F, HW, hw_shift, contig, sB, sC, sF, groups, cg, slabs, (int)bpb);
    else
      gn_stats_vec_kernel<float><<<vgrid, kSlabThreads, 0, st>>>(
          (const float*)x, partials, F, HW, hw_shift, contig, sB, sC, sF, groups, cg, slabs, (int)bpb);
  } else {
    const unsigned pgrid = (unsigned)(grid < resident ? grid : resident);
    if (dtype == VST_BF16)
      gn_stats_kernel<__nv_bfloat16><<<pgrid, kSlabThreads, 0, st>>>((const __nv_bfloat16*)x, partials, C, F, HW,
                                                                      sB, sC, sF, groups, slabs, (int)grid);
    else
      gn_stats_kernel<float><<<pgrid, kSlabThreads, 0, st>>>((const float*)x, partials, C, F, HW, sB, sC, sF,
                                                             groups, slabs, (int)grid);
  }
  VST_LAUNCH_OK();
  return VST_OK;
}

extern "C" int vst_groupnorm5d_finalize(const float* partials, int nparts, int B, int groups,
                                        float eps, float* mean, float* rstd, void* stream) {
  VST_REQUIRE(partials && mean && rstd && nparts > 0 && B > 0 && groups > 0, VST_E_BADARG,
              "groupnorm finalize: bad argument");
  gn_finalize_kernel<<<B * groups, 32, 0, (cudaStream_t)stream>>>(partials, nparts, eps, mean, rstd);
  VST_LAUNCH_OK();
  return VST_OK;
}

extern "C" int vst_layernorm(const void* x, void* y, const void* gamma, const void* beta,
                             const float* pe, int64_t rows, int C, int F, float eps, int in_dtype,
                             int out_dtype, int param_dtype, void* stream) {
  VST_REQUIRE(rows >= 0 && C > 0, VST_E_BADARG, "layernorm: bad shape");
  if (rows == 0) return VST_OK;
  VST_REQUIRE(x && y, VST_E_BADARG, "layernorm: null pointer");
  VST_REQUIRE(C % 8 == 0 && C <= 2048, VST_E_UNSUPPORTED, "layernorm: C must be a multiple of 8, <= 2048");
  VST_REQUIRE(pe == nullptr || F > 0, VST_E_BADARG, "layernorm: F must be > 0 with pe");
  cudaStream_t st = (cudaStream_t)stream;
  const int key = in_dtype * 4 + out_dtype * 2 + param_dtype;
  switch (key) {
    case 0: return launch_ln<float, float, float>(x, y, gamma, beta, pe, rows, C, F, eps, st);
    case 2: return launch_ln<float, __nv_bfloat16, float>(x, y, gamma, beta, pe, rows, C, F, eps, st);
    case 3: return launch_ln<float, __nv_bfloat16, __nv_bfloat16>(x, y, gamma, beta, pe, rows, C, F, eps, st);
    case 6: return launch_ln<__nv_bfloat16, __nv_bfloat16, float>(x, y, gamma, beta, pe, rows, C, F, eps, st);
    case 7: return launch_ln<__nv_bfloat16, __nv_bfloat16, __nv_bfloat16>(x, y, gamma, beta, pe, rows, C, F, eps, st);
    default:
      vst_set_error("layernorm: unsupported dtype combination in=%d out=%d param=%d", in_dtype,
                    out_dtype, param_dtype);
      return VST_E_UNSUPPORTED;
  }
}
