// Backward kernels of the memory-bound motion-module pieces (config 4, train_animatediff.py): the
// LayerNorms (K9), the GEGLU gate (§8f-1) and the GroupNorm in front of the motion module (a10).
// The reference leaves all of them to autograd (diffusers BasicTransformerBlock /
// TransformerTemporalModel, un-vendored; in-tree restatement animatediff/temporal_transformer.py
// :44-59,69-75).  Everything here is one or two passes over the activation: HBM-bound, 16-byte
// accesses, fp32 statistics, parameter gradients summed in a fixed order (no atomics).
#include "common.cuh"

namespace {

__device__ __forceinline__ void load8bf(const __nv_bfloat16* p, float* v) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 e = unpack_bf16(w[i]);
    v[2 * i] = e.x;
    v[2 * i + 1] = e.y;
  }
}
__device__ __forceinline__ void store8bf(__nv_bfloat16* p, const float* v) {
  uint4 w;
  w.x = pack_bf16(v[0], v[1]); w.y = pack_bf16(v[2], v[3]);
  w.z = pack_bf16(v[4], v[5]); w.w = pack_bf16(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = w;
}
template <typename TP>
__device__ __forceinline__ float ldparam(const TP* p, int i) { return p ? (float)p[i] : 1.f; }
// eight consecutive parameters (16-byte aligned offset) in one or two 16-byte loads
__device__ __forceinline__ void ldparam8(const float* p, int off, float* o) {
  if (p == nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = 1.f;
    return;
  }
  const float4 a = __ldg(reinterpret_cast<const float4*>(p + off)), b = __ldg(reinterpret_cast<const float4*>(p + off) + 1);
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
__device__ __forceinline__ void ldparam8(const __nv_bfloat16* p, int off, float* o) {
  if (p == nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = 1.f;
    return;
  }
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(p + off));
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 e = unpack_bf16(w[i]);
    o[2 * i] = e.x;
    o[2 * i + 1] = e.y;
  }
}

// ---- LayerNorm backward -----------------------------------------------------------------------
// y = xhat * gamma + beta (+ pe: additive, no gradient).  Per row, with g = dy * gamma:
//   dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  dgamma += dy * xhat,  dbeta += dy.
// A warp owns a row (held in registers, statistics recomputed); each warp keeps running dgamma /
// dbeta for its columns over all the rows it visits; the 8 warps of a block are summed through
// shared memory and every block writes one partial row, reduced by ln_bwd_reduce_kernel.
constexpr int kLnWarps = 8;
template <typename TP, int MAXV>
__global__ void __launch_bounds__(kLnWarps * 32)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
              const TP* __restrict__ gamma, __nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
              int64_t rows, int C, float eps) {
  extern __shared__ float red[];   // [kLnWarps][2][C]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = C >> 3;
  float dgam[MAXV][8], dbet[MAXV][8];
#pragma unroll
  for (int i = 0; i < MAXV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dgam[i][j] = dbet[i][j] = 0.f;
  const float invC = 1.f / (float)C;
  // the NEXT row's 16-byte loads are in flight (raw, in registers) while this row goes through its three warp
  // reductions: one resident block of 8 warps per SM has nothing else to hide the load latency with
  const int64_t rstep = (int64_t)gridDim.x * kLnWarps;
  int64_t row = (int64_t)blockIdx.x * kLnWarps + warp;
  uint4 nx[MAXV], ng[MAXV];
  auto fetch = [&](int64_t r) {
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        nx[i] = ld_stream16(x + r * C + vi * 8);
        ng[i] = ld_stream16(dy + r * C + vi * 8);
      }
    }
  };
  auto unpack = [](const uint4& u, float* o) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 e = unpack_bf16(w[q]);
      o[2 * q] = e.x;
      o[2 * q + 1] = e.y;
    }
  };
  if (row < rows) fetch(row);
  for (; row < rows; row += rstep) {
    float v[MAXV][8], g[MAXV][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        unpack(nx[i], v[i]);
        unpack(ng[i], g[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += v[i][j];
      }
    }
    if (row + rstep < rows) fetch(row + rstep);
    const float mean = warp_sum(sum) * invC;
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i)
      if (lane + 32 * i < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[i][j] - mean;
          m2 = fmaf(d, d, m2);
        }
      }
    const float rstd = rsqrtf(warp_sum(m2) * invC + eps);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float gam8[8];
        ldparam8(gamma, vi * 8, gam8);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (v[i][j] - mean) * rstd;
          const float d = g[i][j];
          dgam[i][j] = fmaf(d, xh, dgam[i][j]);
          dbet[i][j] += d;
          const float gg = d * gam8[j];
          v[i][j] = xh;
          g[i][j] = gg;
          s1 += gg;
          s2 = fmaf(gg, xh, s2);
        }
      }
    }
    s1 = warp_sum(s1) * invC;
    s2 = warp_sum(s2) * invC;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - s1 - v[i][j] * s2);
        store8bf(dx + row * C + vi * 8, o);
      }
    }
  }
  // block partial: fixed-order sum over the 8 warps
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int vi = lane + 32 * i;
    if (vi < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        red[(warp * 2 + 0) * C + vi * 8 + j] = dgam[i][j];
        red[(warp * 2 + 1) * C + vi * 8 + j] = dbet[i][j];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    const int which = c / C, col = c % C;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kLnWarps; ++w) s += red[(w * 2 + which) * C + col];
    partial[((int64_t)blockIdx.x * 2 + which) * C + col] = s;
  }
}
__global__ void ln_bwd_reduce_kernel(const float* __restrict__ partial, int blocks, int C,
                                     float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 2 * C) return;
  const int which = c / C, col = c % C;
  float s = 0.f;
  for (int b = 0; b < blocks; ++b) s += partial[((int64_t)b * 2 + which) * C + col];
  float* out = which == 0 ? dgamma : dbeta;
  if (out) out[col] = s;
}
// Blocks = what is resident at once: the C > 512 instantiation holds 160 accumulators + 80 row values per lane
// (191 registers) and 8 x 2 x C floats of shared memory, ONE block per SM; four waves of such blocks (r01: a cap of
// 4 x SMs) left each block 28 rows to amortise its 80 KB cross-warp reduction over (S = 1024 level: 126 us, 0.15 of
// HBM).
int ln_bwd_blocks(int64_t rows, int C) {
  int64_t b = (rows + kLnWarps - 1) / kLnWarps;
  const int64_t cap = (C <= 512 ? 2 : 1) * (int64_t)vst_sm_count();
  return (int)(b < cap ? b : cap);
}

// ---- GEGLU backward: out = a * gelu(b), a = hg[:, :inner], b = hg[:, inner:] ------------------
__global__ void geglu_bwd_kernel(const __nv_bfloat16* __restrict__ hg, const __nv_bfloat16* __restrict__ dout,
                                 __nv_bfloat16* __restrict__ dhg, int64_t rows, int inner, int64_t ld_in,
                                 int64_t ld_do, int64_t ld_out) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int per_row = inner >> 3;
  if (gid >= rows * per_row) return;
  const int64_t row = gid / per_row;
  const int c = (int)(gid % per_row) * 8;
  float a[8], b[8], d[8], da[8], db[8];
  load8bf(hg + row * ld_in + c, a);
  load8bf(hg + row * ld_in + inner + c, b);
  load8bf(dout + row * ld_do + c, d);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float cdf = 0.5f * (1.f + erff(b[j] * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * __expf(-0.5f * b[j] * b[j]);
    da[j] = d[j] * b[j] * cdf;
    db[j] = d[j] * a[j] * (cdf + b[j] * pdf);
  }
  store8bf(dhg + row * ld_out + c, da);
  store8bf(dhg + row * ld_out + inner + c, db);
}

// dz = alpha * dy * gelu'(z) (exact erf GELU; temporal_transformer.py:53-55 feed-forward)
__global__ void gelu_bwd_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ dy,
                                __nv_bfloat16* __restrict__ dz, int64_t rows, int N, int64_t ld_z,
                                int64_t ld_dy, int64_t ld_dz, float alpha) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int per_row = N >> 3;
  if (gid >= rows * per_row) return;
  const int64_t row = gid / per_row;
  const int c = (int)(gid % per_row) * 8;
  float b[8], d[8], o[8];
  load8bf(z + row * ld_z + c, b);
  load8bf(dy + row * ld_dy + c, d);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float cdf = 0.5f * (1.f + erff(b[j] * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * __expf(-0.5f * b[j] * b[j]);
    o[j] = alpha * d[j] * (cdf + b[j] * pdf);
  }
  store8bf(dz + row * ld_dz + c, o);
}

// ---- GroupNorm5D backward (frames-major x and dy: element (b, f, c, hw)) ----------------------------
// plane sums: for every (b, f, c) plane of HW contiguous elements, (sum dy, sum dy*x).
__global__ void gn_bwd_plane_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                    float* __restrict__ plane, int HW) {
  const int64_t p = blockIdx.x;   // plane index (b*F + f)*C + c
  const __nv_bfloat16* xp = x + p * HW;
  const __nv_bfloat16* dp = dy + p * HW;
  float s0 = 0.f, s1 = 0.f;
  for (int i = threadIdx.x * 8; i < HW; i += blockDim.x * 8) {
    float a[8], g[8];
    load8bf(xp + i, a);
    load8bf(dp + i, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s0 += g[j];
      s1 = fmaf(g[j], a[j], s1);
    }
  }
  __shared__ float r0[8], r1[8];
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  if ((threadIdx.x & 31) == 0) { r0[threadIdx.x >> 5] = s0; r1[threadIdx.x >> 5] = s1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t0 = 0.f, t1 = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { t0 += r0[w]; t1 += r1[w]; }
    plane[2 * p] = t0;
    plane[2 * p + 1] = t1;
  }
}
// one block per group g: channel sums over (b, f), dgamma / dbeta, and the per-(b, g) coefficients
//   c1 = mean(dy * gamma), c2 = mean(dy * gamma * xhat) over the group.
template <typename TP>
__global__ void gn_bwd_group_kernel(const float* __restrict__ plane, const float* __restrict__ mean,
                                    const float* __restrict__ rstd, const TP* __restrict__ gamma,
                                    float* __restrict__ dgamma, float* __restrict__ dbeta,
                                    float* __restrict__ coef, int B, int C, int F, int HW, int groups) {
  const int g = blockIdx.x, cpg = C / groups;
  __shared__ float sA[64], sB[64];   // per channel of the group (cpg <= 64)
  const float n = (float)cpg * F * HW;
  for (int b = 0; b < B; ++b) {
    const float mu = mean[b * groups + g], rs = rstd[b * groups + g];
    if ((int)threadIdx.x < cpg) {
      const int c = g * cpg + threadIdx.x;
      float a = 0.f, bx = 0.f;
      for (int f = 0; f < F; ++f) {
        const int64_t p = ((int64_t)(b * F + f) * C + c);
        a += plane[2 * p];
        bx += plane[2 * p + 1];
      }
      const float bh = rs * (bx - mu * a);   // sum dy * xhat over the channel
      sA[threadIdx.x] = a;
      sB[threadIdx.x] = bh;
      if (b == 0) {
        if (dbeta) dbeta[c] = a;
        if (dgamma) dgamma[c] = bh;
      } else {
        if (dbeta) dbeta[c] += a;
        if (dgamma) dgamma[c] += bh;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      float s1 = 0.f, s2 = 0.f;
      for (int i = 0; i < cpg; ++i) {
        const float gm = ldparam(gamma, g * cpg + i);
        s1 = fmaf(gm, sA[i], s1);
        s2 = fmaf(gm, sB[i], s2);
      }
      coef[2 * (b * groups + g)] = s1 / n;
      coef[2 * (b * groups + g) + 1] = s2 / n;
    }
    __syncthreads();
  }
}
// dx = rstd * (dy * gamma - c1 - xhat * c2) (+ add)
template <typename TP>
__global__ void gn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                    const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ dx,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ coef, const TP* __restrict__ gamma,
                                    int64_t total8, int C, int F, int HW, int groups) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total8) return;
  const int64_t e = gid * 8;
  const int64_t plane = e / HW;            // (b*F + f)*C + c
  const int c = (int)(plane % C);
  const int b = (int)(plane / ((int64_t)C * F));
  const int bg = b * groups + c / (C / groups);
  const float mu = mean[bg], rs = rstd[bg], c1 = coef[2 * bg], c2 = coef[2 * bg + 1];
  const float gm = ldparam(gamma, c);
  float a[8], g[8], o[8];
  load8bf(x + e, a);
  load8bf(dy + e, g);
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j] = rs * (g[j] * gm - c1 - (a[j] - mu) * rs * c2);
  if (add) {
    float r[8];
    load8bf(add + e, r);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] += r[j];
  }
  store8bf(dx + e, o);
}

}  // namespace

extern "C" int64_t vst_layernorm_bwd_workspace_bytes(int64_t rows, int C) {
  if (rows <= 0 || C <= 0) return 0;
  return (int64_t)ln_bwd_blocks(rows, C) * 2 * C * 4;
}

extern "C" int vst_layernorm_bwd(const void* x, const void* dy, const void* gamma, void* dx,
                                 float* dgamma, float* dbeta, int64_t rows, int C, float eps,
                                 int param_dtype, void* workspace, int64_t workspace_bytes,
                                 void* stream) {
  VST_REQUIRE(rows >= 0 && C > 0, VST_E_BADARG, "layernorm_bwd: bad shape");
  if (rows == 0) return VST_OK;
  VST_REQUIRE(x && dy && dx, VST_E_BADARG, "layernorm_bwd: null pointer");
  VST_REQUIRE(C % 8 == 0 && C <= 1280, VST_E_UNSUPPORTED,
              "layernorm_bwd: C must be a multiple of 8, <= 1280 (motion-module widths)");
  const int blocks = ln_bwd_blocks(rows, C);
  VST_REQUIRE(workspace && workspace_bytes >= (int64_t)blocks * 2 * C * 4, VST_E_BADARG,
              "layernorm_bwd: workspace of vst_layernorm_bwd_workspace_bytes() bytes required");
  cudaStream_t st = (cudaStream_t)stream;
  const int smem = kLnWarps * 2 * C * 4;
  const int nvec = C / 8;
#define LN_BWD(TP, MV)                                                                              \
  do {                                                                                              \
    auto kern = ln_bwd_kernel<TP, MV>;                                                              \
    VST_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));     \
    kern<<<blocks, kLnWarps * 32, smem, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,    \
                                              (const TP*)gamma, (__nv_bfloat16*)dx,                 \
                                              (float*)workspace, rows, C, eps);                     \
  } while (0)
  if (param_dtype == VST_BF16) {
    if (nvec <= 64) LN_BWD(__nv_bfloat16, 2); else if (nvec <= 96) LN_BWD(__nv_bfloat16, 3); else LN_BWD(__nv_bfloat16, 5);
  } else if (param_dtype == VST_F32) {
    if (nvec <= 64) LN_BWD(float, 2); else if (nvec <= 96) LN_BWD(float, 3); else LN_BWD(float, 5);
  } else {
    vst_set_error("layernorm_bwd: bad param dtype %d", param_dtype);
    return VST_E_BADARG;
  }
#undef LN_BWD
  VST_LAUNCH_OK();
  if (dgamma || dbeta) {
    ln_bwd_reduce_kernel<<<(2 * C + 255) / 256, 256, 0, st>>>((const float*)workspace, blocks, C, dgamma, dbeta);
    VST_LAUNCH_OK();
  }
  return VST_OK;
}

extern "C" int vst_geglu_bwd(const void* hg, const void* dout, void* dhg, int64_t rows, int inner,
                             int64_t ld_in, int64_t ld_do, int64_t ld_out, void* stream) {
  VST_REQUIRE(rows >= 0 && inner > 0, VST_E_BADARG, "geglu_bwd: bad shape");
  if (rows == 0) return VST_OK;
  VST_REQUIRE(hg && dout && dhg, VST_E_BADARG, "geglu_bwd: null pointer");
  VST_REQUIRE(inner % 8 == 0 && ld_in % 8 == 0 && ld_do % 8 == 0 && ld_out % 8 == 0, VST_E_UNSUPPORTED,
              "geglu_bwd: inner and row strides must be multiples of 8");
  const int64_t threads = rows * (inner / 8);
  geglu_bwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)hg, (const __nv_bfloat16*)dout, (__nv_bfloat16*)dhg, rows, inner, ld_in,
      ld_do, ld_out);
  VST_LAUNCH_OK();
  return VST_OK;
}

extern "C" int vst_gelu_bwd(const void* z, const void* dy, void* dz, int64_t rows, int N, int64_t ld_z,
                            int64_t ld_dy, int64_t ld_dz, float alpha, void* stream) {
  VST_REQUIRE(rows >= 0 && N > 0, VST_E_BADARG, "gelu_bwd: bad shape");
  if (rows == 0) return VST_OK;
  VST_REQUIRE(z && dy && dz, VST_E_BADARG, "gelu_bwd: null pointer");
  VST_REQUIRE(N % 8 == 0 && ld_z % 8 == 0 && ld_dy % 8 == 0 && ld_dz % 8 == 0, VST_E_UNSUPPORTED,
              "gelu_bwd: N and row strides must be multiples of 8");
  const int64_t threads = rows * (N / 8);
  gelu_bwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)z, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dz, rows, N, ld_z, ld_dy, ld_dz, alpha);
  VST_LAUNCH_OK();
  return VST_OK;
}

extern "C" int64_t vst_groupnorm5d_bwd_workspace_bytes(int B, int C, int F, int groups) {
  if (B <= 0 || C <= 0 || F <= 0 || groups <= 0) return 0;
  return ((int64_t)B * F * C * 2 + (int64_t)B * groups * 2) * 4;
}

extern "C" int vst_groupnorm5d_bwd(const void* x, const void* dy, const void* add, void* dx,
                                   const float* mean, const float* rstd, const void* gamma,
                                   float* dgamma, float* dbeta, int B, int C, int F, int HW,
                                   int groups, int param_dtype, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
  VST_REQUIRE(B > 0 && C > 0 && F > 0 && HW > 0 && groups > 0 && C % groups == 0, VST_E_BADARG,
              "groupnorm5d_bwd: bad shape");
  VST_REQUIRE(x && dy && dx && mean && rstd, VST_E_BADARG, "groupnorm5d_bwd: null pointer");
  VST_REQUIRE(HW % 8 == 0 && C / groups <= 64, VST_E_UNSUPPORTED,
              "groupnorm5d_bwd: HW must be a multiple of 8 and C/groups <= 64");
  VST_REQUIRE(workspace && workspace_bytes >= vst_groupnorm5d_bwd_workspace_bytes(B, C, F, groups),
              VST_E_BADARG, "groupnorm5d_bwd: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  float* plane = (float*)workspace;
  float* coef = plane + (int64_t)B * F * C * 2;
  const int64_t planes = (int64_t)B * F * C;
  VST_REQUIRE(planes < (1ll << 31), VST_E_UNSUPPORTED, "groupnorm5d_bwd: too many planes");
  gn_bwd_plane_kernel<<<(unsigned)planes, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, plane, HW);
  VST_LAUNCH_OK();
  const int64_t total8 = planes * HW / 8;
  const unsigned grid = (unsigned)((total8 + 255) / 256);
  if (param_dtype == VST_BF16) {
    gn_bwd_group_kernel<__nv_bfloat16><<<groups, 64, 0, st>>>(plane, mean, rstd, (const __nv_bfloat16*)gamma,
                                                              dgamma, dbeta, coef, B, C, F, HW, groups);
    VST_LAUNCH_OK();
    gn_bwd_apply_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)add, (__nv_bfloat16*)dx,
        mean, rstd, coef, (const __nv_bfloat16*)gamma, total8, C, F, HW, groups);
  } else if (param_dtype == VST_F32) {
    gn_bwd_group_kernel<float><<<groups, 64, 0, st>>>(plane, mean, rstd, (const float*)gamma, dgamma, dbeta,
                                                      coef, B, C, F, HW, groups);
    VST_LAUNCH_OK();
    gn_bwd_apply_kernel<float><<<grid, 256, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)add, (__nv_bfloat16*)dx,
        mean, rstd, coef, (const float*)gamma, total8, C, F, HW, groups);
  } else {
    vst_set_error("groupnorm5d_bwd: bad param dtype %d", param_dtype);
    return VST_E_BADARG;
  }
  VST_LAUNCH_OK();
  return VST_OK;
}
