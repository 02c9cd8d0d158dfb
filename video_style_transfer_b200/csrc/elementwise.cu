// GEGLU gate of the motion-module feed-forward: out[r, j] =
// hg[r, j] * gelu(hg[r, inner + j]) with exact (erf) GELU, bf16 in/out, fp32 math.
// diffusers FeedForward(activation_fn="geglu") (un-vendored); GELU as temporal_transformer.py:55.
// HBM-bound: 3 * rows * inner * 2 bytes; 16-byte vectors, grid-stride.
#include "common.cuh"

namespace {
__global__ void __launch_bounds__(256)
geglu_kernel(const __nv_bfloat16* __restrict__ hg, __nv_bfloat16* __restrict__ out, int64_t rows,
             int inner, int64_t ld_in, int64_t ld_out) {
  const int vec_per_row = inner >> 3;
  const int64_t total = rows * vec_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / vec_per_row;
    const int j = (int)(i % vec_per_row) * 8;
    const uint4 h = ld_stream16(hg + r * ld_in + j);
    const uint4 g = ld_stream16(hg + r * ld_in + inner + j);
    const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, gw[4] = {g.x, g.y, g.z, g.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 a = unpack_bf16(hw[k]), b = unpack_bf16(gw[k]);
      const float g0 = 0.5f * b.x * (1.f + erff(b.x * 0.70710678118654752f));
      const float g1 = 0.5f * b.y * (1.f + erff(b.y * 0.70710678118654752f));
      // torch: gelu(gate) is rounded to bf16 before the product
      o[k] = pack_bf16(a.x * __bfloat162float(__float2bfloat16_rn(g0)),
                       a.y * __bfloat162float(__float2bfloat16_rn(g1)));
    }
    st_stream16(out + r * ld_out + j, make_uint4(o[0], o[1], o[2], o[3]));
  }
}
}  // namespace

extern "C" int vst_geglu(const void* hg, void* out, int64_t rows, int inner, int64_t ld_in,
                         int64_t ld_out, void* stream) {
  VST_REQUIRE(rows >= 0 && inner > 0, VST_E_BADARG, "geglu: bad shape");
  if (rows == 0) return VST_OK;
  VST_REQUIRE(hg && out, VST_E_BADARG, "geglu: null pointer");
  VST_REQUIRE(inner % 8 == 0 && ld_in % 8 == 0 && ld_out % 8 == 0, VST_E_UNSUPPORTED,
              "geglu: inner and row strides must be multiples of 8");
  const int64_t total = rows * (inner / 8);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)vst_sm_count() * 16;
  if (blocks > cap) blocks = cap;
  geglu_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)hg, (__nv_bfloat16*)out, rows, inner, ld_in, ld_out);
  VST_LAUNCH_OK();
  return VST_OK;
}

// ---- §8f-4: classifier-free-guidance combine + Euler update of the sampler ----------------------
// inference_animatediff.py:122,131 with diffusers' EulerDiscreteScheduler.step (epsilon prediction,
// no churn; un-vendored): eps = eps_u + g (eps_c - eps_u); x' = x + (sigma_next - sigma) eps, and
// optionally the next step's scheduler.scale_model_input, x' / sqrt(sigma_next^2 + 1), in the model
// dtype.  One pass over four / five arrays instead of the sampler's six elementwise launches.
namespace {
template <typename TE, typename TX>
__global__ void __launch_bounds__(256)
cfg_euler_kernel(const TE* __restrict__ eu, const TE* __restrict__ ec, const TX* __restrict__ x,
                 TX* __restrict__ x_out, TE* __restrict__ scaled_next, int64_t n, float guidance,
                 float dt, float next_scale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float u = (float)eu[i], c = (float)ec[i];
    // the reference forms noise_pred in the UNet's output dtype before the fp32 scheduler step
    const float eps = (float)(TE)(u + guidance * (float)(TE)(c - u));
    const float xn = (float)x[i] + dt * eps;
    x_out[i] = (TX)xn;
    if (scaled_next != nullptr) scaled_next[i] = (TE)((float)(TX)xn * next_scale);
  }
}
}  // namespace

extern "C" int vst_cfg_euler_step(const void* eps_uncond, const void* eps_cond, const void* latents,
                                  void* latents_out, void* scaled_next, int64_t n, float guidance,
                                  float sigma, float sigma_next, int eps_dtype, int latent_dtype,
                                  void* stream) {
  VST_REQUIRE(n >= 0, VST_E_BADARG, "cfg_euler_step: bad size");
  if (n == 0) return VST_OK;
  VST_REQUIRE(eps_uncond && eps_cond && latents && latents_out, VST_E_BADARG, "cfg_euler_step: null pointer");
  int64_t blocks = (n + 255) / 256;
  const int64_t cap = (int64_t)vst_sm_count() * 16;
  if (blocks > cap) blocks = cap;
  const float dt = sigma_next - sigma, ns = 1.f / sqrtf(sigma_next * sigma_next + 1.f);
  cudaStream_t st = (cudaStream_t)stream;
#define GO(TE, TX)                                                                               \
  cfg_euler_kernel<TE, TX><<<(unsigned)blocks, 256, 0, st>>>(                                    \
      (const TE*)eps_uncond, (const TE*)eps_cond, (const TX*)latents, (TX*)latents_out,          \
      (TE*)scaled_next, n, guidance, dt, ns)
  if (eps_dtype == VST_BF16 && latent_dtype == VST_F32) GO(__nv_bfloat16, float);
  else if (eps_dtype == VST_BF16 && latent_dtype == VST_BF16) GO(__nv_bfloat16, __nv_bfloat16);
  else if (eps_dtype == VST_F32 && latent_dtype == VST_F32) GO(float, float);
  else {
    vst_set_error("cfg_euler_step: unsupported dtypes eps=%d latents=%d", eps_dtype, latent_dtype);
    return VST_E_UNSUPPORTED;
  }
#undef GO
  VST_LAUNCH_OK();
  return VST_OK;
}
