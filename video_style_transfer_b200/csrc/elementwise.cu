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
