// Library-level plumbing of libvst_b200.so: version, error text, device info, TMA map encode.
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

static thread_local char g_err[512] = "";

void vst_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" int vst_version(void) { return 100; /* 0.1.0 */ }
extern "C" const char* vst_last_error(void) { return g_err; }

// one entry point that answers "how much caller-provided workspace does op need"
// for every op that takes one; the per-op functions stay exported too.
extern "C" int64_t vst_workspace_bytes(int op, const int64_t* dims, int ndims) {
  auto need = [&](int n) {
    if (dims != nullptr && ndims >= n) return true;
    vst_set_error("vst_workspace_bytes: op %d takes %d dims", op, n);
    return false;
  };
  switch (op) {
    case VST_OP_GEMM: return need(4) ? vst_gemm_workspace_bytes(dims[0], dims[1], dims[2], dims[3]) : VST_E_BADARG;
    case VST_OP_COLSUM: return need(2) ? vst_colsum_workspace_bytes(dims[0], (int)dims[1]) : VST_E_BADARG;
    case VST_OP_ATTN_BWD:
      return need(5) ? vst_attn_bwd_workspace_bytes((int)dims[0], (int)dims[1], (int)dims[2], (int)dims[3], (int)dims[4])
                     : VST_E_BADARG;
    case VST_OP_LORA_GEMM_BWD_PARAMS:
      return need(5) ? vst_lora_gemm_bwd_params_workspace_bytes(dims[0], (int)dims[1], (int)dims[2], (int)dims[3],
                                                                (int)dims[4])
                     : VST_E_BADARG;
    case VST_OP_LAYERNORM_BWD: return need(2) ? vst_layernorm_bwd_workspace_bytes(dims[0], (int)dims[1]) : VST_E_BADARG;
    case VST_OP_GROUPNORM5D_BWD:
      return need(4) ? vst_groupnorm5d_bwd_workspace_bytes((int)dims[0], (int)dims[1], (int)dims[2], (int)dims[3])
                     : VST_E_BADARG;
    case VST_OP_REARRANGE: case VST_OP_GROUPNORM5D: case VST_OP_LAYERNORM: case VST_OP_LORA_GEMM:
    case VST_OP_ATTN_FWD: case VST_OP_TEMPORAL_ATTN_FWD: case VST_OP_TEMPORAL_ATTN_BWD: case VST_OP_GEGLU:
    case VST_OP_LORA_GEMM_BWD_DX:
      return 0;    // these ops take no workspace
    default:
      vst_set_error("vst_workspace_bytes: unknown op %d", op);
      return VST_E_BADARG;
  }
}

// ---- SURVEY.md §8b entry points of the fused LoRA linear: forward, dX, parameter gradients ------
// Module-level calls: a C caller runs one LoRA linear's forward and backward from the RAW parameters
// (A, B, mergers, masks), without the Python glue (video_style_transfer_b200/_autograd.py, fused.py).
static int64_t al256(int64_t x) { return (x + 255) / 256 * 256; }

extern "C" int64_t vst_lora_gemm_fwd_workspace_bytes(int64_t M, int N, int Klow, int rank, int branches,
                                                     int have_u_out) {
  if (M < 0 || N <= 0 || Klow <= 0 || rank <= 0 || branches < 1 || branches > 2) return VST_E_BADARG;
  const int64_t k2 = (int64_t)branches * ((rank + 7) / 8 * 8);
  return al256(k2 * Klow * 2) + al256((int64_t)N * k2 * 2) + (have_u_out ? 0 : al256(M * k2 * 2));
}

// y = (x W^T + bias + [(x_c A_c^T)(scale m_c.mask_c.B_c)^T + (x_s A_s^T)(scale m_s.mask_s.B_s)^T] + residual) * out_scale
//   lora_linear.py:74-81 (x_c = x_s = x), lora_unzip.py:66-75 (three streams), unziplora_linear_layer.py:221-263
//   (hard masks multiply), :298-346 (a block-masked branch is absent: pass its A = NULL), :324-345 (modes
//   content / style: one branch, no merger).  u = [x_c A_c^T | x_s A_s^T] is computed INSIDE the main kernel
//   (vst_lora_gemm_fused) when the streams coincide and the shape allows, else by a pre-pass; it is written to
//   u_out when given (the backward needs it).
extern "C" int vst_lora_gemm_fwd(const void* x, int64_t ldx, const void* x_c, int64_t ldxc, const void* x_s,
                                 int64_t ldxs, const void* W, int64_t ldw, const void* bias, int bias_dtype,
                                 const void* A_c, const void* B_c, const void* m_c, const uint8_t* mask_c,
                                 const void* A_s, const void* B_s, const void* m_s, const uint8_t* mask_s,
                                 int rank, int param_dtype, float scale, int mode, const void* residual,
                                 int64_t ldr, float out_scale, void* u_out, int64_t ldu, void* D, int64_t ldd,
                                 int64_t M, int N, int K, int Klow, int out_dtype, void* workspace,
                                 int64_t workspace_bytes, void* stream) {
  VST_REQUIRE(x && W && D && M >= 0 && N > 0 && K > 0, VST_E_BADARG, "vst_lora_gemm_fwd: x, W and D are required");
  VST_REQUIRE(mode >= VST_LORA_BOTH && mode <= VST_LORA_STYLE, VST_E_BADARG, "vst_lora_gemm_fwd: bad mode %d", mode);
  if (M == 0) return VST_OK;
  const bool use_c = mode != VST_LORA_STYLE && A_c != nullptr && B_c != nullptr;
  const bool use_s = mode != VST_LORA_CONTENT && A_s != nullptr && B_s != nullptr;
  const int nbr = (use_c ? 1 : 0) + (use_s ? 1 : 0);
  if (nbr == 0)      // every branch block-masked: the plain linear (unziplora_linear_layer.py:308-309,316-317)
    return vst_lora_gemm(x, ldx, W, ldw, K, nullptr, 0, nullptr, 0, 0, 0, bias, bias_dtype, residual, ldr, out_scale,
                         VST_ACT_NONE, D, ldd, M, N, out_dtype, stream);
  VST_REQUIRE(rank > 0 && rank % 8 == 0, VST_E_UNSUPPORTED, "vst_lora_gemm_fwd: rank=%d must be a multiple of 8", rank);
  if (Klow <= 0) Klow = K;
  if (x_c == nullptr) { x_c = x; ldxc = ldx; VST_REQUIRE(Klow == K, VST_E_BADARG, "vst_lora_gemm_fwd: Klow != K needs x_c"); }
  if (x_s == nullptr) { x_s = x_c; ldxs = ldxc; }
  const int k2 = nbr * rank;
  const int64_t need = vst_lora_gemm_fwd_workspace_bytes(M, N, Klow, rank, nbr, u_out != nullptr);
  VST_REQUIRE(workspace && workspace_bytes >= need && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0, VST_E_BADARG,
              "vst_lora_gemm_fwd: a 256-byte aligned workspace of %lld bytes is required", (long long)need);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  void* a_pack = ws;
  ws += al256((int64_t)k2 * Klow * 2);
  void* b_pack = ws;
  ws += al256((int64_t)N * k2 * 2);
  void* u = u_out ? u_out : (void*)ws;
  const int64_t ldu_ = u_out ? ldu : k2;
  const bool merger = mode == VST_LORA_BOTH;                       // :331,:343: single-branch modes skip the merger
  int rc = vst_lora_pack_down(use_c ? A_c : nullptr, use_s ? A_s : nullptr, a_pack, rank, rank, Klow, param_dtype, stream);
  if (rc) return rc;
  rc = vst_lora_pack_up(use_c ? B_c : nullptr, use_c && merger ? m_c : nullptr, use_c ? mask_c : nullptr,
                        use_s ? B_s : nullptr, use_s && merger ? m_s : nullptr, use_s ? mask_s : nullptr, scale, b_pack,
                        N, rank, param_dtype, stream);
  if (rc) return rc;
  const bool one_stream = (!use_c || x_c == x) && (!use_s || x_s == x) && Klow == K;
  if (one_stream && residual == nullptr && out_dtype == VST_BF16 && vst_lora_gemm_fused_supported(M, N, K, k2, 0))
    return vst_lora_gemm_fused(x, ldx, W, ldw, K, a_pack, Klow, b_pack, k2, k2, 0, bias, bias_dtype, out_scale,
                               VST_ACT_NONE, u_out, ldu, D, ldd, M, N, stream);
  // pre-pass: u[:, branch] = x_branch A_branch^T (one GEMM when both branches read the same stream)
  const uint8_t* ap = reinterpret_cast<const uint8_t*>(a_pack);
  uint8_t* up = reinterpret_cast<uint8_t*>(u);
  if (nbr == 2 && x_s == x_c) {
    rc = vst_lora_gemm(x_c, ldxc, a_pack, Klow, Klow, nullptr, 0, nullptr, 0, 0, 0, nullptr, VST_F32, nullptr, 0, 1.0f,
                       VST_ACT_NONE, u, ldu_, M, k2, VST_BF16, stream);
    if (rc) return rc;
  } else {
    int col = 0;
    if (use_c) {
      rc = vst_lora_gemm(x_c, ldxc, ap, Klow, Klow, nullptr, 0, nullptr, 0, 0, 0, nullptr, VST_F32, nullptr, 0, 1.0f,
                         VST_ACT_NONE, up, ldu_, M, rank, VST_BF16, stream);
      if (rc) return rc;
      col = rank;
    }
    if (use_s) {
      rc = vst_lora_gemm(x_s, ldxs, ap + (int64_t)col * Klow * 2, Klow, Klow, nullptr, 0, nullptr, 0, 0, 0, nullptr, VST_F32,
                         nullptr, 0, 1.0f, VST_ACT_NONE, up + (int64_t)col * 2, ldu_, M, rank, VST_BF16, stream);
      if (rc) return rc;
    }
  }
  return vst_lora_gemm(x, ldx, W, ldw, K, u, ldu_, b_pack, k2, k2, 0, bias, bias_dtype, residual, ldr, out_scale,
                       VST_ACT_NONE, D, ldd, M, N, out_dtype, stream);
}

extern "C" int vst_lora_gemm_bwd_dx(const void* dY, int64_t lddy, const void* W, int64_t ldw, const void* Bp,
                                    int64_t ldbp, const void* A, int64_t lda, int R, void* dU, int64_t lddu,
                                    void* dX, int64_t lddx, int64_t M, int N, int K, float alpha, int accumulate,
                                    int out_dtype, void* stream) {
  VST_REQUIRE(dY && W && dX, VST_E_BADARG, "vst_lora_gemm_bwd_dx: dY, W and dX are required");
  const bool low = (R > 0);
  if (low) {
    VST_REQUIRE(Bp && A && dU, VST_E_BADARG, "vst_lora_gemm_bwd_dx: R > 0 needs Bp, A and dU");
    // dU[M,R] = dY[M,N] * B'[N,R]   (B' stored [contraction, out] -> transB = 1)
    int rc = vst_gemm(dY, lddy, Bp, ldbp, N, nullptr, 0, nullptr, 0, 0, 0, 1, dU, lddu, M, R, 1.0f, 0, VST_BF16,
                      nullptr, 0, stream);
    if (rc != VST_OK) return rc;
  }
  // dX[M,K] (+)= alpha * (dY[M,N] * W[N,K] + dU[M,R] * A[R,K])
  return vst_gemm(dY, lddy, W, ldw, N, low ? dU : nullptr, low ? lddu : 0, low ? A : nullptr, low ? lda : 0,
                  low ? R : 0, 0, 1, dX, lddx, M, K, alpha, accumulate, out_dtype, nullptr, 0, stream);
}

extern "C" int64_t vst_lora_gemm_bwd_params_workspace_bytes(int64_t M, int N, int K, int Klow, int R) {
  int64_t a = vst_gemm_workspace_bytes(N, K, M, 0);          // dW  [N, K]    over M tokens
  int64_t b = R > 0 ? vst_gemm_workspace_bytes(R, Klow, M, 0) : 0;   // dA  [R, Klow]
  int64_t c = R > 0 ? vst_gemm_workspace_bytes(N, R, M, 0) : 0;      // dB' [N, R]
  return a > b ? (a > c ? a : c) : (b > c ? b : c);
}

extern "C" int vst_lora_gemm_bwd_params(const void* dY, int64_t lddy, const void* X, int64_t ldx,
                                        const void* Xlow, int64_t ldxl, const void* U, int64_t ldu,
                                        const void* dU, int64_t lddu, float* dW, float* dA, float* dBp,
                                        int64_t M, int N, int K, int Klow, int R, int accumulate,
                                        void* workspace, int64_t workspace_bytes, void* stream) {
  VST_REQUIRE(dY != nullptr, VST_E_BADARG, "vst_lora_gemm_bwd_params: dY is required");
  int rc = VST_OK;
  if (dW) {      // dW[N,K] = dY^T X
    VST_REQUIRE(X != nullptr, VST_E_BADARG, "vst_lora_gemm_bwd_params: dW needs X");
    rc = vst_gemm(dY, lddy, X, ldx, M, nullptr, 0, nullptr, 0, 0, 1, 1, dW, K, N, K, 1.0f, accumulate, VST_F32,
                  workspace, workspace_bytes, stream);
    if (rc != VST_OK) return rc;
  }
  if (dA) {      // dA[R,Klow] = dU^T Xlow   (Xlow: the stream the adapter reads; X when single-stream)
    VST_REQUIRE(dU && (Xlow || X) && R > 0, VST_E_BADARG, "vst_lora_gemm_bwd_params: dA needs dU and Xlow");
    rc = vst_gemm(dU, lddu, Xlow ? Xlow : X, Xlow ? ldxl : ldx, M, nullptr, 0, nullptr, 0, 0, 1, 1, dA, Klow, R,
                  Klow, 1.0f, accumulate, VST_F32, workspace, workspace_bytes, stream);
    if (rc != VST_OK) return rc;
  }
  if (dBp) {     // dB'[N,R] = dY^T U
    VST_REQUIRE(U && R > 0, VST_E_BADARG, "vst_lora_gemm_bwd_params: dB' needs U");
    rc = vst_gemm(dY, lddy, U, ldu, M, nullptr, 0, nullptr, 0, 0, 1, 1, dBp, R, N, R, 1.0f, accumulate, VST_F32,
                  workspace, workspace_bytes, stream);
  }
  return rc;
}

int vst_sm_count() {
  static int cached = 0;
  if (cached) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 148;
  cached = n;
  return n;
}

extern "C" int vst_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  VST_CUDA_OK(cudaGetDevice(&dev));
  if (sm_count) VST_CUDA_OK(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
  if (cc_major) VST_CUDA_OK(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  if (cc_minor) VST_CUDA_OK(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
  return VST_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int vst_make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz,
                  CUtensorMapL2promotion l2) {
  return vst_make_tmap_dt(out, base, rank, dims, strides_bytes, box, swz, l2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
}

int vst_make_tmap_dt(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz,
                     CUtensorMapL2promotion l2, CUtensorMapDataType dt) {
  EncodeTiledFn enc = get_encode();
  VST_REQUIRE(enc != nullptr, VST_E_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  VST_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, VST_E_BADARG,
              "TMA base pointer must be 16-byte aligned");
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) {
    d[i] = dims[i];
    b[i] = box[i];
    es[i] = 1;
    if (i > 0) {
      s[i - 1] = strides_bytes[i];
      VST_REQUIRE((s[i - 1] & 15) == 0, VST_E_BADARG,
                  "TMA stride %llu (dim %d) must be a multiple of 16 bytes",
                  (unsigned long long)s[i - 1], i);
    }
  }
  CUresult r = enc(out, dt, (cuuint32_t)rank, const_cast<void*>(base),
                   d, s, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, l2,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // a thread that has made no runtime call yet (torch's autograd worker entering a backward whose first
    // device action is this driver call) has no current context: bind the primary one and retry
    cudaFree(0);
    r = enc(out, dt, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, swz, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  VST_REQUIRE(r == CUDA_SUCCESS, VST_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return VST_OK;
}
