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
// Thin compositions of vst_lora_gemm / vst_gemm so a C caller can run one LoRA linear's forward and
// backward without the Python autograd glue (video_style_transfer_b200/_autograd.py does the same).
extern "C" int vst_lora_gemm_fwd(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1,
                                 const void* A2, int64_t lda2, const void* B2, int64_t ldb2, int K2,
                                 int k2_group_n, const void* bias, int bias_dtype, const void* residual,
                                 int64_t ldr, float out_scale, int act, void* D, int64_t ldd, int64_t M,
                                 int N, int out_dtype, void* stream) {
  return vst_lora_gemm(A1, lda1, B1, ldb1, K1, A2, lda2, B2, ldb2, K2, k2_group_n, bias, bias_dtype, residual,
                       ldr, out_scale, act, D, ldd, M, N, out_dtype, stream);
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
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                   d, s, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, l2,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // a thread that has made no runtime call yet (torch's autograd worker entering a backward whose first
    // device action is this driver call) has no current context: bind the primary one and retry
    cudaFree(0);
    r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, swz, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  VST_REQUIRE(r == CUDA_SUCCESS, VST_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return VST_OK;
}
