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
    case VST_OP_LAYERNORM_BWD: return need(2) ? vst_layernorm_bwd_workspace_bytes(dims[0], (int)dims[1]) : VST_E_BADARG;
    case VST_OP_GROUPNORM5D_BWD:
      return need(4) ? vst_groupnorm5d_bwd_workspace_bytes((int)dims[0], (int)dims[1], (int)dims[2], (int)dims[3])
                     : VST_E_BADARG;
    case VST_OP_REARRANGE: case VST_OP_GROUPNORM5D: case VST_OP_LAYERNORM: case VST_OP_LORA_GEMM:
    case VST_OP_ATTN_FWD: case VST_OP_TEMPORAL_ATTN_FWD: case VST_OP_TEMPORAL_ATTN_BWD: case VST_OP_GEGLU:
      return 0;    // these ops take no workspace
    default:
      vst_set_error("vst_workspace_bytes: unknown op %d", op);
      return VST_E_BADARG;
  }
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
  VST_REQUIRE(r == CUDA_SUCCESS, VST_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return VST_OK;
}
