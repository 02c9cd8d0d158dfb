"""ctypes binding of libvst_b200.so (the C ABI declared in include/vst_b200.h).

There is no CPU or eager-PyTorch fallback: if the shared library is missing or a symbol cannot
be resolved, importing this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# VST_LIB_PATH: load another build of the same library (A/B timing of two builds in one GPU call)
LIB_PATH = os.environ.get("VST_LIB_PATH") or os.path.join(_HERE, "libvst_b200.so")

VST_F32, VST_BF16 = 0, 1

_p, _i, _l, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> (restype, argtypes); mirrors include/vst_b200.h one to one
SIGNATURES = {
    "vst_version": (_i, []),
    "vst_last_error": (C.c_char_p, []),
    "vst_device_info": (_i, [C.POINTER(_i)] * 3),
    "vst_rearrange_fwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _i, _i, _p]),
    "vst_rearrange_bwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _i, _i, _p]),
    "vst_groupnorm5d_partials": (_i, [_i, _i, _i, _i]),
    "vst_groupnorm5d_stats": (_i, [_p, _p, _i, _i, _i, _i, _l, _l, _l, _i, _i, _p]),
    "vst_groupnorm5d_finalize": (_i, [_p, _i, _i, _i, _f, _p, _p, _p]),
    "vst_groupnorm5d_apply_rearrange": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _l, _l, _l,
                                             _i, _i, _i, _i, _p]),
    "vst_layernorm": (_i, [_p, _p, _p, _p, _p, _l, _i, _i, _f, _i, _i, _i, _p]),
    "vst_temporal_attn_fwd": (_i, [_p, _p, _p, _p, _l, _i, _i, _i, _l, _l, _l, _l, _f, _p]),
    "vst_lora_gemm": (_i, [_p, _l, _p, _l, _i, _p, _l, _p, _l, _i, _i, _p, _i, _p, _l, _f, _i, _p, _l,
                           _l, _i, _i, _p]),
    "vst_lora_gemm_plan": (_i, [_l, _i, _i, _i, _i]),
    "vst_lora_gemm_fused_supported": (_i, [_l, _i, _i, _i, _i]),
    "vst_lora_gemm_fused": (_i, [_p, _l, _p, _l, _i, _p, _l, _p, _l, _i, _i, _p, _i, _f, _i, _p, _l, _p, _l, _l, _i, _p]),
    "vst_lora_pack_down": (_i, [_p, _p, _p, _i, _i, _i, _i, _p]),
    "vst_lora_gemm_fwd_workspace_bytes": (_l, [_l, _i, _i, _i, _i, _i]),
    # x ldx x_c ldxc x_s ldxs W ldw bias bias_dtype A_c B_c m_c mask_c A_s B_s m_s mask_s rank param_dtype scale mode
    # residual ldr out_scale u_out ldu D ldd M N K Klow out_dtype workspace bytes stream
    "vst_lora_gemm_fwd": (_i, [_p, _l, _p, _l, _p, _l, _p, _l, _p, _i] + [_p] * 8 + [_i, _i, _f, _i, _p, _l, _f, _p, _l,
                               _p, _l, _l, _i, _i, _i, _i, _p, _l, _p]),
    # dY lddy W ldw Bp ldbp A lda R dU lddu dX lddx M N K alpha accumulate out_dtype stream
    "vst_lora_gemm_bwd_dx": (_i, [_p, _l, _p, _l, _p, _l, _p, _l, _i, _p, _l, _p, _l, _l, _i, _i, _f, _i, _i, _p]),
    "vst_lora_gemm_bwd_params_workspace_bytes": (_l, [_l, _i, _i, _i, _i]),
    # dY lddy X ldx Xlow ldxl U ldu dU lddu dW dA dBp M N K Klow R accumulate workspace bytes stream
    "vst_lora_gemm_bwd_params": (_i, [_p, _l, _p, _l, _p, _l, _p, _l, _p, _l, _p, _p, _p, _l, _i, _i, _i, _i, _i,
                                      _p, _l, _p]),
    "vst_lora_pack_up": (_i, [_p, _p, _p, _p, _p, _p, _f, _p, _i, _i, _i, _p]),
    "vst_gemm_workspace_bytes": (_l, [_l, _i, _l, _l]),
    "vst_gemm": (_i, [_p, _l, _p, _l, _l, _p, _l, _p, _l, _l, _i, _i, _p, _l, _l, _i, _f, _i, _i, _p,
                      _l, _p]),
    "vst_lora_pack_up_bwd": (_i, [_p, _l, _i, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _i, _i, _i, _p]),
    "vst_colsum_workspace_bytes": (_l, [_l, _i]),
    "vst_colsum": (_i, [_p, _l, _l, _i, _f, _p, _p, _l, _p]),
    "vst_colsum_noalloc": (_i, [_p, _l, _l, _i, _f, _p, _p]),
    "vst_geglu": (_i, [_p, _p, _l, _i, _l, _l, _p]),
    "vst_cfg_euler_step": (_i, [_p, _p, _p, _p, _p, _l, _f, _f, _f, _i, _i, _p]),
    "vst_attn_bwd_workspace_bytes": (_l, [_i, _i, _i, _i, _i]),
    "vst_attn_bwd": (_i, [_p] * 9 + [_i] * 5 + [_l] * 8 + [_i, _f, _p, _l, _l, _l, _p, _l, _p]),
    "vst_temporal_attn_bwd": (_i, [_p] * 7 + [_l, _i, _i, _i] + [_l] * 7 + [_f, _p]),
    "vst_layernorm_bwd_workspace_bytes": (_l, [_l, _i]),
    "vst_layernorm_bwd": (_i, [_p, _p, _p, _p, _p, _p, _l, _i, _f, _i, _p, _l, _p]),
    "vst_geglu_bwd": (_i, [_p, _p, _p, _l, _i, _l, _l, _l, _p]),
    "vst_gelu_bwd": (_i, [_p, _p, _p, _l, _i, _l, _l, _l, _f, _p]),
    "vst_groupnorm5d_bwd_workspace_bytes": (_l, [_i, _i, _i, _i]),
    "vst_groupnorm5d_bwd": (_i, [_p] * 9 + [_i] * 6 + [_p, _l, _p]),
    "vst_comm_alloc": (_i, [_l, C.POINTER(_p)]),
    "vst_comm_free": (_i, [_p]),
    "vst_comm_ipc_get": (_i, [_p, _p]),
    "vst_comm_ipc_open": (_i, [_p, C.POINTER(_p)]),
    "vst_comm_ipc_close": (_i, [_p]),
    "vst_comm_init": (_i, [_i, _i, C.POINTER(_p), C.POINTER(_p), C.POINTER(_p)]),
    "vst_comm_fini": (_i, [_p]),
    "vst_comm_barrier": (_i, [_p, _p]),
    "vst_workspace_bytes": (_l, [_i, _p, _i]),
    "vst_comm_allgather_partials": (_i, [_p, _p, _i, _i, _l, _p]),
    "vst_a2a_frames_to_hw": (_i, [_p, _p, _l, _p, _p, _p, _p, _i, _i, _i, _i, _i, _l, _l, _l, _i, _i, _p]),
    "vst_a2a_hw_to_frames": (_i, [_p, _l, _p, _p, _i, _i, _i, _i, _i, _l, _l, _l, _p]),
    "vst_groupnorm_tokens_apply": (_i, [_p, _p, _p, _p, _p, _i, _l, _i, _i, _i, _p]),
    "vst_attn_fwd": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _l, _l, _l, _l, _i, _f, _p, _l, _l, _l, _p]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make` (or __graft_entry__.build()). "
            "video_style_transfer_b200 has no CPU / eager fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


class VstError(RuntimeError):
    pass


launch_count = 0   # kernels launched through the C ABI (each compute entry point launches one)


def check(rc: int, what: str):
    global launch_count
    launch_count += 1
    if rc != 0:
        msg = lib.vst_last_error()
        raise VstError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
