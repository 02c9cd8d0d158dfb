"""One small invocation of every kernel family through the C ABI, for compute-sanitizer
(scripts/sanitize.sh).  Shapes are small on purpose (the tools slow kernels down 10-100x) but hit the
ragged edges: rows / columns that do not fill a tile, Sk = 77, residual / bias / low-rank operands."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_style_transfer_b200 import ops  # noqa: E402

bf = torch.bfloat16
torch.manual_seed(0)
dev = "cuda"
rn = lambda *s: torch.randn(*s, device=dev).to(bf)


def case(name, fn):
    fn()
    torch.cuda.synchronize()
    print("ok", name, flush=True)


pe = torch.randn(32, 64, device=dev)
x5 = rn(1, 8, 64, 6, 5).permute(0, 2, 1, 3, 4)                     # (B,C,F,H,W) view of frames-major storage
case("rearrange_fwd+pe", lambda: ops.rearrange_fwd(x5, pe))
case("rearrange_bwd+res", lambda: ops.rearrange_bwd(rn(30, 8, 64), 1, 6, 5, residual5=x5, frames_major=True))
case("groupnorm5d_rearrange", lambda: ops.groupnorm5d_rearrange(x5, torch.ones(64, device=dev), torch.zeros(64, device=dev)))
case("layernorm+pe", lambda: ops.layernorm(rn(240, 64), torch.ones(64, device=dev), torch.zeros(64, device=dev), pe=pe, frames=8))
case("layernorm flat 320", lambda: ops.layernorm(rn(96, 320), torch.ones(320, device=dev), torch.zeros(320, device=dev)))
case("temporal_attn", lambda: ops.temporal_attn(rn(9, 16, 320), rn(9, 16, 320), rn(9, 16, 320), 8))
case("lora_gemm single-CTA ragged", lambda: ops.lora_gemm(rn(300, 320), rn(136, 320), rn(300, 32), rn(136, 32),
                                                         bias=torch.randn(136, device=dev), residual=rn(300, 136)))
case("lora_gemm pair + residual", lambda: ops.lora_gemm(rn(4100, 320), rn(320, 320), bias=torch.randn(320, device=dev),
                                                       residual=rn(4100, 320)))
case("lora_gemm pair grouped", lambda: ops.lora_gemm(rn(4096, 640), rn(1920, 640), rn(4096, 384), rn(1920, 128), k2_group_n=640))
case("lora_gemm geglu", lambda: ops.lora_gemm(rn(2048, 320), rn(2560, 320), bias=torch.randn(2560, device=dev), act="geglu"))
case("lora_gemm fp32 out", lambda: ops.lora_gemm(rn(77, 2048), rn(64, 2048), out_dtype=torch.float32))
case("gemm TN split-K", lambda: ops.gemm(rn(4096, 128), rn(4096, 72), trans_a=True, trans_b=True, out_dtype=torch.float32))
case("gemm NN two pairs", lambda: ops.gemm(rn(520, 640), rn(640, 320), rn(520, 128), rn(128, 320), trans_b=True))
q, k, v = rn(2, 200, 128), rn(2, 136, 128), rn(2, 136, 128)
case("attn_fwd stream (2 key blocks)", lambda: ops.attn_fwd_raw(q, k, v, 2))
case("attn_fwd text cross Sk=77", lambda: ops.attn_fwd_raw(rn(4, 256, 128), rn(1, 77, 128), rn(1, 77, 128), 2))
case("attn_fwd exact kernel (mask)", lambda: ops.attn_fwd_raw(q, k, v, 2, mask=torch.zeros(1, 1, 1, 136, device=dev)))
o, lse = ops.attn_fwd_raw(q, k, v, 2, return_lse=True)
case("attn_bwd", lambda: ops.attn_bwd(q, k, v, o, rn(2, 200, 128), lse, 2))
case("temporal_attn_bwd", lambda: ops.temporal_attn_bwd(rn(9, 16, 320), rn(9, 16, 320), rn(9, 16, 320), rn(9, 16, 320), 8))
case("layernorm_bwd", lambda: ops.layernorm_bwd(rn(240, 320), rn(240, 320), torch.ones(320, device=dev), 1e-5))
case("geglu fwd/bwd", lambda: ops.geglu_bwd(rn(64, 256), rn(64, 128)))
case("cfg_euler", lambda: ops.cfg_euler_step(rn(1000), rn(1000), torch.randn(1000, device=dev), 7.5, 2.0, 1.5))
print("all cases ran")
