"""Context line (not a bench arm): the SAME cfg3 hot-path workload through the reference's own
modules on `cuda` in bf16 -- torch eager, i.e. what the reference runs on a B200 today:
F.scaled_dot_product_attention + cuBLAS GEMMs, the dense 3-GEMM UnZipLoRA form of
unziplora_linear_layer.py:298-346 materialised on every call, three F.linear per temporal LoRA,
ATen GroupNorm / LayerNorm / permute copies.  Modules come from oracle/_ref (oracle/build_ref.py);
the motion block is oracle/ref_cpu.RefMotionModule (diffusers-shaped, un-vendored).  Full shapes,
full call counts (70 self + 70 cross + 15 motion per UNet call, two B=1 calls per step as
inference_animatediff.py:110-121), distinct weights per module.  CUDA-event timing.

    python scripts/torch_eager_context.py [--steps 5] [--warmup 2] > profiles/torch_eager_context_r02.json
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_cpu  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--latent", type=int, default=128)
    args = ap.parse_args()
    ns = ref_cpu.load_reference()
    if ns is None:
        print(json.dumps({"unavailable": "oracle/_ref is not in place (run python -m oracle.build_ref)"}))
        return
    dev, bf = torch.device("cuda", 0), torch.bfloat16
    F, S = args.frames, args.latent
    torch.manual_seed(0)
    # SDXL plan (video_style_transfer_b200/sdxl_hotpath.py): per UNet call 10 L1 + 60 L2 blocks, 5 motion per level
    plan = []
    for c, heads, side, nblk in ((640, 10, S // 2, 10), (1280, 20, S // 4, 60)):
        for _ in range(nblk):
            a_s, p = ref_cpu.ref_spatial_attn(ns, c, heads, c)
            a_c, _ = ref_cpu.ref_spatial_attn(ns, c, heads, 2048)
            for a in (a_s, a_c):
                for lin in (a.to_q, a.to_k, a.to_v, a.to_out):
                    lin.to(dev, bf)
            plan.append(("self", c, side, a_s, p))
            plan.append(("cross", c, side, a_c, p))
    for c, side in ((320, S), (640, S // 2), (1280, S // 4)):
        for _ in range(5):
            plan.append(("motion", c, side, ref_cpu.RefMotionModule(ns, c).to(dev, bf).eval(), None))
    x3 = {c: torch.randn(F, side * side, c, device=dev, dtype=bf) for c, side in ((640, S // 2), (1280, S // 4))}
    x4 = {c: torch.randn(F, c, side, side, device=dev, dtype=bf) for c, side in ((320, S), (640, S // 2), (1280, S // 4))}
    texts = [torch.randn(1, 77, 2048, device=dev, dtype=bf) for _ in range(2)]

    @torch.no_grad()
    def step():
        for text in texts:
            for kind, c, side, mod, proc in plan:
                if kind == "self":
                    proc(mod, x3[c], scale=1.0)
                elif kind == "cross":
                    proc(mod, x3[c], encoder_hidden_states=text, scale=1.0)
                else:
                    mod(x4[c], F)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    print(json.dumps({"what": "cfg3 hot-path step through the reference's modules, torch eager bf16 on cuda "
                              "(SDPA + cuBLAS, dense 3-GEMM UnZipLoRA form): context, not a bench arm",
                      "ms_per_step": ms, "steps_per_s": 1e3 / ms, "steps": args.steps, "warmup": args.warmup,
                      "frames": F, "latent": S, "torch": torch.__version__,
                      "sdpa_backends": {"flash": torch.backends.cuda.flash_sdp_enabled(),
                                        "cudnn": torch.backends.cuda.cudnn_sdp_enabled(),
                                        "mem_efficient": torch.backends.cuda.mem_efficient_sdp_enabled()},
                      "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1)}))


if __name__ == "__main__":
    main()
