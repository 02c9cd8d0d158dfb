"""Per-kernel microbenchmarks at the config-3 shapes (16 frames, 1024^2 latents): CUDA-event timing,
algorithmic bytes / FLOPs per SURVEY.md §8(d), fraction of the measured peaks.  Development tool;
bench.py is the judged entry point."""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_style_transfer_b200 import ops  # noqa: E402

PEAKS = {"hbm_gbs": 6550.4, "bf16_tflops": 1659.1}
try:
    PEAKS.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
except Exception:
    pass

_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    _flush.zero_()


def timeit(fn, iters=None, warmup=None, flush=True):
    iters = iters or int(os.environ.get("KBENCH_ITERS", "10"))      # 1 / 1 under ncu --set full
    warmup = warmup or int(os.environ.get("KBENCH_WARMUP", "3"))
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def report(name, ms, best, nbytes=None, flops=None):
    s = f"{name:58s} med {ms*1e3:9.1f} us  best {best*1e3:9.1f} us"
    if nbytes:
        gbs = nbytes / (ms * 1e-3) / 1e9
        s += f"  {gbs:8.0f} GB/s ({gbs / PEAKS['hbm_gbs']:.2f} of HBM)"
    if flops:
        tf = flops / (ms * 1e-3) / 1e12
        s += f"  {tf:8.1f} TFLOP/s ({tf / PEAKS['bf16_tflops']:.2f} of bf16 burst)"
    print(s, flush=True)


def main():
    which = sys.argv[1:] or ["layout", "norm", "tattn", "gemm", "attn"]
    bf = torch.bfloat16
    F = 16
    levels = [("L0", 320, 128), ("L1", 640, 64), ("L2", 1280, 32)]
    if "layout" in which:
        for tag, c, hw in levels:
            x = torch.randn(1, F, c, hw, hw, device="cuda").to(bf).permute(0, 2, 1, 3, 4)
            pe = torch.randn(32, c, device="cuda")
            nb = 2 * x.numel() * 2
            ms, b = timeit(lambda: ops.rearrange_fwd(x, pe))
            report(f"rearrange_fwd+pe {tag} {tuple(x.shape)}", ms, b, nb)
            y = ops.rearrange_fwd(x, pe)
            ms, b = timeit(lambda: ops.rearrange_bwd(y, 1, hw, hw, residual5=x, frames_major=True))
            report(f"rearrange_bwd+res {tag}", ms, b, 3 * x.numel() * 2)
            ms, b = timeit(lambda: torch.empty_like(y).copy_(y))
            report(f"  (torch copy_ same bytes) {tag}", ms, b, nb)
    if "norm" in which:
        for tag, c, hw in levels:
            x = torch.randn(1, F, c, hw, hw, device="cuda").to(bf).permute(0, 2, 1, 3, 4)
            g, be = torch.randn(c, device="cuda").to(bf), torch.randn(c, device="cuda").to(bf)
            ms, b = timeit(lambda: ops.groupnorm5d_stats(x))
            report(f"groupnorm5d stats+finalize {tag}", ms, b, x.numel() * 2)
            ms, b = timeit(lambda: ops.groupnorm5d_rearrange(x, g, be))
            report(f"groupnorm5d stats+apply+rearrange {tag}", ms, b, 3 * x.numel() * 2)
            t = torch.randn(hw * hw * F, c, device="cuda").to(bf)
            pe = torch.randn(32, c, device="cuda")
            ms, b = timeit(lambda: ops.layernorm(t, g, be, 1e-5, pe=pe, frames=F))
            report(f"layernorm+pe {tag} rows={t.shape[0]}", ms, b, 2 * t.numel() * 2)
    if "tattn" in which:
        for tag, c, hw in levels:
            n = hw * hw
            qkv = torch.randn(n, F, 3 * c, device="cuda").to(bf)
            q, k, v = qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:]
            ms, b = timeit(lambda: ops.temporal_attn(q, k, v, 8))
            report(f"temporal_attn {tag} N={n} d={c // 8}", ms, b, 4 * n * F * c * 2,
                   4 * n * F * F * c)
    if "gemm" in which:
        cases = [("plain 8192^3", 8192, 8192, 8192, 0),
                 ("L1 self proj M=65536 C=640 r=64x2", 65536, 640, 640, 128),
                 ("L2 self proj M=16384 C=1280 r=64x2", 16384, 1280, 1280, 128),
                 ("L1 u = x A^T  N=128", 65536, 128, 640, 0),
                 ("L2 cross kv M=1232 K=2048 N=1280", 1232, 1280, 2048, 128),
                 ("L2 text k|v folded M=77 K=2048 N=2560", 77, 2560, 2048, 0),
                 ("L1 text k|v folded M=77 K=2048 N=1280", 77, 1280, 2048, 0),
                 ("L0 temporal proj M=262144 C=320 r=32", 262144, 320, 320, 32),
                 ("L2 ff M=16384 1280->10240", 16384, 10240, 1280, 0)]
        for tag, m, n, k1, k2 in cases:
            a1 = torch.randn(m, k1, device="cuda").to(bf)
            b1 = torch.randn(n, k1, device="cuda").to(bf)
            a2 = torch.randn(m, k2, device="cuda").to(bf) if k2 else None
            b2 = torch.randn(n, k2, device="cuda").to(bf) if k2 else None
            bias = torch.randn(n, device="cuda").to(bf)
            out = torch.empty(m, n, device="cuda", dtype=bf)
            ms, b = timeit(lambda: ops.lora_gemm(a1, b1, a2, b2, bias=bias, out=out))
            report(f"lora_gemm {tag}", ms, b, None, 2 * m * n * (k1 + k2))
            ms, b = timeit(lambda: torch.nn.functional.linear(a1, b1, bias))
            report(f"  (cuBLAS F.linear, K1 only) {tag}", ms, b, None, 2 * m * n * k1)
    if "dom" in which:
        # the dominant launches of the batched-CFG step (lora_gemm2_kernel<256,0>, adapters folded): M = 2 x 16 frames
        for tag, m, n, k in [("L2 q|k|v folded M=32768 N=3840 K=1280", 32768, 3840, 1280),
                             ("L2 to_q / to_out folded M=32768 N=1280 K=1280", 32768, 1280, 1280),
                             ("L1 q|k|v folded M=131072 N=1920 K=640", 131072, 1920, 640),
                             ("L2 motion q|k|v M=32768 N=3840 K=1280 (same shape, temporal)", 32768, 3840, 1280)][:3]:
            x = torch.randn(m, k, device="cuda").to(torch.bfloat16)
            w = (torch.randn(n, k, device="cuda") / math.sqrt(k)).to(torch.bfloat16)
            ms, b = timeit(lambda: ops.lora_gemm(x, w))
            report(f"lora_gemm {tag}", ms, b, 2 * (m * k + n * k + m * n), 2 * m * n * k)
    if "geglu" in which:
        # motion-module / spatial feed-forward, first GEMM with the GEGLU epilogue (batched-CFG M), per level
        for tag, m, c in [("L0 M=524288 C=320", 524288, 320), ("L1 M=131072 C=640", 131072, 640),
                          ("L2 M=32768 C=1280", 32768, 1280)]:
            x = torch.randn(m, c, device="cuda").to(torch.bfloat16)
            w = (torch.randn(8 * c, c, device="cuda") / math.sqrt(c)).to(torch.bfloat16)
            bias = torch.randn(8 * c, device="cuda").to(torch.bfloat16)
            ms, b = timeit(lambda: ops.lora_gemm(x, w, bias=bias, act="geglu"))
            report(f"lora_gemm GEGLU {tag} -> 4C", ms, b, 2 * (m * c + 8 * c * c + m * 4 * c), 2 * m * c * 8 * c)
            ms, b = timeit(lambda: ops.lora_gemm(x, w[:4 * c], bias=bias[:4 * c]))
            report(f"  (plain GEMM, N = 4C: half the FLOPs, same output bytes) {tag}", ms, b,
                   2 * (m * c + 4 * c * c + m * 4 * c), 2 * m * c * 4 * c)
    if "fused" in which:
        # single-launch fused LoRA projection against the two-launch form (u = x A^T pre-pass + extra K blocks)
        # and against the plain GEMM with the adapters folded into W
        for tag, m, n, k1, k2, gn in [("L2 q|k|v M=32768 C=1280 r=64x2", 32768, 3840, 1280, 128, 1280),
                                      ("L2 to_q   M=32768 C=1280 r=64x2", 32768, 1280, 1280, 128, 0),
                                      ("L1 q|k|v M=131072 C=640 r=64x2", 131072, 1920, 640, 128, 640),
                                      ("L1 to_q   M=131072 C=640 r=64x2", 131072, 640, 640, 128, 0),
                                      ("L2 q|k|v M=16384 (two-pass CFG)", 16384, 3840, 1280, 128, 1280)]:
            g = n // gn if gn else 1
            x = torch.randn(m, k1, device="cuda").to(torch.bfloat16)
            w = (torch.randn(n, k1, device="cuda") / math.sqrt(k1)).to(torch.bfloat16)
            a = (torch.randn(g * k2, k1, device="cuda") / math.sqrt(k1)).to(torch.bfloat16)
            bp = (torch.randn(n, k2, device="cuda") / math.sqrt(k2)).to(torch.bfloat16)
            fl = 2 * m * n * (k1 + k2) + 2 * m * g * k2 * k1
            ms, b = timeit(lambda: ops.lora_gemm_fused(x, w, a, bp, None, 1.0, gn))
            report(f"lora_gemm_fused {tag}", ms, b, None, fl)
            ms, b = timeit(lambda: ops.lora_gemm(x, w, ops.lora_gemm(x, a), bp, k2_group_n=gn))
            report(f"  (two launches) {tag}", ms, b, None, fl)
            ms, b = timeit(lambda: ops.lora_gemm(x, w))
            report(f"  (folded: plain GEMM) {tag}", ms, b, None, 2 * m * n * k1)
    if "gemmres" in which:
        # memory-bound GEMMs with a fused residual add (motion-module to_out / FF2, spatial to_out)
        for tag, m, n, k in [("L0 to_out +res M=262144 320x320", 262144, 320, 320),
                             ("L0 ff2 +res M=262144 N=320 K=1280", 262144, 320, 1280),
                             ("L1 to_out +res M=65536 640x640", 65536, 640, 640),
                             ("L1 ff2 +res M=65536 N=640 K=2560", 65536, 640, 2560),
                             ("L2 to_out +res M=16384 1280x1280", 16384, 1280, 1280)]:
            a1 = torch.randn(m, k, device="cuda").to(bf)
            b1 = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
            bias = torch.randn(n, device="cuda").to(bf)
            res = torch.randn(m, n, device="cuda").to(bf)
            out = torch.empty(m, n, device="cuda", dtype=bf)
            nb = (a1.numel() + b1.numel() + 2 * res.numel()) * 2
            ms, b = timeit(lambda: ops.lora_gemm(a1, b1, bias=bias, residual=res, out=out))
            report(f"lora_gemm {tag}", ms, b, nb, 2 * m * n * k)
            ms, b = timeit(lambda: ops.lora_gemm(a1, b1, bias=bias, out=out))
            report(f"  (same without residual) {tag}", ms, b, nb - res.numel() * 2, 2 * m * n * k)
    if "attn" in which:
        for tag, bt, h, sq, sk in [("L1 self", 16, 10, 4096, 4096), ("L2 self", 16, 20, 1024, 1024),
                                   ("L1 cross", 16, 10, 4096, 77), ("L2 cross", 16, 20, 1024, 77)]:
            c = h * 64
            q = torch.randn(bt, sq, c, device="cuda").to(bf)
            bk = bt if "self" in tag else 1
            k = torch.randn(bk, sk, c, device="cuda").to(bf)
            v = torch.randn(bk, sk, c, device="cuda").to(bf)
            fl = 4 * bt * h * sq * sk * 64
            nb = (2 * q.numel() + 2 * k.numel()) * 2
            ms, b = timeit(lambda: ops.attn_fwd(q, k, v, h))
            report(f"attn_fwd {tag} B={bt} h={h} Sq={sq} Sk={sk}", ms, b, nb, fl)
            qh = q.view(bt, sq, h, 64).transpose(1, 2)
            kh = k.view(bk, sk, h, 64).transpose(1, 2).expand(bt, h, sk, 64) if bk == 1 else \
                k.view(bk, sk, h, 64).transpose(1, 2)
            vh = v.view(bk, sk, h, 64).transpose(1, 2).expand(bt, h, sk, 64) if bk == 1 else \
                v.view(bk, sk, h, 64).transpose(1, 2)
            ms, b = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(qh, kh, vh))
            report(f"  (torch SDPA) {tag}", ms, b, nb, fl)
    if "bwd" in which:
        # the training-step (config 2 / 4) backward kernels at the same shapes
        for tag, m, n, k in [("L1 dx = dy W  M=65536 640x640", 65536, 640, 640),
                             ("L2 dx = dy W  M=16384 1280x1280", 16384, 1280, 1280),
                             ("L2 ff dx M=16384 N=1280 K=10240", 16384, 1280, 10240)]:
            dy = torch.randn(m, k, device="cuda").to(bf)
            w = torch.randn(k, n, device="cuda").to(bf)          # stored [out, in]: opB = W^T via trans_b
            ms, b = timeit(lambda: ops.gemm(dy, w, trans_b=True))
            report(f"vst_gemm NN {tag}", ms, b, None, 2 * m * n * k)
            ms, b = timeit(lambda: dy @ w)
            report(f"  (cuBLAS matmul) {tag}", ms, b, None, 2 * m * n * k)
        for tag, m, n, k in [("L1 dA = u^T x  128x640 K=65536", 128, 640, 65536),
                             ("L1 dW = dy^T x 640x640 K=65536", 640, 640, 65536),
                             ("L2 dW = dy^T x 1280x1280 K=16384", 1280, 1280, 16384),
                             ("L0 temporal dB 320x32 K=262144", 320, 32, 262144)]:
            a = torch.randn(k, m, device="cuda").to(bf)
            x = torch.randn(k, n, device="cuda").to(bf)
            ms, b = timeit(lambda: ops.gemm(a, x, trans_a=True, trans_b=True, out_dtype=torch.float32))
            report(f"vst_gemm TN split-K {tag}", ms, b, (a.numel() + x.numel()) * 2, 2 * m * n * k)
            ms, b = timeit(lambda: a.t() @ x)
            report(f"  (cuBLAS matmul) {tag}", ms, b, (a.numel() + x.numel()) * 2, 2 * m * n * k)
    if "bwd" in which or "colsum" in which:
        for tag, m, n in [("L0 FF bias grad M=262144 N=2560", 262144, 2560), ("L0 proj bias grad M=262144 N=320", 262144, 320),
                          ("L2 FF bias grad M=16384 N=10240", 16384, 10240)]:
            dy = torch.randn(m, n, device="cuda").to(bf)
            ms, b = timeit(lambda: ops.colsum(dy))
            report(f"colsum {tag}", ms, b, m * n * 2)
    if "bwd" in which or "abwd" in which:
        for tag, bt, h, sq, sk in [("L1 self", 16, 10, 4096, 4096), ("L2 self", 16, 20, 1024, 1024),
                                   ("L1 cross", 16, 10, 4096, 77), ("L2 cross", 16, 20, 1024, 77)]:
            if "bwd" not in which and "cross" in tag:
                continue
            c = h * 64
            bk = bt if "self" in tag else 1
            q = torch.randn(bt, sq, c, device="cuda").to(bf)
            k = torch.randn(bk, sk, c, device="cuda").to(bf)
            v = torch.randn(bk, sk, c, device="cuda").to(bf)
            do = torch.randn(bt, sq, c, device="cuda").to(bf)
            o, lse = ops.attn_fwd_raw(q, k, v, h, return_lse=True)
            fl = 10 * bt * h * sq * sk * 64
            ms, b = timeit(lambda: ops.attn_bwd(q, k, v, o, do, lse, h))
            report(f"attn_bwd {tag} B={bt} h={h} Sq={sq} Sk={sk}", ms, b, None, fl)
            if bk == bt and "bwd" in which:
                qh, kh, vh = (t.view(bt, -1, h, 64).transpose(1, 2).detach().requires_grad_() for t in (q, k, v))
                with torch.enable_grad():
                    oh = torch.nn.functional.scaled_dot_product_attention(qh, kh, vh)
                doh = do.view(bt, sq, h, 64).transpose(1, 2)
                ms, b = timeit(lambda: torch.autograd.grad(oh, (qh, kh, vh), doh, retain_graph=True))
                report(f"  (torch SDPA backward) {tag}", ms, b, None, fl)
        for tag, c, hw in levels:
            n = hw * hw
            qkv = torch.randn(n, F, 3 * c, device="cuda").to(bf)
            q, k, v = qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:]
            do = torch.randn(n, F, c, device="cuda").to(bf)
            ms, b = timeit(lambda: ops.temporal_attn_bwd(q, k, v, do, 8))
            report(f"temporal_attn_bwd {tag} N={n} d={c // 8}", ms, b, 7 * n * F * c * 2, 10 * n * F * F * c)
            t = torch.randn(n * F, c, device="cuda").to(bf)
            dy = torch.randn(n * F, c, device="cuda").to(bf)
            g = torch.randn(c, device="cuda").to(bf)
            ms, b = timeit(lambda: ops.layernorm_bwd(t, dy, g, 1e-5))
            report(f"layernorm_bwd {tag} rows={n * F}", ms, b, 3 * t.numel() * 2)
            hg = torch.randn(n * F, 8 * c, device="cuda").to(bf)
            dg = torch.randn(n * F, 4 * c, device="cuda").to(bf)
            ms, b = timeit(lambda: ops.geglu_bwd(hg, dg))
            report(f"geglu_bwd {tag} rows={n * F} inner={4 * c}", ms, b, (2 * hg.numel() + dg.numel()) * 2)
            del hg, dg
            x = torch.randn(F, c, hw, hw, device="cuda").to(bf)
            dyx = torch.randn(F, c, hw, hw, device="cuda").to(bf)
            mean, rstd = ops.groupnorm5d_stats(x.view(1, F, c, hw, hw).permute(0, 2, 1, 3, 4))
            ms, b = timeit(lambda: ops.groupnorm5d_bwd(x, dyx, mean, rstd, g, 1, F, 32))
            report(f"groupnorm5d_bwd {tag}", ms, b, 5 * x.numel() * 2)


if __name__ == "__main__":
    main()
