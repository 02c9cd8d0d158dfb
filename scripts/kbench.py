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


def timeit(fn, iters=10, warmup=3, flush=True):
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


if __name__ == "__main__":
    main()
