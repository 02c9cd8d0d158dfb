"""Kernel table of one config-2 step (self-attention processor, Stage-1 trainable set) under torch.profiler:
which launches make up the fwd+bwd time (tuning aid)."""
import os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench_extra
from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0

kind, stage = (sys.argv[1] if len(sys.argv) > 1 else "self"), int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda", 0)
torch.manual_seed(0)
bf = torch.bfloat16
proc = AnimateDiffAttnProcessor2_0()
F, S, C = 16, 1024, 1280
attn = bench_extra._midblock_attn(kind, stage, dev)
x = torch.randn(F, S, C, device=dev).to(bf).requires_grad_(True)
text = None if kind == "self" else torch.randn(1, 77, 2048, device=dev).to(bf).requires_grad_(True)
dy = torch.randn(F, S, C, device=dev).to(bf)


def fwd_bwd():
    x.grad = None
    y = proc(attn, x, text)
    y.backward(dy)


for _ in range(4):
    fwd_bwd()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        fwd_bwd()
    torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total / 3.0) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[2])
tot = sum(r[2] for r in rows)
print(f"{kind} stage {stage}: device time per step {tot:.1f} us over {sum(r[1] for r in rows) // 3} launches")
for k, n, t in rows[:40]:
    print(f"{t:9.1f} us {n // 3:4d}  {k[:110]}")
