"""LayerNorm flat kernel: how much of its time is the gamma / beta / position-table traffic through L1?"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = [sys.argv[0]]
from video_style_transfer_b200 import ops
from scripts.kbench import timeit, report
bf = torch.bfloat16
torch.manual_seed(0)
for tag, rows, c in (("L0", 524288, 320), ("L1", 131072, 640), ("L2", 32768, 1280)):
    x = torch.randn(rows, c, device="cuda").to(bf)
    g, b = torch.randn(c, device="cuda").to(bf), torch.randn(c, device="cuda").to(bf)
    pe = torch.randn(32, c, device="cuda")
    nb = 2 * rows * c * 2
    for name, fn in (("gamma+beta+pe", lambda: ops.layernorm(x, g, b, 1e-5, pe=pe, frames=16)),
                     ("gamma+beta", lambda: ops.layernorm(x, g, b, 1e-5)),
                     ("no params", lambda: ops.layernorm(x, None, None, 1e-5)),
                     ("torch copy_", lambda: torch.empty_like(x).copy_(x))):
        y = fn()
        if name != "torch copy_":     # bit-level checksum: equal across VST_LN_PERS=0/1 and across builds
            print(f"  checksum {tag} [{name}]: {int(y.view(torch.int16).to(torch.int64).sum())} "
                  f"{int((y.view(torch.int16).to(torch.int64) * torch.arange(y.numel(), device='cuda').view_as(y) % 65521).sum())}")
        ms, best = timeit(fn)
        report(f"layernorm {tag} rows={rows} C={c} [{name}]", ms, best, nb)
