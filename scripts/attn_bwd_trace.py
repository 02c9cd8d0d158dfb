"""Pipeline trace of the attention backward kernel (tuning aid, not part of the product).
Build (here, no GPU needed):  python scripts/attn_bwd_trace.py build
Run (GPU box):                python scripts/attn_bwd_trace.py run [Sq] [heads] [Bt]
CTA (0,0,0) stamps clock64() per (query tile, key tile) pair i at: 0 sdp_full seen, 1 S / dP in registers,
2 arithmetic done, 3 pds_free seen, 4 P / dS stored (softmax warp 0); 5 S / dP MMA issue, 6 dV / dK / dQ MMA issue
(MMA warp); 7 dq_full seen, 8 dQ atomics issued (drain warp 10); 9 TMA warp sees the Q / dO stage of pair i free;
10 MMA warp sees pds_full(i); 11 MMA warp starts waiting for Q / dO of pair i."""
import ctypes, glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CSRC = os.path.join(ROOT, "video_style_transfer_b200", "csrc")
LIB = os.environ.get("TRACE_LIB", os.path.join(ROOT, "video_style_transfer_b200", "libvst_b200_bwdtrace.so"))
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]


def build():
    obj = "/tmp/attn_bwd_trace.o"
    subprocess.check_call(["nvcc"] + FLAGS + ["-DVST_BWD_TRACE"] + os.environ.get("TRACE_FLAGS", "").split() + ["-c", os.path.join(CSRC, "attn_bwd.cu"), "-o", obj])
    others = [o for o in glob.glob(os.path.join(CSRC, "*.o")) if not o.endswith("/attn_bwd.o")]
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB, obj] + others)


def run(sq=4096, heads=10, bt=16):
    os.environ["VST_LIB_PATH"] = LIB
    import numpy as np
    import torch
    from video_style_transfer_b200 import _lib, ops
    c = heads * 64
    q, k, v, do = (torch.randn(bt, sq, c, device="cuda").to(torch.bfloat16) for _ in range(4))
    o, lse = ops.attn_fwd_raw(q, k, v, heads, return_lse=True)
    for _ in range(3):
        ops.attn_bwd(q, k, v, o, do, lse, heads)
        torch.cuda.synchronize()
    raw = ctypes.CDLL(LIB)
    buf = np.zeros(16 * 64, dtype=np.int64)
    assert raw.vst_attn_bwd_debug_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
    t = buf.reshape(16, 64)
    n = min(sq // 128, 64)
    t0 = t[5, 0]
    names = ["sdp_full", "in_regs", "math_done", "pds_free", "stored", "SdP_issue", "grad_issue", "dq_full", "dq_added",
             "q_stage_free", "pds_full@mma", "q_wait@mma"]
    print("pair " + " ".join(f"{x:>10s}" for x in names) + "     period")
    for i in range(n):
        row = [int(t[s, i] - t0) for s in range(12)]
        per = int(t[4, i] - t[4, i - 1]) if i else 0
        print(f"{i:4d} " + " ".join(f"{x:10d}" for x in row) + f" {per:10d}")
    life = t[12]
    print("CTA lifetime (clk from entry): set-up done %d, K/V landed %d, first S/dP issue %d, last P/dS stored %d, "
          "dV/dK stored %d, exit %d" % tuple(int(x - life[0]) for x in (life[1], life[2], t[5, 0], t[4, n - 1], life[3], life[4])))
    lo, hi = min(4, n // 2), n - 2
    def mean(a, b):
        return float(np.mean([t[b, i] - t[a, i] for i in range(lo, hi)]))
    print("steady state (pairs %d..%d), clk:" % (lo, hi - 1))
    print("  period                         %8.0f" % float(np.mean([t[4, i + 1] - t[4, i] for i in range(lo, hi)])))
    print("  sdp_full -> S/dP in registers  %8.0f" % mean(0, 1))
    print("  arithmetic (second chunk)      %8.0f" % mean(1, 2))
    print("  wait pds_free                  %8.0f" % mean(2, 3))
    print("  P/dS stores + fence + arrive   %8.0f" % mean(3, 4))
    print("  stored(i) -> sdp_full(i+1)     %8.0f" % float(np.mean([t[0, i + 1] - t[4, i] for i in range(lo, hi)])))
    print("  S/dP issue(i+1) after in_regs(i) %6.0f" % float(np.mean([t[5, i + 1] - t[1, i] for i in range(lo, hi)])))
    print("  S/dP issue(i+1) -> sdp_full(i+1) %6.0f" % float(np.mean([t[0, i + 1] - t[5, i + 1] for i in range(lo, hi)])))
    print("  grad issue(i) after stored(i)  %8.0f" % mean(4, 6))
    print("  grad issue(i) -> dq_full(i)    %8.0f" % mean(6, 7))
    print("  grad issue(i) -> pds_free seen by softmax(i+1) %6.0f" % float(np.mean([t[3, i + 1] - t[6, i] for i in range(lo, hi)])))
    print("  dq_full -> atomics issued      %8.0f" % mean(7, 8))


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build()
    else:
        run(*[int(x) for x in sys.argv[2:]])
