"""Pipeline trace of the persistent attention kernel (tuning aid, not part of the product).
Build (here, no GPU needed):  python scripts/attn_trace.py build
Run (GPU box):                python scripts/attn_trace.py run [Sq] [heads] [Bt]
CTA 0 stamps clock64() at: 0 s_full seen, 1 S in registers, 2 exp loop done, 3 o_full seen,
4 P stored (softmax warp 0); 5 S MMA issue, 6 PV MMA issue (MMA warp)."""
import ctypes, glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "video_style_transfer_b200", "csrc")
LIB = os.path.join(ROOT, "scripts", "ubench", "libvst_attn_trace.so")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]

def build():
    obj = "/tmp/attn_trace.o"
    subprocess.check_call(["nvcc"] + FLAGS + os.environ.get("TRACE_FLAGS", "").split() + ["-DVST_ATTN_TRACE", "-c", os.path.join(CSRC, "attn.cu"), "-o", obj])
    others = [o for o in glob.glob(os.path.join(CSRC, "*.o")) if not o.endswith("/attn.o")]
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB, obj] + others)

def run(sq=1024, heads=20, bt=16):
    import torch, numpy as np
    lib = ctypes.CDLL(LIB)
    c = heads * 64
    q, k, v = (torch.randn(bt, sq, c, device="cuda").to(torch.bfloat16) for _ in range(3))
    o = torch.empty_like(q)
    P, L, I, F = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float
    lib.vst_attn_fwd.argtypes = [P, P, P, P, P, I, I, I, I, I, L, L, L, L, I, F, P, L, L, L, P]
    for _ in range(3):
        rc = lib.vst_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), None, bt, heads, sq, sq, 64,
                              c, c, c, c, 1, 0.125, None, 0, 0, 0, None)
        assert rc == 0
        torch.cuda.synchronize()
    buf = np.zeros(4 * 12 * 1024, dtype=np.int64)
    assert lib.vst_attn_debug_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
    full = buf.reshape(4, 12, 1024)
    # the MMA warp (warp 5) and the TMA warp land in banks 1 and 0: fold the MMA stamps into bank 0's view
    w = int(os.environ.get("TRACE_WARP", "0"))
    t = full[w].copy()
    t[5], t[6] = full[1][5], full[1][6]     # the MMA warp (warp 5) stamps into bank 1
    nb = sq // 128
    print("math phase per softmax warp (mean over item 2):",
          [float(np.mean([full[x][2][i] - full[x][1][i] for i in range(nb + 1, 2 * nb)])) for x in range(4)])
    print("block period per softmax warp:",
          [float(np.mean([full[x][4][i + 1] - full[x][4][i] for i in range(nb + 1, 2 * nb - 1)])) for x in range(4)])
    print("P_stored skew vs warp 0 (mean):",
          [float(np.mean([full[x][4][i] - full[0][4][i] for i in range(nb + 1, 2 * nb)])) for x in range(4)])
    nblk = sq // 128
    n = min(3 * nblk, 1024)
    t0 = t[5, 0]
    names = ["s_full", "S_in_reg", "exp_done", "o_full", "P_stored", "S_issue", "PV_issue"]
    print("blk  " + "  ".join(f"{x:>9s}" for x in names) + "   softmax_period")
    prev = None
    for b in range(n):
        row = [int(t[s, b] - t0) if t[s, b] else -1 for s in range(7)]
        per = row[4] - prev if prev is not None else 0
        prev = row[4]
        print(f"{b:3d}  " + "  ".join(f"{x:9d}" for x in row) + f"   {per:9d}" + ("   <- item start" if b % nblk == 0 else ""))
    print("deferred epilogue of item i inside block 0 of item i+1: exp_done -> o_full seen -> O staged -> P stored")
    for i in range(3):
        nxt = (i + 1) * nblk
        print(i, [int(t[s_, i] - t[2, nxt]) for s_ in (7, 8)], int(t[4, nxt] - t[2, nxt]))
    # steady state summary over the second item
    b0, b1 = nblk + 1, 2 * nblk
    d = lambda a, b_: np.mean([t[a, i] - t[b_, i] for i in range(b0, b1)])
    print("mean ld (S_in_reg - s_full)      ", d(1, 0))
    print("mean math (exp_done - S_in_reg)  ", d(2, 1))
    print("mean o_full wait                 ", d(3, 2))
    print("mean P store (P_stored - o_full) ", d(4, 3))
    print("mean s_full(j+1) - S_issue(j+1)  ", np.mean([t[0, i] - t[5, i] for i in range(b0, b1)]))
    print("mean S_issue(j+1) - S_in_reg(j)  ", np.mean([t[5, i + 1] - t[1, i] for i in range(b0, b1 - 1)]))
    print("mean PV_issue(j) - P_stored(j)   ", np.mean([t[6, i] - t[4, i] for i in range(b0, b1)]))
    print("mean block period                ", np.mean([t[4, i + 1] - t[4, i] for i in range(b0, b1 - 1)]))

if __name__ == "__main__":
    if sys.argv[1] == "build":
        build()
    else:
        run(*[int(x) for x in sys.argv[2:]])
