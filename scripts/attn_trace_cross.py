"""Pipeline trace of the persistent attention kernel on the TEXT CROSS-ATTENTION shape (Sk = 77: one 80-key block per
item, so item index == block index).  Tuning aid; build first with `python scripts/attn_trace.py build`.
Stamps (softmax warp 0 unless noted): 0 s_full seen, 1 S in registers, 2 exp loop done, 4 P stored, 7 o_full seen in the
deferred epilogue of the PREVIOUS item, 8 its O staged; MMA warp: 5 S MMA issue, 6 PV MMA issue."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "scripts", "ubench", "libvst_attn_trace.so")


def run(sq=1024, heads=20, bt=32, bk=2, sk=77):
    lib = ctypes.CDLL(LIB)
    c = heads * 64
    q = torch.randn(bt, sq, c, device="cuda").to(torch.bfloat16)
    k, v = (torch.randn(bk, sk, c, device="cuda").to(torch.bfloat16) for _ in range(2))
    o = torch.empty_like(q)
    P, L, I, F = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float
    lib.vst_attn_fwd.argtypes = [P, P, P, P, P, I, I, I, I, I, L, L, L, L, I, F, P, L, L, L, P]
    for _ in range(3):
        rc = lib.vst_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), None, bt, heads, sq, sk, 64,
                              c, c, c, c, bt // bk, 0.125, None, 0, 0, 0, None)
        assert rc == 0, rc
        torch.cuda.synchronize()
    buf = np.zeros(4 * 12 * 1024, dtype=np.int64)
    assert lib.vst_attn_debug_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
    full = buf.reshape(4, 12, 1024)
    t = full[0].copy()
    t[5], t[6] = full[1][5], full[1][6]
    n_items = int((t[4] != 0).sum())
    print("items traced on CTA 0:", n_items)
    t0 = t[5, 0]
    names = ["S_issue", "s_full", "S_in_reg", "exp_done", "epi_ofull", "epi_staged", "P_stored", "PV_issue"]
    slots = [5, 0, 1, 2, None, None, 4, 6]
    print("item " + " ".join(f"{x:>10s}" for x in names) + "     period")
    prev = None
    for i in range(min(n_items, 24)):
        row = []
        for nm, s in zip(names, slots):
            if nm == "epi_ofull":
                row.append(int(t[7, i - 1] - t0) if i > 0 and t[7, i - 1] else -1)
            elif nm == "epi_staged":
                row.append(int(t[8, i - 1] - t0) if i > 0 and t[8, i - 1] else -1)
            else:
                row.append(int(t[s, i] - t0) if t[s, i] else -1)
        per = row[6] - prev if prev is not None else 0
        prev = row[6]
        print(f"{i:4d} " + " ".join(f"{x:10d}" for x in row) + f" {per:10d}")
    a, b = 4, min(n_items, 1000) - 2
    m = lambda f: float(np.mean([f(i) for i in range(a, b)]))
    print("steady state (mean over items %d..%d), clocks:" % (a, b))
    print("  item period (P_stored to P_stored)   ", m(lambda i: t[4, i + 1] - t[4, i]))
    print("  s_full - S_issue                      ", m(lambda i: t[0, i] - t[5, i]))
    print("  S load (S_in_reg - s_full)            ", m(lambda i: t[1, i] - t[0, i]))
    print("  math (exp_done - S_in_reg)            ", m(lambda i: t[2, i] - t[1, i]))
    print("  epilogue wait (epi_ofull - exp_done)  ", m(lambda i: t[7, i - 1] - t[2, i]))
    print("  epilogue body (epi_staged - epi_ofull)", m(lambda i: t[8, i - 1] - t[7, i - 1]))
    print("  P store (P_stored - epi_staged)       ", m(lambda i: t[4, i] - t[8, i - 1]))
    print("  next s_full seen - P_stored           ", m(lambda i: t[0, i + 1] - t[4, i]))
    print("  S_issue(i+1) - S_in_reg(i)            ", m(lambda i: t[5, i + 1] - t[1, i]))
    print("  PV_issue(i) - P_stored(i)             ", m(lambda i: t[6, i] - t[4, i]))
    print("  PV_issue(i) - S_in_reg(i+1)           ", m(lambda i: t[6, i] - t[1, i + 1]))
    mm = full[1]                      # the MMA warp's bank: 9 loop top of item i+1's S, 10 q_full seen, 11 s_free seen
    print("  MMA warp, item i+1's S: loop top - S_in_reg(i)", m(lambda i: mm[9, i + 1] - t[1, i]))
    print("     q_full wait                          ", m(lambda i: mm[10, i + 1] - mm[9, i + 1]))
    print("     s_free wait                          ", m(lambda i: mm[11, i + 1] - mm[10, i + 1]))
    print("     k_full wait                          ", m(lambda i: t[5, i + 1] - mm[11, i + 1]))
    print("  other softmax warps: S_in_reg skew vs warp 0", [m(lambda i, x=x: full[x][1, i] - full[0][1, i]) for x in range(4)])


if __name__ == "__main__":
    run(*[int(x) for x in sys.argv[1:]])
