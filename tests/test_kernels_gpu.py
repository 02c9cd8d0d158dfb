"""-m gpu parity tests of the individual CUDA kernels, called through the C ABI
(video_style_transfer_b200.ops -> ctypes -> libvst_b200.so) and checked against the CPU oracle
(oracle/hotpath.py) on the same seeded inputs.

Tolerances (BASELINE.json north_star): layout kernels bit-exact; bf16 attention / GEMM outputs
within 1e-2 max relative error vs the fp32 oracle, defined per tensor as max|a-b| / max|b|.
"""
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
REL_TOL = 1e-2


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def ops():
    from video_style_transfer_b200 import ops as o
    return o


@pytest.fixture(scope="module")
def hp():
    from oracle import hotpath
    return hotpath


# ---- a9: layout (bit-exact) ----------------------------------------------------------------------
def test_rearrange_golden_fp32(ops):
    g = np.load(os.path.join(GOLD, "layout.npz"))
    x = torch.from_numpy(g["rearr_x"]).cuda()
    from oracle.hotpath import positional_table
    pe = positional_table(24).cuda()
    y = ops.rearrange_fwd(x, pe)
    assert torch.equal(y.cpu(), torch.from_numpy(g["rearr_y"]))
    back = ops.rearrange_bwd(y, 2, 3, 5)
    assert torch.equal(back.cpu(), torch.from_numpy(g["rearr_back"]))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("shape", [(1, 320, 16, 32, 32), (2, 64, 8, 8, 8), (1, 72, 5, 6, 4),
                                   (1, 640, 16, 16, 16), (1, 8, 1, 1, 8)])
@pytest.mark.parametrize("frames_major", [False, True])
def test_rearrange_bit_exact(ops, hp, dtype, shape, frames_major):
    torch.manual_seed(0)
    b, c, f, h, w = shape
    if frames_major:   # live layout: (B*F,C,H,W) storage viewed as (B,C,F,H,W)
        x = torch.randn(b, f, c, h, w).to(dtype).cuda().permute(0, 2, 1, 3, 4)
    else:
        x = torch.randn(b, c, f, h, w).to(dtype).cuda()
    pe = hp.positional_table(c).cuda()
    y = ops.rearrange_fwd(x, None)
    assert torch.equal(y.cpu(), hp.rearrange_fwd(x.cpu()))
    y_pe = ops.rearrange_fwd(x, pe)
    want = (hp.rearrange_fwd(x.cpu().float()) + pe.cpu()[:, :f]).to(dtype)   # one fp32 add, one rounding
    assert torch.equal(y_pe.cpu(), want)
    back = ops.rearrange_bwd(y, b, h, w, frames_major=frames_major)
    assert torch.equal(back.cpu(), x.cpu())                                  # round trip
    res = torch.randn(x.shape).to(dtype).cuda()
    if frames_major:
        res = torch.randn(b, f, c, h, w).to(dtype).cuda().permute(0, 2, 1, 3, 4)
    back_r = ops.rearrange_bwd(y, b, h, w, residual5=res, frames_major=frames_major)
    assert torch.equal(back_r.cpu(), (x.cpu().float() + res.cpu().float()).to(dtype))


def test_rearrange_empty(ops):
    x = torch.empty(0, 64, 16, 8, 8, dtype=torch.bfloat16, device="cuda")
    assert ops.rearrange_fwd(x).shape == (0, 16, 64)


# ---- a10 / K9: norms ---------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(1, 320, 16, 32, 32), (2, 64, 4, 8, 8), (1, 96, 3, 5, 3)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_groupnorm5d_rearrange(ops, hp, shape, dtype):
    torch.manual_seed(1)
    b, c, f, h, w = shape
    x = (torch.randn(b, f, c, h, w) * 2 + 0.5).to(dtype).cuda().permute(0, 2, 1, 3, 4)
    gamma, beta = torch.randn(c).cuda(), torch.randn(c).cuda()
    y, mean, rstd = ops.groupnorm5d_rearrange(x, gamma, beta, 32, 1e-6)
    xr = x.cpu().float()
    want = hp.rearrange_fwd(hp.groupnorm5d(xr, gamma.cpu(), beta.cpu(), 32, 1e-6))
    mu = xr.reshape(b, 32, -1).mean(-1).reshape(-1)
    assert torch.allclose(mean.cpu(), mu, atol=1e-5, rtol=1e-5)
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
    assert rel_err(y, want) < tol


@pytest.mark.parametrize("layout", ["frames", "channels"])
@pytest.mark.parametrize("shape", [(2, 320, 5, 32, 32),     # HW a power of two, 2 slabs per (group, frame), ragged last
                                   (1, 64, 3, 6, 4),        # HW = 24: 16-byte friendly but not a power of two
                                   (3, 96, 2, 16, 24),      # HW = 384, three batches
                                   (1, 32, 4, 5, 3)])       # HW = 15: the scalar kernel
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_groupnorm5d_stats_addressing(ops, shape, dtype, layout):
    """Every addressing branch of the statistics kernel (csrc/norm.cu: channels of a group contiguous /
    shift-and-mask / division / scalar) against fp64 statistics, on frames-major (B*F, C, H, W) activations --
    the path's layout -- and on a contiguous (B, C, F, H, W) tensor (the reference's view, attention_processor.py
    :33-36 / GroupNorm over the 5-D view)."""
    torch.manual_seed(5)
    b, c, f, h, w = shape
    if layout == "frames":
        x = (torch.randn(b, f, c, h, w) * 1.7 - 0.4).to(dtype).cuda().permute(0, 2, 1, 3, 4)
    else:
        x = (torch.randn(b, c, f, h, w) * 1.7 - 0.4).to(dtype).cuda()
    groups = 32 if c % 32 == 0 and c >= 64 else 8
    mean, rstd = ops.groupnorm5d_stats(x, groups, 1e-6)
    xr = x.double().cpu().reshape(b, groups, -1)
    assert torch.allclose(mean.cpu().double(), xr.mean(-1).reshape(-1), atol=2e-6, rtol=1e-5)
    want_rstd = (xr.var(-1, unbiased=False) + 1e-6).rsqrt().reshape(-1)
    assert torch.allclose(rstd.cpu().double(), want_rstd, atol=0, rtol=2e-5)


@pytest.mark.parametrize("c", [320, 640, 1280, 64])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_layernorm(ops, hp, c, dtype):
    torch.manual_seed(2)
    f = 16
    x = (torch.randn(37 * f, c) * 1.5 + 0.3).to(dtype).cuda()
    g, b = torch.randn(c).cuda(), torch.randn(c).cuda()
    y = ops.layernorm(x, g, b, 1e-5)
    want = hp.layernorm(x.cpu().float(), g.cpu(), b.cpu(), 1e-5)
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
    assert rel_err(y, want) < tol
    pe = hp.positional_table(c).cuda()
    y2 = ops.layernorm(x, g, b, 1e-5, pe=pe, frames=f)
    want2 = want.reshape(37, f, c) + pe.cpu()[:, :f]
    assert rel_err(y2, want2.reshape(-1, c)) < tol


@pytest.mark.parametrize("pdtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("c,f,extra", [(320, 16, 0), (320, 5, 3), (320, 7, 1), (640, 16, 1), (640, 3, 0),
                                       (1280, 16, 0), (1280, 6, 5), (320, 32, 2)])
def test_layernorm_persistent_path(ops, hp, c, f, extra, pdtype):
    """Row counts large enough for the persistent flat kernel (>= 2 groups per resident warp), ragged last
    group, frame counts that do not divide the rows a warp holds (the grid is then trimmed so that a warp's
    position-table rows stay fixed): bit-exact against the torch restatement with the same two roundings."""
    torch.manual_seed(12)
    rows = 6000 * (160 // (c // 8)) + extra          # > 2 * 8 warps * 2 blocks * 148 SMs groups of 160 vectors
    x = (torch.randn(rows, c) * 1.5 + 0.3).to(torch.bfloat16).cuda()
    # (bf16 parameters: the inference stack; fp32 parameters: the training stack's master weights)
    g, b = torch.randn(c).to(pdtype).cuda(), torch.randn(c).to(pdtype).cuda()
    pe = hp.positional_table(c).cuda()
    got = ops.layernorm(x, g, b, 1e-5, pe=pe, frames=f)
    xf = x.float()
    mu = xf.mean(-1, keepdim=True)
    var = ((xf - mu) ** 2).mean(-1, keepdim=True)
    ln = ((xf - mu) * torch.rsqrt(var + 1e-5) * g.float() + b.float())
    fr = torch.arange(rows, device="cuda") % f
    want = (ln.to(torch.bfloat16).float() + pe[0, fr]).to(torch.bfloat16)
    # fp32 statistics are summed in a different order than torch's: allow one bf16 ulp on a few elements
    diff = (got.float() - want.float()).abs()
    assert float(diff.max()) <= 2.0 ** -7 * float(want.float().abs().max())
    assert float((diff > 0).float().mean()) < 0.02
    plain = ops.layernorm(x, g, b, 1e-5)
    assert rel_err(plain, ln) < 1e-2
    bare = ops.layernorm(x, None, None, 1e-5)                # no parameters at all
    assert rel_err(bare, (xf - mu) * torch.rsqrt(var + 1e-5)) < 1e-2
    # the last (ragged) rows and the first rows of the second sweep of the grid
    assert torch.isfinite(got.float()).all()


# ---- a8: temporal attention ------------------------------------------------------------------------
@pytest.mark.parametrize("f,heads,d", [(16, 8, 40), (16, 8, 80), (16, 8, 160), (32, 8, 40),
                                       (32, 8, 160), (8, 8, 8), (5, 4, 40), (24, 2, 64)])
def test_temporal_attn(ops, hp, f, heads, d):
    torch.manual_seed(3)
    n, c = 67, heads * d
    qkv = torch.randn(n, f, 3 * c).to(torch.bfloat16).cuda()       # packed projection output
    q, k, v = qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:]
    o = ops.temporal_attn(q, k, v, heads)

    def sp(t):
        return t.cpu().float().reshape(n, f, heads, d).transpose(1, 2)

    want = hp.sdpa(sp(q), sp(k), sp(v)).transpose(1, 2).reshape(n, f, c)
    assert rel_err(o, want) < REL_TOL


# ---- a3/a7: fused LoRA GEMM --------------------------------------------------------------------------
@pytest.mark.parametrize("m,n,k1,k2", [(128, 128, 64, 0), (128, 64, 128, 0), (256, 256, 256, 64),
                                       (1000, 640, 640, 128), (300, 320, 320, 32),
                                       (77, 1280, 2048, 128), (21, 40, 48, 16), (513, 960, 320, 96),
                                       # large enough for the CTA-pair (cta_group::2) kernel
                                       (19200, 1280, 1280, 128), (19000, 640, 640, 128),
                                       (37888, 384, 320, 32)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_lora_gemm(ops, m, n, k1, k2, out_dtype):
    torch.manual_seed(4)
    a1 = torch.randn(m, k1).to(torch.bfloat16).cuda()
    b1 = (torch.randn(n, k1) / math.sqrt(k1)).to(torch.bfloat16).cuda()
    a2 = b2 = None
    if k2:
        a2 = torch.randn(m, k2).to(torch.bfloat16).cuda()
        b2 = (torch.randn(n, k2) / math.sqrt(k2)).to(torch.bfloat16).cuda()
    bias = torch.randn(n).cuda()
    res = torch.randn(m, n).to(out_dtype).cuda()
    out = ops.lora_gemm(a1, b1, a2, b2, bias=bias, residual=res, out_scale=0.5, out_dtype=out_dtype)
    want = a1.cpu().float() @ b1.cpu().float().t() + bias.cpu() + res.cpu().float()
    if k2:
        want = want + a2.cpu().float() @ b2.cpu().float().t()
    want = want * 0.5
    assert rel_err(out, want) < (REL_TOL if out_dtype == torch.bfloat16 else 1e-4)
    plain = ops.lora_gemm(a1, b1, out_dtype=out_dtype)
    assert rel_err(plain, a1.cpu().float() @ b1.cpu().float().t()) < (
        REL_TOL if out_dtype == torch.bfloat16 else 1e-4)


@pytest.mark.parametrize("m,n,k", [(262144, 320, 320), (65536, 640, 640), (16384, 1280, 1280), (4100, 200, 72)])
def test_lora_gemm_residual_tma_epilogue(ops, m, n, k):
    """bf16 output + residual (to_out / ff.net.2 / proj_out of every module): the residual tile is
    TMA-loaded into the output staging buffers and added in place.  Full config-3 shapes, a strided
    residual view (row stride != N), and the exact identity D = 0 * A + R == R."""
    torch.manual_seed(m % 97)
    a = torch.randn(m, k).to(torch.bfloat16).cuda()
    w = (torch.randn(n, k) / math.sqrt(k)).to(torch.bfloat16).cuda()
    wide = torch.randn(m, n + 64).to(torch.bfloat16).cuda()
    res = wide[:, 8:8 + n]                                      # row stride n + 64, 16-byte aligned
    bias = torch.randn(n).cuda()
    out = ops.lora_gemm(a, w, bias=bias, residual=res, out_scale=0.25)
    want = (a.float() @ w.float().t() + bias + res.float()) * 0.25
    assert rel_err(out, want.cpu()) < REL_TOL
    zero = torch.zeros_like(a)
    assert torch.equal(ops.lora_gemm(zero, w, residual=res), res)


@pytest.mark.parametrize("m,n,k1,k2,gn", [(4096, 1280, 1280, 128, 0), (4100, 3840, 1280, 128, 1280),
                                          (8192, 1920, 640, 128, 640), (2048, 640, 640, 64, 0),
                                          (16384, 960, 320, 64, 320), (32768, 1280, 1280, 128, 0)])
def test_lora_gemm_fused_single_launch(ops, m, n, k1, k2, gn):
    """vst_lora_gemm_fused (u = x A^T computed inside the main kernel) against the two-launch form it replaces
    (bit-identical: same bf16 u, same accumulation order), against fp32, and its u_out against the u GEMM."""
    torch.manual_seed(m + n + k2)
    groups = n // gn if gn else 1
    x = torch.randn(m, k1).to(torch.bfloat16).cuda()
    w = (torch.randn(n, k1) / math.sqrt(k1)).to(torch.bfloat16).cuda()
    a = (torch.randn(groups * k2, k1) / math.sqrt(k1)).to(torch.bfloat16).cuda()
    bp = (torch.randn(n, k2) / math.sqrt(k2)).to(torch.bfloat16).cuda()
    bias = torch.randn(n).cuda()
    assert ops.lora_gemm_fused_supported(m, n, k1, k2, gn)
    y, u = ops.lora_gemm_fused_raw(x, w, a, bp, bias, 0.5, gn, save_u=True)
    u_ref = ops.lora_gemm(x, a)
    y_ref = ops.lora_gemm(x, w, u_ref, bp, bias=bias, out_scale=0.5, k2_group_n=gn)
    assert torch.equal(u, u_ref)
    assert torch.equal(y, y_ref)
    y_nou = ops.lora_gemm_fused(x, w, a, bp, bias, 0.5, gn)
    assert torch.equal(y_nou, y)
    sel = slice(0, 512)                                          # fp32 check on a slab
    uf = (x[sel].float() @ a.float().t()).to(torch.bfloat16).float()
    want = x[sel].float() @ w.float().t() + bias
    for g_ in range(groups):
        cols = slice(g_ * gn, (g_ + 1) * gn) if gn else slice(0, n)
        want[:, cols] += uf[:, g_ * k2:(g_ + 1) * k2] @ bp[cols].float().t()
    assert rel_err(y[sel], (want * 0.5).cpu()) < REL_TOL
    assert not ops.lora_gemm_fused_supported(1000, n, k1, k2, gn)       # small M: the two-launch form stays


def test_lora_gemm_fused_autograd_matches_two_launch_form(ops):
    """FusedLoraFn (one forward launch, u saved by the kernel) gives the gradients of the two-launch autograd
    graph (LoraGemmFn twice) and of fp32 autograd."""
    torch.manual_seed(8)
    m, n, k1, k2, gn = 4096, 1920, 640, 128, 640
    mk = lambda *s_, sc=1.0: (torch.randn(*s_) * sc).to(torch.bfloat16).cuda()
    x0, w0, a0, bp0 = mk(m, k1), mk(n, k1, sc=k1 ** -0.5), mk(3 * k2, k1, sc=k1 ** -0.5), mk(n, k2, sc=k2 ** -0.5)
    b0 = torch.randn(n).cuda()
    dy = mk(m, n, sc=0.1)

    def run(fused):
        ts = [t.clone().requires_grad_(True) for t in (x0, w0, a0, bp0, b0)]
        x, w, a, bp, b = ts
        if fused:
            y = ops.lora_gemm_fused(x, w, a, bp, b, 1.0, gn)
        else:
            y = ops.lora_gemm(x, w, ops.lora_gemm(x, a), bp, bias=b, k2_group_n=gn)
        y.backward(dy)
        return y.detach(), [t.grad for t in ts]

    y1, g1 = run(True)
    y2, g2 = run(False)
    assert torch.equal(y1, y2)
    for a_, b_, nm in zip(g1, g2, ("dx", "dw", "da", "dbp", "dbias")):
        # (the two-launch graph sums dy W and du A as two bf16 tensors; the fused backward accumulates both
        # operand pairs in one fp32 accumulator: they differ by bf16 rounding)
        assert rel_err(a_, b_.float().cpu()) < 1.5e-2, nm
    xf, wf, af, bpf = (t.float().cpu().requires_grad_(True) for t in (x0, w0, a0, bp0))
    uf = xf @ af.t()
    yf = xf @ wf.t() + b0.cpu()
    yf = yf + torch.cat([uf[:, i * k2:(i + 1) * k2] @ bpf[i * gn:(i + 1) * gn].t() for i in range(3)], dim=1)
    yf.backward(dy.float().cpu())
    for got, ref, nm in zip(g1[:4], (xf.grad, wf.grad, af.grad, bpf.grad), ("dx", "dw", "da", "dbp")):
        assert rel_err(got, ref) < 1.5e-2, nm


def test_lora_gemm_linearity(ops):
    """Size-independent property at a full config-3 shape (M=16384, C=1280): exact identity
    columns of B pick out columns of A, and doubling A doubles D (power-of-two scaling is exact)."""
    torch.manual_seed(5)
    m, k = 16384, 1280
    a = torch.randn(m, k).to(torch.bfloat16).cuda()
    eye = torch.eye(k).to(torch.bfloat16).cuda()
    d = ops.lora_gemm(a, eye)
    assert torch.equal(d, a)
    w = (torch.randn(k, k) / 36).to(torch.bfloat16).cuda()
    d1, d2 = ops.lora_gemm(a, w), ops.lora_gemm(a * 2, w)
    assert torch.equal(d1 * 2, d2)


def test_lora_pack_up(ops):
    torch.manual_seed(6)
    n, r = 40, 8
    bc, bs = torch.randn(n, r).cuda(), torch.randn(n, r).cuda()
    mc, ms = torch.rand(n).cuda(), torch.rand(n).cuda()
    mk = (torch.rand(n) > 0.5).cuda()
    out = ops.lora_pack_up(bc, mc, mk, bs, ms, None, 0.7)
    want = torch.cat([bc * (0.7 * mc * mk)[:, None], bs * (0.7 * ms)[:, None]], 1)
    assert rel_err(out, want) < 5e-3
    only_s = ops.lora_pack_up(None, None, None, bs, None, None, 1.0)
    assert torch.equal(only_s.cpu(), bs.cpu().to(torch.bfloat16))


# ---- a2: spatial / cross attention ------------------------------------------------------------------
@pytest.mark.parametrize("bt,bk,heads,sq,sk", [(2, 2, 2, 128, 128), (2, 2, 3, 256, 384),
                                               (4, 2, 2, 200, 77), (1, 1, 10, 1024, 1024),
                                               (3, 1, 1, 100, 300), (1, 1, 2, 64, 1),
                                               # persistent form, more items than resident CTAs:
                                               # several items per CTA, one / two key blocks each
                                               (16, 1, 4, 1024, 77), (16, 16, 5, 512, 200),
                                               # the 80-key cross-attention instantiation at its key-count
                                               # limits (65, 80), 5+ items per CTA (Q ring and O stage wrap),
                                               # ragged last query tile
                                               (40, 2, 8, 640, 65), (24, 3, 12, 520, 80)])
def test_attn_fwd(ops, hp, bt, bk, heads, sq, sk):
    torch.manual_seed(7)
    c = heads * 64
    q = torch.randn(bt, sq, c).to(torch.bfloat16).cuda()
    k = torch.randn(bk, sk, c).to(torch.bfloat16).cuda()
    v = torch.randn(bk, sk, c).to(torch.bfloat16).cuda()
    o, lse = ops.attn_fwd(q, k, v, heads, return_lse=True)
    g = bt // bk

    def sp(t, b, s):
        return t.cpu().float().reshape(b, s, heads, 64).transpose(1, 2)

    kk = sp(k, bk, sk).repeat_interleave(g, dim=0)
    vv = sp(v, bk, sk).repeat_interleave(g, dim=0)
    qq = sp(q, bt, sq)
    want = hp.sdpa(qq, kk, vv).transpose(1, 2).reshape(bt, sq, c)
    assert rel_err(o, want) < REL_TOL
    want_lse = torch.logsumexp(qq @ kk.transpose(-1, -2) / 8.0, dim=-1)
    assert torch.allclose(lse.cpu(), want_lse, atol=2e-2, rtol=1e-3)


@pytest.mark.parametrize("bt,bk,heads,sq,sk", [(16, 16, 20, 1024, 1024), (16, 1, 20, 1024, 77),
                                               (16, 16, 10, 4096, 4096)])
def test_attn_fwd_full_size_schedule_independence(ops, bt, bk, heads, sq, sk):
    """Config-3 shapes, where the oracle is too slow: size-independent properties of the persistent
    kernel.  (1) An item's result does not depend on which CTA walked it or on what that CTA did
    before (Q ring slot, deferred epilogue, barrier phases): one (batch, head) slice computed alone
    is BIT-identical to the same slice of the full launch.  (2) Attention is invariant under a
    joint permutation of keys and values (different block composition, same softmax): the outputs
    agree to bf16 rounding.  (3) Rows of P sum to one: with V = 1 the output is exactly 1."""
    torch.manual_seed(17)
    c = heads * 64
    q = torch.randn(bt, sq, c).to(torch.bfloat16).cuda()
    k = torch.randn(bk, sk, c).to(torch.bfloat16).cuda()
    v = torch.randn(bk, sk, c).to(torch.bfloat16).cuda()
    o, lse = ops.attn_fwd(q, k, v, heads, return_lse=True)
    b, h = bt - 3, heads - 2
    kb = b // (bt // bk)
    cols = slice(h * 64, (h + 1) * 64)
    o1, lse1 = ops.attn_fwd(q[b:b + 1, :, cols].contiguous(), k[kb:kb + 1, :, cols].contiguous(),
                            v[kb:kb + 1, :, cols].contiguous(), 1, return_lse=True)
    assert torch.equal(o1[0], o[b, :, cols])
    assert torch.equal(lse1.reshape(-1), lse[b, h].reshape(-1))
    perm = torch.randperm(sk, device="cuda")
    o2 = ops.attn_fwd(q, k[:, perm].contiguous(), v[:, perm].contiguous(), heads)
    assert rel_err(o2, o.float().cpu()) < REL_TOL
    ones = torch.ones_like(v)
    o3 = ops.attn_fwd(q, k, ones, heads)
    assert (o3.float() - 1.0).abs().max().item() <= 2 ** -7


@pytest.mark.parametrize("m,c,k1,k2,groups", [(300, 128, 64, 16, 3), (1000, 640, 640, 128, 3),
                                              (6400, 1280, 1280, 128, 3), (16384, 640, 640, 128, 3),
                                              (77, 1280, 2048, 128, 2), (513, 320, 320, 32, 3)])
def test_lora_gemm_grouped_low_rank(ops, m, c, k1, k2, groups):
    """q|k|v fused along N: output columns of group g read only their own K2 columns of A2."""
    torch.manual_seed(8)
    n = groups * c
    a1 = torch.randn(m, k1).to(torch.bfloat16).cuda()
    b1 = (torch.randn(n, k1) / math.sqrt(k1)).to(torch.bfloat16).cuda()
    a2 = torch.randn(m, groups * k2).to(torch.bfloat16).cuda()
    b2 = (torch.randn(n, k2) / math.sqrt(k2)).to(torch.bfloat16).cuda()
    out = ops.lora_gemm(a1, b1, a2, b2, k2_group_n=c)
    want = a1.cpu().float() @ b1.cpu().float().t()
    for g in range(groups):
        want[:, g * c:(g + 1) * c] += (a2.cpu().float()[:, g * k2:(g + 1) * k2]
                                       @ b2.cpu().float()[g * c:(g + 1) * c].t())
    assert rel_err(out, want) < REL_TOL


@pytest.mark.parametrize("m,n", [(128, 128), (300, 192), (4096, 256)])
def test_lora_gemm_identity_row_col_mapping(ops, m, n):
    """D = A I must reproduce A bit for bit: pins the TMEM-lane -> row and chunk -> column mapping
    of the TMA-store epilogue (values <= 256 are exact in bf16)."""
    eye = torch.eye(n).to(torch.bfloat16).cuda()
    rows = (torch.arange(m) % 251).float()[:, None].expand(m, n).to(torch.bfloat16).cuda().contiguous()
    cols = torch.arange(n).float()[None, :].expand(m, n).to(torch.bfloat16).cuda().contiguous()
    assert torch.equal(ops.lora_gemm(rows, eye), rows)
    assert torch.equal(ops.lora_gemm(cols, eye), cols)


# ---- a2 with the optional additive mask (attention_processor.py:46-48) ---------------------------------
@pytest.mark.parametrize("bt,bk,heads,sq,sk,mshape", [
    (2, 2, 2, 128, 128, (2, 1, 1, 128)),        # key-padding mask, broadcast over heads and queries
    (4, 2, 2, 200, 77, (4, 2, 1, 77)),          # per head, ragged Sk (scalar-load tail path)
    (2, 2, 3, 256, 384, (2, 3, 256, 384)),      # full (b, h, q, k) mask, 16-byte aligned rows
    (1, 1, 2, 64, 300, (1, 1, 64, 300)),        # per query, several key blocks + tail
])
def test_attn_fwd_additive_mask(ops, hp, bt, bk, heads, sq, sk, mshape):
    torch.manual_seed(11)
    c = heads * 64
    q = torch.randn(bt, sq, c).to(torch.bfloat16).cuda()
    k = torch.randn(bk, sk, c).to(torch.bfloat16).cuda()
    v = torch.randn(bk, sk, c).to(torch.bfloat16).cuda()
    mask = torch.randn(mshape) * 2.0
    mask[..., ::5] = -1e4                      # some keys (nearly) masked out
    mask[..., 0] = 0.0                         # never a fully masked row
    o, lse = ops.attn_fwd(q, k, v, heads, return_lse=True, mask=mask.cuda())
    g = bt // bk
    sp = lambda t, b, s: t.cpu().float().reshape(b, s, heads, 64).transpose(1, 2)
    kk, vv, qq = sp(k, bk, sk).repeat_interleave(g, 0), sp(v, bk, sk).repeat_interleave(g, 0), sp(q, bt, sq)
    want = hp.sdpa(qq, kk, vv, mask).transpose(1, 2).reshape(bt, sq, c)
    assert rel_err(o, want) < REL_TOL
    want_lse = torch.logsumexp(qq @ kk.transpose(-1, -2) / 8.0 + mask, dim=-1)
    assert torch.allclose(lse.cpu(), want_lse, atol=2e-2, rtol=1e-3)
    # bool masks follow F.scaled_dot_product_attention: False = excluded
    keep = torch.rand(bt, 1, 1, sk) > 0.3
    keep[..., 0] = True
    o2 = ops.attn_fwd(q, k, v, heads, mask=keep.cuda())
    add = torch.zeros(keep.shape).masked_fill(~keep, float("-inf"))
    want2 = hp.sdpa(qq, kk, vv, add).transpose(1, 2).reshape(bt, sq, c)
    assert rel_err(o2, want2) < REL_TOL


def test_attn_bwd_additive_mask(ops):
    torch.manual_seed(12)
    bt, heads, sq, sk = 2, 2, 200, 136
    c = heads * 64
    q, k, v, d_o = (torch.randn(bt, s, c).to(torch.bfloat16).cuda().requires_grad_(True)
                    for s in (sq, sk, sk, sq))
    mask = (torch.randn(bt, 1, 1, sk) * 2.0).cuda()
    o = ops.attn_fwd(q, k, v, heads, mask=mask)
    o.backward(d_o.detach())
    qr, kr, vr = (t.detach().float().cpu().requires_grad_(True) for t in (q, k, v))
    sp = lambda t, s: t.reshape(bt, s, heads, 64).transpose(1, 2)
    ref = torch.nn.functional.scaled_dot_product_attention(sp(qr, sq), sp(kr, sk), sp(vr, sk),
                                                           attn_mask=mask.cpu())
    ref.transpose(1, 2).reshape(bt, sq, c).backward(d_o.detach().float().cpu())
    for got, want, nm in ((q.grad, qr.grad, "dq"), (k.grad, kr.grad, "dk"), (v.grad, vr.grad, "dv")):
        assert rel_err(got, want) < 2e-2, nm


# ---- §8f-1: GEGLU in the GEMM epilogue (VST_ACT_GEGLU) ------------------------------------------------
@pytest.mark.parametrize("m,k,inner", [(300, 320, 1280), (4096, 640, 2560), (1000, 64, 128), (16384, 1280, 5120)])
@pytest.mark.parametrize("bias_dtype", [torch.bfloat16, torch.float32, None])
def test_lora_gemm_geglu_epilogue(ops, m, k, inner, bias_dtype):
    torch.manual_seed(13)
    x = torch.randn(m, k).to(torch.bfloat16).cuda()
    w = (torch.randn(2 * inner, k) / k ** 0.5).to(torch.bfloat16).cuda()
    b = None if bias_dtype is None else torch.randn(2 * inner).to(bias_dtype).cuda()
    got = ops.lora_gemm(x, w, bias=b, act="geglu")
    assert got.shape == (m, inner)
    z = x.float() @ w.float().t()
    if b is not None:
        z = z + b.float()
    want = z[:, :inner] * torch.nn.functional.gelu(z[:, inner:])
    assert rel_err(got, want) < REL_TOL
    # and against the two-kernel path it replaces (bf16-rounded pre-activation)
    two = ops.geglu(ops.lora_gemm(x, w, bias=b))
    assert rel_err(got, two) < REL_TOL


# ---- §8f-4: CFG combine + Euler update ----------------------------------------------------------------
@pytest.mark.parametrize("edt,xdt", [(torch.bfloat16, torch.float32), (torch.bfloat16, torch.bfloat16),
                                     (torch.float32, torch.float32)])
def test_cfg_euler_step(ops, edt, xdt):
    torch.manual_seed(3)
    shape = (1, 4, 16, 32, 32)
    eu, ec = torch.randn(shape).to(edt), torch.randn(shape).to(edt)
    x = torch.randn(shape).to(xdt)
    g, s, sn = 7.5, 3.1, 2.2
    out, nxt = ops.cfg_euler_step(eu.cuda(), ec.cuda(), x.cuda(), g, s, sn, want_scaled_next=True)
    # inference_animatediff.py:122 in the UNet dtype, then EulerDiscreteScheduler.step in fp32
    noise = eu + g * (ec - eu)
    want = (x.float() + (sn - s) * noise.float()).to(xdt)
    assert out.dtype == xdt and nxt.dtype == edt
    tol = 0 if edt == torch.float32 else 1e-2
    assert rel_err(out, want) <= max(tol, 1e-6)
    assert rel_err(nxt, want.float() / (sn * sn + 1) ** 0.5) < 1e-2
    assert torch.equal(ops.cfg_euler_step(eu.cuda(), ec.cuda(), x.cuda(), g, s, sn), out)
