"""-m gpu parity tests of the backward kernels (vst_gemm = lora_gemm_bwd_dx /
_bwd_params, vst_attn_bwd, vst_temporal_attn_bwd, norm / GEGLU gradients), through the C ABI,
against fp32 torch-CPU autograd of the oracle functions on the same seeded inputs.
Tolerance: bf16 operands, fp32 accumulation -> max|a-b| / max|b| <= 1e-2 per tensor."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
REL_TOL = 1e-2


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def ops():
    from video_style_transfer_b200 import ops as o
    return o


def _bf(t):
    return t.to(torch.bfloat16)


# ---- vst_gemm: every transpose combination, ragged edges, split-K, two operand pairs -------------
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (200, 136, 328), (1000, 264, 72), (136, 128, 16384),
                                   (64, 1288, 4104),
                                   (392, 320, 200), (256, 640, 2056),       # tile widths 160 and 224 (N = 320, 640)
                                   (5000, 1000, 136)])     # 80 pair tiles: more than one per SM pair (persistent walk)
def test_gemm_transposes(ops, ta, tb, m, n, k):
    if ta and m % 8:
        pytest.skip("MN-major A needs M % 8 == 0")
    if (not ta or not tb) and k % 8:
        pytest.skip("K-major operands need K % 8 == 0")
    torch.manual_seed(m + n + k + 2 * ta + tb)
    a = _bf(torch.randn(m, k))
    b = _bf(torch.randn(n, k))
    want = a.float() @ b.float().T
    ad = (a.T.contiguous() if ta else a).cuda()
    bd = (b.T.contiguous() if tb else b).cuda()
    got = ops.gemm(ad, bd, trans_a=bool(ta), trans_b=bool(tb), out_dtype=torch.float32)
    assert rel_err(got, want) < 2e-3      # fp32 output of bf16 operands: only summation order differs
    got16 = ops.gemm(ad, bd, trans_a=bool(ta), trans_b=bool(tb), alpha=0.5)
    assert rel_err(got16, 0.5 * want) < REL_TOL


def test_gemm_two_pairs_accumulate(ops):
    """dx = dy W + du A (transB = 1, two operand pairs) and D += (accumulate)."""
    torch.manual_seed(3)
    m, n, k1, k2 = 520, 640, 640, 128
    dy, w = _bf(torch.randn(m, k1)), _bf(torch.randn(k1, n))       # W stored [out=k1, in=n]
    du, a = _bf(torch.randn(m, k2)), _bf(torch.randn(k2, n))
    want = dy.float() @ w.float() + du.float() @ a.float()
    got = ops.gemm(dy.cuda(), w.cuda(), du.cuda(), a.cuda(), trans_b=True, out_dtype=torch.float32)
    assert rel_err(got, want) < 2e-3
    base = torch.randn(m, n)
    out = base.clone().cuda()
    ops.gemm(dy.cuda(), w.cuda(), du.cuda(), a.cuda(), trans_b=True, out=out, accumulate=True, alpha=2.0)
    assert rel_err(out, base + 2.0 * want) < 2e-3


def test_gemm_split_k_is_deterministic(ops):
    torch.manual_seed(4)
    tokens, n_out, r2 = 16384, 1280, 128
    dy, u = _bf(torch.randn(tokens, n_out)).cuda(), _bf(torch.randn(tokens, r2)).cuda()
    a = ops.gemm(dy, u, trans_a=True, trans_b=True, out_dtype=torch.float32)
    b = ops.gemm(dy, u, trans_a=True, trans_b=True, out_dtype=torch.float32)
    assert torch.equal(a, b)
    want = dy.float().T @ u.float()
    assert rel_err(a, want) < 2e-3


# ---- a4 / a7: parameter and input gradients of the LoRA layers against the reference's own autograd
#      results (tests/golden/*.npz, written by oracle/make_golden.py from the reference modules) ----
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _sub(g, prefix):
    return {k[len(prefix):]: torch.from_numpy(np.asarray(g[k])) for k in g.files if k.startswith(prefix)}


def _close(got, want, what, tol=2e-2):
    assert got is not None, f"{what}: no gradient"
    assert got.shape == want.shape, (what, got.shape, want.shape)
    e = rel_err(got, want)
    assert e < tol, f"{what}: rel err {e}"


@pytest.mark.parametrize("ft", ["both", "content", "style"])
@pytest.mark.parametrize("masked", [0, 1])
def test_unziplora_train_gradients_golden(ft, masked):
    """unziplora_linear_layer.py:221-263 forward + backward: dx (both streams), dA, dB and the merger
    gradients of both branches, hard column masks on and off."""
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    g = np.load(os.path.join(GOLD, "unziplora.npz"))
    p, t = _sub(g, "infer/"), _sub(g, "train/")
    lay = UnZipLoRALinearLayer(48, 40, 8, ["content", "style"], lora_matrix_num=2)
    with torch.no_grad():
        for k in ("content_down", "content_up", "style_down", "style_up"):
            lay.lora_matrix_dic[k].weight.copy_(t[k])
        lay.merge_content.copy_(t["merge_content"])
        lay.merge_style.copy_(t["merge_style"])
    lay = lay.cuda()
    lay.mask_content, lay.mask_style = t["mask_content"].cuda(), t["mask_style"].cuda()
    lay.set_forward(ft)
    lay.set_layer_mask("content", bool(masked))
    lay.set_layer_mask("style", bool(masked))
    x = p["x"].cuda().requires_grad_(True)
    x2 = p["x2"].cuda().requires_grad_(True)
    y = lay(x, x2)
    w = _sub(g, f"train/{ft}_{masked}/")
    _close(y, w["y"], "y")
    y.backward(t["dy"].cuda())
    d = lay.lora_matrix_dic
    if "dx" in w:
        _close(x.grad, w["dx"], "dx")
    if "dx2" in w:
        _close(x2.grad, w["dx2"], "dx2")
    for key in ("content", "style"):
        if f"d_{key}_down" in w:
            _close(d[f"{key}_down"].weight.grad, w[f"d_{key}_down"], f"d_{key}_down")
            _close(d[f"{key}_up"].weight.grad, w[f"d_{key}_up"], f"d_{key}_up")
    if "dmc" in w:
        _close(lay.merge_content.grad, w["dmc"], "dmc")
        _close(lay.merge_style.grad, w["dms"], "dms")


def test_temporal_lora_gradients_golden():
    """temporal_lora.py:29-32: dx, dA, dB with the base frozen (its grads must stay None)."""
    from video_style_transfer_b200.animatediff.temporal_lora import TemporalLoRALinear
    g = np.load(os.path.join(GOLD, "lora_linear.npz"))
    T = lambda a: torch.from_numpy(np.asarray(a))
    base = torch.nn.Linear(48, 40)
    with torch.no_grad():
        base.weight.copy_(T(g["t_weight"]))
        base.bias.copy_(T(g["t_bias"]))
    m = TemporalLoRALinear(base, rank=4, alpha=2.0)
    with torch.no_grad():
        m.lora_A.copy_(T(g["t_A"]))
        m.lora_B.copy_(T(g["t_B"]))
    m = m.cuda()
    x = T(g["t_x"]).cuda().requires_grad_(True)
    y = m(x)
    _close(y, T(g["t_y"]), "y")
    y.backward(T(g["t_dy"]).cuda())
    _close(x.grad, T(g["t_dx"]), "dx")
    _close(m.lora_A.grad, T(g["t_dA"]), "dA")
    _close(m.lora_B.grad, T(g["t_dB"]), "dB")
    assert m.base.weight.grad is None and m.base.bias.grad is None


def test_plain_linear_weight_and_bias_gradients(ops):
    """Trainable projections of the motion module (proj_in / proj_out / feed-forward keep their
    weights trainable, animatediff/utils.py:66-96): dW = dy^T x, db = colsum(dy), dx = dy W."""
    from video_style_transfer_b200._host import fused_linear
    torch.manual_seed(5)
    m, k, n = 1000, 320, 640
    x = torch.randn(m, k).to(torch.bfloat16)
    w = torch.nn.Parameter(torch.randn(n, k) * 0.05)
    b = torch.nn.Parameter(torch.randn(n))
    res = torch.randn(m, n).to(torch.bfloat16)
    dy = torch.randn(m, n).to(torch.bfloat16)
    xr, rr = x.float().requires_grad_(True), res.float().requires_grad_(True)
    want = (torch.nn.functional.linear(xr, w.detach().to(torch.bfloat16).float().requires_grad_(True), b.detach()) + rr) * 0.5
    wref = [t for t in (xr, rr)]
    # reference gradients in fp32 on the bf16-rounded operands
    w32 = w.detach().to(torch.bfloat16).float().requires_grad_(True)
    b32 = b.detach().clone().requires_grad_(True)
    y32 = (torch.nn.functional.linear(xr, w32, b32) + rr) * 0.5
    y32.backward(dy.float())
    wd, bd = torch.nn.Parameter(w.detach().cuda()), torch.nn.Parameter(b.detach().cuda())
    xd, rd = x.cuda().requires_grad_(True), res.cuda().requires_grad_(True)
    y = fused_linear(xd, wd, bd, None, rd, 0.5)
    _close(y, y32.detach(), "y")
    y.backward(dy.cuda())
    _close(xd.grad, xr.grad, "dx")
    _close(rd.grad, rr.grad, "dres")
    _close(wd.grad, w32.grad, "dW")
    _close(bd.grad, b32.grad, "db")


# ---- a2 backward: flash-attention gradients vs fp32 autograd of the same bf16 inputs -----------------
def _sdpa_ref(q, k, v, heads, d_o):
    bt, sq, c = q.shape
    bk, sk, _ = k.shape
    grp = bt // bk
    qf, kf, vf = (t.float().requires_grad_(True) for t in (q, k, v))
    qh = qf.view(bt, sq, heads, 64).transpose(1, 2)
    kh = kf.repeat_interleave(grp, dim=0).view(bt, sk, heads, 64).transpose(1, 2)
    vh = vf.repeat_interleave(grp, dim=0).view(bt, sk, heads, 64).transpose(1, 2)
    o = torch.nn.functional.scaled_dot_product_attention(qh, kh, vh).transpose(1, 2).reshape(bt, sq, c)
    o.backward(d_o.float())
    return o.detach(), qf.grad, kf.grad, vf.grad


@pytest.mark.parametrize("bt,bk,heads,sq,sk", [(2, 2, 2, 128, 128), (2, 2, 3, 256, 384), (3, 3, 2, 200, 136),
                                               (4, 1, 2, 256, 77), (2, 2, 1, 64, 520), (6, 2, 2, 130, 77),
                                               # shared K / V with a long ragged query walk: the launch splits the walk
                                               # over three CTAs per key tile (9 query tiles, the last one 76 rows)
                                               (4, 1, 2, 1100, 77),
                                               # 11 query tiles against 3 key tiles: the three-stage Q / dO ring wraps,
                                               # dQ tiles leave by bulk tensor reduction with a clipped last tile
                                               (1, 1, 1, 1300, 300)])
def test_attn_bwd_vs_autograd(ops, bt, bk, heads, sq, sk):
    torch.manual_seed(bt * 1000 + sq + sk)
    c = heads * 64
    q, k, v = _bf(torch.randn(bt, sq, c)), _bf(torch.randn(bk, sk, c)), _bf(torch.randn(bk, sk, c))
    d_o = _bf(torch.randn(bt, sq, c))
    o_ref, dq_ref, dk_ref, dv_ref = _sdpa_ref(q, k, v, heads, d_o)
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    o = ops.attn_fwd(qd, kd, vd, heads)
    assert rel_err(o, o_ref) < REL_TOL
    o.backward(d_o.cuda())
    assert rel_err(qd.grad, dq_ref) < REL_TOL, "dq"
    assert rel_err(kd.grad, dk_ref) < REL_TOL, "dk"
    assert rel_err(vd.grad, dv_ref) < REL_TOL, "dv"


def test_attn_bwd_packed_qkv_views(ops):
    """q | k | v as column views of one packed projection (row stride 3C), as the processor feeds them."""
    torch.manual_seed(11)
    bt, heads, s = 2, 2, 256
    c = heads * 64
    qkv = _bf(torch.randn(bt, s, 3 * c))
    d_o = _bf(torch.randn(bt, s, c))
    o_ref, dq_ref, dk_ref, dv_ref = _sdpa_ref(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], heads, d_o)
    x = qkv.cuda().requires_grad_(True)
    o = ops.attn_fwd(x[..., :c], x[..., c:2 * c], x[..., 2 * c:], heads)
    o.backward(d_o.cuda())
    want = torch.cat([dq_ref, dk_ref, dv_ref], dim=-1)
    assert rel_err(x.grad, want) < REL_TOL


# (config 2 at its full (16, 1024, 1280) size: tests/test_parity_full_gpu.py)


# ---- a8 / K9 / §8f-1 / a10 backward kernels vs fp32 autograd --------------------------------------------


@pytest.mark.parametrize("m,n,k,r", [(300, 136, 72, 16), (1024, 640, 640, 128), (77, 64, 2048, 0), (4096, 640, 640, 64)])
def test_lora_linear_c_abi_fwd_bwd(m, n, k, r):
    """SURVEY §8b names: vst_lora_gemm_fwd, vst_lora_gemm_bwd_dx, vst_lora_gemm_bwd_params called raw
    through ctypes, against fp32 autograd of y = x W^T + (x A^T) B'^T on the same bf16 operands."""
    from video_style_transfer_b200 import _lib
    lib = _lib.lib
    torch.manual_seed(0)
    dev = "cuda"
    bf = lambda *s_, sc=1.0: (torch.randn(*s_, device=dev) * sc).to(torch.bfloat16)
    x, w, dy = bf(m, k), bf(n, k, sc=k ** -0.5), bf(m, n)
    a = bf(r, k, sc=k ** -0.5) if r else None
    bp = bf(n, r, sc=r ** -0.5) if r else None
    st = torch.cuda.current_stream().cuda_stream
    P = lambda t: 0 if t is None else t.data_ptr()
    # forward through the MODULE-level entry: raw A / B in, u saved for the backward (one branch, no merger,
    # bf16 parameters so the packed operands are A and B themselves)
    u = torch.empty(m, r, dtype=torch.bfloat16, device=dev) if r else None
    y = torch.empty(m, n, dtype=torch.bfloat16, device=dev)
    wsb = lib.vst_lora_gemm_fwd_workspace_bytes(m, n, k, max(r, 8), 1, 1)
    wsf = torch.empty(wsb + 256, dtype=torch.uint8, device=dev)
    wsp = wsf.data_ptr() + (-wsf.data_ptr()) % 256
    _lib.check(lib.vst_lora_gemm_fwd(P(x), k, 0, 0, 0, 0, P(w), k, 0, 0, P(a), P(bp), 0, 0, 0, 0, 0, 0, r, 1, 1.0, 0,
                                     0, 0, 1.0, P(u), r, P(y), n, m, n, k, k, 1, wsp, wsb, st), "vst_lora_gemm_fwd")
    # fp32 reference
    xf, wf = x.float().requires_grad_(True), w.float().requires_grad_(True)
    af = a.float().requires_grad_(True) if r else None
    bpf = bp.float().requires_grad_(True) if r else None
    yf = xf @ wf.t()
    if r:
        yf = yf + (xf @ af.t()) @ bpf.t()
    assert rel_err(y, yf) < 1e-2
    yf.backward(dy.float())
    # backward
    du = torch.empty(m, r, dtype=torch.bfloat16, device=dev) if r else None
    dx = torch.empty(m, k, dtype=torch.bfloat16, device=dev)
    _lib.check(lib.vst_lora_gemm_bwd_dx(P(dy), n, P(w), k, P(bp), r, P(a), k, r, P(du), r, P(dx), k, m, n, k, 1.0, 0,
                                        1, st), "dx")
    assert rel_err(dx, xf.grad) < 1e-2
    wsb = lib.vst_lora_gemm_bwd_params_workspace_bytes(m, n, k, k, r)
    assert wsb >= 0
    dims = (ctypes.c_int64 * 5)(m, n, k, k, r)
    assert lib.vst_workspace_bytes(14, dims, 5) == wsb
    ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=dev)
    dw = torch.empty(n, k, dtype=torch.float32, device=dev)
    da = torch.empty(r, k, dtype=torch.float32, device=dev) if r else None
    dbp = torch.empty(n, r, dtype=torch.float32, device=dev) if r else None
    _lib.check(lib.vst_lora_gemm_bwd_params(P(dy), n, P(x), k, 0, 0, P(u), r, P(du), r, P(dw), P(da), P(dbp), m, n, k,
                                            k, r, 0, P(ws), wsb, st), "params")
    assert rel_err(dw, wf.grad) < 1e-2
    if r:
        assert rel_err(da, af.grad) < 1.5e-2
        assert rel_err(dbp, bpf.grad) < 1e-2
    # bad arguments are refused with a message, not executed
    assert lib.vst_lora_gemm_bwd_dx(0, n, P(w), k, 0, 0, 0, 0, 0, 0, 0, P(dx), k, m, n, k, 1.0, 0, 1, st) != 0
    assert b"required" in lib.vst_last_error()


@pytest.mark.parametrize("kind", ["self", "cross"])
def test_stage1_processor_golden_fwd_bwd(kind):
    """a6: the Stage-1 three-stream processor drop-in (unzip_attention_processor.py:664-759) against the
    fixture made by the reference's own processor: output, d hidden_states, d of the three text streams,
    and dA / dB / d-merger of every projection (hard column masks on the train layers)."""
    import types
    from video_style_transfer_b200.unziplora_unet.unzip_attention_processor import AttnProcessor2_0
    from video_style_transfer_b200.unziplora_unet.lora_unzip import LoRACompatibleLinear
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    g = np.load(os.path.join(GOLD, "stage1_processor.npz"))

    def mk(nm):
        f = _sub(g, f"{kind}/{nm}/")
        out_f, in_f = f["weight"].shape
        lin = LoRACompatibleLinear(in_f, out_f, bias="bias" in f)
        lay = UnZipLoRALinearLayer(in_f, out_f, f["a_c"].shape[0], ["content", "style"], lora_matrix_num=2)
        with torch.no_grad():
            lin.weight.copy_(f["weight"])
            if "bias" in f:
                lin.bias.copy_(f["bias"])
            d = lay.lora_matrix_dic
            d["content_down"].weight.copy_(f["a_c"]); d["content_up"].weight.copy_(f["b_c"])
            d["style_down"].weight.copy_(f["a_s"]); d["style_up"].weight.copy_(f["b_s"])
            lay.merge_content.copy_(f["m_c"]); lay.merge_style.copy_(f["m_s"])
        lin.set_lora_layer(lay)
        lin = lin.cuda()
        lay.mask_content, lay.mask_style = f["mask_c"].cuda(), f["mask_s"].cuda()
        return lin, f

    lins = {nm: mk(nm) for nm in ("to_q", "to_k", "to_v", "to_out")}
    heads, res = [int(v) for v in g[f"{kind}/meta"]]
    attn = types.SimpleNamespace(
        spatial_norm=None, group_norm=None, norm_cross=False, heads=heads, to_q=lins["to_q"][0],
        to_k=lins["to_k"][0], to_v=lins["to_v"][0],
        to_out=torch.nn.ModuleList([lins["to_out"][0], torch.nn.Dropout(0.0)]),
        residual_connection=bool(res), rescale_output_factor=float(g[f"{kind}/rescale"]))
    T = lambda k: torch.from_numpy(np.asarray(g[k])).cuda()
    hs = T(f"{kind}/hs").requires_grad_(True)
    txt = {n: T(f"cross/{n}").requires_grad_(True) for n in ("e", "ec", "es")} if kind == "cross" else {}
    y = AttnProcessor2_0()(attn, hs, encoder_hidden_states=txt.get("e"),
                           encoder_hidden_states_content=txt.get("ec"),
                           encoder_hidden_states_style=txt.get("es"), scale=0.5)   # scale is ignored (:675)
    _close(y, T(f"{kind}/y").cpu(), "y")
    y.backward(T(f"{kind}/dy"))
    _close(hs.grad, T(f"{kind}/d_hs").cpu(), "d hidden_states", 3e-2)
    for n, t in txt.items():
        _close(t.grad, T(f"cross/d_{n}").cpu(), f"d {n}", 3e-2)
    for nm, (lin, f) in lins.items():
        d, lay = lin.lora_layer.lora_matrix_dic, lin.lora_layer
        for fld, t in (("a_c", d["content_down"].weight), ("b_c", d["content_up"].weight),
                       ("a_s", d["style_down"].weight), ("b_s", d["style_up"].weight),
                       ("m_c", lay.merge_content), ("m_s", lay.merge_style)):
            _close(t.grad, f["d_" + fld], f"{nm}.{fld}", 3e-2)


@pytest.mark.parametrize("n,f,heads,d", [(64, 16, 8, 40), (33, 16, 8, 80), (9, 32, 8, 160), (17, 5, 2, 16)])
def test_temporal_attn_bwd_vs_autograd(ops, n, f, heads, d):
    torch.manual_seed(n + f + d)
    c = heads * d
    qkv = _bf(torch.randn(n, f, 3 * c))
    d_o = _bf(torch.randn(n, f, c))
    ref = qkv.float().requires_grad_(True)
    sp = lambda t: t.reshape(n, f, heads, d).transpose(1, 2)
    o_ref = torch.nn.functional.scaled_dot_product_attention(
        sp(ref[..., :c]), sp(ref[..., c:2 * c]), sp(ref[..., 2 * c:])).transpose(1, 2).reshape(n, f, c)
    o_ref.backward(d_o.float())
    x = qkv.cuda().requires_grad_(True)
    o = ops.temporal_attn(x[..., :c], x[..., c:2 * c], x[..., 2 * c:], heads)
    assert rel_err(o, o_ref.detach()) < REL_TOL
    o.backward(d_o.cuda())
    for i, nm in enumerate("qkv"):
        assert rel_err(x.grad[..., i * c:(i + 1) * c], ref.grad[..., i * c:(i + 1) * c]) < REL_TOL, nm


@pytest.mark.parametrize("rows,c,with_pe", [(100, 320, True), (77, 640, False), (300, 1280, True), (5, 64, False)])
def test_layernorm_bwd_vs_autograd(ops, rows, c, with_pe):
    torch.manual_seed(rows + c)
    f = 5
    x = _bf(torch.randn(rows * f, c) * 2 + 0.5)
    g, b = torch.randn(c), torch.randn(c)
    pe = torch.randn(f, c) if with_pe else None
    dy = _bf(torch.randn(rows * f, c))
    xr, gr, br = x.float().requires_grad_(True), g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y_ref = torch.nn.functional.layer_norm(xr, (c,), gr, br, 1e-5)
    if pe is not None:
        y_ref = y_ref + pe.repeat(rows, 1)
    y_ref.backward(dy.float())
    xd = x.cuda().requires_grad_(True)
    gd, bd = torch.nn.Parameter(g.cuda()), torch.nn.Parameter(b.cuda())
    y = ops.layernorm(xd, gd, bd, 1e-5, pe=None if pe is None else pe.cuda(), frames=f)
    assert rel_err(y, y_ref.detach()) < REL_TOL
    y.backward(dy.cuda())
    assert rel_err(xd.grad, xr.grad) < REL_TOL
    assert rel_err(gd.grad, gr.grad) < REL_TOL
    assert rel_err(bd.grad, br.grad) < REL_TOL


def test_geglu_bwd_vs_autograd(ops):
    torch.manual_seed(8)
    rows, inner = 333, 640
    hg = _bf(torch.randn(rows, 2 * inner))
    dy = _bf(torch.randn(rows, inner))
    r = hg.float().requires_grad_(True)
    (r[:, :inner] * torch.nn.functional.gelu(r[:, inner:])).backward(dy.float())
    x = hg.cuda().requires_grad_(True)
    ops.geglu(x).backward(dy.cuda())
    assert rel_err(x.grad, r.grad) < REL_TOL


@pytest.mark.parametrize("b,c,f,h,w", [(1, 320, 16, 8, 8), (2, 64, 5, 4, 6), (1, 1280, 4, 4, 4)])
def test_groupnorm5d_rearrange_bwd_vs_autograd(ops, b, c, f, h, w):
    """GroupNorm(32, C, 1e-6) over the (B,C,F,H,W) view + rearrange to tokens: dx, dgamma, dbeta."""
    torch.manual_seed(c + f)
    x = _bf(torch.randn(b * f, c, h, w) * 1.5 + 0.3)
    g, be = torch.randn(c), torch.randn(c)
    dtok = _bf(torch.randn(b * h * w, f, c))
    xr, gr, br = x.float().requires_grad_(True), g.clone().requires_grad_(True), be.clone().requires_grad_(True)
    x5r = xr.view(b, f, c, h, w).permute(0, 2, 1, 3, 4)
    y5 = torch.nn.functional.group_norm(x5r, 32, gr, br, 1e-6)
    tok_ref = y5.permute(0, 3, 4, 2, 1).reshape(b * h * w, f, c)
    tok_ref.backward(dtok.float())
    xd = x.cuda().requires_grad_(True)
    gd, bd = torch.nn.Parameter(g.cuda()), torch.nn.Parameter(be.cuda())
    x5 = xd.view(b, f, c, h, w).permute(0, 2, 1, 3, 4)
    tok, _, _ = ops.groupnorm5d_rearrange(x5, gd, bd, 32, 1e-6)
    assert rel_err(tok, tok_ref.detach()) < REL_TOL
    tok.backward(dtok.cuda())
    assert rel_err(xd.grad, xr.grad) < REL_TOL
    assert rel_err(gd.grad, gr.grad) < REL_TOL
    assert rel_err(bd.grad, br.grad) < REL_TOL


# ---- config 4 building block: one motion module with temporal LoRA, forward + backward -----------------
@pytest.mark.parametrize("c,hw,f", [(320, 8, 16), (640, 4, 16)])
def test_motion_module_fwd_bwd_vs_oracle(c, hw, f):
    """Trainable set of train_animatediff.py inside a motion module (animatediff/utils.py:66-96):
    lora_A / lora_B of the eight wrapped projections, the norms, proj_in / proj_out and the
    feed-forward; base projections frozen.  Gradients vs fp32 autograd of oracle/processor.py."""
    from oracle import processor as op
    from video_style_transfer_b200.motion_module import MotionModule
    from video_style_transfer_b200.animatediff.temporal_lora import inject_temporal_lora, TemporalLoRALinear
    torch.manual_seed(0)
    holder = torch.nn.Module()
    holder.motion_modules = torch.nn.ModuleList([MotionModule(c, 8)])
    with torch.no_grad():
        for n, p in holder.named_parameters():
            if "norm" in n or n.endswith("bias"):
                p.add_(0.1 * torch.randn_like(p))
    assert inject_temporal_lora(holder, rank=32, alpha=1.0) == 8
    with torch.no_grad():
        for m in holder.modules():
            if isinstance(m, TemporalLoRALinear):
                m.lora_B.normal_(0, 0.2)
    mm = holder.motion_modules[0]
    sd = mm.temporal_transformer.state_dict()
    frozen = lambda k: ".base." in k or k.endswith(".pe")
    w = {k: v.float().clone().requires_grad_(not frozen(k)) for k, v in sd.items()}
    w["pos_embed.pe"] = w["transformer_blocks.0.pos_embed.pe"]
    x = _bf(torch.randn(f, c, hw, hw))
    dy = _bf(torch.randn(f, c, hw, hw))
    xr = x.float().requires_grad_(True)
    want = op.motion_module(xr, w, f, 8, 32, {"rank": 32, "scale": 1.0 / 32})
    want.backward(dy.float())

    mm = mm.cuda()
    xd = x.cuda().requires_grad_(True)
    y = mm(xd, num_frames=f)
    assert rel_err(y, want.detach()) < 2e-2
    y.backward(dy.cuda())
    _close(xd.grad, xr.grad, "dx", 3e-2)
    checked = 0
    for k, p in mm.temporal_transformer.named_parameters():
        if frozen(k):
            assert p.grad is None, k
            continue
        _close(p.grad, w[k].grad, k, 3e-2)
        checked += 1
    assert checked == 8 * 2 + 4 * 2 + 2 * 2 + 2 * 2   # lora A/B, 4 norms, proj_in/out, the two ff linears


def test_temporal_transformer_fwd_bwd_vs_reference_golden_weights():
    """The reference's own smoke test back-propagates a loss through TemporalTransformer
    (animatediff/temporal_transformer.py:149-176).  Same here: forward + backward of the drop-in with
    the golden 'tiny' weights, gradients vs fp32 torch autograd of the same module restated with
    torch ops (nn.MultiheadAttention / nn.LayerNorm / exact GELU are what the reference itself uses)."""
    from video_style_transfer_b200.animatediff.temporal_transformer import TemporalTransformer
    torch.manual_seed(2)
    c, f, hw, heads = 64, 8, 4, 8
    m = TemporalTransformer(c, 2, heads)
    x = torch.randn(1, c, f, hw, hw)
    dy = torch.randn(1, c, f, hw, hw)

    # fp32 torch restatement sharing the parameters
    def ref_forward(xin):
        b, cc, ff, h, w = xin.shape
        t = xin.permute(0, 3, 4, 2, 1).reshape(b * h * w, ff, cc) + m.pos_encoding.pe[:, :ff]
        for blk in m.blocks:
            n1 = blk.norm1(t)
            t = t + blk.attn(n1, n1, n1, need_weights=False)[0]
            t = t + blk.ffn(blk.norm2(t))
        t = m.norm(t)
        return t.reshape(b, h, w, ff, cc).permute(0, 4, 3, 1, 2)

    xr = x.clone().requires_grad_(True)
    want = ref_forward(xr)
    want.backward(dy)
    ref_grads = {k: p.grad.clone() for k, p in m.named_parameters()}
    m.zero_grad()

    m = m.cuda()
    xd = x.cuda().requires_grad_(True)
    y = m(xd, num_frames=f)
    assert rel_err(y, want.detach()) < 2e-2
    y.backward(dy.cuda())
    _close(xd.grad, xr.grad, "dx", 3e-2)
    for k, p in m.named_parameters():
        _close(p.grad, ref_grads[k], k, 3e-2)


@pytest.mark.parametrize("m", [200, 4096])
@pytest.mark.parametrize("mode,three", [(0, False), (0, True), (1, False), (2, True)])
def test_lora_gemm_fwd_module_level_vs_oracle(m, mode, three):
    """vst_lora_gemm_fwd from RAW parameters (fp32 A / B / mergers, hard column masks, scale, forward type,
    one or three activation streams) against the oracle's UnZipLoRALinearLayer + LoRACompatibleLinear
    (unziplora_linear_layer.py:221-263, lora_unzip.py:66-75).  m = 4096 single-stream takes the single-launch
    kernel, the rest the pre-pass form; u_out is what the backward consumes."""
    from oracle import hotpath as hp
    from video_style_transfer_b200 import _lib
    lib = _lib.lib
    torch.manual_seed(m + 10 * mode + three)
    n, k, r = 640, 640, 64
    f32 = lambda *s_, sc=1.0: (torch.randn(*s_) * sc)
    x, x1, x2 = (f32(m, k).to(torch.bfloat16) for _ in range(3))
    w, bias = f32(n, k, sc=k ** -0.5).to(torch.bfloat16), f32(n)
    a_c, b_c, a_s, b_s = f32(r, k, sc=1 / r), f32(n, r, sc=1 / r), f32(r, k, sc=1 / r), f32(n, r, sc=1 / r)
    m_c, m_s = torch.rand(n), torch.rand(n)
    mk_c, mk_s = torch.rand(n) > 0.3, torch.rand(n) > 0.3
    ft = ("both", "content", "style")[mode]
    xc, xs = (x1, x2) if three else (x, x)
    delta = hp.unziplora_delta(xc.float(), xs.float(), a_c, b_c, a_s, b_s, m_c, m_s, ft, True, True, infer=False,
                               mask_c=mk_c, mask_s=mk_s)
    want = hp.lora_compatible_linear(x.float(), w.float(), bias, delta, 0.7)
    d = lambda t: t.cuda().contiguous()
    X, X1, X2, W, Bi = d(x), d(x1), d(x2), d(w), d(bias)
    Ac, Bc, As, Bs, Mc, Ms = d(a_c), d(b_c), d(a_s), d(b_s), d(m_c), d(m_s)
    Kc, Ks = d(mk_c.to(torch.uint8)), d(mk_s.to(torch.uint8))
    nbr = 2 if mode == 0 else 1
    y = torch.empty(m, n, dtype=torch.bfloat16, device="cuda")
    u = torch.empty(m, nbr * r, dtype=torch.bfloat16, device="cuda")
    wsb = lib.vst_lora_gemm_fwd_workspace_bytes(m, n, k, r, nbr, 1)
    ws = torch.empty(wsb + 256, dtype=torch.uint8, device="cuda")
    wsp = ws.data_ptr() + (-ws.data_ptr()) % 256
    P = lambda t: t.data_ptr()
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.vst_lora_gemm_fwd(P(X), k, P(X1) if three else 0, k, P(X2) if three else 0, k, P(W), k, P(Bi), 0,
                                     P(Ac), P(Bc), P(Mc), P(Kc), P(As), P(Bs), P(Ms), P(Ks), r, 0, 0.7, mode,
                                     0, 0, 1.0, P(u), nbr * r, P(y), n, m, n, k, k, 1, wsp, wsb, st), "vst_lora_gemm_fwd")
    assert rel_err(y, want) < 1e-2
    u_want = []
    if mode != 2:
        u_want.append(xc.float() @ a_c.to(torch.bfloat16).float().t())
    if mode != 1:
        u_want.append(xs.float() @ a_s.to(torch.bfloat16).float().t())
    assert rel_err(u, torch.cat(u_want, dim=1)) < 1e-2
    # every branch block-masked (inference layer): A = NULL -> the plain linear
    _lib.check(lib.vst_lora_gemm_fwd(P(X), k, 0, 0, 0, 0, P(W), k, P(Bi), 0, 0, 0, 0, 0, 0, 0, 0, 0, r, 0, 0.7, mode,
                                     0, 0, 1.0, 0, 0, P(y), n, m, n, k, k, 1, 0, 0, st), "vst_lora_gemm_fwd plain")
    assert rel_err(y, x.float() @ w.float().t() + bias) < 1e-2
