"""CPU suite (-m "not gpu"): the oracle against the committed golden vectors (generated from the
reference's own modules by oracle/make_golden.py), the host-side logic of the drop-in modules,
and that the C-ABI library loads and exports every symbol include/vst_b200.h declares.
No compute entry point is called here."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def T(a):
    return torch.from_numpy(np.asarray(a))


def sub(g, prefix):
    return {k[len(prefix):]: T(g[k]) for k in g.files if k.startswith(prefix)}


def close(a, b, tol):
    err = (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)
    assert err <= tol, err


# ---- oracle vs golden -------------------------------------------------------------------------------
def test_oracle_layout_golden():
    from oracle import hotpath as hp
    g = np.load(os.path.join(GOLD, "layout.npz"))
    for c in (40, 64):
        assert torch.equal(hp.positional_table(c, 32), T(g[f"pe_{c}"]))
    y = hp.rearrange_fwd(T(g["rearr_x"]), hp.positional_table(24))
    assert torch.equal(y, T(g["rearr_y"]))
    assert torch.equal(hp.rearrange_bwd(y, 2, 3, 5), T(g["rearr_back"]))


@pytest.mark.parametrize("tag", ["tiny", "d40"])
def test_oracle_temporal_transformer_golden(tag):
    from oracle import hotpath as hp
    g = np.load(os.path.join(GOLD, "temporal_transformer.npz"))
    b, c, f, h, w, layers, heads = [int(v) for v in g[f"{tag}/meta"]]
    y = hp.temporal_transformer(T(g[f"{tag}/x"]), sub(g, f"{tag}/sd/"), layers, heads)
    close(y, T(g[f"{tag}/y"]), 2e-5)


def test_oracle_unziplora_golden():
    from oracle import hotpath as hp
    g = np.load(os.path.join(GOLD, "unziplora.npz"))
    p = sub(g, "infer/")
    for ft in ("both", "content", "style"):
        for mc in (0, 1):
            for ms in (0, 1):
                for two in (0, 1):
                    y = hp.unziplora_delta(p["x"], p["x2"] if two else None, p["content_down"],
                                           p["content_up"], p["style_down"], p["style_up"],
                                           p["merge_content"], p["merge_style"], ft, bool(mc),
                                           bool(ms), infer=True)
                    close(y, p[f"y_{ft}_{mc}{ms}_{two}"], 1e-6)
    t = sub(g, "train/")
    for ft in ("both", "content", "style"):
        for masked in (0, 1):
            y = hp.unziplora_delta(p["x"], p["x2"], t["content_down"], t["content_up"],
                                   t["style_down"], t["style_up"], t["merge_content"],
                                   t["merge_style"], ft, bool(masked), bool(masked), infer=False,
                                   mask_c=t["mask_content"], mask_s=t["mask_style"])
            close(y, t[f"{ft}_{masked}/y"], 1e-6)


def test_oracle_lora_linear_golden():
    from oracle import hotpath as hp
    g = np.load(os.path.join(GOLD, "lora_linear.npz"))
    u = sub(np.load(os.path.join(GOLD, "unziplora.npz")), "infer/")
    for x1, x2, key in ((None, None, "y"), (T(g["x1"]), T(g["x2"]), "y3")):
        delta = hp.unziplora_delta(T(g["x"]) if x1 is None else x1, x2, u["content_down"],
                                   u["content_up"], u["style_down"], u["style_up"],
                                   u["merge_content"], u["merge_style"])
        y = hp.lora_compatible_linear(T(g["x"]), T(g["weight"]), T(g["bias"]), delta,
                                      float(g["scale"]))
        close(y, T(g[key]), 1e-6)
    y = hp.temporal_lora_linear(T(g["t_x"]), T(g["t_weight"]), T(g["t_bias"]), T(g["t_A"]),
                                T(g["t_B"]), float(g["t_scale"]))
    close(y, T(g["t_y"]), 1e-6)


def _params(g, kind):
    from oracle import processor as op

    def proj(nm):
        f = sub(g, f"{kind}/{nm}/")
        return op.UnzipProj(f["weight"], f.get("bias"), f["a_c"], f["b_c"], f["a_s"], f["b_s"],
                            f["m_c"], f["m_s"])

    heads, res = [int(v) for v in g[f"{kind}/meta"]]
    return op.SpatialAttnParams(heads, proj("to_q"), proj("to_k"), proj("to_v"), proj("to_out"),
                                bool(res), float(g[f"{kind}/rescale"]))


@pytest.mark.parametrize("kind", ["self", "cross"])
def test_oracle_processor_golden(kind):
    from oracle import processor as op
    g = np.load(os.path.join(GOLD, "attn_processor.npz"))
    p = _params(g, kind)
    ehs = T(g["cross/ehs"]) if kind == "cross" else None
    close(op.animatediff_attn_processor(p, T(g[f"{kind}/hs"]), ehs, scale=0.9), T(g[f"{kind}/y"]), 2e-5)
    if kind == "self":
        close(op.animatediff_attn_processor(p, T(g["self/hs4"]), None, scale=0.9), T(g["self/y4"]), 2e-5)


@pytest.mark.parametrize("kind", ["self", "cross"])
def test_oracle_stage1_processor_golden(kind):
    """a6: oracle vs the fixture made by the reference's own Stage-1 AttnProcessor2_0
    (unzip_attention_processor.py:664-759) on lora_unzip linears + train-mode UnZipLoRA layers."""
    from oracle import processor as op
    g = np.load(os.path.join(GOLD, "stage1_processor.npz"))

    def proj(nm):
        f = sub(g, f"{kind}/{nm}/")
        return op.UnzipProj(f["weight"], f.get("bias"), f["a_c"], f["b_c"], f["a_s"], f["b_s"],
                            f["m_c"], f["m_s"], infer=False, mask_c=f["mask_c"], mask_s=f["mask_s"])

    heads, res = [int(v) for v in g[f"{kind}/meta"]]
    p = op.SpatialAttnParams(heads, proj("to_q"), proj("to_k"), proj("to_v"), proj("to_out"),
                             bool(res), float(g[f"{kind}/rescale"]))
    hs = T(g[f"{kind}/hs"]).requires_grad_(True)
    txt = [T(g[f"cross/{n}"]).requires_grad_(True) for n in ("e", "ec", "es")] if kind == "cross" else [None] * 3
    y = op.stage1_attn_processor(p, hs, *txt)
    close(y, T(g[f"{kind}/y"]), 2e-5)
    y.backward(T(g[f"{kind}/dy"]))
    close(hs.grad, T(g[f"{kind}/d_hs"]), 2e-5)
    if kind == "cross":
        for t, n in zip(txt, ("e", "ec", "es")):
            close(t.grad, T(g[f"cross/d_{n}"]), 2e-5)


def test_oracle_motion_module_shapes():
    """The diffusers-owned motion block is 'parity unpinned' (oracle/__init__.py); check the
    restatement is self-consistent: zero-init proj_out => identity (+residual)."""
    from oracle import processor as op
    from oracle import hotpath as hp
    torch.manual_seed(0)
    c, f = 64, 4
    w = {"norm.weight": torch.ones(c), "norm.bias": torch.zeros(c),
         "proj_in.weight": torch.randn(c, c) * 0.1, "proj_in.bias": torch.zeros(c),
         "proj_out.weight": torch.zeros(c, c), "proj_out.bias": torch.zeros(c),
         "pos_embed.pe": hp.positional_table(c)}
    tb = "transformer_blocks.0."
    for n in ("norm1", "norm2", "norm3"):
        w[tb + n + ".weight"], w[tb + n + ".bias"] = torch.ones(c), torch.zeros(c)
    for a in ("attn1", "attn2"):
        for p in ("to_q", "to_k", "to_v", "to_out.0"):
            w[tb + f"{a}.{p}.weight"] = torch.randn(c, c) * 0.1
        w[tb + f"{a}.to_out.0.bias"] = torch.zeros(c)
    w[tb + "ff.net.0.proj.weight"], w[tb + "ff.net.0.proj.bias"] = torch.randn(8 * c, c) * 0.1, torch.zeros(8 * c)
    w[tb + "ff.net.2.weight"], w[tb + "ff.net.2.bias"] = torch.randn(c, 4 * c) * 0.1, torch.zeros(c)
    x = torch.randn(2 * f, c, 3, 3)
    assert torch.allclose(op.motion_module(x, w, f), x)
    w["proj_out.weight"] = torch.randn(c, c) * 0.1
    assert op.motion_module(x, w, f).shape == x.shape


# ---- the C-ABI library ---------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "vst_b200.h")).read()
    declared = set(re.findall(r"\b(vst_[a-z0-9_]+)\s*\(", hdr))
    assert {"vst_lora_gemm", "vst_attn_fwd", "vst_rearrange_fwd"} <= declared
    so = os.path.join(ROOT, "video_style_transfer_b200", "libvst_b200.so")
    assert os.path.exists(so), "build the library first: make"
    lib = ctypes.CDLL(so)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/vst_b200.h but not exported"
    from video_style_transfer_b200 import _lib
    assert set(_lib.SIGNATURES) == declared
    assert _lib.lib.vst_version() >= 100


def test_ops_refuse_cpu_tensors():
    from video_style_transfer_b200 import ops
    with pytest.raises(RuntimeError):
        ops.rearrange_fwd(torch.zeros(1, 8, 2, 2, 2))
    with pytest.raises(RuntimeError):
        ops.lora_gemm(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16))


# ---- host logic of the drop-ins -------------------------------------------------------------------------
def test_unziplora_layer_api_and_bookkeeping_golden():
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import (
        UnZipLoRALinearLayer, UnZipLoRALinearLayerInfer)
    g = np.load(os.path.join(GOLD, "unziplora.npz"))
    t, bk = sub(g, "train/"), sub(g, "book/")
    fin, fout, r = 48, 40, 8
    inf = UnZipLoRALinearLayerInfer(fin, fout, r, ["content", "style"])
    assert set(inf.state_dict()) == {"merge_content", "merge_style",
                                     "lora_matrix_dic.content_down.weight",
                                     "lora_matrix_dic.content_up.weight",
                                     "lora_matrix_dic.style_down.weight",
                                     "lora_matrix_dic.style_up.weight"}
    with pytest.raises(AssertionError):
        inf.set_forward("nope")
    inf.set_layer_mask("content", True)
    assert [k for k, _, _ in inf._active()] == ["style"]
    lay = UnZipLoRALinearLayer(fin, fout, r, ["content", "style"], lora_matrix_num=2)
    with torch.no_grad():
        for k in ("content_down", "content_up", "style_down", "style_up"):
            lay.lora_matrix_dic[k].weight.copy_(t[k])
            lay.lora_matrix_dic[k].weight.grad = bk[f"grad_{k}"].clone()
        lay.merge_content.copy_(t["merge_content"])
        lay.merge_style.copy_(t["merge_style"])
    lay.merge_content.grad = bk["grad_merge_content"].clone()
    lay.merge_style.grad = bk["grad_merge_style"].clone()
    close(lay.compute_mergers_similarity().detach(), bk["similarity_unmasked"], 1e-6)
    for key in ("content", "style"):
        close(lay.get_unziplora_norm(key, "L2").detach(), bk[f"norm_L2_{key}"], 1e-5)
        close(lay.get_unziplora_norm(key, "L1", multiple=False).detach(), bk[f"norm_L1_{key}"], 1e-5)
        close(lay.get_unziplora_weight(key)[1].detach(), bk[f"weight_up_{key}"], 1e-6)
        lay.set_cone_score(key)
        lay.get_unziplora_cone(key, accumulate=True)
        close(getattr(lay, f"column_score_{key}").detach(), bk[f"cone_acc_{key}"], 1e-5)
        lay.get_unziplora_cone(key, accumulate=False)
        assert torch.equal(getattr(lay, f"column_score_{key}"), bk[f"cone_sparsity_{key}"])
    lay.column_score_content, lay.column_score_style = bk["score_content"].clone(), bk["score_style"].clone()
    lay.mask_updated_elements(None, step_ratio=0.25, avoid=True)
    assert torch.equal(lay.mask_content, bk["mask_content_after"])
    assert torch.equal(lay.mask_style, bk["mask_style_after"])
    lay.set_layer_mask("content", True)
    lay.set_layer_mask("style", True)
    close(lay.compute_mergers_similarity().detach(), bk["similarity_masked"], 1e-6)
    lay.set_gradient_mask(False)
    close(lay.merge_content.grad, bk["grad_merge_content_overlap"], 1e-6)
    lay.mask_updated_elements("style", step_ratio=0.5)
    assert torch.equal(lay.mask_content, bk["mask_content_keyed"])
    assert torch.equal(lay.mask_style, bk["mask_style_keyed"])
    close(lay.log_selected_mask("style"), bk["log_selected_style"], 1e-6)


def _toy_unet(g):
    from video_style_transfer_b200.unziplora_unet.lora_linear import LoRACompatibleLinear
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayerInfer
    nn = torch.nn

    class Attn(nn.Module):
        def __init__(self, dim, spatial):
            super().__init__()
            mk = (lambda: LoRACompatibleLinear(dim, dim)) if spatial else (lambda: nn.Linear(dim, dim))
            self.to_q, self.to_k, self.to_v = mk(), mk(), mk()
            self.to_out = nn.ModuleList([mk(), nn.Dropout(0.0)])

    dim, r = 16, 4
    unet, blk, sp, mm = nn.Module(), nn.Module(), nn.Module(), nn.Module()
    sp.transformer_blocks = nn.ModuleList([nn.Module()])
    sp.transformer_blocks[0].attn1 = Attn(dim, True)
    mm.temporal_transformer = nn.Module()
    mm.temporal_transformer.transformer_blocks = nn.ModuleList([nn.Module()])
    mm.temporal_transformer.transformer_blocks[0].attn1 = Attn(dim, False)
    blk.attentions, blk.motion_modules = nn.ModuleList([sp]), nn.ModuleList([mm])
    unet.down_blocks = nn.ModuleList([blk])
    a = sp.transformer_blocks[0].attn1
    for lin in (a.to_q, a.to_k, a.to_v, a.to_out[0]):
        lin.set_lora_layer(UnZipLoRALinearLayerInfer(dim, dim, r, ["content", "style"]))
    unet.load_state_dict(sub(g, "toy/sd/"))
    return unet


def test_temporal_lora_host_logic_golden():
    from video_style_transfer_b200.animatediff import temporal_lora as tl
    g = np.load(os.path.join(GOLD, "lora_linear.npz"))
    unet = _toy_unet(g)
    assert tl.inject_temporal_lora(unet, rank=4, alpha=2.0) == int(g["toy/n_wrapped"]) == 4
    assert tl.inject_temporal_lora(unet, rank=4, alpha=2.0) == 0            # idempotent
    lora = sub(g, "toy/lora/")
    with torch.no_grad():
        for n, m in unet.named_modules():
            if isinstance(m, tl.TemporalLoRALinear):
                m.lora_A.copy_(lora[n + ".lora_A"])
                m.lora_B.copy_(lora[n + ".lora_B"])
                assert not m.base.weight.requires_grad and m.lora_A.requires_grad
    names = [n for n, _ in unet.named_parameters()]
    assert any(".base.weight" in n for n in names) and any(n.endswith(".lora_A") for n in names)
    idx = tl.build_spatial_lora_index(unet)
    assert sorted(idx) == [str(k) for k in g["toy/index_keys"]]
    loss = tl.compute_orth_loss(unet, idx, 1e-2)
    close(loss.detach(), T(g["toy/orth_loss"]), 1e-4)
    loss.backward()                                                          # grads reach A and B
    m = unet.down_blocks[0].motion_modules[0].temporal_transformer.transformer_blocks[0].attn1.to_q
    assert m.lora_A.grad is not None and m.lora_B.grad is not None
    merged = tl.get_merged_motion_state_dict(unet)
    want = sub(g, "toy/merged/")
    assert set(merged) == set(want)
    for k in want:
        close(merged[k], want[k], 1e-6)
    assert tl.compute_orth_loss(unet, {}, 1e-2).item() == 0.0


def test_temporal_lora_linear_golden_host_side():
    from video_style_transfer_b200.animatediff.temporal_lora import TemporalLoRALinear
    g = np.load(os.path.join(GOLD, "lora_linear.npz"))
    base = torch.nn.Linear(48, 40)
    with torch.no_grad():
        base.weight.copy_(T(g["t_weight"]))
        base.bias.copy_(T(g["t_bias"]))
    m = TemporalLoRALinear(base, rank=4, alpha=2.0)
    assert m.lora_B.abs().max() == 0 and abs(m.scale - float(g["t_scale"])) < 1e-12
    with torch.no_grad():
        m.lora_A.copy_(T(g["t_A"]))
        m.lora_B.copy_(T(g["t_B"]))
    close(m.get_delta().detach(), T(g["t_delta"]), 1e-6)
    close(m.merged_weight(), T(g["t_merged"]), 1e-6)


def test_positional_encoding_matches_reference_table():
    from video_style_transfer_b200.animatediff.temporal_transformer import PositionalEncoding, TemporalTransformer
    g = np.load(os.path.join(GOLD, "layout.npz"))
    assert torch.equal(PositionalEncoding(64, 32).pe, T(g["pe_64"]))
    gt = np.load(os.path.join(GOLD, "temporal_transformer.npz"))
    m = TemporalTransformer(64, 2, 8)
    m.load_state_dict(sub(gt, "tiny/sd/"))                                   # same parameter names
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 64, 2, 2, 2))                                       # no CPU fallback


def test_bf16_cache_invalidates_on_inplace_update():
    from video_style_transfer_b200._host import bf16_cached
    w = torch.nn.Parameter(torch.randn(8, 8))
    with torch.no_grad():
        a = bf16_cached(w)
        assert bf16_cached(w) is a
    # with autograd recording, the same cached storage comes back wrapped in a graph node whose
    # backward hands the operand gradient to the parameter (cast to its dtype)
    t = bf16_cached(w)
    assert t.requires_grad and t.data_ptr() == a.data_ptr()
    t.backward(torch.ones_like(t))
    assert w.grad is not None and w.grad.dtype == w.dtype and torch.equal(w.grad, torch.ones_like(w))
    with torch.no_grad():
        w.add_(1.0)
        b = bf16_cached(w)
    assert b is not a and torch.equal(b, w.detach().to(torch.bfloat16))


# ---- §8e: the sequence-parallel split, world_size 2 over gloo on CPU ------------------------------------
def _sp_gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    from oracle import processor as op
    from oracle import sp
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                       # same weights and video on every rank
        c, f, hw = 64, 8, 4
        w = _motion_weights(c)
        x = torch.randn(f, c, hw, hw)
        want = op.motion_module(x, w, f, 8, 32, None)
        (f0, f1), (p0, p1) = sp.sp_plan(f, hw * hw, world, rank)
        assert (f1 - f0) * world == f and (p1 - p0) * world == hw * hw
        got = sp.motion_module_sp(x[f0:f1].contiguous(), w, f, 8, 32, None)
        err = ((got - want[f0:f1]).abs().max() / want.abs().max()).item()
        q.put((rank, err))
    finally:
        dist.destroy_process_group()


def _motion_weights(c):
    g = torch.Generator().manual_seed(7)
    r = lambda *s: torch.randn(*s, generator=g) * 0.1
    w = {"norm.weight": 1 + r(c), "norm.bias": r(c), "proj_in.weight": r(c, c), "proj_in.bias": r(c),
         "proj_out.weight": r(c, c), "proj_out.bias": r(c)}
    tb = "transformer_blocks.0."
    for n in ("norm1", "norm2", "norm3"):
        w[tb + n + ".weight"], w[tb + n + ".bias"] = 1 + r(c), r(c)
    for a in ("attn1", "attn2"):
        for p in ("to_q", "to_k", "to_v"):
            w[tb + a + "." + p + ".weight"] = r(c, c)
        w[tb + a + ".to_out.0.weight"], w[tb + a + ".to_out.0.bias"] = r(c, c), r(c)
    w[tb + "ff.net.0.proj.weight"], w[tb + "ff.net.0.proj.bias"] = r(8 * c, c), r(8 * c)
    w[tb + "ff.net.2.weight"], w[tb + "ff.net.2.bias"] = r(c, 4 * c), r(c)
    from oracle import hotpath as hp
    w["pos_embed.pe"] = hp.positional_table(c, 32)
    return w


def test_sequence_parallel_split_world2_gloo():
    """Frames sharded over 2 ranks, GroupNorm statistics all-reduced, frames<->positions exchanges:
    the sharded oracle equals the unsharded one on every rank's frames."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29611
    procs = [ctx.Process(target=_sp_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    errs = dict(q.get(timeout=5) for _ in range(2))
    assert set(errs) == {0, 1} and max(errs.values()) < 1e-5, errs


def test_generic_workspace_entry_matches_per_op_functions():
    """vst_workspace_bytes(op, dims) is pure host arithmetic: no GPU needed."""
    import ctypes as C
    from video_style_transfer_b200._lib import lib

    def generic(op, *dims):
        arr = (C.c_int64 * len(dims))(*dims)
        return lib.vst_workspace_bytes(op, arr, len(dims))

    assert generic(7, 640, 640, 65536, 0) == lib.vst_gemm_workspace_bytes(640, 640, 65536, 0) > 0
    assert generic(8, 65536, 640) == lib.vst_colsum_workspace_bytes(65536, 640)
    assert generic(9, 16, 20, 1024, 77, 16) == lib.vst_attn_bwd_workspace_bytes(16, 20, 1024, 77, 16) > 0
    assert generic(11, 4096, 320) == lib.vst_layernorm_bwd_workspace_bytes(4096, 320)
    assert generic(12, 1, 320, 16, 32) == lib.vst_groupnorm5d_bwd_workspace_bytes(1, 320, 16, 32)
    assert generic(4) == 0 and generic(3, 1, 2, 3) == 0          # ops without a workspace
    assert generic(99) < 0 and generic(7, 1, 2) < 0              # unknown op / too few dims


# ---- §8f-3: on-disk formats round-trip through the drop-in wiring (host logic, no GPU) --------------------
class _ToyAttn(torch.nn.Module):
    def __init__(self, c, kv):
        super().__init__()
        self.to_q = torch.nn.Linear(c, c, bias=False)
        self.to_k = torch.nn.Linear(kv, c, bias=False)
        self.to_v = torch.nn.Linear(kv, c, bias=False)
        self.to_out = torch.nn.ModuleList([torch.nn.Linear(c, c), torch.nn.Dropout(0.0)])


class _ToyUNet(torch.nn.Module):
    def __init__(self, c=16, ctx=24):
        super().__init__()
        blk = lambda: torch.nn.ModuleDict({"attn1": _ToyAttn(c, c), "attn2": _ToyAttn(c, ctx)})
        self.down_blocks = torch.nn.ModuleList([torch.nn.ModuleDict({"transformer_blocks": torch.nn.ModuleList([blk()])})])
        self.motion_modules = torch.nn.ModuleList([torch.nn.ModuleDict({"attn1": _ToyAttn(c, c)})])

    @property
    def attn_processors(self):      # diffusers naming: "<attention module>.processor"
        return {n + ".processor": None for n, m in self.named_modules() if isinstance(m, _ToyAttn)}


def test_unziplora_formats_round_trip(tmp_path):
    """LoRA safetensors (unet.unet.{attn}.{part}.lora.{down,up}.weight) + merger .pth
    (unet.{attn}.{part}.lora.merge_*) -> insert_unziplora_to_unet -> layers hold exactly those tensors,
    base Parameters are shared, motion modules are skipped; Stage-2 mergers save in the format the
    loader reads (reference unziplora_unet/utils.py:324-483, animatediff/utils.py:137-163)."""
    from safetensors.torch import save_file
    from video_style_transfer_b200.unziplora_unet import utils as U
    from video_style_transfer_b200.animatediff import utils as AU
    torch.manual_seed(0)
    unet, r = _ToyUNet(), 4
    names = U.spatial_attention_names(unet)
    assert len(names) == 2 and all("motion_modules" not in n for n in names)
    files, mergers = {}, {}
    for branch in ("content", "style"):
        sd, md = {}, {}
        for n in names:
            attn = unet.get_submodule(n)
            for part in U.PARTS:
                lin = attn.get_submodule(part)
                sd[f"unet.unet.{n}.{part}.lora.down.weight"] = torch.randn(r, lin.in_features)
                sd[f"unet.unet.{n}.{part}.lora.up.weight"] = torch.randn(lin.out_features, r)
                md[f"unet.{n}.{part}.lora.merge_{branch}"] = torch.rand(lin.out_features)
        d = tmp_path / branch
        d.mkdir()
        save_file(sd, str(d / U.LORA_WEIGHT_NAME_SAFE))
        torch.save(md, str(tmp_path / f"merger_{branch}.pth"))
        files[branch], mergers[branch] = sd, md
    w_q = unet.down_blocks[0]["transformer_blocks"][0]["attn1"].to_q.weight
    U.insert_unziplora_to_unet(unet, str(tmp_path / "content"), str(tmp_path / "style"),
                               str(tmp_path / "merger_content.pth"), str(tmp_path / "merger_style.pth"), rank=r)
    assert unet.down_blocks[0]["transformer_blocks"][0]["attn1"].to_q.weight is w_q      # Parameters shared
    assert not hasattr(unet.motion_modules[0]["attn1"].to_q, "lora_layer")               # motion skipped
    for n in names:
        for part in U.PARTS:
            lay = unet.get_submodule(n).get_submodule(part).lora_layer
            for branch in ("content", "style"):
                for half in ("down", "up"):
                    assert torch.equal(lay.lora_matrix_dic[f"{branch}_{half}"].weight,
                                       files[branch][f"unet.unet.{n}.{part}.lora.{half}.weight"])
                assert torch.equal(getattr(lay, f"merge_{branch}"), mergers[branch][f"unet.{n}.{part}.lora.merge_{branch}"])
    U.unziplora_set_forward_type(unet, "style")
    assert all(l.forward_type == "style" for _, l in U._lora_sites(unet))
    # Stage-2 merger files: same keys as the files loaded above
    with torch.no_grad():
        for _, l in U._lora_sites(unet):
            l.merge_content.mul_(0.5)
    counts = AU.freeze_spatial_layers(unet, unfreeze_mergers=True)
    assert counts["mergers"] == sum(l.merge_content.numel() + l.merge_style.numel() for _, l in U._lora_sites(unet))
    assert all(p.requires_grad == ("motion_modules" in k or "merge_" in k) for k, p in unet.named_parameters())
    path = AU.save_checkpoint(unet, str(tmp_path), 7, save_mergers=True)
    mc = torch.load(os.path.join(path, "merger_content_stage2.pth"))
    assert set(mc) == set(mergers["content"])
    k0 = next(iter(mc))
    assert torch.equal(mc[k0], mergers["content"][k0] * 0.5)
    motion = torch.load(os.path.join(path, "motion_modules.pth"))
    assert set(motion) == {k for k in unet.state_dict() if "motion_modules" in k}
    # and the merger files load straight back through the inference wiring
    unet2 = _ToyUNet()
    U.insert_unziplora_to_unet(unet2, files["content"], files["style"], os.path.join(path, "merger_content_stage2.pth"),
                               os.path.join(path, "merger_style_stage2.pth"), rank=r)
    lay2 = unet2.get_submodule(names[0]).to_q.lora_layer
    assert torch.equal(lay2.merge_content, mc[f"unet.{names[0]}.to_q.lora.merge_content"])


def test_stage1_export_key_format():
    """unet_inverse_ziplora_state_dict (reference unziplora_unet/utils.py:288-310): merger-folded up
    matrices and the merger itself under ``unet.{linear}.lora.*`` keys."""
    from video_style_transfer_b200.unziplora_unet import utils as U
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    torch.manual_seed(1)
    unet = _ToyUNet()
    attn = unet.down_blocks[0]["transformer_blocks"][0]["attn2"]
    U._make_lora_compatible(attn, "to_k")
    lay = UnZipLoRALinearLayer(24, 16, 4, ["content", "style"], lora_matrix_num=2)
    with torch.no_grad():
        lay.merge_style.uniform_(0, 1)
    attn.to_k.set_lora_layer(lay)
    lora_sd, mask_sd = U.unet_inverse_ziplora_state_dict(unet, "style")
    stem = "unet.down_blocks.0.transformer_blocks.0.attn2.to_k.lora"
    assert set(lora_sd) == {stem + ".up.weight", stem + ".down.weight"} and set(mask_sd) == {stem + ".merge_style"}
    assert torch.equal(lora_sd[stem + ".down.weight"], lay.lora_matrix_dic["style_down"].weight)
    assert torch.allclose(lora_sd[stem + ".up.weight"], lay.lora_matrix_dic["style_up"].weight * lay.merge_style[:, None])
    assert torch.equal(mask_sd[stem + ".merge_style"], lay.merge_style)


# ---- the oracle's library-shaped pieces against torch.nn.functional -------------------
def test_oracle_pieces_match_torch_functional():
    """GroupNorm5D / LayerNorm / GELU / sdpa of oracle/hotpath.py are the pieces the (unpinned) motion
    block restatement is assembled from; each is checked against the torch library op diffusers calls
    at that point (F.group_norm on the (B,C,F,H,W) view, F.layer_norm, F.gelu, F.scaled_dot_product_attention)."""
    import torch.nn.functional as F
    from oracle import hotpath as hp
    torch.manual_seed(0)
    x5 = torch.randn(2, 64, 5, 6, 7) * 3 + 1
    w, b = torch.randn(64), torch.randn(64)
    close(hp.groupnorm5d(x5, w, b, 32, 1e-6), F.group_norm(x5, 32, w, b, 1e-6), 2e-6)
    x = torch.randn(37, 11, 96)
    w, b = torch.randn(96), torch.randn(96)
    close(hp.layernorm(x, w, b, 1e-5), F.layer_norm(x, (96,), w, b, 1e-5), 2e-6)
    close(hp.gelu(x), F.gelu(x), 1e-6)
    q, k, v = torch.randn(3, 4, 50, 24), torch.randn(3, 4, 77, 24), torch.randn(3, 4, 77, 24)
    close(hp.sdpa(q, k, v), F.scaled_dot_product_attention(q, k, v), 2e-6)
    m = torch.zeros(3, 1, 1, 77)
    m[..., 40:] = -1e4
    close(hp.sdpa(q, k, v, m), F.scaled_dot_product_attention(q, k, v, attn_mask=m), 2e-6)
    hg = torch.randn(9, 64)
    hid, gate = hg.chunk(2, dim=-1)
    close(hid * hp.gelu(gate), hid * F.gelu(gate), 1e-6)                 # diffusers GEGLU


def test_oracle_orth_loss_dense_matches_reference_golden():
    """oracle.hotpath.orth_loss_dense restates animatediff/temporal_lora.py:126-166; the fixture value was
    produced by the reference's own compute_orth_loss (oracle/make_golden.py)."""
    from oracle import hotpath as hp
    from video_style_transfer_b200.animatediff import temporal_lora as tl
    g = np.load(os.path.join(GOLD, "lora_linear.npz"))
    unet = _toy_unet(g)
    tl.inject_temporal_lora(unet, rank=4, alpha=2.0)
    lora = sub(g, "toy/lora/")
    with torch.no_grad():
        for n, m in unet.named_modules():
            if isinstance(m, tl.TemporalLoRALinear):
                m.lora_A.copy_(lora[n + ".lora_A"])
                m.lora_B.copy_(lora[n + ".lora_B"])
    idx = tl.build_spatial_lora_index(unet)
    pairs = []
    for n, m in unet.named_modules():
        if n in idx and isinstance(m, tl.TemporalLoRALinear):
            d = idx[n].lora_matrix_dic
            pairs.append((m.lora_A, m.lora_B, m.scale, d["content_down"].weight, d["content_up"].weight,
                          d["style_down"].weight, d["style_up"].weight))
    close(hp.orth_loss_dense(pairs, 1e-2).detach(), T(g["toy/orth_loss"]), 1e-5)


def test_cpu_arm_uses_reference_modules_when_placed():
    """oracle/ref_cpu.py: with oracle/_ref in place (oracle/build_ref.py) the timed CPU arm is the
    reference's own code; its spatial path agrees with the oracle restatement on the same weights."""
    from oracle import build_ref, ref_cpu, processor as op
    have = build_ref.build()
    ns = ref_cpu.load_reference()
    if not have:
        assert ns is None
        pytest.skip("no /root/reference and no oracle/_ref on this machine")
    assert ns is not None
    torch.manual_seed(0)
    attn, proc = ref_cpu.ref_spatial_attn(ns, 64, 2, 96, rank=8)
    x, text = torch.randn(4, 10, 64), torch.randn(1, 77, 96)

    def oproj(l):
        ll, d = l.lora_layer, l.lora_layer.lora_matrix_dic
        f = lambda t: None if t is None else t.detach()
        return op.UnzipProj(f(l.weight), f(l.bias), f(d["content_down"].weight), f(d["content_up"].weight),
                            f(d["style_down"].weight), f(d["style_up"].weight), f(ll.merge_content), f(ll.merge_style))
    P = op.SpatialAttnParams(2, oproj(attn.to_q), oproj(attn.to_k), oproj(attn.to_v), oproj(attn.to_out[0]))
    with torch.no_grad():
        close(op.animatediff_attn_processor(P, x, text), proc(attn, x, encoder_hidden_states=text, scale=1.0), 1e-5)
        mm = ref_cpu.RefMotionModule(ns, 64).eval()
        xm = torch.randn(8, 64, 3, 3)
        w = {"norm.weight": mm.norm.weight, "norm.bias": mm.norm.bias, "pos_embed.pe": mm.pe,
             "proj_in.weight": mm.proj_in.weight, "proj_in.bias": mm.proj_in.bias,
             "proj_out.weight": mm.proj_out.weight, "proj_out.bias": mm.proj_out.bias}
        tb = "transformer_blocks.0."
        for i, nm in enumerate(("norm1", "norm2", "norm3")):
            w[tb + nm + ".weight"], w[tb + nm + ".bias"] = mm.norms[i].weight, mm.norms[i].bias
        for ai, a in enumerate(("attn1", "attn2")):
            for pi, p_ in enumerate(("to_q", "to_k", "to_v", "to_out.0")):
                lin = mm.attns[ai][pi]
                w[tb + f"{a}.{p_}.base.weight"] = lin.base.weight
                if lin.base.bias is not None:
                    w[tb + f"{a}.{p_}.base.bias"] = lin.base.bias
                w[tb + f"{a}.{p_}.lora_A"], w[tb + f"{a}.{p_}.lora_B"] = lin.lora_A, lin.lora_B
        w[tb + "ff.net.0.proj.weight"], w[tb + "ff.net.0.proj.bias"] = mm.ff_in.weight, mm.ff_in.bias
        w[tb + "ff.net.2.weight"], w[tb + "ff.net.2.bias"] = mm.ff_out.weight, mm.ff_out.bias
        close(op.motion_module(xm, w, 8, 8, 32, {"scale": 1.0 / 32}), mm(xm, 8), 2e-5)
    arm = ref_cpu.CpuArm(8, 32)
    assert arm.kind == "reference" and set(arm.units) == set(arm.CALLS)


# ---- on-disk formats against files the REFERENCE's own helpers wrote / read (formats were only
# ---- checked against this repository's own loader) ------------------------------------------------------------
def test_formats_match_reference_written_files(tmp_path):
    """tests/golden/formats.npz holds what the reference's insert_unziplora_to_unet / freeze_spatial_layers /
    save_checkpoint / unet_inverse_ziplora_state_dict produced on the toy UNet of oracle/toy_unet.py
    (oracle/make_golden_formats.py; reference unziplora_unet/utils.py:289-310,388-483,
    animatediff/utils.py:66-163, temporal_lora.py:86-192).  The drop-in helpers must read the same input
    files into the same state dict and write files with the same keys and tensors."""
    from safetensors.torch import save_file
    from oracle.toy_unet import ToyUNet
    from video_style_transfer_b200.unziplora_unet import utils as U
    from video_style_transfer_b200.animatediff import utils as AU
    from video_style_transfer_b200.animatediff.temporal_lora import inject_temporal_lora, TemporalLoRALinear
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    g = np.load(os.path.join(GOLD, "formats.npz"))
    sub = lambda prefix: {k[len(prefix):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix)}
    assert str(g["lora_file_name"]) == U.LORA_WEIGHT_NAME_SAFE
    unet = ToyUNet()
    unet.load_state_dict(sub("toy/"))
    for b in ("content", "style"):
        (tmp_path / b).mkdir()
        save_file(sub(f"in/lora_{b}/"), str(tmp_path / b / U.LORA_WEIGHT_NAME_SAFE))
        torch.save(sub(f"in/merger_{b}/"), str(tmp_path / f"merger_{b}.pth"))
    U.insert_unziplora_to_unet(unet, str(tmp_path / "content"), str(tmp_path / "style"),
                               str(tmp_path / "merger_content.pth"), str(tmp_path / "merger_style.pth"), rank=4)
    want, got = sub("inserted/"), unet.state_dict()
    assert set(got) == set(want)
    assert all(torch.equal(got[k], want[k]) for k in want)
    # Stage 2: temporal LoRA injection count, the same parameter updates, the freeze map, the checkpoint files
    assert inject_temporal_lora(unet, rank=2, alpha=1.0) == int(g["n_wrapped"])
    with torch.no_grad():
        for n, m in unet.named_modules():
            if isinstance(m, TemporalLoRALinear):
                m.lora_A.copy_(torch.from_numpy(g[f"lora_A/{n}"]))
                m.lora_B.copy_(torch.from_numpy(g[f"lora_B/{n}"]))
        for _, ll in U._lora_sites(unet):
            ll.merge_content.mul_(0.5)
            ll.merge_style.add_(0.25)
    AU.freeze_spatial_layers(unet, unfreeze_mergers=True)
    rg = {k[len("requires_grad/"):]: bool(g[k]) for k in g.files if k.startswith("requires_grad/")}
    assert {n: p.requires_grad for n, p in unet.named_parameters()} == rg
    path = AU.save_checkpoint(unet, str(tmp_path), 7, save_mergers=True)
    assert sorted(os.listdir(path)) == [str(s) for s in g["checkpoint_files"]]
    for fn in os.listdir(path):
        want, got = sub(f"ckpt/{fn}/"), torch.load(os.path.join(path, fn))
        assert set(got) == set(want), fn
        for k in want:
            assert got[k].dtype == want[k].dtype and got[k].shape == want[k].shape, (fn, k)
            assert torch.allclose(got[k], want[k], rtol=0, atol=1e-6), (fn, k)    # merged W + scale*B A: fp32 order
    # the reference-written Stage-2 merger files load straight back through the drop-in's inference wiring
    unet2 = ToyUNet()
    for fn in ("merger_content_stage2.pth", "merger_style_stage2.pth"):
        torch.save(sub(f"ckpt/{fn}/"), str(tmp_path / fn))
    U.insert_unziplora_to_unet(unet2, sub("in/lora_content/"), sub("in/lora_style/"),
                               str(tmp_path / "merger_content_stage2.pth"), str(tmp_path / "merger_style_stage2.pth"), rank=4)
    for name, ll in U._lora_sites(unet2):
        assert torch.equal(ll.merge_content, torch.from_numpy(g[f"ckpt/merger_content_stage2.pth/unet.{name}.lora.merge_content"]))
        assert torch.equal(ll.merge_style, torch.from_numpy(g[f"ckpt/merger_style_stage2.pth/unet.{name}.lora.merge_style"]))
    # Stage-1 export of trained layers
    unet1 = ToyUNet()
    unet1.load_state_dict(sub("toy/"))
    for n in U.spatial_attention_names(unet1):
        attn = unet1.get_submodule(n)
        for parent, leaf, part in ((attn, "to_q", "to_q"), (attn, "to_k", "to_k"), (attn, "to_v", "to_v"),
                                   (attn.to_out, "0", "to_out.0")):
            U._make_lora_compatible(parent, leaf)
            lin = getattr(parent, leaf)
            lay = UnZipLoRALinearLayer(lin.in_features, lin.out_features, 4, ["content", "style"], lora_matrix_num=2)
            missing = lay.load_state_dict(sub(f"stage1/layer/{n}.{part}/"), strict=True)
            assert not missing.missing_keys and not missing.unexpected_keys
            lin.set_lora_layer(lay)
    for key in ("content", "style"):
        lora_sd, mask_sd = U.unet_inverse_ziplora_state_dict(unet1, key)
        for kind, got in (("lora", lora_sd), ("mask", mask_sd)):
            want = sub(f"stage1/export_{key}/{kind}/")
            assert set(got) == set(want)
            assert all(torch.allclose(got[k].detach(), want[k], rtol=0, atol=1e-7) for k in want), (key, kind)
