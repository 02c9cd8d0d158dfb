"""Generate tests/golden/*.npz by running the REFERENCE's own modules (imported from
/root/reference in the build container) on seeded inputs, and check the oracle restatement
(oracle/hotpath.py, oracle/processor.py) against them while doing so.

Run here only:  python -m oracle.make_golden
The GPU box has no /root/reference; tests read only the committed .npz files.
"""
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("VST_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _np(d):
    return {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
            for k, v in d.items()}


def _close(a, b, tol, what):
    err = (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)
    assert err <= tol, f"oracle != reference for {what}: rel err {err:.3e} > {tol}"
    return err


def _import_stage1_processor():
    """unziplora_unet/unzip_attention_processor.py imports diffusers (absent here) for the base
    class of its ``Attention`` and a few helpers.  ``AttnProcessor2_0.__call__`` (:664-759) itself
    only touches ``attn.*`` and F.scaled_dot_product_attention, so the import lines are satisfied
    with empty stand-ins: no diffusers code runs, the processor body that runs is the reference's."""
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules.setdefault(name, m)

    class _Log:
        def __getattr__(self, k):
            return lambda *a, **kw: None

    stub("diffusers")
    stub("diffusers.utils", USE_PEFT_BACKEND=False, deprecate=lambda *a, **kw: None,
         logging=types.SimpleNamespace(get_logger=lambda n: _Log()))
    stub("diffusers.models")
    stub("diffusers.models.attention_processor", Attention=torch.nn.Module)
    stub("diffusers.utils.import_utils", is_xformers_available=lambda: False)
    stub("diffusers.utils.torch_utils", maybe_allow_in_graph=lambda c: c)
    import unziplora_unet.unzip_attention_processor as uap
    return uap.AttnProcessor2_0


def stage1_golden(report):
    """a6: the Stage-1 three-stream processor (unzip_attention_processor.py:664-759) run on
    lora_unzip.LoRACompatibleLinear + UnZipLoRALinearLayer (train, hard masks).  The processor
    spells the two extra streams ``hidden_states_content`` / ``hidden_states_style`` while the
    linear it is wired to spells them ``hidden_states_1`` / ``hidden_states_2`` (lora_unzip.py:66);
    ``_Kw`` below is that renaming and nothing else."""
    from unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayer
    import unziplora_unet.lora_unzip as lora_unzip
    from oracle import processor as op
    Proc = _import_stage1_processor()

    class _Kw(torch.nn.Module):
        def __init__(self, lin):
            super().__init__()
            self.lin = lin

        def forward(self, hidden_states, hidden_states_content=None, hidden_states_style=None):
            return self.lin(hidden_states, 1.0, hidden_states_content, hidden_states_style)

    C, heads, ctx, rr = 128, 2, 24, 8
    torch.manual_seed(11)

    def mk(i, o, bias):
        l = lora_unzip.LoRACompatibleLinear(i, o, bias=bias)
        ll = UnZipLoRALinearLayer(i, o, rr, ["content", "style"], lora_matrix_num=2)
        with torch.no_grad():
            ll.merge_content.uniform_(0, 1)
            ll.merge_style.uniform_(0, 1)
        ll.mask_content = torch.rand(o) > 0.3
        ll.mask_style = torch.rand(o) > 0.3
        l.set_lora_layer(ll)
        return l

    def oproj(l, grad=False):
        ll, d = l.lora_layer, l.lora_layer.lora_matrix_dic
        f = (lambda t: t.detach().clone().requires_grad_(True)) if grad else (lambda t: t.detach())
        return op.UnzipProj(l.weight.detach(), None if l.bias is None else l.bias.detach(),
                            f(d["content_down"].weight), f(d["content_up"].weight),
                            f(d["style_down"].weight), f(d["style_up"].weight),
                            f(ll.merge_content), f(ll.merge_style), ll.forward_type,
                            ll.masked_matrix["content"], ll.masked_matrix["style"], infer=False,
                            mask_c=ll.mask_content, mask_s=ll.mask_style)

    g = {}
    proc = Proc()
    for kind in ("self", "cross"):
        kv_in = C if kind == "self" else ctx
        lins = dict(to_q=mk(C, C, False), to_k=mk(kv_in, C, False), to_v=mk(kv_in, C, False),
                    to_out=mk(C, C, True))
        attn = types.SimpleNamespace(
            spatial_norm=None, group_norm=None, norm_cross=False, heads=heads,
            to_q=_Kw(lins["to_q"]), to_k=_Kw(lins["to_k"]), to_v=_Kw(lins["to_v"]),
            to_out=torch.nn.ModuleList([_Kw(lins["to_out"]), torch.nn.Dropout(0.0)]),
            residual_connection=(kind == "self"), rescale_output_factor=1.5 if kind == "self" else 1.0)
        hs = torch.randn(3, 20, C, requires_grad=True)
        if kind == "self":
            e = ec = es = None
        else:
            e, ec, es = (torch.randn(3, 11, ctx, requires_grad=True) for _ in range(3))
        y = proc(attn, hs, encoder_hidden_states=e, encoder_hidden_states_content=ec,
                 encoder_hidden_states_style=es)
        dy = torch.randn_like(y)
        y.backward(dy)
        # oracle restatement, forward and gradients
        hs_o = hs.detach().clone().requires_grad_(True)
        eo = [None if t is None else t.detach().clone().requires_grad_(True) for t in (e, ec, es)]
        P = op.SpatialAttnParams(heads, oproj(lins["to_q"], True), oproj(lins["to_k"], True),
                                 oproj(lins["to_v"], True), oproj(lins["to_out"], True),
                                 attn.residual_connection, attn.rescale_output_factor)
        yo = op.stage1_attn_processor(P, hs_o, *eo)
        report[f"stage1_{kind}"] = _close(yo, y.detach(), 2e-5, f"stage-1 processor {kind}")
        yo.backward(dy)
        _close(hs_o.grad, hs.grad, 2e-5, f"stage-1 {kind} d hidden_states")
        g[f"{kind}/hs"], g[f"{kind}/y"], g[f"{kind}/dy"], g[f"{kind}/d_hs"] = hs, y, dy, hs.grad
        for nm, t, to in (("e", e, eo[0]), ("ec", ec, eo[1]), ("es", es, eo[2])):
            if t is not None:
                _close(to.grad, t.grad, 2e-5, f"stage-1 {kind} d {nm}")
                g[f"{kind}/{nm}"], g[f"{kind}/d_{nm}"] = t, t.grad
        g[f"{kind}/meta"] = np.array([heads, int(attn.residual_connection)])
        g[f"{kind}/rescale"] = attn.rescale_output_factor
        for nm, l in lins.items():
            ll, d, po = l.lora_layer, l.lora_layer.lora_matrix_dic, getattr(P, nm)
            g[f"{kind}/{nm}/weight"] = l.weight
            if l.bias is not None:
                g[f"{kind}/{nm}/bias"] = l.bias
            for fld, t, to in (("a_c", d["content_down"].weight, po.a_c), ("b_c", d["content_up"].weight, po.b_c),
                               ("a_s", d["style_down"].weight, po.a_s), ("b_s", d["style_up"].weight, po.b_s),
                               ("m_c", ll.merge_content, po.m_c), ("m_s", ll.merge_style, po.m_s)):
                g[f"{kind}/{nm}/{fld}"] = t
                _close(to.grad, t.grad, 5e-5, f"stage-1 {kind} d {nm}.{fld}")
                g[f"{kind}/{nm}/d_{fld}"] = t.grad
            g[f"{kind}/{nm}/mask_c"], g[f"{kind}/{nm}/mask_s"] = ll.mask_content, ll.mask_style
    np.savez_compressed(os.path.join(OUT, "stage1_processor.npz"), **_np(g))


def main():
    sys.path.insert(0, REF)
    if sys.argv[1:] == ["stage1"]:          # only the Stage-1 fixture (the others are unchanged)
        rep = {}
        os.makedirs(OUT, exist_ok=True)
        stage1_golden(rep)
        print(rep)
        return
    from animatediff.temporal_transformer import TemporalTransformer, PositionalEncoding
    from animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    from animatediff.temporal_lora import TemporalLoRALinear
    from unziplora_unet.unziplora_linear_layer import (UnZipLoRALinearLayer,
                                                       UnZipLoRALinearLayerInfer)
    from unziplora_unet.lora_linear import LoRACompatibleLinear
    import unziplora_unet.lora_unzip as lora_unzip
    from oracle import hotpath as hp
    from oracle import processor as op

    os.makedirs(OUT, exist_ok=True)
    report = {}

    # ---- a9: positional table + rearranges -------------------------------------------
    g = {}
    for c in (40, 64, 320):
        pe = PositionalEncoding(c, 32).pe
        assert torch.equal(pe, hp.positional_table(c, 32))
        if c != 320:
            g[f"pe_{c}"] = pe
    torch.manual_seed(0)
    x = torch.randn(2, 24, 8, 3, 5)
    y_ref = x.permute(0, 3, 4, 2, 1).reshape(-1, 8, 24) + PositionalEncoding(24, 32).pe[:, :8]
    assert torch.equal(y_ref, hp.rearrange_fwd(x, hp.positional_table(24)))
    back_ref = y_ref.view(2, 3, 5, 8, 24).permute(0, 4, 3, 1, 2).contiguous()
    assert torch.equal(back_ref, hp.rearrange_bwd(y_ref, 2, 3, 5))
    g.update(rearr_x=x, rearr_y=y_ref, rearr_back=back_ref)
    np.savez_compressed(os.path.join(OUT, "layout.npz"), **_np(g))

    # ---- a11: TemporalTransformer (config-1 family) -----------------------------------
    g = {}
    for tag, (b, c, f, h, w, layers, heads) in {
        "tiny": (2, 64, 8, 3, 4, 2, 8),
        "d40": (1, 160, 16, 4, 4, 1, 4),
    }.items():
        torch.manual_seed(1)
        m = TemporalTransformer(c, layers, heads).eval()
        # default LN/bias init is trivial (1/0): randomise so the affine terms are exercised
        with torch.no_grad():
            for n_, p_ in m.named_parameters():
                if "norm" in n_ or n_.endswith("bias"):
                    p_.add_(0.1 * torch.randn_like(p_))
        x = torch.randn(b, c, f, h, w)
        with torch.no_grad():
            y = m(x, num_frames=f)
        sd = {k: v for k, v in m.state_dict().items()}
        yo = hp.temporal_transformer(x, sd, layers, heads)
        report[f"tt_{tag}"] = _close(yo, y, 2e-5, f"TemporalTransformer {tag}")
        g[f"{tag}/x"] = x
        g[f"{tag}/y"] = y
        g[f"{tag}/meta"] = np.array([b, c, f, h, w, layers, heads])
        for k, v in sd.items():
            g[f"{tag}/sd/{k}"] = v
    np.savez_compressed(os.path.join(OUT, "temporal_transformer.npz"), **_np(g))
    # config 1 at full size: checked here, not stored (21 MB)
    torch.manual_seed(0)
    m = TemporalTransformer(320, 2, 8).eval()
    x = torch.randn(1, 320, 16, 32, 32)
    with torch.no_grad():
        y = m(x, 16)
    report["tt_config1"] = _close(hp.temporal_transformer(x, dict(m.state_dict()), 2, 8), y, 2e-5,
                                  "TemporalTransformer config1")

    # ---- a3/a4: UnZipLoRA layers -------------------------------------------------------
    g = {}
    fin, fout, r = 48, 40, 8
    torch.manual_seed(2)
    lay = UnZipLoRALinearLayerInfer(fin, fout, r, ["content", "style"])
    with torch.no_grad():
        lay.merge_content.uniform_(0, 1)
        lay.merge_style.uniform_(0, 1)
    W = {k: lay.lora_matrix_dic[k].weight.detach() for k in
         ("content_down", "content_up", "style_down", "style_up")}
    x = torch.randn(3, 7, fin)
    x2 = torch.randn(3, 7, fin)
    g.update({"infer/" + k: v for k, v in W.items()})
    g.update({"infer/merge_content": lay.merge_content, "infer/merge_style": lay.merge_style,
              "infer/x": x, "infer/x2": x2})
    for ft in ("both", "content", "style"):
        for mc in (False, True):
            for ms in (False, True):
                lay.set_forward(ft)
                lay.set_layer_mask("content", mc)
                lay.set_layer_mask("style", ms)
                for two in (False, True):
                    with torch.no_grad():
                        y = lay(x, x2 if two else None)
                    yo = hp.unziplora_delta(x, x2 if two else None, W["content_down"],
                                            W["content_up"], W["style_down"], W["style_up"],
                                            lay.merge_content.detach(), lay.merge_style.detach(),
                                            ft, mc, ms, infer=True)
                    _close(yo, y, 1e-6, f"infer {ft} {mc} {ms} {two}")
                    g[f"infer/y_{ft}_{int(mc)}{int(ms)}_{int(two)}"] = y
    # training layer: hard masks multiply; grads through A, B, mergers, x
    torch.manual_seed(3)
    tl = UnZipLoRALinearLayer(fin, fout, r, ["content", "style"], lora_matrix_num=2)
    with torch.no_grad():
        tl.merge_content.uniform_(0, 1)
        tl.merge_style.uniform_(0, 1)
    tl.mask_content = torch.rand(fout) > 0.5
    tl.mask_style = torch.rand(fout) > 0.5
    g.update({"train/" + k: tl.lora_matrix_dic[k].weight.detach().clone() for k in W})
    g.update({"train/merge_content": tl.merge_content.detach().clone(),
              "train/merge_style": tl.merge_style.detach().clone(),
              "train/mask_content": tl.mask_content, "train/mask_style": tl.mask_style})
    dy = torch.randn(3, 7, fout)
    g["train/dy"] = dy
    for ft in ("both", "content", "style"):
        for masked in (False, True):
            tl.set_forward(ft)
            tl.set_layer_mask("content", masked)
            tl.set_layer_mask("style", masked)
            xa = x.clone().requires_grad_(True)
            xb = x2.clone().requires_grad_(True)
            tl.zero_grad()
            y = tl(xa, xb)
            y.backward(dy)
            args = [tl.lora_matrix_dic[k].weight.detach().clone().requires_grad_(True) for k in
                    ("content_down", "content_up", "style_down", "style_up")]
            mco = tl.merge_content.detach().clone().requires_grad_(True)
            mso = tl.merge_style.detach().clone().requires_grad_(True)
            xo, xo2 = x.clone().requires_grad_(True), x2.clone().requires_grad_(True)
            yo = hp.unziplora_delta(xo, xo2, *args, mco, mso, ft, masked, masked, infer=False,
                                    mask_c=tl.mask_content, mask_s=tl.mask_style)
            _close(yo, y, 1e-6, f"train {ft} {masked}")
            yo.backward(dy)
            tag = f"train/{ft}_{int(masked)}"
            g[tag + "/y"] = y
            pairs = {"dx": (xa.grad, xo.grad), "dx2": (xb.grad, xo2.grad),
                     "dmc": (tl.merge_content.grad, mco.grad),
                     "dms": (tl.merge_style.grad, mso.grad)}
            for k, a in zip(("content_down", "content_up", "style_down", "style_up"), args):
                pairs["d_" + k] = (tl.lora_matrix_dic[k].weight.grad, a.grad)
            for k, (gr, go) in pairs.items():
                if gr is None:
                    assert go is None or go.abs().max() == 0, k
                    continue
                _close(go, gr, 1e-5, tag + k)
                g[f"{tag}/{k}"] = gr.clone()
    # host-side bookkeeping of the training layer (reference :53-220), pinned for the drop-in
    tl.set_forward("both"); tl.set_layer_mask("content", False); tl.set_layer_mask("style", False)
    tl.zero_grad()
    tl(x.clone(), x2.clone()).backward(dy)
    for k in W:
        g[f"book/grad_{k}"] = tl.lora_matrix_dic[k].weight.grad.clone()
    g["book/grad_merge_content"] = tl.merge_content.grad.clone()
    g["book/grad_merge_style"] = tl.merge_style.grad.clone()
    g["book/similarity_unmasked"] = tl.compute_mergers_similarity()
    for key in ("content", "style"):
        g[f"book/norm_L2_{key}"] = tl.get_unziplora_norm(key, "L2")
        g[f"book/norm_L1_{key}"] = tl.get_unziplora_norm(key, "L1", multiple=False)
        d_, u_ = tl.get_unziplora_weight(key)
        g[f"book/weight_up_{key}"] = u_.clone()
        tl.set_cone_score(key)
        tl.get_unziplora_cone(key, accumulate=True)
        g[f"book/cone_acc_{key}"] = getattr(tl, f"column_score_{key}").clone()
        tl.get_unziplora_cone(key, accumulate=False)
        g[f"book/cone_sparsity_{key}"] = getattr(tl, f"column_score_{key}").clone()
    tl.column_score_content = torch.rand(fout)
    tl.column_score_style = torch.rand(fout)
    g["book/score_content"], g["book/score_style"] = tl.column_score_content.clone(), tl.column_score_style.clone()
    tl.mask_content = torch.zeros(fout, dtype=torch.bool)
    tl.mask_style = torch.zeros(fout, dtype=torch.bool)
    tl.mask_updated_elements(None, step_ratio=0.25, avoid=True)
    g["book/mask_content_after"], g["book/mask_style_after"] = tl.mask_content.clone(), tl.mask_style.clone()
    tl.set_layer_mask("content", True); tl.set_layer_mask("style", True)
    g["book/similarity_masked"] = tl.compute_mergers_similarity()
    tl.set_gradient_mask(False)
    g["book/grad_merge_content_overlap"] = tl.merge_content.grad.clone()
    tl.mask_updated_elements("style", step_ratio=0.5)
    g["book/mask_content_keyed"], g["book/mask_style_keyed"] = tl.mask_content.clone(), tl.mask_style.clone()
    g["book/log_selected_style"] = tl.log_selected_mask("style")
    np.savez_compressed(os.path.join(OUT, "unziplora.npz"), **_np(g))

    # ---- a3/a5: LoRACompatibleLinear (both variants) + a7 TemporalLoRALinear ----------
    g = {}
    torch.manual_seed(4)
    lin = LoRACompatibleLinear(fin, fout)
    lay.set_forward("both"); lay.set_layer_mask("content", False); lay.set_layer_mask("style", False)
    lin.set_lora_layer(lay)
    with torch.no_grad():
        y = lin(x, 0.7)
    yo = hp.lora_compatible_linear(x, lin.weight.detach(), lin.bias.detach(),
                                   hp.unziplora_delta(x, None, W["content_down"], W["content_up"],
                                                      W["style_down"], W["style_up"],
                                                      lay.merge_content.detach(),
                                                      lay.merge_style.detach()), 0.7)
    _close(yo, y, 1e-6, "LoRACompatibleLinear")
    lin3 = lora_unzip.LoRACompatibleLinear(fin, fout)
    with torch.no_grad():
        lin3.weight.copy_(lin.weight)
        lin3.bias.copy_(lin.bias)
    lin3.set_lora_layer(lay)
    x1 = torch.randn(3, 7, fin)
    with torch.no_grad():
        y3 = lin3(x, 0.7, x1, x2)
    yo3 = hp.lora_compatible_linear(x, lin.weight.detach(), lin.bias.detach(),
                                    hp.unziplora_delta(x1, x2, W["content_down"], W["content_up"],
                                                       W["style_down"], W["style_up"],
                                                       lay.merge_content.detach(),
                                                       lay.merge_style.detach()), 0.7)
    _close(yo3, y3, 1e-6, "lora_unzip.LoRACompatibleLinear")
    g.update(weight=lin.weight, bias=lin.bias, x=x, x1=x1, x2=x2, y=y, y3=y3, scale=0.7)
    torch.manual_seed(5)
    base = torch.nn.Linear(fin, fout)
    tlo = TemporalLoRALinear(base, rank=4, alpha=2.0)
    with torch.no_grad():
        tlo.lora_B.normal_(0, 0.25)
    xt = torch.randn(5, 6, fin, requires_grad=True)
    yt = tlo(xt)
    dyt = torch.randn_like(yt)
    yt.backward(dyt)
    yo = hp.temporal_lora_linear(xt.detach(), base.weight.detach(), base.bias.detach(),
                                 tlo.lora_A.detach(), tlo.lora_B.detach(), tlo.scale)
    _close(yo, yt.detach(), 1e-6, "TemporalLoRALinear")
    g.update(t_weight=base.weight, t_bias=base.bias, t_A=tlo.lora_A, t_B=tlo.lora_B,
             t_scale=tlo.scale, t_x=xt, t_y=yt, t_dy=dyt, t_dx=xt.grad, t_dA=tlo.lora_A.grad,
             t_dB=tlo.lora_B.grad, t_delta=tlo.get_delta(), t_merged=tlo.merged_weight())
    # compute_orth_loss / merged checkpoint on a two-module toy "unet" (temporal_lora.py:126-192)
    from animatediff.temporal_lora import (compute_orth_loss, build_spatial_lora_index,
                                           inject_temporal_lora, get_merged_motion_state_dict)
    torch.manual_seed(7)

    class Attn(torch.nn.Module):
        def __init__(self, dim, spatial):
            super().__init__()
            mk_ = (lambda: LoRACompatibleLinear(dim, dim)) if spatial else (lambda: torch.nn.Linear(dim, dim))
            self.to_q, self.to_k, self.to_v = mk_(), mk_(), mk_()
            self.to_out = torch.nn.ModuleList([mk_(), torch.nn.Dropout(0.0)])

    def toy(dim=16, r=4):
        unet = torch.nn.Module()
        blk = torch.nn.Module()
        sp = torch.nn.Module(); sp.transformer_blocks = torch.nn.ModuleList([torch.nn.Module()])
        sp.transformer_blocks[0].attn1 = Attn(dim, True)
        mm = torch.nn.Module(); mm.temporal_transformer = torch.nn.Module()
        mm.temporal_transformer.transformer_blocks = torch.nn.ModuleList([torch.nn.Module()])
        mm.temporal_transformer.transformer_blocks[0].attn1 = Attn(dim, False)
        blk.attentions = torch.nn.ModuleList([sp]); blk.motion_modules = torch.nn.ModuleList([mm])
        unet.down_blocks = torch.nn.ModuleList([blk])
        for lin in (sp.transformer_blocks[0].attn1.to_q, sp.transformer_blocks[0].attn1.to_k,
                    sp.transformer_blocks[0].attn1.to_v, sp.transformer_blocks[0].attn1.to_out[0]):
            lin.set_lora_layer(UnZipLoRALinearLayerInfer(dim, dim, r, ["content", "style"]))
        return unet

    unet = toy()
    g["toy/state"] = np.array(0)
    for k, v in unet.state_dict().items():
        g[f"toy/sd/{k}"] = v.clone()
    n_wrapped = inject_temporal_lora(unet, rank=4, alpha=2.0)
    with torch.no_grad():
        for n_, m_ in unet.named_modules():
            if isinstance(m_, TemporalLoRALinear):
                m_.lora_B.normal_(0, 0.3)
                g[f"toy/lora/{n_}.lora_A"] = m_.lora_A.clone()
                g[f"toy/lora/{n_}.lora_B"] = m_.lora_B.clone()
    idx = build_spatial_lora_index(unet)
    g["toy/n_wrapped"] = n_wrapped
    g["toy/index_keys"] = np.array(sorted(idx.keys()))
    g["toy/orth_loss"] = compute_orth_loss(unet, idx, 1e-2).detach()
    for k, v in get_merged_motion_state_dict(unet).items():
        g[f"toy/merged/{k}"] = v
    np.savez_compressed(os.path.join(OUT, "lora_linear.npz"), **_np(g))

    # ---- a1: AnimateDiffAttnProcessor2_0 on a duck-typed attn ---------------------------
    g = {}
    C, heads, ctx, rr = 128, 2, 24, 8
    torch.manual_seed(6)

    def mk(i, o, bias):
        l = LoRACompatibleLinear(i, o, bias=bias)
        ll = UnZipLoRALinearLayerInfer(i, o, rr, ["content", "style"])
        with torch.no_grad():
            ll.merge_content.uniform_(0, 1)
            ll.merge_style.uniform_(0, 1)
        l.set_lora_layer(ll)
        return l

    def oproj(l):
        ll = l.lora_layer
        d = ll.lora_matrix_dic
        return op.UnzipProj(l.weight.detach(), None if l.bias is None else l.bias.detach(),
                            d["content_down"].weight.detach(), d["content_up"].weight.detach(),
                            d["style_down"].weight.detach(), d["style_up"].weight.detach(),
                            ll.merge_content.detach(), ll.merge_style.detach(), ll.forward_type,
                            ll.masked_matrix["content"], ll.masked_matrix["style"])

    proc = AnimateDiffAttnProcessor2_0()
    for kind in ("self", "cross"):
        kv_in = C if kind == "self" else ctx
        attn = types.SimpleNamespace(
            spatial_norm=None, group_norm=None, norm_cross=False, heads=heads,
            to_q=mk(C, C, False), to_k=mk(kv_in, C, False), to_v=mk(kv_in, C, False),
            to_out=torch.nn.ModuleList([mk(C, C, True), torch.nn.Dropout(0.0)]),
            residual_connection=(kind == "self"), rescale_output_factor=1.5 if kind == "self" else 1.0)
        hs = torch.randn(6, 20, C)                         # B*F = 6 (B=2, F=3)
        ehs = None if kind == "self" else torch.randn(2, 11, ctx)
        with torch.no_grad():
            y = proc(attn, hs, ehs, scale=0.9)
        P = op.SpatialAttnParams(heads, oproj(attn.to_q), oproj(attn.to_k), oproj(attn.to_v),
                                 oproj(attn.to_out[0]), attn.residual_connection,
                                 attn.rescale_output_factor)
        report[f"proc_{kind}"] = _close(op.animatediff_attn_processor(P, hs, ehs, scale=0.9), y,
                                        2e-5, f"processor {kind}")
        g[f"{kind}/hs"] = hs
        g[f"{kind}/y"] = y
        if ehs is not None:
            g[f"{kind}/ehs"] = ehs
        g[f"{kind}/meta"] = np.array([heads, int(attn.residual_connection)])
        g[f"{kind}/rescale"] = attn.rescale_output_factor
        for nm, l in (("to_q", attn.to_q), ("to_k", attn.to_k), ("to_v", attn.to_v),
                      ("to_out", attn.to_out[0])):
            pr = oproj(l)
            for fld in ("weight", "bias", "a_c", "b_c", "a_s", "b_s", "m_c", "m_s"):
                v = getattr(pr, fld)
                if v is not None:
                    g[f"{kind}/{nm}/{fld}"] = v
        if kind == "self":                                  # 4-D input branch :33-36,88-89
            hs4 = torch.randn(6, C, 4, 5)
            with torch.no_grad():
                y4 = proc(attn, hs4, None, scale=0.9)
            _close(op.animatediff_attn_processor(P, hs4, None, scale=0.9), y4, 2e-5, "4-D")
            g["self/hs4"], g["self/y4"] = hs4, y4
    np.savez_compressed(os.path.join(OUT, "attn_processor.npz"), **_np(g))

    stage1_golden(report)

    for k, v in report.items():
        print(f"{k}: oracle-vs-reference rel err {v:.2e}")
    print("golden written to", OUT)


if __name__ == "__main__":
    main()
