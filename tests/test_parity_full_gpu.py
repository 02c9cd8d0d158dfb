"""-m gpu parity at the BASELINE configs' real sizes (round-1 VERDICT item 1):
  * config 2 at (16, 1024, 1280) tokens, text (1, 77, 2048), rank 64: forward + every gradient;
  * the 25-step CFG + Euler denoise at the L2 level geometry (C=1280, 20 heads, 32^2, 16 frames),
    EulerDiscrete's SDXL sigma schedule, update undamped, per-step latent cosine >= 0.999;
  * gradient checkpointing: recomputed forward gives bit-identical gradients;
  * config 4: one Stage-2 optimiser step (loss terms, gradients, updated parameters) against
    oracle autograd + torch.optim.AdamW on the CPU;
  * compute_orth_loss on device tensors; folded vs low-rank adapters at small magnitudes;
    batched CFG == two-pass CFG; text K/V cache invalidation; operand-cache growth.
Tolerances: bf16 path vs fp32 oracle, max|a-b| / max|b| <= 1e-2 per GEMM/attention tensor, 2e-2 for
whole modules fed bf16 inputs (north_star); stated per assertion."""
import os

import numpy as np
import pytest
import torch

from .helpers import (attn_linears, attn_to, make_attn, make_motion, motion_weights, oracle_attn, rel,
                      sdxl_euler_sigmas)

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
bf = torch.bfloat16


# ---- config 2 at full size --------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["self", "cross"])
def test_config2_full_size_fwd_bwd_vs_oracle(kind):
    """BASELINE config 2 exactly: SDXL mid-block attention (C=1280, 20 heads, ctx 2048), rank-64
    content+style UnZipLoRA, hidden states (16, 1024, 1280) bf16, text (1, 77, 2048); forward and the
    gradients w.r.t. the input, the text and A / B / merger of all four projections (Stage-1 trains
    A, B; Stage-2 the mergers) against fp32 torch-CPU autograd of oracle/processor.py (dense 3-GEMM
    form of the reference, unziplora_linear_layer.py:221-263)."""
    from oracle import processor as op
    from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    torch.manual_seed(2)
    C, heads, ctx = 1280, 20, 2048
    attn = make_attn(C, heads, C if kind == "self" else ctx, 64, train=True, residual=True, rescale=2.0)
    P = oracle_attn(attn, grad=True, infer=False)
    hs = torch.randn(16, 1024, C).to(bf)
    ehs = None if kind == "self" else torch.randn(1, 77, ctx).to(bf)
    dy = (torch.randn(16, 1024, C) * 0.1).to(bf)
    hs_r = hs.float().requires_grad_(True)
    ehs_r = None if ehs is None else ehs.float().requires_grad_(True)
    want = op.animatediff_attn_processor(P, hs_r, ehs_r, scale=1.0)
    want.backward(dy.float())

    attn_to(attn, "cuda")                      # fp32 parameters, bf16 activations (autocast-like)
    hs_d = hs.cuda().requires_grad_(True)
    ehs_d = None if ehs is None else ehs.cuda().requires_grad_(True)
    y = AnimateDiffAttnProcessor2_0()(attn, hs_d, ehs_d)
    assert y.shape == (16, 1024, C) and y.dtype == bf
    assert rel(y, want) < 2e-2
    cos = torch.nn.functional.cosine_similarity(y.float().cpu().flatten(), want.detach().flatten(), dim=0)
    assert cos > 0.9999
    y.backward(dy.cuda())
    assert rel(hs_d.grad, hs_r.grad) < 2e-2, "d hidden_states"
    if ehs is not None:
        assert rel(ehs_d.grad, ehs_r.grad) < 2e-2, "d encoder_hidden_states"
    for n, pr in zip(("to_q", "to_k", "to_v", "to_out"), (P.to_q, P.to_k, P.to_v, P.to_out)):
        lay = attn_linears(attn)[n].lora_layer
        d = lay.lora_matrix_dic
        for got, ref, what in ((d["content_down"].weight.grad, pr.a_c.grad, "dA_c"),
                               (d["content_up"].weight.grad, pr.b_c.grad, "dB_c"),
                               (d["style_down"].weight.grad, pr.a_s.grad, "dA_s"),
                               (d["style_up"].weight.grad, pr.b_s.grad, "dB_s"),
                               (lay.merge_content.grad, pr.m_c.grad, "dm_c"),
                               (lay.merge_style.grad, pr.m_s.grad, "dm_s")):
            assert rel(got, ref) < 2e-2, f"{kind} {n} {what}"


# ---- 25-step denoise at the L2 geometry ---------------------------------------------------------------
def test_denoise_25_steps_l2_geometry_undamped():
    """north_star's acceptance bar at a real level geometry: C=1280, 20 heads, 32^2 positions, 16
    frames; eps-network = spatial self-attention -> text cross-attention (UnZipLoRA rank 64,
    content+style, residual connections) -> motion module (temporal LoRA rank 32); CFG 7.5 as two
    calls (inference_animatediff.py:110-122); EulerDiscrete with SDXL's sigma schedule (:67,:95,:106,
    :131), the update x <- x + (sigma_next - sigma) eps UNDAMPED.  bf16 sm_100a path against the fp32
    oracle from the same initial latent; per-step latent cosine >= 0.999.

    The residual branches (to_out of both attentions, proj_out of the motion module) carry a gain of
    0.3, which makes |eps| ~ 1.3 for unit-variance scaled input, i.e. the synthetic network behaves
    like a trained epsilon-predictor (diffusers zero-initialises the motion proj_out for the same
    reason).  At gain 1.0 a random-init stack is a chaotic map: the fp32 oracle and the SAME oracle
    with activations rounded to bf16 decorrelate by step 5 (cos 0.987, 0.76 at step 25;
    scripts/denoise_dynamics.py, profiles/denoise_dynamics_r02.log), so no bf16 implementation can
    meet the bar there and the test would measure the map, not the kernels.

    The oracle (oracle/processor.py, plain torch) is evaluated in fp32 on the device for the 50
    network evaluations (TF32 off); its first evaluation is pinned against its CPU execution."""
    from oracle import processor as op
    from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    torch.manual_seed(7)
    C, heads, ctx, F_, hw = 1280, 20, 2048, 16, 32
    steps, guidance, gain = 25, 7.5, 0.3
    a_self = make_attn(C, heads, C, 64, residual=True, out_gain=gain)
    a_cross = make_attn(C, heads, ctx, 64, residual=True, out_gain=gain)
    _, mm = make_motion(C, 32, lora_b_std=0.1, out_gain=gain)
    text_u, text_c = torch.randn(1, 77, ctx), torch.randn(1, 77, ctx)
    sig, init_sigma = sdxl_euler_sigmas(steps)
    x0 = torch.randn(F_, C, hw, hw) * init_sigma
    lora_cfg = {"rank": 32, "scale": 1.0 / 32}
    tok = lambda x4: x4.flatten(2).transpose(1, 2)
    untok = lambda x3: x3.transpose(1, 2).reshape(F_, C, hw, hw)

    def eps_oracle(P_s, P_c, w, x4, text):
        h = op.animatediff_attn_processor(P_s, tok(x4))
        h = op.animatediff_attn_processor(P_c, h, text)
        return op.motion_module(untok(h), w, F_, 8, 32, lora_cfg)

    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            # pin the device-executed oracle to its CPU execution (one evaluation, uncond text)
            cpu_eps = eps_oracle(oracle_attn(a_self), oracle_attn(a_cross), motion_weights(mm),
                                 x0 / (float(sig[0]) ** 2 + 1) ** 0.5, text_u)
            Ps, Pc = oracle_attn(a_self, device="cuda"), oracle_attn(a_cross, device="cuda")
            w = motion_weights(mm, device="cuda")
            dev_eps = eps_oracle(Ps, Pc, w, x0.cuda() / (float(sig[0]) ** 2 + 1) ** 0.5, text_u.cuda())
            assert rel(dev_eps, cpu_eps) < 1e-4, "fp32 oracle on the device != fp32 oracle on the CPU"

            attn_to(a_self, "cuda", bf), attn_to(a_cross, "cuda", bf)
            mm.cuda().to(bf)
            proc = AnimateDiffAttnProcessor2_0()

            def eps_ours(x4, text):
                h = proc(a_self, tok(x4).contiguous())
                h = proc(a_cross, h, text)
                return mm(untok(h).contiguous(), num_frames=F_)

            assert rel(eps_ours((x0 / (float(sig[0]) ** 2 + 1) ** 0.5).cuda().to(bf), text_u.cuda().to(bf)),
                       cpu_eps) < 3e-2
            xo, xg = x0.clone().cuda(), x0.clone().cuda()
            tu32, tc32 = text_u.cuda(), text_c.cuda()
            tu, tc = tu32.to(bf), tc32.to(bf)
            worst, eps_std = 1.0, []
            for i in range(steps):
                s, sn = float(sig[i]), float(sig[i + 1])
                c_in = 1.0 / (s * s + 1.0) ** 0.5                  # scheduler.scale_model_input
                eu, ec = eps_oracle(Ps, Pc, w, xo * c_in, tu32), eps_oracle(Ps, Pc, w, xo * c_in, tc32)
                e = eu + guidance * (ec - eu)
                eps_std.append(e.std().item())
                xo = xo + (sn - s) * e
                gin = (xg * c_in).to(bf)
                gu, gc = eps_ours(gin, tu).float(), eps_ours(gin, tc).float()
                xg = xg + (sn - s) * (gu + guidance * (gc - gu))
                cos = torch.nn.functional.cosine_similarity(xg.flatten(), xo.flatten(), dim=0).item()
                worst = min(worst, cos)
                assert cos >= 0.999, (i, cos)
            # the synthetic network is in the regime the docstring claims, and the latent was denoised
            assert 0.5 < min(eps_std) and max(eps_std) < 4.0, eps_std
            assert xo.std().item() < 0.5 * init_sigma
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    assert worst >= 0.999


# ---- gradient checkpointing ---------------------------------------------------------------------------
@pytest.mark.parametrize("what", ["motion", "processor"])
def test_gradient_checkpointing_bit_exact(what):
    """The reference trains with gradient checkpointing on (train_animatediff.py:100-101): autograd
    re-enters the forward during backward.  Every op is deterministic (no atomics in the forward,
    fixed-order reductions), so the recomputed forward is bit-identical and so are the gradients
    wherever the backward is deterministic too (the whole motion module)."""
    from torch.utils.checkpoint import checkpoint
    from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    torch.manual_seed(3)
    if what == "motion":
        c, f, hw = 320, 8, 8
        _, mod = make_motion(c, 32)
        mod.cuda()
        for n, p in mod.named_parameters():
            p.requires_grad_(".base." not in n)
        x = torch.randn(2 * f, c, hw, hw, device="cuda").to(bf)
        fn = lambda t: mod(t, num_frames=f)
        params = [p for p in mod.parameters() if p.requires_grad]
    else:
        attn = attn_to(make_attn(640, 10, 2048, 64, train=True, residual=True), "cuda")
        proc = AnimateDiffAttnProcessor2_0()
        x = torch.randn(4, 256, 640, device="cuda").to(bf)
        text = torch.randn(1, 77, 2048, device="cuda").to(bf)
        fn = lambda t: proc(attn, t, text)
        params = [p for l in attn_linears(attn).values() for p in l.parameters() if p.requires_grad]
    dy = torch.randn_like(fn(x.detach()))

    def grads(use_ckpt):
        for p in params:
            p.grad = None
        xi = x.detach().clone().requires_grad_(True)
        y = checkpoint(fn, xi, use_reentrant=False) if use_ckpt else fn(xi)
        y.backward(dy)
        return y.detach().clone(), xi.grad.clone(), [p.grad.clone() for p in params]

    y0, dx0, g0 = grads(False)
    y1, dx1, g1 = grads(True)
    assert torch.equal(y0, y1), "recomputed forward differs"
    assert len(g0) == len(g1) > 0
    if what == "motion":
        assert torch.equal(dx0, dx1)
        for a, b in zip(g0, g1):
            assert torch.equal(a, b)
    else:
        # the spatial attention backward adds its per-key-tile dQ contributions with fp32 atomics
        # (csrc/attn_bwd.cu), so dQ is reproducible only up to fp32 summation order, with or without
        # checkpointing -- and dQ is then rounded to bf16, where a different order can flip a rounding:
        # one bf16 ulp (2^-8 of an element) on isolated elements, run to run.  The bound asserted is
        # that ulp against the tensor's max (observed: 0 ... 6e-5); the bf16 parity tolerance is 1e-2.
        ULP = 2.0 ** -8
        assert rel(dx1, dx0) < ULP
        for a, b in zip(g0, g1):
            assert rel(b, a) < ULP
    # reentrant flavour (what diffusers' _gradient_checkpointing_func uses on older versions)
    xi = x.detach().clone().requires_grad_(True)
    for p in params:
        p.grad = None
    checkpoint(fn, xi, use_reentrant=True).backward(dy)
    assert torch.equal(xi.grad, dx0) if what == "motion" else rel(xi.grad, dx0) < 2.0 ** -8


# ---- config 4: one Stage-2 optimiser step ----------------------------------------------------------------
def test_config4_train_step_vs_oracle_adamw():
    """One Stage-2 step of train_animatediff.py:212-317 over a depth-scaled hot-path stack (SDXL widths
    and module order, one transformer block per attention, 4 frames, latent 16): forward of all 37
    modules, backward from synthetic upstream gradients plus the orthogonality loss (:307-308), global
    gradient-norm clipping (:316), AdamW (:165,:317).  Oracle: the same stack through oracle/*.py in
    fp32 on the CPU under torch autograd, torch.nn.utils.clip_grad_norm_ and torch.optim.AdamW.
    Compared: orth loss, pre-clip gradient norm, every trainable tensor's gradient, every updated
    trainable tensor."""
    from oracle import hotpath as hp, processor as op
    from video_style_transfer_b200 import sdxl_hotpath as hpth
    from video_style_transfer_b200.animatediff.temporal_lora import TemporalLoRALinear, build_spatial_lora_index
    torch.manual_seed(11)
    F_, S, lam, lr, wd, clip = 4, 16, 1e-4, 1e-3, 1e-2, 1.0
    model = hpth.SDXLHotPath(temporal_rank=32, depth_scale=0.1).train()
    with torch.no_grad():      # Stage-1-like LoRA magnitudes so the orth loss and its gradient are not ~0
        for n, p in model.named_parameters():
            if n.endswith("lora_B"):
                p.normal_(0, 0.05)
    plan = model.forward_plan()
    names = {id(m): n for n, m in model.named_modules()}
    hpth.freeze_for_stage2(model, unfreeze_mergers=True)
    # ---- oracle side: fp32 CPU leaves for every parameter --------------------------------------------
    cpu = {n: p.detach().float().clone().requires_grad_(p.requires_grad) for n, p in model.named_parameters()}
    bufs = {n: b.detach().float().clone() for n, b in model.named_buffers()}
    shapes = hpth.level_shapes(F_, S)
    g = torch.Generator().manual_seed(0)
    xs4 = [torch.randn(s, generator=g).to(bf) for s in shapes]
    xs3 = [None] + [torch.randn(s[0], s[2] * s[3], s[1], generator=g).to(bf) for s in shapes[1:]]
    text = torch.randn(1, 77, 2048, generator=g).to(bf)
    up = {"motion": [(1e-2 * torch.randn(s, generator=g)).to(bf) for s in shapes],
          "self": [None] + [(1e-2 * torch.randn(t.shape, generator=g)).to(bf) for t in xs3[1:]],
          "cross": [None] + [(1e-2 * torch.randn(t.shape, generator=g)).to(bf) for t in xs3[1:]]}

    def oproj(prefix):
        ll = prefix + ".lora_layer."
        return op.UnzipProj(cpu[prefix + ".weight"], cpu.get(prefix + ".bias"),
                            cpu[ll + "lora_matrix_dic.content_down.weight"], cpu[ll + "lora_matrix_dic.content_up.weight"],
                            cpu[ll + "lora_matrix_dic.style_down.weight"], cpu[ll + "lora_matrix_dic.style_up.weight"],
                            cpu[ll + "merge_content"], cpu[ll + "merge_style"], infer=True)

    outs, grads = [], []
    for kind, lvl, mod in plan:
        n = names[id(mod)]
        if kind == "motion":
            pre = n + ".temporal_transformer."
            w = {k[len(pre):]: v for k, v in cpu.items() if k.startswith(pre)}
            w["pos_embed.pe"] = bufs[pre + "transformer_blocks.0.pos_embed.pe"]
            y = op.motion_module(xs4[lvl].float(), w, F_, 8, 32, {"rank": 32, "scale": 1.0 / 32})
        else:
            P = op.SpatialAttnParams(mod.heads, oproj(n + ".to_q"), oproj(n + ".to_k"), oproj(n + ".to_v"),
                                     oproj(n + ".to_out.0"))
            y = op.animatediff_attn_processor(P, xs3[lvl].float(), None if kind == "self" else text.float())
        outs.append(y)
        grads.append(up[kind][lvl].float())
    index = build_spatial_lora_index(model)
    assert len(index) > 0
    pairs = []
    for n, m in model.named_modules():
        if n in index and isinstance(m, TemporalLoRALinear):
            sp = [k for k, v in model.named_modules() if v is index[n]][0] + ".lora_matrix_dic."
            pairs.append((cpu[n + ".lora_A"], cpu[n + ".lora_B"], m.scale, cpu[sp + "content_down.weight"],
                          cpu[sp + "content_up.weight"], cpu[sp + "style_down.weight"], cpu[sp + "style_up.weight"]))
    orth_ref = hp.orth_loss_dense(pairs, lam)
    torch.autograd.backward(outs + [orth_ref], grads + [torch.ones_like(orth_ref)])
    train_names = [n for n, p in cpu.items() if p.requires_grad]
    grad_ref = {n: cpu[n].grad.clone() for n in train_names}
    norm_ref = torch.nn.utils.clip_grad_norm_([cpu[n] for n in train_names], clip)
    opt_ref = torch.optim.AdamW([cpu[n] for n in train_names], lr=lr, weight_decay=wd)
    before = {n: cpu[n].detach().clone() for n in train_names}
    opt_ref.step()

    # ---- sm_100a side: fp32 parameters on the device, bf16 activations -----------------------------------
    model.cuda()
    params = [p for p in model.parameters() if p.requires_grad]
    assert sorted(n for n, p in model.named_parameters() if p.requires_grad) == sorted(train_names)
    opt = torch.optim.AdamW(params, lr=lr, weight_decay=wd)
    cu = lambda t: None if t is None else t.cuda()
    orth = hpth.train_step_hotpath(model, plan, [cu(t) for t in xs4], [cu(t) for t in xs3], cu(text),
                                   {k: [cu(t) for t in v] for k, v in up.items()}, F_, opt, index, lam, clip)
    assert abs(orth.item() - orth_ref.item()) <= 2e-3 * abs(orth_ref.item()) + 1e-12
    got = dict(model.named_parameters())
    # gradients (clip_grad_norm_ scaled both sides by clip / norm in place): per tensor 2e-2, and the
    # global norm through the concatenation of all of them
    scale_ref = min(1.0, clip / (norm_ref.item() + 1e-6))
    sq = 0.0
    for n in train_names:
        gr = got[n].grad.float().cpu()
        sq += float((gr.double() ** 2).sum())
        want = grad_ref[n] * scale_ref
        if want.abs().max() > 1e-7 * scale_ref:       # tensors whose gradient is pure rounding noise are skipped
            assert rel(gr, want) < 3e-2, n
    clipped_norm = sq ** 0.5
    assert abs(clipped_norm - min(norm_ref.item(), clip)) < 2e-2 * min(norm_ref.item(), clip)
    # updated parameters: the step itself (Adam's first step is lr * g / (|g| + eps): sign-like, so a
    # bf16-noise-sized gradient may legitimately flip) -- compare where the oracle gradient is clear of
    # that noise, and the whole update through its cosine
    num = den_a = den_b = 0.0
    for n in train_names:
        d_ref = cpu[n].detach() - before[n]
        d_got = got[n].detach().float().cpu() - before[n]
        num += float((d_ref.double() * d_got.double()).sum())
        den_a += float((d_ref.double() ** 2).sum())
        den_b += float((d_got.double() ** 2).sum())
        gref = grad_ref[n] * scale_ref
        clear = gref.abs() > 0.05 * gref.abs().max()
        if clear.any():
            assert (d_got[clear] - d_ref[clear]).abs().max() <= 0.05 * lr, n
    assert num / (den_a ** 0.5 * den_b ** 0.5) > 0.97


# ---- §8f-2 on the device --------------------------------------------------------------------------------
def test_compute_orth_loss_on_device_matches_reference_golden():
    """compute_orth_loss (r x r core algebra) on CUDA tensors against the value the REFERENCE's
    compute_orth_loss produced for the same toy UNet (tests/golden/lora_linear.npz, written by
    oracle/make_golden.py from animatediff/temporal_lora.py:126-166), and its gradients against
    fp32 autograd of the dense restatement."""
    from oracle import hotpath as hp
    from video_style_transfer_b200.animatediff import temporal_lora as tl
    from .test_cpu import _toy_unet, sub, T
    g = np.load(os.path.join(GOLD, "lora_linear.npz"))
    unet = _toy_unet(g)
    tl.inject_temporal_lora(unet, rank=4, alpha=2.0)
    lora = sub(g, "toy/lora/")
    with torch.no_grad():
        for n, m in unet.named_modules():
            if isinstance(m, tl.TemporalLoRALinear):
                m.lora_A.copy_(lora[n + ".lora_A"])
                m.lora_B.copy_(lora[n + ".lora_B"])
    idx_cpu = tl.build_spatial_lora_index(unet)
    pairs = []
    for n, m in unet.named_modules():
        if n in idx_cpu and isinstance(m, tl.TemporalLoRALinear):
            d = idx_cpu[n].lora_matrix_dic
            pairs.append((m.lora_A.detach().clone().requires_grad_(True), m.lora_B.detach().clone().requires_grad_(True),
                          m.scale, d["content_down"].weight.detach(), d["content_up"].weight.detach(),
                          d["style_down"].weight.detach(), d["style_up"].weight.detach()))
    ref = hp.orth_loss_dense(pairs, 1e-2)
    assert abs(ref.item() - float(g["toy/orth_loss"])) <= 1e-5 * abs(float(g["toy/orth_loss"]))
    ref.backward()
    unet.cuda()
    idx = tl.build_spatial_lora_index(unet)
    loss = tl.compute_orth_loss(unet, idx, 1e-2)
    assert loss.is_cuda
    assert abs(loss.item() - float(g["toy/orth_loss"])) <= 1e-4 * abs(float(g["toy/orth_loss"]))
    loss.backward()
    i = 0
    for n, m in unet.named_modules():
        if n in idx and isinstance(m, tl.TemporalLoRALinear):
            assert rel(m.lora_A.grad, pairs[i][0].grad) < 1e-4
            assert rel(m.lora_B.grad, pairs[i][1].grad) < 1e-4
            i += 1
    assert i == len(pairs) > 0


# ---- folded vs low-rank adapters at small magnitudes (round-1 advisor finding) -----------------------------
def test_folded_weight_keeps_small_adapters(monkeypatch):
    """W_eff = W + B'A is accumulated in fp32 and rounded to bf16 once.  With adapter deltas far
    below one bf16 ulp of W (lora_B ~ 1e-3, mergers ~ 0.05: an early Stage-2 checkpoint) the folded
    projection still has to agree with the low-rank form and with the fp32 oracle."""
    from oracle import hotpath as hp
    from video_style_transfer_b200 import fused
    from video_style_transfer_b200.unziplora_unet.lora_linear import LoRACompatibleLinear
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import UnZipLoRALinearLayerInfer
    torch.manual_seed(4)
    i, o, r = 640, 640, 64
    lin = LoRACompatibleLinear(i, o, bias=True)
    lay = UnZipLoRALinearLayerInfer(i, o, r, ["content", "style"])
    with torch.no_grad():
        for k in ("content_up", "style_up"):
            lay.lora_matrix_dic[k].weight.normal_(0, 1e-3)
        lay.merge_content.fill_(0.05)
        lay.merge_style.fill_(0.05)
    lin.set_lora_layer(lay)
    x = torch.randn(512, i)
    d = lay.lora_matrix_dic
    delta = hp.unziplora_delta(x, None, d["content_down"].weight, d["content_up"].weight, d["style_down"].weight,
                               d["style_up"].weight, lay.merge_content, lay.merge_style)
    want = hp.lora_compatible_linear(x, lin.weight, lin.bias, delta, 1.0).detach()
    lin.cuda()
    xd = x.cuda().to(bf)
    monkeypatch.setattr(fused, "MERGE_LORA", True)
    y_fold = lin(xd)
    monkeypatch.setattr(fused, "MERGE_LORA", False)
    y_low = lin(xd)
    assert rel(y_fold, want) < 1e-2 and rel(y_low, want) < 1e-2
    assert rel(y_fold, y_low.float().cpu()) < 1e-2
    # the fold happens in fp32: W_eff is the correctly rounded W + B'A, not bf16(W) + bf16 noise
    w_eff = fused.merged_weight([lin.weight], [lin.site_downs()], [lin.site_up(1.0)])
    bprime = torch.cat([d["content_up"].weight * 0.05, d["style_up"].weight * 0.05], dim=1).float()
    a = torch.cat([d["content_down"].weight, d["style_down"].weight], dim=0).float()
    exact = (lin.weight.float() + bprime.to(bf).float() @ a.to(bf).float()).to(bf)
    assert (w_eff != exact).float().mean().item() < 1e-3          # identical up to fp32 summation order


def test_operand_cache_does_not_grow_across_parameter_updates():
    """Round-1 advisor finding: every merger / B update used to leave one more folded W_eff in
    _host._cache.  Entries derived from rebuilt operands now overwrite their predecessors."""
    from video_style_transfer_b200 import _host
    torch.manual_seed(0)
    attn = attn_to(make_attn(640, 10, 640, 64), "cuda", bf)
    from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0
    proc = AnimateDiffAttnProcessor2_0()
    x = torch.randn(2, 64, 640, device="cuda").to(bf)
    sizes, outs = [], []
    for it in range(6):
        with torch.no_grad():
            for lin in attn_linears(attn).values():
                lin.lora_layer.merge_content.mul_(0.9)
                lin.lora_layer.lora_matrix_dic["style_up"].weight.mul_(1.01)
            outs.append(proc(attn, x).float().cpu())
        sizes.append(_host.cache_stats()["bytes"])
    assert sizes[-1] == sizes[1], sizes             # steady state after the first rebuild
    assert rel(outs[-1], outs[0]) > 1e-4            # and the updates did reach the output
    # writes through .data bump no version: documented, with an explicit way out
    with torch.no_grad():
        y0 = proc(attn, x).float().cpu()
        attn.to_q.lora_layer.merge_style.data.zero_()
        attn.to_q.lora_layer.merge_content.data.zero_()
        _host.invalidate_caches()
        y1 = proc(attn, x).float().cpu()
    assert rel(y1, y0) > 1e-4


# ---- step-invariant text K/V, batched CFG -------------------------------------------------------------------
def test_text_kv_cache_tracks_text_and_parameters():
    from oracle import processor as op
    from video_style_transfer_b200.animatediff import attention_processor as ap
    torch.manual_seed(9)
    attn = make_attn(640, 10, 2048, 64)
    x, t1, t2 = torch.randn(4, 256, 640), torch.randn(1, 77, 2048), torch.randn(1, 77, 2048)
    P = oracle_attn(attn)
    want1, want2 = op.animatediff_attn_processor(P, x, t1), op.animatediff_attn_processor(P, x, t2)
    attn_to(attn, "cuda", bf)
    proc = ap.AnimateDiffAttnProcessor2_0()
    xd, text = x.cuda().to(bf), t1.cuda().to(bf)
    with torch.no_grad():
        n0 = len(ap._text_kv_cache)
        y1 = proc(attn, xd, text)
        assert len(ap._text_kv_cache) == n0 + 1
        from video_style_transfer_b200 import _lib
        l0 = _lib.launch_count
        y1b = proc(attn, xd, text)
        launches_cached = _lib.launch_count - l0
        assert torch.equal(y1, y1b) and rel(y1, want1) < 2e-2
        text.copy_(t2.cuda())                                   # new prompt in the same buffer
        l0 = _lib.launch_count
        y2 = proc(attn, xd, text)
        assert _lib.launch_count - l0 > launches_cached         # the K/V projection ran again
        assert rel(y2, want2) < 2e-2 and len(ap._text_kv_cache) == n0 + 1
        attn.to_k.lora_layer.merge_content.zero_()              # a merger update invalidates too
        attn.to_k.lora_layer.merge_style.zero_()
        P.to_k.m_c.zero_(), P.to_k.m_s.zero_()
        assert rel(proc(attn, xd, text), op.animatediff_attn_processor(P, x, t2)) < 2e-2
    # with autograd on the text (Stage-1 trains through it) nothing is cached
    text_g = t1.cuda().to(bf).requires_grad_(True)
    attn2 = attn_to(make_attn(640, 10, 2048, 64, train=True), "cuda")
    n1 = len(ap._text_kv_cache)
    proc(attn2, xd, text_g).float().sum().backward()
    assert text_g.grad is not None and len(ap._text_kv_cache) == n1


def test_batched_cfg_equals_two_pass_and_graph_set_text():
    """uncond | cond stacked along the batch (one B=2 pass, pipeline_animatediff_xl.py:353,373-375)
    returns exactly what the two B=1 passes of inference_animatediff.py:110-121 return, per half;
    GraphedDenoiseStep.set_text swaps the prompt under a captured graph."""
    from video_style_transfer_b200 import sdxl_hotpath as hpth
    torch.manual_seed(13)
    F_, S = 4, 16
    with torch.device("cuda"):
        model = hpth.SDXLHotPath(temporal_rank=8, depth_scale=0.1)
    model = model.to(bf).eval()
    plan = model.forward_plan()
    g = torch.Generator().manual_seed(5)
    sh2 = hpth.level_shapes(F_, S, batch=2)
    x4 = [torch.randn(s, generator=g).to(bf).cuda() for s in sh2]
    x3 = [None] + [torch.randn(s[0], s[2] * s[3], s[1], generator=g).to(bf).cuda() for s in sh2[1:]]
    tu, tc = (torch.randn(1, 77, 2048, generator=g).to(bf).cuda() for _ in range(2))
    both = hpth.run_unet_hotpath(plan, x4, x3, torch.cat([tu, tc]), F_)
    for half, text in ((0, tu), (1, tc)):
        sl = slice(half * F_, (half + 1) * F_)
        one = hpth.run_unet_hotpath(plan, [t[sl].contiguous() for t in x4],
                                    [None] + [t[sl].contiguous() for t in x3[1:]], text, F_)
        for a, b in zip(both, one):
            assert torch.equal(a[sl], b), "batched CFG differs from the two-pass form"
    # a plan whose last module per level is a text cross-attention, so the outputs depend on the prompt
    plan_t = plan + [next(p for p in plan if p[0] == "cross" and p[1] == lvl) for lvl in (1, 2)]
    both = hpth.run_unet_hotpath(plan_t, x4, x3, torch.cat([tu, tc]), F_)
    graphed = hpth.GraphedDenoiseStep(plan_t, x4, x3, [torch.cat([tu, tc])], F_)
    for a, b in zip(graphed.replay(), both):
        assert torch.equal(a, b)
    tu2, tc2 = (torch.randn(1, 77, 2048, generator=g).to(bf).cuda() for _ in range(2))
    want = hpth.run_unet_hotpath(plan_t, x4, x3, torch.cat([tu2, tc2]), F_)
    graphed.set_text([torch.cat([tu2, tc2])])
    got = graphed.replay()
    assert not torch.equal(got[2], both[2]) and not torch.equal(got[1], both[1])
    for a, b in zip(got, want):
        assert torch.equal(a, b)


# ---- §8f-1, spatial half: the full BasicTransformerBlock ----------------------------------------------------
def _block_and_oracle(c, heads, ctx, rank, train=False):
    from video_style_transfer_b200.unziplora_unet.unzip_attention import BasicTransformerBlock
    from video_style_transfer_b200 import sdxl_hotpath as hpth

    class _A(torch.nn.Module):                       # diffusers-protocol Attention duck type as an nn.Module
        def __init__(self, ns):
            super().__init__()
            self.__dict__.update({k: v for k, v in vars(ns).items() if not isinstance(v, torch.nn.Module)})
            self.to_q, self.to_k, self.to_v, self.to_out = ns.to_q, ns.to_k, ns.to_v, ns.to_out
            self.processor = None

        def forward(self, x, encoder_hidden_states=None, **kw):
            from video_style_transfer_b200.animatediff.attention_processor import AnimateDiffAttnProcessor2_0
            if self.processor is None:
                self.processor = AnimateDiffAttnProcessor2_0()
            return self.processor(self, x, encoder_hidden_states=encoder_hidden_states, **kw)

    a1, a2 = make_attn(c, heads, c, rank, train=train), make_attn(c, heads, ctx, rank, train=train)
    blk = BasicTransformerBlock(c, heads, 64, cross_attention_dim=ctx, attn1=_A(a1), attn2=_A(a2))
    with torch.no_grad():
        for n, p in blk.named_parameters():
            if n.startswith("norm") or (n.startswith("ff") and n.endswith("bias")):
                p.add_(0.1 * torch.randn_like(p))
    return blk, a1, a2


def _block_weights(blk, grad=False):
    return {k: v.detach().float().cpu().clone().requires_grad_(grad)
            for k, v in blk.named_parameters() if k.startswith(("norm", "ff."))}


def test_spatial_transformer_block_vs_oracle_l2_geometry():
    """SDXL L2 spatial BasicTransformerBlock (unzip_attention.py:113-239: LN -> self-attn -> +h; LN ->
    text cross-attn -> +h; LN -> GEGLU FF -> +h) at C=1280, 20 heads, 1024 tokens, 4 frames, text
    (1, 77, 2048), UnZipLoRA rank 64, against the fp32 oracle."""
    from oracle import processor as op
    torch.manual_seed(21)
    C, heads, ctx = 1280, 20, 2048
    blk, a1, a2 = _block_and_oracle(C, heads, ctx, 64)
    x, text = torch.randn(4, 1024, C), torch.randn(1, 77, ctx)
    want = op.basic_transformer_block(oracle_attn(a1), oracle_attn(a2), _block_weights(blk), x, text)
    blk = blk.cuda().to(bf)
    with torch.no_grad():
        y = blk(x.cuda().to(bf), encoder_hidden_states=text.cuda().to(bf))
    assert y.shape == x.shape and y.dtype == bf
    assert rel(y, want) < 2e-2
    assert torch.nn.functional.cosine_similarity(y.float().cpu().flatten(), want.flatten(), dim=0) > 0.9999


def test_spatial_transformer_block_backward_vs_oracle():
    """Gradients of the full block (input, text, LayerNorm / FF parameters, UnZipLoRA A / B / mergers) against
    fp32 autograd of the oracle at a size the CPU differentiates in seconds."""
    from oracle import processor as op
    torch.manual_seed(22)
    C, heads, ctx = 320, 5, 256
    blk, a1, a2 = _block_and_oracle(C, heads, ctx, 16, train=True)
    x, text = torch.randn(3, 200, C).to(bf), torch.randn(1, 77, ctx).to(bf)
    dy = (torch.randn(3, 200, C) * 0.1).to(bf)
    P1, P2 = oracle_attn(a1, grad=True, infer=False), oracle_attn(a2, grad=True, infer=False)
    w = _block_weights(blk, grad=True)
    xr, tr = x.float().requires_grad_(True), text.float().requires_grad_(True)
    want = op.basic_transformer_block(P1, P2, w, xr, tr)
    want.backward(dy.float())
    blk = blk.cuda()
    xd, td = x.cuda().requires_grad_(True), text.cuda().requires_grad_(True)
    y = blk(xd, encoder_hidden_states=td)
    assert rel(y, want.detach()) < 2e-2
    y.backward(dy.cuda())
    assert rel(xd.grad, xr.grad) < 3e-2 and rel(td.grad, tr.grad) < 3e-2
    got = dict(blk.named_parameters())
    for k, v in w.items():
        if v.requires_grad:
            assert rel(got[k].grad, v.grad) < 3e-2, k
    for attn, P in ((a1, P1), (a2, P2)):
        for n, pr in zip(("to_q", "to_k", "to_v", "to_out"), (P.to_q, P.to_k, P.to_v, P.to_out)):
            lay = attn_linears(attn)[n].lora_layer
            assert rel(lay.merge_content.grad, pr.m_c.grad) < 3e-2, n
            assert rel(lay.lora_matrix_dic["style_up"].weight.grad, pr.b_s.grad) < 3e-2, n
            assert rel(lay.lora_matrix_dic["content_down"].weight.grad, pr.a_c.grad) < 3e-2, n
