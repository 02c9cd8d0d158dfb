"""Builders shared by the -m gpu parity tests: a diffusers-protocol ``attn`` duck type on the
drop-in linears, its oracle twin (oracle/processor.py), a motion module and its oracle weights."""
import types

import torch


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def make_attn(c, heads, kv, rank=64, train=False, residual=False, rescale=1.0, out_gain=1.0,
              trainable=("lora", "merger")):
    """to_q / to_k / to_v / to_out = LoRACompatibleLinear + UnZipLoRALinearLayer(Infer), fp32 on the
    CPU (move with ``attn_to``).  out_gain scales to_out (weight, bias, up-projections): the
    residual-branch gain of the synthetic denoiser in the 25-step test."""
    from video_style_transfer_b200.unziplora_unet.lora_linear import LoRACompatibleLinear
    from video_style_transfer_b200.unziplora_unet.unziplora_linear_layer import (UnZipLoRALinearLayer,
                                                                                 UnZipLoRALinearLayerInfer)

    def mk(i, o, bias, gain=1.0):
        lin = LoRACompatibleLinear(i, o, bias=bias)
        lay = (UnZipLoRALinearLayer(i, o, rank, ["content", "style"], lora_matrix_num=2) if train
               else UnZipLoRALinearLayerInfer(i, o, rank, ["content", "style"]))
        with torch.no_grad():
            lay.merge_content.uniform_(0, 1)
            lay.merge_style.uniform_(0, 1)
            if gain != 1.0:
                lin.weight.mul_(gain)
                if bias:
                    lin.bias.mul_(gain)
                for k in ("content_up", "style_up"):
                    lay.lora_matrix_dic[k].weight.mul_(gain)
        lin.set_lora_layer(lay)
        for n, p in lin.named_parameters():
            p.requires_grad_(train and (("merge_" in n and "merger" in trainable) or
                                        ("lora_matrix_dic" in n and "lora" in trainable)))
        return lin

    return types.SimpleNamespace(
        spatial_norm=None, group_norm=None, norm_cross=False, heads=heads,
        to_q=mk(c, c, False), to_k=mk(kv, c, False), to_v=mk(kv, c, False),
        to_out=torch.nn.ModuleList([mk(c, c, True, out_gain), torch.nn.Dropout(0.0)]),
        residual_connection=residual, rescale_output_factor=rescale)


def attn_linears(attn):
    return {"to_q": attn.to_q, "to_k": attn.to_k, "to_v": attn.to_v, "to_out": attn.to_out[0]}


def attn_to(attn, device=None, dtype=None):
    for lin in attn_linears(attn).values():
        lin.to(device=device, dtype=dtype)
    return attn


def oracle_attn(attn, grad=False, infer=True, device="cpu"):
    """oracle.processor.SpatialAttnParams holding fp32 clones of the attn's tensors."""
    from oracle import processor as op

    def f(t, g):
        return None if t is None else t.detach().float().to(device).clone().requires_grad_(g)

    def oproj(l):
        ll, d = l.lora_layer, l.lora_layer.lora_matrix_dic
        return op.UnzipProj(f(l.weight, False), f(l.bias, False), f(d["content_down"].weight, grad),
                            f(d["content_up"].weight, grad), f(d["style_down"].weight, grad),
                            f(d["style_up"].weight, grad), f(ll.merge_content, grad), f(ll.merge_style, grad),
                            infer=infer)

    L = attn_linears(attn)
    return op.SpatialAttnParams(attn.heads, oproj(L["to_q"]), oproj(L["to_k"]), oproj(L["to_v"]),
                                oproj(L["to_out"]), residual_connection=attn.residual_connection,
                                rescale_output_factor=attn.rescale_output_factor)


def make_motion(c, rank=32, lora_b_std=0.1, out_gain=1.0, randomize_affine=True):
    """(holder, motion module): MotionModule(c, 8) with temporal LoRA injected (non-zero B)."""
    from video_style_transfer_b200.animatediff.temporal_lora import TemporalLoRALinear, inject_temporal_lora
    from video_style_transfer_b200.motion_module import MotionModule
    holder = torch.nn.Module()
    holder.motion_modules = torch.nn.ModuleList([MotionModule(c, 8)])
    with torch.no_grad():
        if randomize_affine:
            for n, p in holder.named_parameters():
                if "norm" in n or n.endswith("bias"):
                    p.add_(0.1 * torch.randn_like(p))
        tt = holder.motion_modules[0].temporal_transformer
        tt.proj_out.weight.mul_(out_gain)
        tt.proj_out.bias.mul_(out_gain)
    inject_temporal_lora(holder, rank=rank, alpha=1.0)
    with torch.no_grad():
        for m in holder.modules():
            if isinstance(m, TemporalLoRALinear):
                m.lora_B.normal_(0, lora_b_std)
    return holder, holder.motion_modules[0]


def motion_weights(mm, device="cpu", grad_names=()):
    """Oracle weight dict (oracle/processor.py:motion_module key names) from a MotionModule."""
    w = {}
    for k, v in mm.temporal_transformer.state_dict().items():
        t = v.detach().float().to(device).clone()
        if any(s in k for s in grad_names) and ".base." not in k:
            t.requires_grad_(True)
        w[k] = t
    w["pos_embed.pe"] = w["transformer_blocks.0.pos_embed.pe"]
    return w


def sdxl_euler_sigmas(steps):
    """EulerDiscreteScheduler as SDXL configures it (diffusers, un-vendored -- recalled:
    scaled_linear betas 0.00085..0.012 over 1000 train steps, 'leading' timestep spacing,
    steps_offset 1; sigma = sqrt((1 - abar) / abar), a final 0 appended).  The reference drives it
    at inference_animatediff.py:67 (set_timesteps), :95 (init_noise_sigma), :106 (scale_model_input),
    :131 (step, epsilon prediction: x <- x + (sigma_next - sigma) * eps)."""
    betas = torch.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000, dtype=torch.float64) ** 2
    abar = torch.cumprod(1.0 - betas, 0)
    sig_all = ((1.0 - abar) / abar) ** 0.5
    ts = (torch.arange(steps) * (1000 // steps)).flip(0) + 1
    sig = torch.cat([sig_all[ts], torch.zeros(1, dtype=torch.float64)]).float()
    init_noise_sigma = float((sig[0] ** 2 + 1) ** 0.5)
    return sig, init_noise_sigma
