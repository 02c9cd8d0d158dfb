"""Oracle restatement of the attention processors and the motion block (test infrastructure).

``SpatialAttnParams`` is the plain-tensor state of one diffusers ``Attention`` module wired
the way the live path wires it (unziplora_unet/utils.py:435-483): to_q/k/v/out are
``LoRACompatibleLinear`` whose ``lora_layer`` is a ``UnZipLoRALinearLayerInfer``.
"""
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from . import hotpath as hp


@dataclass
class UnzipProj:
    """One LoRACompatibleLinear + UnZipLoRA layer as raw tensors."""
    weight: torch.Tensor
    bias: Optional[torch.Tensor]
    a_c: torch.Tensor
    b_c: torch.Tensor
    a_s: torch.Tensor
    b_s: torch.Tensor
    m_c: torch.Tensor
    m_s: torch.Tensor
    forward_type: str = "both"
    masked_content: bool = False
    masked_style: bool = False
    infer: bool = True
    mask_c: Optional[torch.Tensor] = None
    mask_s: Optional[torch.Tensor] = None

    def __call__(self, x, scale=1.0, x1=None, x2=None):
        # lora_linear.py:74-81: lora_layer(hidden_states) (single stream)
        # lora_unzip.py:66-75: lora_layer(hidden_states_1, hidden_states_2) (three streams)
        xc = x if x1 is None else x1
        xs = x2
        delta = hp.unziplora_delta(xc, xs, self.a_c, self.b_c, self.a_s, self.b_s, self.m_c,
                                   self.m_s, self.forward_type, self.masked_content,
                                   self.masked_style, self.infer, self.mask_c, self.mask_s)
        return hp.lora_compatible_linear(x, self.weight, self.bias, delta, scale)


@dataclass
class SpatialAttnParams:
    heads: int
    to_q: UnzipProj
    to_k: UnzipProj
    to_v: UnzipProj
    to_out: UnzipProj
    residual_connection: bool = False
    rescale_output_factor: float = 1.0
    group_norm: Optional[Dict[str, torch.Tensor]] = None   # {"weight","bias","groups","eps"}


def animatediff_attn_processor(p: SpatialAttnParams, hidden_states, encoder_hidden_states=None,
                               attention_mask=None, scale: float = 1.0):
    """AnimateDiffAttnProcessor2_0.__call__.  animatediff/attention_processor.py:18-96.
    (spatial_norm / norm_cross are None on the SDXL path and are not restated.)"""
    residual = hidden_states                                             # :28
    input_ndim = hidden_states.ndim
    if input_ndim == 4:                                                  # :33-36
        bsz, ch, hh, ww = hidden_states.shape
        hidden_states = hidden_states.view(bsz, ch, hh * ww).transpose(1, 2)
    batch = hidden_states.shape[0]                                       # :39
    if p.group_norm is not None:                                         # :50-51
        g = p.group_norm
        hidden_states = hp.groupnorm5d(hidden_states.transpose(1, 2), g["weight"], g["bias"],
                                       g["groups"], g["eps"]).transpose(1, 2)
    q = p.to_q(hidden_states, scale)                                     # :54-55
    ehs = hidden_states if encoder_hidden_states is None else encoder_hidden_states
    if ehs.shape[0] != batch:                                            # :62-66
        ehs = ehs.repeat_interleave(batch // ehs.shape[0], dim=0)
    k = p.to_k(ehs, scale)                                               # :68-69
    v = p.to_v(ehs, scale)
    d = k.shape[-1] // p.heads                                           # :71-72

    def sp(t):
        return t.reshape(batch, -1, p.heads, d).transpose(1, 2)          # :74-76

    o = hp.sdpa(sp(q), sp(k), sp(v), attention_mask)                     # :78-80
    o = o.transpose(1, 2).reshape(batch, -1, p.heads * d)                # :82
    o = p.to_out(o, scale)                                               # :85 (dropout p=0 :86)
    if input_ndim == 4:                                                  # :88-89
        o = o.transpose(-1, -2).reshape(bsz, ch, hh, ww)
    if p.residual_connection:                                            # :91-92
        o = o + residual
    return o / p.rescale_output_factor                                   # :94


def stage1_attn_processor(p: SpatialAttnParams, hidden_states, ehs=None, ehs_content=None,
                          ehs_style=None, attention_mask=None):
    """Stage-1 AttnProcessor2_0 with three text streams.
    unziplora_unet/unzip_attention_processor.py:664-759 executed against
    lora_unzip.py:66-75 semantics (hidden_states_content/_style -> hidden_states_1/_2);
    self-attention aliases all three streams to hidden_states (:711-714).  ``scale`` is accepted
    by the reference (:675) but never forwarded: every projection runs at its default 1.0.
    (spatial_norm / group_norm / norm_cross are None on the SDXL path and are not restated.)"""
    residual = hidden_states                                             # :677
    input_ndim = hidden_states.ndim
    if input_ndim == 4:                                                  # :683-685
        bsz, ch, hh, ww = hidden_states.shape
        hidden_states = hidden_states.view(bsz, ch, hh * ww).transpose(1, 2)
    batch = hidden_states.shape[0] if ehs is None else ehs.shape[0]      # :687-689
    q = p.to_q(hidden_states, 1.0, hidden_states, hidden_states)         # :706-708
    if ehs is None:                                                      # :711-714
        ehs = ehs_content = ehs_style = hidden_states
    k = p.to_k(ehs, 1.0, ehs_content, ehs_style)                         # :720-725
    v = p.to_v(ehs, 1.0, ehs_content, ehs_style)
    d = k.shape[-1] // p.heads                                           # :727-728

    def sp(t):
        return t.reshape(batch, -1, p.heads, d).transpose(1, 2)          # :730-733

    o = hp.sdpa(sp(q), sp(k), sp(v), attention_mask)                     # :737-739
    o = o.transpose(1, 2).reshape(batch, -1, p.heads * d)                # :741
    o = p.to_out(o, 1.0, o, o)                                           # :745-747
    if input_ndim == 4:                                                  # :751-752
        o = o.transpose(-1, -2).reshape(bsz, ch, hh, ww)
    if p.residual_connection:                                            # :754-755
        o = o + residual
    return o / p.rescale_output_factor                                   # :757


def basic_transformer_block(p1: SpatialAttnParams, p2: SpatialAttnParams, w: Dict[str, torch.Tensor],
                            hidden_states, encoder_hidden_states=None, scale: float = 1.0):
    """Spatial BasicTransformerBlock.forward for SDXL (norm_type "layer_norm", GEGLU feed-forward).
    unziplora_unet/unzip_attention.py:113-239: h = attn1(norm1(h)) + h (:128-166);
    h = attn2(norm2(h), text) + h (:176-203); h = ff(norm3(h)) + h (:206-235).  The class subclasses
    diffusers' block (un-vendored), so the wrapper itself cannot be imported here: PARITY UNPINNED for
    the composition, every piece is pinned (processor: tests/golden/attn_processor.npz; LayerNorm / GELU:
    tests/test_cpu.py against torch.nn.functional; GEGLU = hidden * gelu(gate), diffusers recalled).
    ``w``: norm{1,2,3}.{weight,bias}, ff.net.0.proj.{weight,bias}, ff.net.2.{weight,bias}."""
    h = hidden_states
    h = animatediff_attn_processor(p1, hp.layernorm(h, w["norm1.weight"], w["norm1.bias"]), None, scale=scale) + h
    h = animatediff_attn_processor(p2, hp.layernorm(h, w["norm2.weight"], w["norm2.bias"]),
                                   encoder_hidden_states, scale=scale) + h
    hn = hp.layernorm(h, w["norm3.weight"], w["norm3.bias"])
    hg = hn @ w["ff.net.0.proj.weight"].t() + w["ff.net.0.proj.bias"]
    hid, gate = hg.chunk(2, dim=-1)
    return (hid * hp.gelu(gate)) @ w["ff.net.2.weight"].t() + w["ff.net.2.bias"] + h


# --------------------------------------------------------------------------------------
# Motion block (diffusers TransformerTemporalModel, un-vendored: recalled, PARITY UNPINNED)
# --------------------------------------------------------------------------------------


def temporal_attention(x, w: Dict[str, torch.Tensor], prefix: str, heads: int, lora=None):
    """Motion-module ``Attention`` with the default processor: to_q/k/v (no bias),
    frame-axis SDPA, to_out.0 (bias).  Projections are TemporalLoRALinear when
    ``lora`` = {"rank","scale"} (temporal_lora.py:29-32, keys ``{proj}.base.weight``,
    ``{proj}.lora_A``, ``{proj}.lora_B``)."""
    n, f, c = x.shape

    def proj(name, t):
        if lora is None:
            return hp.lora_compatible_linear(t, w[prefix + name + ".weight"],
                                             w.get(prefix + name + ".bias"))
        return hp.temporal_lora_linear(t, w[prefix + name + ".base.weight"],
                                       w.get(prefix + name + ".base.bias"),
                                       w[prefix + name + ".lora_A"], w[prefix + name + ".lora_B"],
                                       lora["scale"])

    q, k, v = proj("to_q", x), proj("to_k", x), proj("to_v", x)
    d = c // heads

    def sp(t):
        return t.reshape(n, f, heads, d).transpose(1, 2)

    o = hp.sdpa(sp(q), sp(k), sp(v)).transpose(1, 2).reshape(n, f, c)
    return proj("to_out.0", o)


def motion_module_tokens(y, w: Dict[str, torch.Tensor], num_frames: int, heads: int = 8, lora=None):
    """The position-wise body of the motion module on token rows (N, F, C): proj_in ->
    [LN,+PE,attn1,+res; LN,+PE,attn2,+res; LN,GEGLU FF,+res] -> proj_out.  Every row of N is
    independent, which is what the sequence-parallel split over positions relies on."""
    y = y @ w["proj_in.weight"].t() + w["proj_in.bias"]
    pe = w["pos_embed.pe"]
    tb = "transformer_blocks.0."
    for attn, norm in (("attn1", "norm1"), ("attn2", "norm2")):
        yn = hp.layernorm(y, w[tb + norm + ".weight"], w[tb + norm + ".bias"])
        yn = yn + pe[:, :num_frames, :]
        y = y + temporal_attention(yn, w, tb + attn + ".", heads, lora)
    yn = hp.layernorm(y, w[tb + "norm3.weight"], w[tb + "norm3.bias"])
    hg = yn @ w[tb + "ff.net.0.proj.weight"].t() + w[tb + "ff.net.0.proj.bias"]
    hid, gate = hg.chunk(2, dim=-1)
    y = y + ((hid * hp.gelu(gate)) @ w[tb + "ff.net.2.weight"].t() + w[tb + "ff.net.2.bias"])
    y = y @ w["proj_out.weight"].t() + w["proj_out.bias"]
    return y


def motion_module(x, w: Dict[str, torch.Tensor], num_frames: int, heads: int = 8,
                  groups: int = 32, lora=None):
    """x (B*F, C, H, W) -> same shape.  GroupNorm5D -> (BHW,F,C) -> proj_in ->
    [LN,+PE,attn1,+res; LN,+PE,attn2,+res; LN,GEGLU FF,+res] -> proj_out -> back -> +res.
    Key names follow diffusers (temporal_lora.py:76-84,100-111 show the naming)."""
    bf, c, h, wd = x.shape
    b = bf // num_frames
    residual = x
    x5 = x.reshape(b, num_frames, c, h, wd).permute(0, 2, 1, 3, 4)
    x5 = hp.groupnorm5d(x5, w["norm.weight"], w["norm.bias"], groups, 1e-6)
    y = hp.rearrange_fwd(x5)
    y = motion_module_tokens(y, w, num_frames, heads, lora)
    y5 = hp.rearrange_bwd(y, b, h, wd)                     # (B,C,F,H,W)
    out = y5.permute(0, 2, 1, 3, 4).reshape(bf, c, h, wd)
    return out + residual
