"""Plain torch-CPU fp32 restatement of the reference hot path (oracle; test infrastructure).

Each function cites the reference file:line (relative to /root/reference) it follows.  The
restatements deliberately avoid the fused library ops the reference calls
(``nn.MultiheadAttention``, ``F.scaled_dot_product_attention``) so that they state the
arithmetic explicitly; ``oracle/make_golden.py`` checks them against the reference's own
modules.
"""
import math
from typing import Dict, Optional

import torch

# --------------------------------------------------------------------------------------
# a9: layout rearrange + sinusoidal positional add
# --------------------------------------------------------------------------------------


def positional_table(channels: int, max_len: int = 32) -> torch.Tensor:
    """pe[0,p,2i]=sin(p*w_i), pe[0,p,2i+1]=cos(p*w_i), w_i=exp(2i*(-ln 1e4/C)).

    animatediff/temporal_transformer.py:11-18 (identical to diffusers'
    SinusoidalPositionalEmbedding, un-vendored)."""
    position = torch.arange(max_len).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, channels, 2) * (-math.log(10000.0) / channels))
    pe = torch.zeros(1, max_len, channels)
    pe[0, :, 0::2] = torch.sin(position * div_term)
    pe[0, :, 1::2] = torch.cos(position * div_term)
    return pe


def rearrange_fwd(x: torch.Tensor, pe: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(B,C,F,H,W) -> (B*H*W, F, C) [+ pe[:, :F]].  temporal_transformer.py:129-133."""
    b, c, f, h, w = x.shape
    y = x.permute(0, 3, 4, 2, 1).reshape(-1, f, c)
    if pe is not None:
        y = y + pe[:, :f, :]
    return y


def rearrange_bwd(y: torch.Tensor, b: int, h: int, w: int) -> torch.Tensor:
    """(B*H*W, F, C) -> (B,C,F,H,W) contiguous.  temporal_transformer.py:143-144."""
    n, f, c = y.shape
    return y.reshape(b, h, w, f, c).permute(0, 4, 3, 1, 2).contiguous()


# --------------------------------------------------------------------------------------
# norms
# --------------------------------------------------------------------------------------


def layernorm(x, weight, bias, eps: float = 1e-5):
    """LayerNorm over the last dim (biased variance).  temporal_transformer.py:44,52,111."""
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    y = (x - mu) / torch.sqrt(var + eps)
    if weight is not None:
        y = y * weight
    if bias is not None:
        y = y + bias
    return y


def groupnorm5d(x, weight, bias, groups: int = 32, eps: float = 1e-6):
    """GroupNorm(groups, C) on a (B,C,F,H,W) view: statistics over (C/groups, F, H, W).

    Live path: diffusers TransformerTemporalModel.norm (un-vendored, recalled); in-tree
    analogue call animatediff/attention_processor.py:50-51."""
    b, c = x.shape[:2]
    xr = x.reshape(b, groups, -1)
    mu = xr.mean(dim=-1, keepdim=True)
    var = ((xr - mu) ** 2).mean(dim=-1, keepdim=True)
    y = ((xr - mu) / torch.sqrt(var + eps)).reshape(x.shape)
    shape = [1, c] + [1] * (x.dim() - 2)
    return y * weight.reshape(shape) + bias.reshape(shape)


# --------------------------------------------------------------------------------------
# a2 / a8: attention core
# --------------------------------------------------------------------------------------


def sdpa(q, k, v, mask=None):
    """softmax(q k^T / sqrt(d) + mask) v, non-causal, no dropout.
    q (B,h,Sq,d), k/v (B,h,Sk,d).  attention_processor.py:78-80."""
    d = q.shape[-1]
    s = (q @ k.transpose(-1, -2)) / math.sqrt(d)
    if mask is not None:
        s = s + mask
    p = torch.softmax(s, dim=-1)
    return p @ v


def mha_self(x, in_w, in_b, out_w, out_b, heads: int):
    """nn.MultiheadAttention(batch_first=True)(x,x,x) restated.
    temporal_transformer.py:45-50,70: packed in_proj (3C,C) rows = [q;k;v]."""
    n, f, c = x.shape
    d = c // heads
    qkv = x @ in_w.t() + in_b
    q, k, v = qkv.split(c, dim=-1)

    def sp(t):
        return t.reshape(n, f, heads, d).transpose(1, 2)

    o = sdpa(sp(q), sp(k), sp(v)).transpose(1, 2).reshape(n, f, c)
    return o @ out_w.t() + out_b


def gelu(x):
    """Exact (erf) GELU, nn.GELU() default.  temporal_transformer.py:55."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


# --------------------------------------------------------------------------------------
# a11: standalone temporal transformer (config 1)
# --------------------------------------------------------------------------------------


def temporal_block(x, p: Dict[str, torch.Tensor], prefix: str, heads: int):
    """TemporalTransformerBlock.forward.  temporal_transformer.py:61-76."""
    g = lambda k: p[prefix + k]
    xn = layernorm(x, g("norm1.weight"), g("norm1.bias"))
    x = x + mha_self(xn, g("attn.in_proj_weight"), g("attn.in_proj_bias"),
                     g("attn.out_proj.weight"), g("attn.out_proj.bias"), heads)
    xn = layernorm(x, g("norm2.weight"), g("norm2.bias"))
    hdn = gelu(xn @ g("ffn.0.weight").t() + g("ffn.0.bias"))
    x = x + (hdn @ g("ffn.3.weight").t() + g("ffn.3.bias"))
    return x


def temporal_transformer(x, p: Dict[str, torch.Tensor], num_layers: int, heads: int):
    """TemporalTransformer.forward on (B,C,F,H,W).  temporal_transformer.py:113-146.
    ``p`` is the module's state_dict (keys ``pos_encoding.pe``, ``blocks.{i}.*``, ``norm.*``)."""
    b, c, f, h, w = x.shape
    y = rearrange_fwd(x, p["pos_encoding.pe"])
    for i in range(num_layers):
        y = temporal_block(y, p, f"blocks.{i}.", heads)
    y = layernorm(y, p["norm.weight"], p["norm.bias"])
    return rearrange_bwd(y, b, h, w)


# --------------------------------------------------------------------------------------
# a3 / a4 / a5 / a7: LoRA projections
# --------------------------------------------------------------------------------------


def unziplora_delta(x_content, x_style, a_c, b_c, a_s, b_s, m_c, m_s,
                    forward_type: str = "both", masked_content: bool = False,
                    masked_style: bool = False, infer: bool = True,
                    mask_c: Optional[torch.Tensor] = None,
                    mask_s: Optional[torch.Tensor] = None):
    """UnZipLoRALinearLayer(Infer).forward in the reference's dense 3-GEMM form.

    a_* = ``{key}_down.weight`` (r,in); b_* = ``{key}_up.weight`` (out,r); m_* mergers (out,).
    infer=True : unziplora_linear_layer.py:298-346 (block-masked branch -> zeros :308,316;
                 content/style modes skip the merger :331,343).
    infer=False: :221-263 (hard bool column mask multiplies :236-237,243-244; content/style
                 modes skip the merger :249-261)."""
    if x_style is None:
        x_style = x_content

    def branch(x, a, b, m, masked, mask, use_merger):
        if infer and masked:
            return torch.zeros(x.shape[:-1] + (b.shape[0],), dtype=x.dtype)
        wmat = a.t() @ b.t()                      # (in,out) materialised every call
        if use_merger:
            wmat = wmat * m
        if (not infer) and masked:
            wmat = wmat * mask
        return x @ wmat

    if forward_type == "both":
        return (branch(x_style, a_s, b_s, m_s, masked_style, mask_s, True)
                + branch(x_content, a_c, b_c, m_c, masked_content, mask_c, True))
    if forward_type == "content":
        return branch(x_content, a_c, b_c, m_c, masked_content, mask_c, False)
    if forward_type == "style":
        return branch(x_style, a_s, b_s, m_s, masked_style, mask_s, False)
    raise AssertionError(forward_type)


def lora_compatible_linear(x, weight, bias, delta=None, scale: float = 1.0):
    """Linear(x) + scale * lora_layer(...).  unziplora_unet/lora_linear.py:74-81 and
    lora_unzip.py:66-75 (there ``delta`` is computed from hidden_states_1/_2)."""
    out = x @ weight.t()
    if bias is not None:
        out = out + bias
    if delta is not None:
        out = out + scale * delta
    return out


def temporal_lora_linear(x, weight, bias, lora_a, lora_b, scale: float):
    """base(x) + scale * (x A^T) B^T.  animatediff/temporal_lora.py:29-32."""
    out = x @ weight.t()
    if bias is not None:
        out = out + bias
    return out + ((x @ lora_a.t()) @ lora_b.t()) * scale


# --------------------------------------------------------------------------------------
# §8f-2: temporal / spatial orthogonality loss
# --------------------------------------------------------------------------------------


def orth_loss_dense(pairs, lambda_orth: float):
    """lambda / N * sum_l ||D_t^T D_c||_F^2 + ||D_t^T D_s||_F^2 in the reference's dense form.
    animatediff/temporal_lora.py:126-166.  ``pairs``: [(A_t (r,in), B_t (out,r), scale_t,
    A_c, B_c, A_s, B_s)] per paired layer; D_t = scale_t * B_t A_t carries the gradient (:147),
    the spatial deltas are constants (:148-155)."""
    if lambda_orth == 0.0 or not pairs:
        return torch.tensor(0.0)
    total = None
    for a_t, b_t, scale_t, a_c, b_c, a_s, b_s in pairs:
        w = (b_t @ a_t) * scale_t                                   # (out, in)   :147
        d_c = (b_c @ a_c).detach()                                  # :153
        d_s = (b_s @ a_s).detach()                                  # :154
        contrib = torch.sum((w.T @ d_c) ** 2) + torch.sum((w.T @ d_s) ** 2)   # :156-158
        total = contrib if total is None else total + contrib
    return lambda_orth * total / len(pairs)                         # :164
