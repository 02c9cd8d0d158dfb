"""Drop-in for unziplora_unet/unzip_attention.py (reference :13-239): the spatial
``BasicTransformerBlock`` of the SDXL UNet, whose LayerNorms and GEGLU feed-forward are the other
half of SURVEY.md §8f-1 (~45 % of the UNet's FLOPs sit in ``ff``, reference :224-235).

The reference subclasses diffusers' block (un-vendored) only to wire its own ``Attention`` in
(:66-111); the ``forward`` it runs for SDXL (``norm_type="layer_norm"``) is :113-239:

    h = attn1(norm1(h)) + h                       :128-166
    h = attn2(norm2(h), text...) + h              :176-203
    h = ff(norm3(h)) + h                          :206-235      ff = Linear(C, 8C) GEGLU -> Linear(4C, C)

Here every step is on the sm_100a kernels: LayerNorm = ``vst_layernorm`` (flat, one pass), the
attentions go through the plugged-in processor with the residual carried by the to_out epilogue
(``fused_residual`` kwarg, swallowed by processors that do not know it), GEGLU lives in the first
FF GEMM's epilogue and the closing residual in the second's.  Sub-module names follow diffusers
(``norm1/attn1/norm2/attn2/norm3/ff.net.0.proj/ff.net.2``) so UNet state dicts load.
AdaLayerNorm variants, GLIGEN and chunked feed-forward are not on the SDXL path and raise.
"""
from typing import Any, Dict, Optional

import torch
from torch import nn

from .. import ops
from ..motion_module import FeedForward


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim: int, num_attention_heads: int, attention_head_dim: int, dropout=0.0,
                 cross_attention_dim: Optional[int] = None, activation_fn: str = "geglu",
                 norm_type: str = "layer_norm", norm_eps: float = 1e-5, norm_elementwise_affine: bool = True,
                 attn1: Optional[nn.Module] = None, attn2: Optional[nn.Module] = None, **unused):
        super().__init__()
        if norm_type != "layer_norm" or activation_fn != "geglu" or dropout != 0.0:
            raise NotImplementedError("BasicTransformerBlock (B200): the SDXL configuration only "
                                      "(layer_norm, geglu, no dropout)")
        if attn1 is None or attn2 is None:
            raise ValueError("pass the block's attention modules (diffusers-protocol `Attention` with a "
                             "video_style_transfer_b200 processor) as attn1 / attn2")
        self.norm1 = nn.LayerNorm(dim, norm_eps, norm_elementwise_affine)
        self.attn1 = attn1
        self.norm2 = nn.LayerNorm(dim, norm_eps, norm_elementwise_affine)
        self.attn2 = attn2
        self.norm3 = nn.LayerNorm(dim, norm_eps, norm_elementwise_affine)
        self.ff = FeedForward(dim)

    def forward(self, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                encoder_hidden_states: Optional[torch.Tensor] = None,
                encoder_attention_mask: Optional[torch.Tensor] = None,
                cross_attention_kwargs: Optional[Dict[str, Any]] = None, **kwargs) -> torch.Tensor:
        if not hidden_states.is_cuda:
            raise RuntimeError("BasicTransformerBlock (B200) needs CUDA tensors; there is no CPU fallback")
        kw = dict(cross_attention_kwargs or {})
        if kw.pop("gligen", None) is not None:
            raise NotImplementedError("GLIGEN is not on the SDXL-AnimateDiff path")
        in_dtype = hidden_states.dtype
        h = hidden_states if in_dtype == torch.bfloat16 else hidden_states.to(torch.bfloat16)
        h = h.contiguous()
        shape = h.shape
        c = shape[-1]

        def ln(norm, x):
            return ops.layernorm(x, norm.weight, norm.bias, norm.eps)

        # the processors add `fused_residual` in the to_out epilogue; one that ignores the kwarg returns
        # the bare attention output and the add happens here (reference :166, :203)
        def attend(attn, x, **a):
            proc = getattr(attn, "processor", None)
            if proc is None or getattr(proc, "fuses_residual", False):
                return attn(ln(attn_norm[attn], x), fused_residual=x, **a, **kw)
            return attn(ln(attn_norm[attn], x), **a, **kw) + x

        attn_norm = {self.attn1: self.norm1, self.attn2: self.norm2}
        h = attend(self.attn1, h, attention_mask=attention_mask)                       # :128-166
        h = attend(self.attn2, h, encoder_hidden_states=encoder_hidden_states,        # :176-203
                   attention_mask=encoder_attention_mask)
        h2 = h.reshape(-1, c)
        h2 = self.ff.forward_tokens(ln(self.norm3, h2), h2)                            # :206-235
        h = h2.view(shape)
        return h if in_dtype == torch.bfloat16 else h.to(in_dtype)
