"""Drop-in for animatediff/temporal_transformer.py (reference :6-146): the standalone temporal
transformer of BASELINE config 1.  Same classes, constructor arguments and parameter names (so
the reference's state dict loads: ``pos_encoding.pe``, ``blocks.{i}.norm1/attn.in_proj_*/
attn.out_proj/norm2/ffn.0/ffn.3``, ``norm``); forward runs on sm_100a:

  (B,C,F,H,W) -> tokens + pe        vst_rearrange_fwd           (:129-133, one pass)
  LN -> packed QKV projection       vst_layernorm, vst_lora_gemm (:69-70, in_proj 3C x C)
  frame-axis attention              vst_temporal_attn_fwd        (reads q/k/v out of the packed buffer)
  out_proj + residual               vst_lora_gemm epilogue       (:70-71)
  LN -> Linear+GELU -> Linear +res  vst_lora_gemm (GELU / residual epilogues)   (:74)
  final LN, tokens -> (B,C,F,H,W)   vst_layernorm, vst_rearrange_bwd           (:140-144)
"""
import math

import torch
import torch.nn as nn

from .. import ops
from .._host import fused_linear


class PositionalEncoding(nn.Module):
    """Sinusoidal frame-position table, pe[0,p,2i]=sin(p w_i), pe[0,p,2i+1]=cos(p w_i)
    (reference :6-27)."""

    def __init__(self, d_model: int, max_len: int = 32):
        super().__init__()
        pos = torch.arange(max_len).unsqueeze(1)
        freq = torch.exp(torch.arange(0, d_model, 2) * (-math.log(10000.0) / d_model))
        pe = torch.zeros(1, max_len, d_model)
        pe[0, :, 0::2] = torch.sin(pos * freq)
        pe[0, :, 1::2] = torch.cos(pos * freq)
        self.register_buffer("pe", pe)

    def table(self, device) -> torch.Tensor:
        """(max_len, C) fp32 view of the buffer for the kernels' fused position add; cached (the cast used to
        be one ATen copy kernel per motion-module call once the module is bf16)."""
        pe = self.pe
        if pe.dtype == torch.float32 and pe.device == torch.device(device) and pe.is_contiguous():
            return pe.reshape(-1, pe.shape[-1])
        stamp = (pe.data_ptr(), pe._version, pe.dtype, str(device))
        hit = getattr(self, "_table_cache", None)
        if hit is None or hit[0] != stamp:
            hit = (stamp, pe.to(device=device, dtype=torch.float32).reshape(-1, pe.shape[-1]).contiguous())
            object.__setattr__(self, "_table_cache", hit)
        return hit[1]

    def forward(self, x):
        """x (N, F, C) + pe[:, :F].  A (N,F,C) tensor is the planar view (B=N, C, F, HW=1), so the
        add runs through the rearrange kernel's fused position add."""
        if not x.is_cuda:
            raise RuntimeError("PositionalEncoding (B200) needs CUDA tensors; no CPU fallback")
        n, f, c = x.shape
        x5 = x.contiguous().view(n, f, c, 1, 1).permute(0, 2, 1, 3, 4)
        return ops.rearrange_fwd(x5, self.table(x.device))


class TemporalTransformerBlock(nn.Module):
    def __init__(self, channels: int, num_heads: int = 8, dropout: float = 0.0):
        super().__init__()
        if dropout != 0.0:
            raise NotImplementedError("dropout > 0 is not used by the reference configs")
        self.norm1 = nn.LayerNorm(channels)
        self.attn = nn.MultiheadAttention(embed_dim=channels, num_heads=num_heads, dropout=dropout,
                                          batch_first=True)
        self.norm2 = nn.LayerNorm(channels)
        self.ffn = nn.Sequential(nn.Linear(channels, channels * 4), nn.GELU(), nn.Dropout(dropout),
                                 nn.Linear(channels * 4, channels), nn.Dropout(dropout))
        self.num_heads = num_heads

    def forward_tokens(self, x: torch.Tensor) -> torch.Tensor:
        """x: (N, F, C) contiguous residual stream (bf16, or fp32 when the caller runs the
        reference's fp32 configuration); GEMM / attention operands are bf16 either way."""
        n, f, c = x.shape
        sd, bf = x.dtype, torch.bfloat16
        x2 = x.view(n * f, c)
        xn = ops.layernorm(x2, self.norm1.weight, self.norm1.bias, self.norm1.eps, out_dtype=bf)
        from .._autograd import split_cols
        qkv = fused_linear(xn, self.attn.in_proj_weight, self.attn.in_proj_bias)        # (N*F, 3C) packed
        q, k, v = split_cols(qkv, c, 3, (n, f))
        o = ops.temporal_attn(q, k, v, self.num_heads)
        x2 = fused_linear(o.view(n * f, c), self.attn.out_proj.weight, self.attn.out_proj.bias,
                          residual=x2, out_dtype=sd)
        xn = ops.layernorm(x2, self.norm2.weight, self.norm2.bias, self.norm2.eps, out_dtype=bf)
        h = fused_linear(xn, self.ffn[0].weight, self.ffn[0].bias, act="gelu")
        x2 = fused_linear(h, self.ffn[3].weight, self.ffn[3].bias, residual=x2, out_dtype=sd)
        return x2.view(n, f, c)

    def forward(self, x):
        return self.forward_tokens(x.contiguous())


class TemporalTransformer(nn.Module):
    def __init__(self, in_channels: int, num_layers: int = 2, num_heads: int = 8,
                 dropout: float = 0.0):
        super().__init__()
        self.in_channels = in_channels
        self.num_layers = num_layers
        self.pos_encoding = PositionalEncoding(in_channels, max_len=32)
        self.blocks = nn.ModuleList([TemporalTransformerBlock(in_channels, num_heads, dropout)
                                     for _ in range(num_layers)])
        self.norm = nn.LayerNorm(in_channels)

    def forward(self, hidden_states: torch.Tensor, num_frames: int = 1) -> torch.Tensor:
        if not hidden_states.is_cuda:
            raise RuntimeError("TemporalTransformer (B200) needs CUDA tensors; no CPU fallback")
        b, c, f, h, w = hidden_states.shape
        x = hidden_states
        if x.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("TemporalTransformer (B200) takes float32 or bfloat16 latents")
        # the residual stream stays in the caller's dtype (fp32 for BASELINE config 1)
        tok = ops.rearrange_fwd(x, self.pos_encoding.table(x.device))          # (BHW, F, C) + pe
        for block in self.blocks:
            tok = block.forward_tokens(tok)
        tok = ops.layernorm(tok, self.norm.weight, self.norm.bias, self.norm.eps)
        return ops.rearrange_bwd(tok, b, h, w)                                  # (B,C,F,H,W)
