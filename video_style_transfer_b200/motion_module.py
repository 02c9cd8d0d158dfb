"""The AnimateDiff-XL motion module (diffusers ``TransformerTemporalModel`` as instantiated by
``MotionAdapter`` / ``UNetMotionModel``; un-vendored in the reference, whose in-tree restatement is
animatediff/temporal_transformer.py) on the sm_100a kernels.

Sub-module names follow diffusers so a motion-adapter state dict loads and so
``inject_temporal_lora`` / ``freeze_spatial_layers`` (name-substring based, animatediff/utils.py
:66-95, temporal_lora.py:44-69) find what they look for:
  norm, proj_in, transformer_blocks.0.{norm1, attn1.{to_q,to_k,to_v,to_out.0}, norm2, attn2.{...},
  norm3, ff.net.0.proj, ff.net.2, pos_embed}, proj_out

forward((B*F, C, H, W), num_frames):
  GroupNorm(32, C, 1e-6) over the (B,C,F,H,W) view + rearrange   stats, finalize, apply+rearrange
  proj_in                                                          tcgen05 GEMM
  [LN + pe -> q,k,v (temporal-LoRA fused GEMMs) -> frame attention -> to_out + residual] x 2
  LN -> GEGLU feed-forward + residual                              GEMM, geglu, GEMM(+res)
  proj_out, tokens -> (B*F,C,H,W), + input residual                GEMM, rearrange_bwd(+res)
"""
from typing import Optional

import torch
import torch.nn as nn

from . import ops
from ._host import fused_linear
from .fused import project_fused, project_one
from .animatediff.temporal_lora import TemporalLoRALinear
from .animatediff.temporal_transformer import PositionalEncoding


def _project(layer, x2, residual=None):
    return project_one(layer, x2, residual=residual)


class TemporalAttention(nn.Module):
    """Motion-module ``Attention`` (self-attention over frames, 8 heads, no qkv bias)."""

    def __init__(self, dim: int, heads: int = 8, bias: bool = False):
        super().__init__()
        self.heads = heads
        self.to_q = nn.Linear(dim, dim, bias=bias)
        self.to_k = nn.Linear(dim, dim, bias=bias)
        self.to_v = nn.Linear(dim, dim, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim, bias=True), nn.Dropout(0.0)])

    def forward_tokens(self, xn: torch.Tensor, n: int, f: int, residual: torch.Tensor):
        """xn, residual: (N*F, C) bf16 -> residual + to_out(attn(q,k,v))."""
        c = xn.shape[1]
        # one GEMM, N = 3C; q / k / v are (N, F, C) column views with the packed row stride
        q, k, v = project_fused((self.to_q, self.to_k, self.to_v), xn, batch=(n, f))
        o = ops.temporal_attn(q, k, v, self.heads)
        return _project(self.to_out[0], o.view(n * f, c), residual=residual)


class GEGLU(nn.Module):
    def __init__(self, dim: int, inner: int):
        super().__init__()
        self.proj = nn.Linear(dim, inner * 2)


class FeedForward(nn.Module):
    def __init__(self, dim: int, mult: int = 4):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * mult), nn.Dropout(0.0), nn.Linear(dim * mult, dim)])

    def forward_tokens(self, xn: torch.Tensor, residual: torch.Tensor):
        proj = self.net[0].proj
        inner = proj.out_features // 2
        if inner % 128 == 0 and not (torch.is_grad_enabled() and ops._needs_grad(xn, proj.weight, proj.bias)):
            # inference: hidden * gelu(gate) in the GEMM epilogue, the 2*inner-wide tensor never exists
            h = fused_linear(xn, proj.weight, proj.bias, act="geglu")
        else:
            h = ops.geglu(fused_linear(xn, proj.weight, proj.bias))
        return fused_linear(h, self.net[2].weight, self.net[2].bias, residual=residual)


class TemporalBasicTransformerBlock(nn.Module):
    def __init__(self, dim: int, heads: int = 8, max_seq_length: int = 32):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = TemporalAttention(dim, heads)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = TemporalAttention(dim, heads)       # double self-attention
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)
        self.pos_embed = PositionalEncoding(dim, max_seq_length)

    def forward_tokens(self, x2: torch.Tensor, n: int, f: int):
        pe = self.pos_embed.table(x2.device)
        for norm, attn in ((self.norm1, self.attn1), (self.norm2, self.attn2)):
            xn = ops.layernorm(x2, norm.weight, norm.bias, norm.eps, pe=pe, frames=f)
            x2 = attn.forward_tokens(xn, n, f, x2)
        xn = ops.layernorm(x2, self.norm3.weight, self.norm3.bias, self.norm3.eps)
        return self.ff.forward_tokens(xn, x2)


class MotionModule(nn.Module):
    """``motion_modules.{i}`` of UNetMotionModel; holds ``temporal_transformer`` like diffusers."""

    def __init__(self, in_channels: int, heads: int = 8, norm_num_groups: int = 32,
                 max_seq_length: int = 32, num_layers: int = 1):
        super().__init__()
        self.temporal_transformer = TransformerTemporalModel(in_channels, heads, norm_num_groups,
                                                             max_seq_length, num_layers)

    def forward(self, hidden_states, num_frames: int = 1, **kwargs):
        return self.temporal_transformer(hidden_states, num_frames=num_frames)


class TransformerTemporalModel(nn.Module):
    def __init__(self, in_channels: int, heads: int = 8, norm_num_groups: int = 32,
                 max_seq_length: int = 32, num_layers: int = 1):
        super().__init__()
        self.in_channels = in_channels
        self.groups = norm_num_groups
        self.norm = nn.GroupNorm(norm_num_groups, in_channels, eps=1e-6, affine=True)
        self.proj_in = nn.Linear(in_channels, in_channels)
        self.transformer_blocks = nn.ModuleList(
            [TemporalBasicTransformerBlock(in_channels, heads, max_seq_length)
             for _ in range(num_layers)])
        self.proj_out = nn.Linear(in_channels, in_channels)

    def forward(self, hidden_states: torch.Tensor, encoder_hidden_states=None, timestep=None,
                class_labels=None, num_frames: int = 1, cross_attention_kwargs=None,
                return_dict: bool = False):
        if not hidden_states.is_cuda:
            raise RuntimeError("MotionModule (B200) needs CUDA tensors; there is no CPU fallback")
        bf, c, h, w = hidden_states.shape
        b = bf // num_frames
        in_dtype = hidden_states.dtype
        x = hidden_states.contiguous()
        if in_dtype != torch.bfloat16:
            x = x.to(torch.bfloat16)
        x5 = x.view(b, num_frames, c, h, w).permute(0, 2, 1, 3, 4)         # (B,C,F,H,W) view
        tok, _, _ = ops.groupnorm5d_rearrange(x5, self.norm.weight, self.norm.bias, self.groups,
                                              self.norm.eps)
        n = b * h * w
        x2 = fused_linear(tok.view(n * num_frames, c), self.proj_in.weight, self.proj_in.bias)
        for block in self.transformer_blocks:
            x2 = block.forward_tokens(x2, n, num_frames)
        x2 = fused_linear(x2, self.proj_out.weight, self.proj_out.bias)
        out5 = ops.rearrange_bwd(x2.view(n, num_frames, c), b, h, w, residual5=x5, frames_major=True)
        out = out5.permute(0, 2, 1, 3, 4).reshape(bf, c, h, w)              # storage is (B,F,C,H,W)
        return out if in_dtype == torch.bfloat16 else out.to(in_dtype)
