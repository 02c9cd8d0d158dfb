"""Drop-in for unziplora_unet/lora_linear.py (reference :10-81): ``LoRALinearLayer`` and the
``LoRACompatibleLinear`` the live AnimateDiff path wires into every spatial attention
(unziplora_unet/utils.py:712-725).  ``forward`` is ONE fused tcgen05 GEMM:
    y = x W^T + b + scale * [(x A_c^T) (m_c . B_c)^T + (x A_s^T) (m_s . B_s)^T]
instead of the reference's Linear + 3 dense GEMMs (:80-81 -> unziplora_linear_layer.py:311-323).
``LoRACompatibleConv`` / ``_fuse_lora`` are out of scope (never hit on the video path).
"""
from typing import Optional, Union

import torch
from torch import nn

from .. import ops
from .._host import as_rows_bf16, bf16_cached, cached_build, fused_linear, packed_up, pad_branch_cols


class LoRALinearLayer(nn.Module):
    """Plain rank-r adapter (reference :10-38): up(down(x)) * network_alpha / rank."""

    def __init__(self, in_features: int, out_features: int, rank: int = 4,
                 network_alpha: Optional[float] = None,
                 device: Optional[Union[torch.device, str]] = None,
                 dtype: Optional[torch.dtype] = None):
        super().__init__()
        self.down = nn.Linear(in_features, rank, bias=False, device=device, dtype=dtype)
        self.up = nn.Linear(rank, out_features, bias=False, device=device, dtype=dtype)
        self.network_alpha = network_alpha
        self.rank = rank
        self.out_features = out_features
        self.in_features = in_features
        nn.init.normal_(self.down.weight, std=1 / rank)
        nn.init.zeros_(self.up.weight)

    def site_downs(self):
        return [self.down.weight]

    def site_up(self, scale: float = 1.0):
        s = scale if self.network_alpha is None else scale * self.network_alpha / self.rank

        return packed_up(None, None, None, self.up.weight, None, None, float(s), self.rank)

    def fused_operands(self, x_content, x_style=None, scale: float = 1.0):
        u = ops.lora_gemm(x_content, bf16_cached(self.down.weight, pad_rows=True))
        return u, self.site_up(scale)

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        u, bp = self.fused_operands(as_rows_bf16(hidden_states), None, 1.0)
        out = ops.lora_gemm(u, bp)
        return out.reshape(hidden_states.shape[:-1] + (self.out_features,)).to(hidden_states.dtype)


class LoRACompatibleLinear(nn.Linear):
    """nn.Linear with an optional ``lora_layer`` (reference :41-81)."""

    def __init__(self, *args, lora_layer: Optional[nn.Module] = None, **kwargs):
        super().__init__(*args, **kwargs)
        self.lora_layer = lora_layer

    def set_lora_layer(self, lora_layer: Optional[nn.Module]):
        self.lora_layer = lora_layer

    # "LoRA site" protocol (see fused.py): what a projection fused along N needs from each linear
    def site_weight(self):
        return self.weight

    def site_bias(self):
        return self.bias

    def site_downs(self):
        return [] if self.lora_layer is None else self.lora_layer.site_downs()

    def site_up(self, scale: float = 1.0):
        return None if self.lora_layer is None else self.lora_layer.site_up(scale)

    def fused_forward(self, x2: torch.Tensor, scale: float = 1.0,
                      x_content: Optional[torch.Tensor] = None,
                      x_style: Optional[torch.Tensor] = None,
                      residual: Optional[torch.Tensor] = None, out_scale: float = 1.0):
        """2-D bf16 in, 2-D bf16 out; optional residual / output scale fused into the epilogue
        (attention_processor.py:91-94)."""
        low_rank = None
        if self.lora_layer is not None:
            if not hasattr(self.lora_layer, "fused_operands"):
                raise TypeError("lora_layer must be a video_style_transfer_b200 LoRA layer "
                                "(no eager fallback on the sm_100a path)")
            if x_content is None and x_style is None:
                # single stream: frozen adapters -> W_eff = W + B'A and one plain GEMM; otherwise the low-rank
                # form, in ONE launch where the shape allows (fused.project_one)
                from ..fused import project_one
                return project_one(self, x2, scale, residual, out_scale)
            xc = x2 if x_content is None else x_content
            low_rank = self.lora_layer.fused_operands(xc, x_style, scale)
        return fused_linear(x2, self.weight, self.bias, low_rank, residual, out_scale)

    def forward(self, hidden_states: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
        y = self.fused_forward(as_rows_bf16(hidden_states), scale)
        return y.reshape(hidden_states.shape[:-1] + (self.out_features,)).to(hidden_states.dtype)
