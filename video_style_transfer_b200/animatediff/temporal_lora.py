"""Drop-in for animatediff/temporal_lora.py (reference :10-192).

``TemporalLoRALinear`` keeps the load-bearing names (``base``, ``lora_A``, ``lora_B``, ``scale``;
SURVEY.md §8b) and runs y = W x + b + (alpha/r) B (A x) as one fused tcgen05 GEMM (the reference
issues three F.linear calls, :29-32).  Injection / pairing / orthogonality-loss / merged-checkpoint
helpers are host logic re-implemented against the same module-name contract.
"""
from typing import Dict, Optional

import torch
import torch.nn as nn

from .. import ops
from .._host import as_rows_bf16, bf16_cached, cached_build, fused_linear, packed_up, pad_branch_cols


class TemporalLoRALinear(nn.Module):
    def __init__(self, base: nn.Linear, rank: int = 32, alpha: float = 1.0):
        super().__init__()
        self.base = base
        self.base.weight.requires_grad_(False)
        if self.base.bias is not None:
            self.base.bias.requires_grad_(False)
        self.in_features = base.in_features
        self.out_features = base.out_features
        self.rank = rank
        self.scale = alpha / rank
        w = base.weight
        self.lora_A = nn.Parameter(torch.randn(rank, base.in_features, device=w.device,
                                               dtype=w.dtype) * 0.01)
        self.lora_B = nn.Parameter(torch.zeros(base.out_features, rank, device=w.device,
                                               dtype=w.dtype))

    # "LoRA site" protocol shared with LoRACompatibleLinear (see fused.py)
    def site_weight(self):
        return self.base.weight

    def site_bias(self):
        return self.base.bias

    def site_downs(self):
        return [self.lora_A]

    def site_up(self, scale: float = 1.0):
        """(alpha/r) * B as bf16 (out, pad8(r)), cached on lora_B's version."""
        return packed_up(None, None, None, self.lora_B, None, None, float(self.scale), self.rank)

    def fused_forward(self, x2: torch.Tensor, scale: float = 1.0, residual=None,
                      out_scale: float = 1.0, act: str = "none",
                      out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
        # fused.project_one: frozen adapter -> cached W + (alpha/r) B A and one plain GEMM (the
        # reference's own merged_weight, :35-41); trainable adapter -> u = x A^T + fused GEMM
        from ..fused import project_one
        return project_one(self, x2, 1.0, residual, out_scale, act, out_dtype)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = self.fused_forward(as_rows_bf16(x))
        return y.reshape(x.shape[:-1] + (self.out_features,)).to(x.dtype)

    def get_delta(self) -> torch.Tensor:
        return (self.lora_B @ self.lora_A) * self.scale

    def merged_weight(self) -> torch.Tensor:
        with torch.no_grad():
            return (self.base.weight + self.get_delta()).detach()


def _is_attention(module) -> bool:
    return all(hasattr(module, n) for n in ("to_q", "to_k", "to_v", "to_out"))


def inject_temporal_lora(unet, rank: int = 32, alpha: float = 1.0) -> int:
    """Wrap to_q/k/v/to_out[0] of every attention under a ``motion_modules`` path (reference
    :44-69).  Idempotent.  Returns the number of linears wrapped."""
    wrapped = 0
    for name, module in unet.named_modules():
        if "motion_modules" not in name or not _is_attention(module):
            continue
        if isinstance(module.to_q, TemporalLoRALinear):
            continue
        for proj in ("to_q", "to_k", "to_v"):
            lin = getattr(module, proj)
            if isinstance(lin, nn.Linear):
                setattr(module, proj, TemporalLoRALinear(lin, rank, alpha))
                wrapped += 1
        if len(module.to_out) > 0 and isinstance(module.to_out[0], nn.Linear):
            module.to_out[0] = TemporalLoRALinear(module.to_out[0], rank, alpha)
            wrapped += 1
    return wrapped


def build_spatial_lora_index(unet) -> Dict[str, object]:
    """{temporal linear name -> spatial UnZipLoRA layer} (reference :72-123): drop
    ``motion_modules.{i}[.temporal_transformer]`` from the temporal path, insert
    ``attentions.{i}``, and keep the pair when the shapes agree."""
    spatial = {n: m.lora_layer for n, m in unet.named_modules()
               if "motion_modules" not in n and getattr(m, "lora_layer", None) is not None}
    index: Dict[str, object] = {}
    for name, module in unet.named_modules():
        if "motion_modules" not in name or not isinstance(module, TemporalLoRALinear):
            continue
        parts = name.split(".")
        at = parts.index("motion_modules")
        tail = parts[at + 2:]
        if "temporal_transformer" in tail:
            tail = tail[tail.index("temporal_transformer") + 1:]
        path = ".".join(parts[:at]) + ".attentions." + parts[at + 1] + "." + ".".join(tail)
        lora = spatial.get(path)
        if lora is None:
            continue
        try:
            up, down = lora.lora_matrix_dic.content_up.weight, lora.lora_matrix_dic.content_down.weight
        except AttributeError:
            continue
        if up.shape[0] == module.out_features and down.shape[1] == module.in_features:
            index[name] = lora
    return index


def compute_orth_loss(unet, spatial_index: Dict, lambda_orth: float) -> torch.Tensor:
    """lambda/N * sum_l ||D_t^T D_c||_F^2 + ||D_t^T D_s||_F^2 (reference :126-166).
    D_t^T D_x = A_t^T (B_t^T B_x) A_x, so each term goes through the r x r core
    K = B_t^T B_x and ||A_t^T K A_x||_F^2 = tr((A_t A_t^T) K (A_x A_x^T) K^T): O(r^2 (in+out))
    instead of the reference's two dense in x in products."""
    if lambda_orth == 0.0 or not spatial_index:
        return torch.tensor(0.0)
    # Layers of equal shape are stacked and go through batched matmuls: per layer the terms are five r x r products
    # (plus their backward), ~7 us of latency each on the GPU -- 1050 fp32 SGEMM launches, 7.5 ms of a 259 ms
    # config-4 step (profiles/cfg4_kernels_r02.tsv) -- and the SDXL stack has only a handful of distinct shapes.
    groups: Dict[tuple, list] = {}
    for name, module in unet.named_modules():
        if name not in spatial_index or not isinstance(module, TemporalLoRALinear):
            continue
        d = spatial_index[name].lora_matrix_dic
        key = (tuple(module.lora_A.shape), tuple(module.lora_B.shape)) + tuple(
            tuple(d[f"{k}_{w}"].weight.shape) for k in ("content", "style") for w in ("down", "up"))
        groups.setdefault(key, []).append((module, d))
    count = sum(len(v) for v in groups.values())
    if count == 0:
        return torch.tensor(0.0)
    total: Optional[torch.Tensor] = None
    for items in groups.values():
        a_t = torch.stack([m.lora_A.float() for m, _ in items])                  # (n, r_t, in)
        b_t = torch.stack([m.lora_B.float() * m.scale for m, _ in items])        # (n, out, r_t)
        gram_t = a_t @ a_t.transpose(1, 2)
        for key in ("content", "style"):
            with torch.no_grad():
                b_x = torch.stack([d[f"{key}_up"].weight.float() for _, d in items])
                a_x = torch.stack([d[f"{key}_down"].weight.float() for _, d in items])
                gram_x = a_x @ a_x.transpose(1, 2)
            core = b_t.transpose(1, 2) @ b_x                                     # (n, r_t, r_x)
            val = torch.sum((gram_t @ core) * (core @ gram_x))
            total = val if total is None else total + val
    return lambda_orth * total / count


def get_merged_motion_state_dict(unet) -> Dict[str, torch.Tensor]:
    """motion_modules state dict with the temporal LoRA folded into the base weights under the
    original key names (reference :169-192)."""
    merged: Dict[str, torch.Tensor] = {}
    wrapped = []
    for name, module in unet.named_modules():
        if isinstance(module, TemporalLoRALinear) and "motion_modules" in name:
            merged[name + ".weight"] = module.merged_weight().cpu()
            if module.base.bias is not None:
                merged[name + ".bias"] = module.base.bias.detach().cpu()
            wrapped.append(name + ".")
    for key, value in unet.state_dict().items():
        if "motion_modules" in key and not any(key.startswith(w) for w in wrapped):
            merged[key] = value.detach().cpu()
    return merged
