"""Drop-in for the diffusers-free part of animatediff/utils.py (reference :66-163): which parameters
Stage 2 trains, and the checkpoint formats the sampler reads back.

  motion_modules.pth          motion-module state dict under the ORIGINAL key names, temporal LoRA
                              folded into the base weights (reference :115-135, temporal_lora.py:169-192)
  merger_{content,style}_stage2.pth
                              ``unet.{name}.lora.merge_{content,style}`` (reference :137-163), the same
                              format as the Stage-1 merger files ``insert_unziplora_to_unet`` loads
"""
import os
from typing import Dict, List, Tuple

import torch
from torch import nn

from .temporal_lora import TemporalLoRALinear, get_merged_motion_state_dict


def freeze_spatial_layers(unet: nn.Module, unfreeze_mergers: bool = False) -> Dict[str, int]:
    """reference :66-94.  Inside ``motion_modules``: ``.base.weight`` / ``.base.bias`` frozen, everything
    else (lora_A / lora_B, norms, feed-forward, positions) trainable; outside: frozen, except the UnZipLoRA
    mergers when ``unfreeze_mergers``.  Returns the element counts the reference logs."""
    n = dict(trainable=0, frozen=0, mergers=0, temporal_lora=0)
    for name, p in unet.named_parameters():
        if "motion_modules" in name:
            train = not (".base.weight" in name or ".base.bias" in name)
            if train and (".lora_A" in name or ".lora_B" in name):
                n["temporal_lora"] += p.numel()
        else:
            train = unfreeze_mergers and ("merge_content" in name or "merge_style" in name)
            if train:
                n["mergers"] += p.numel()
        p.requires_grad_(train)
        n["trainable" if train else "frozen"] += p.numel()
    return n


def get_trainable_parameters(unet: nn.Module) -> List[nn.Parameter]:
    return [p for p in unet.parameters() if p.requires_grad]


def _extract_merger_state_dicts(unet: nn.Module) -> Tuple[Dict[str, torch.Tensor], Dict[str, torch.Tensor]]:
    """reference :147-163."""
    content, style = {}, {}
    for name, module in unet.named_modules():
        lora = None if "motion_modules" in name else getattr(module, "lora_layer", None)
        mc, ms = getattr(lora, "merge_content", None), getattr(lora, "merge_style", None)
        if mc is None or ms is None:
            continue
        content[f"unet.{name}.lora.merge_content"] = mc.detach().cpu()
        style[f"unet.{name}.lora.merge_style"] = ms.detach().cpu()
    return content, style


def motion_state_dict(unet: nn.Module) -> Dict[str, torch.Tensor]:
    """What goes into motion_modules.pth (reference :115-131)."""
    if any(isinstance(m, TemporalLoRALinear) for m in unet.modules()):
        return get_merged_motion_state_dict(unet)
    return {k: v.cpu() for k, v in unet.state_dict().items() if "motion_modules" in k}


def save_checkpoint(unet: nn.Module, output_dir: str, step, save_full_model: bool = False,
                    save_mergers: bool = False) -> str:
    """reference :100-145 (``save_full_model`` is diffusers' ``save_pretrained``: out of scope)."""
    if save_full_model:
        raise NotImplementedError("save_full_model is UNet.save_pretrained of diffusers (out of scope)")
    path = os.path.join(output_dir, f"checkpoint-{step}")
    os.makedirs(path, exist_ok=True)
    torch.save(motion_state_dict(unet), os.path.join(path, "motion_modules.pth"))
    if save_mergers:
        content, style = _extract_merger_state_dicts(unet)
        if content:
            torch.save(content, os.path.join(path, "merger_content_stage2.pth"))
            torch.save(style, os.path.join(path, "merger_style_stage2.pth"))
    return path
