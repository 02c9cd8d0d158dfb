"""Drop-in for the diffusers-free part of unziplora_unet/utils.py: the on-disk formats and the
wiring that puts UnZipLoRA layers on a UNet's spatial attentions.

The reference functions take a diffusers ``UNet2DConditionModel`` / ``UNetMotionModel``; nothing here
imports diffusers.  A "unet" is any ``nn.Module`` tree whose attention modules are reachable by the
dotted names in ``attn_processors`` (``unet.attn_processors`` when it exists, else every sub-module
that owns ``to_q`` / ``to_k`` / ``to_v`` / ``to_out``).

On-disk formats kept byte-compatible with the reference (file:line into /root/reference):
  LoRA safetensors      ``unet.unet.{attn}.{part}.lora.{down,up}.weight``      utils.py:347-366
                        file name ``pytorch_lora_weights.safetensors``         utils.py:27,131-160
  exported LoRA         ``unet.{linear}.lora.{up,down}.weight``                utils.py:288-310
  merger .pth           ``unet.{attn}.{part}.lora.merge_{content,style}``      utils.py:368-386,
                        (Stage 1 and ``merger_*_stage2.pth``)                   animatediff/utils.py:147-163
  layer state dict      ``lora_matrix_dic.{content,style}_{down,up}.weight``,
                        ``merge_content``, ``merge_style``                     utils.py:324-333
"""
import os
from typing import Dict, Iterable, Optional, Tuple, Union

import torch
from torch import nn

from .lora_linear import LoRACompatibleLinear
from .unziplora_linear_layer import UnZipLoRALinearLayerInfer

LORA_WEIGHT_NAME_SAFE = "pytorch_lora_weights.safetensors"      # reference utils.py:27
PARTS = ("to_q", "to_k", "to_v", "to_out.0")                     # reference utils.py:361, :381


def get_lora_weights(lora_name_or_path: str, subfolder: Optional[str] = None, **kwargs) -> Dict[str, torch.Tensor]:
    """Read a LoRA safetensors file (a file, or a folder holding ``pytorch_lora_weights.safetensors``)
    into a CPU dict.  reference utils.py:131-160.  Hub ids are not resolved: there is no network path."""
    path = lora_name_or_path
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path}: only local LoRA files are supported (no hub download)")
    if subfolder is not None:
        path = os.path.join(path, subfolder)
    if os.path.isdir(path):
        path = os.path.join(path, LORA_WEIGHT_NAME_SAFE)
    if not path.endswith(".safetensors"):
        raise ValueError("Currently only safetensors is supported")
    from safetensors import safe_open
    with safe_open(path, framework="pt", device="cpu") as f:
        return {k: f.get_tensor(k) for k in f.keys()}


def _per_part(tensors: Dict[str, torch.Tensor], stem: str, suffix: str) -> Dict[str, torch.Tensor]:
    return {part: tensors[f"{stem}.{part}.lora.{suffix}"] for part in PARTS}


def use_lora_weights_for_inference(tensors: Dict[str, torch.Tensor], key: str, prefix: str = "unet.unet."):
    """({part: down}, {part: up}) of attention ``key``.  reference utils.py:347-366."""
    return _per_part(tensors, prefix + key, "down.weight"), _per_part(tensors, prefix + key, "up.weight")


def use_lora_mergers_for_inference(tensors_content: Dict[str, torch.Tensor], tensors_style: Dict[str, torch.Tensor],
                                   key: str, prefix: str = "unet.unet."):
    """({part: merge_content}, {part: merge_style}) of attention ``key``.  reference utils.py:368-386."""
    return (_per_part(tensors_content, prefix + key, "merge_content"),
            _per_part(tensors_style, prefix + key, "merge_style"))


def initialize_unziplora_layer_for_inference(state_dict_content_down, state_dict_content_up, state_dict_style_down,
                                             state_dict_style_up, state_dict_merger_content, state_dict_merger_style,
                                             part, **model_kwargs) -> UnZipLoRALinearLayerInfer:
    """One inference layer loaded from the per-part dicts; mergers stay at ones when either merger dict is
    None.  reference utils.py:312-345."""
    layer = UnZipLoRALinearLayerInfer(**model_kwargs)
    sd = {"lora_matrix_dic.content_down.weight": state_dict_content_down[part],
          "lora_matrix_dic.content_up.weight": state_dict_content_up[part],
          "lora_matrix_dic.style_down.weight": state_dict_style_down[part],
          "lora_matrix_dic.style_up.weight": state_dict_style_up[part]}
    if state_dict_merger_content is not None and state_dict_merger_style is not None:
        sd["merge_content"] = state_dict_merger_content[part]
        sd["merge_style"] = state_dict_merger_style[part]
    layer.load_state_dict(sd, strict=False)
    return layer


def _make_lora_compatible(module: nn.Module, attr: str) -> None:
    """nn.Linear -> LoRACompatibleLinear in place, SHARING the weight / bias Parameters.
    reference utils.py:712-725."""
    layer = getattr(module, attr)
    if hasattr(layer, "set_lora_layer"):
        return
    new = LoRACompatibleLinear(layer.in_features, layer.out_features, bias=layer.bias is not None,
                               device=layer.weight.device, dtype=layer.weight.dtype)
    new.weight = layer.weight
    if layer.bias is not None:
        new.bias = layer.bias
    setattr(module, attr, new)


def spatial_attention_names(unet: nn.Module) -> Iterable[str]:
    """Dotted names of the attention modules UnZipLoRA is inserted into: every ``attn_processors`` key
    minus its ``.processor`` leaf, skipping the motion modules (reference utils.py:406-409)."""
    if hasattr(unet, "attn_processors"):
        names = [".".join(k.split(".")[:-1]) for k in unet.attn_processors.keys()]
    else:
        names = [n for n, m in unet.named_modules()
                 if all(hasattr(m, a) for a in ("to_q", "to_k", "to_v", "to_out"))]
    return [n for n in names if "motion_modules" not in n]


def _load(path_or_dict: Union[None, str, Dict[str, torch.Tensor]], loader):
    if path_or_dict is None or isinstance(path_or_dict, dict):
        return path_or_dict
    return loader(path_or_dict)


def insert_unziplora_to_unet(unet: nn.Module, content_lora_path, style_lora_path, weight_content_path=None,
                             weight_style_path=None, rank: int = 64,
                             device: Optional[Union[torch.device, str]] = None, **kwargs) -> nn.Module:
    """Put a loaded ``UnZipLoRALinearLayerInfer`` on to_q / to_k / to_v / to_out[0] of every spatial
    attention.  reference utils.py:387-483.  The four path arguments may also be already-loaded dicts.
    LoRA keys use the ``unet.unet.`` prefix, merger keys ``unet.`` (reference :419)."""
    content = _load(content_lora_path, lambda p: get_lora_weights(p, **kwargs))
    style = _load(style_lora_path, lambda p: get_lora_weights(p, **kwargs))
    w_content = _load(weight_content_path, lambda p: torch.load(p, map_location="cpu"))
    w_style = _load(weight_style_path, lambda p: torch.load(p, map_location="cpu"))
    for attn_name in spatial_attention_names(unet):
        attn = unet.get_submodule(attn_name)
        c_down, c_up = use_lora_weights_for_inference(content, key=attn_name)
        s_down, s_up = use_lora_weights_for_inference(style, key=attn_name)
        m_c = m_s = None
        if w_content is not None and w_style is not None:
            m_c, m_s = use_lora_mergers_for_inference(w_content, w_style, key=attn_name, prefix="unet.")
        for part in PARTS:
            parent, leaf = (attn, part) if "." not in part else (attn.get_submodule(part.rsplit(".", 1)[0]),
                                                                 part.rsplit(".", 1)[1])
            _make_lora_compatible(parent, leaf)
            lin = getattr(parent, leaf)
            lin.set_lora_layer(initialize_unziplora_layer_for_inference(
                c_down, c_up, s_down, s_up, m_c, m_s, part, in_features=lin.in_features,
                out_features=lin.out_features, lora_matrix_key=["content", "style"], rank=rank, device=device))
    return unet


def _lora_sites(unet: nn.Module):
    for name, module in unet.named_modules():
        if hasattr(module, "set_lora_layer") and getattr(module, "lora_layer", None) is not None:
            yield name, module.lora_layer


def unziplora_set_forward_type(unet: nn.Module, type: str = "both") -> nn.Module:
    """reference utils.py:162-174."""
    assert type in ["both", "content", "style"]
    for _, layer in _lora_sites(unet):
        assert hasattr(layer, "set_forward"), layer
        layer.set_forward(type)
    return unet


def unet_inverse_ziplora_state_dict(unet: nn.Module, key: str, quick_release: bool = False
                                    ) -> Tuple[Dict[str, torch.Tensor], Dict[str, torch.Tensor]]:
    """(LoRA state dict, merger state dict) of branch ``key`` in the exported key format
    (reference utils.py:288-310): what Stage 1 writes and ``insert_unziplora_to_unet`` later reads."""
    lora_sd, mask_sd = {}, {}
    for name, layer in _lora_sites(unet):
        assert hasattr(layer, "get_unziplora_weight"), layer
        down, up = layer.get_unziplora_weight(key)
        lora_sd[f"unet.{name}.lora.up.weight"] = up.contiguous()
        lora_sd[f"unet.{name}.lora.down.weight"] = down.contiguous()
        mask_sd[f"unet.{name}.lora.merge_{key}"] = layer.get_merger_mask(key).contiguous()
        if quick_release:
            layer.cpu()
    return lora_sd, mask_sd
