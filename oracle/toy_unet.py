"""TEST INFRASTRUCTURE (never imported by the product path).

A diffusers-free stand-in for the UNet the reference's format helpers walk
(unziplora_unet/utils.py:387-483, animatediff/utils.py:66-163): plain ``nn.Linear`` attentions
reachable by the dotted names diffusers would give them, an ``attn_processors`` mapping keyed
``<attention module>.processor`` (train_animatediff.py:121-125), and one motion module whose names
follow temporal_lora.py:76-84.  Used by oracle/make_golden_formats.py (reference side, in the build
container) and by tests/test_cpu.py (this repository's side) so both walk the same tree.
"""
import torch
from torch import nn


class ToyAttn(nn.Module):
    def __init__(self, c: int, kv: int):
        super().__init__()
        self.to_q = nn.Linear(c, c, bias=False)
        self.to_k = nn.Linear(kv, c, bias=False)
        self.to_v = nn.Linear(kv, c, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(c, c), nn.Dropout(0.0)])


class _Block(nn.Module):
    def __init__(self, c: int, ctx: int):
        super().__init__()
        self.attn1 = ToyAttn(c, c)
        self.attn2 = ToyAttn(c, ctx)
        self.norm1 = nn.LayerNorm(c)


class _Holder(nn.Module):
    pass


class ToyUNet(nn.Module):
    """down_blocks.0.attentions.0.transformer_blocks.0.{attn1,attn2}  (spatial)
    down_blocks.0.motion_modules.0.temporal_transformer.transformer_blocks.0.{attn1,attn2}  (motion)"""

    def __init__(self, c: int = 16, ctx: int = 24):
        super().__init__()
        blk = _Holder()
        sp = _Holder()
        sp.transformer_blocks = nn.ModuleList([_Block(c, ctx)])
        mm = _Holder()
        mm.temporal_transformer = _Holder()
        mm.temporal_transformer.transformer_blocks = nn.ModuleList([_Block(c, c)])
        mm.temporal_transformer.norm = nn.GroupNorm(4, c)
        blk.attentions = nn.ModuleList([sp])
        blk.motion_modules = nn.ModuleList([mm])
        self.down_blocks = nn.ModuleList([blk])
        self.conv_in = nn.Conv2d(4, c, 3, padding=1)

    @property
    def attn_processors(self):
        return {n + ".processor": None for n, m in self.named_modules() if isinstance(m, ToyAttn)}


def seeded_toy(seed: int = 11, c: int = 16, ctx: int = 24) -> ToyUNet:
    torch.manual_seed(seed)
    return ToyUNet(c, ctx)
