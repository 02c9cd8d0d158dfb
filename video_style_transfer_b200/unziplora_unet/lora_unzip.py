"""Drop-in for unziplora_unet/lora_unzip.py (reference :6-75): the three-stream
``LoRACompatibleLinear`` of Stage 1 -- the base projection reads ``hidden_states`` while the
content / style low-rank branches read ``hidden_states_1`` / ``hidden_states_2``.  Still one fused
GEMM: the up-projections ride as extra K-blocks, only the rank-r down-projections see the other
two streams.

The Stage-1 processor of the reference calls these linears with the keywords
``hidden_states_content=`` / ``hidden_states_style=`` (unzip_attention_processor.py:706-708,
720-725, 745-747) while this class spells them ``hidden_states_1`` / ``hidden_states_2`` (:66);
both spellings are accepted here and mean the same streams."""
from typing import Optional

import torch

from .._host import as_rows_bf16
from .lora_linear import LoRACompatibleLinear as _Base
from .lora_linear import LoRALinearLayer  # noqa: F401  (the reference re-exports it via *)


class LoRACompatibleLinear(_Base):
    def forward(self, hidden_states: torch.Tensor, scale: float = 1.0,
                hidden_states_1: torch.Tensor = None,
                hidden_states_2: torch.Tensor = None, *,
                hidden_states_content: torch.Tensor = None,
                hidden_states_style: torch.Tensor = None) -> torch.Tensor:
        if hidden_states_1 is None:
            hidden_states_1 = hidden_states_content
        if hidden_states_2 is None:
            hidden_states_2 = hidden_states_style
        x2 = as_rows_bf16(hidden_states)
        if self.lora_layer is None:
            y = self.fused_forward(x2, scale)
        else:
            x1 = x2 if hidden_states_1 is hidden_states else as_rows_bf16(hidden_states_1)
            if hidden_states_2 is None or hidden_states_2 is hidden_states_1:
                xs = None
            elif hidden_states_2 is hidden_states:
                xs = x2
            else:
                xs = as_rows_bf16(hidden_states_2)
            y = self.fused_forward(x2, scale, x1, xs)
        return y.reshape(hidden_states.shape[:-1] + (self.out_features,)).to(hidden_states.dtype)
