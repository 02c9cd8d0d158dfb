"""Drop-in for animatediff/attention_processor.py (reference :6-96): the diffusers-protocol
spatial attention processor used inside ``UNetMotionModel``.

Same ``__call__(attn, hidden_states, encoder_hidden_states=None, attention_mask=None, temb=None,
scale=1.0, **kwargs)`` contract and the same reads from ``attn``; the body runs on sm_100a:
  to_q / to_k / to_v / to_out[0]  one fused tcgen05 LoRA GEMM each (``fused_forward``)
  head split, SDPA, head merge    one tcgen05 flash-attention kernel (``ops.attn_fwd``); the text
                                  ``repeat_interleave`` (:62-66) is a kv batch divisor, so K/V are
                                  projected once per prompt instead of once per frame
  residual add, / rescale         folded into the to_out epilogue (:91-94)
"""
from typing import Optional

import torch

from .. import ops
from .._host import as_rows_bf16
from ..fused import project_fused


class AnimateDiffAttnProcessor2_0:
    def __init__(self):
        if not torch.cuda.is_available():
            raise ImportError("AnimateDiffAttnProcessor2_0 (B200) requires a CUDA device; "
                              "there is no CPU fallback.")

    @staticmethod
    def _proj(layer, x2, scale, **kw):
        if not hasattr(layer, "fused_forward"):
            raise TypeError(f"{type(layer).__name__} is not a video_style_transfer_b200 "
                            "LoRACompatibleLinear; wrap the attention projections first")
        return layer.fused_forward(x2, scale, **kw)

    def __call__(self, attn, hidden_states: torch.Tensor,
                 encoder_hidden_states: Optional[torch.Tensor] = None,
                 attention_mask: Optional[torch.Tensor] = None,
                 temb: Optional[torch.Tensor] = None, scale: float = 1.0, **kwargs) -> torch.Tensor:
        residual = hidden_states
        in_dtype = hidden_states.dtype
        if attn.spatial_norm is not None:
            hidden_states = attn.spatial_norm(hidden_states, temb)
        input_ndim = hidden_states.ndim
        if input_ndim == 4:
            bsz, channel, height, width = hidden_states.shape
            hidden_states = hidden_states.view(bsz, channel, height * width).transpose(1, 2)
        batch_size, seq_q, _ = hidden_states.shape
        if attention_mask is not None:                                       # reference :46-48
            seq_len = seq_q if encoder_hidden_states is None else encoder_hidden_states.shape[1]
            attention_mask = attn.prepare_attention_mask(attention_mask, seq_len, batch_size)
            attention_mask = attention_mask.view(batch_size, attn.heads, -1, attention_mask.shape[-1])
        if attn.group_norm is not None:
            hidden_states = attn.group_norm(hidden_states.transpose(1, 2)).transpose(1, 2)
        lora_scale = scale if hasattr(attn.to_q, "lora_layer") else 1.0

        x2 = as_rows_bf16(hidden_states)
        if encoder_hidden_states is None:
            # self-attention: q | k | v read the same tokens -> one GEMM, N = 3C
            q, k, v = project_fused((attn.to_q, attn.to_k, attn.to_v), x2, lora_scale, batch=(batch_size, seq_q))
            kv_batch, seq_k = batch_size, seq_q
        else:
            q = self._proj(attn.to_q, x2, lora_scale)
            q = q.view(batch_size, seq_q, q.shape[1])
            if attn.norm_cross:
                encoder_hidden_states = attn.norm_encoder_hidden_states(encoder_hidden_states)
            kv_batch, seq_k = encoder_hidden_states.shape[:2]
            if batch_size % kv_batch != 0:
                raise ValueError("text batch must divide the B*F batch")
            e2 = as_rows_bf16(encoder_hidden_states)       # NOT repeated: kv_group handles :62-66
            k, v = project_fused((attn.to_k, attn.to_v), e2, lora_scale, batch=(kv_batch, seq_k))   # N = 2C
        inner = k.shape[-1]
        # q / k / v: (b, s, inner) column views of the packed projection output (packed row stride)
        o = ops.attn_fwd(q, k, v, attn.heads, mask=attention_mask)
        o2 = o.view(batch_size * seq_q, inner)

        fuse_res = (attn.residual_connection and input_ndim == 3 and residual.dtype == torch.bfloat16
                    and residual.is_contiguous())
        res2 = residual.view(batch_size * seq_q, -1) if fuse_res else None
        post = 1.0 / float(attn.rescale_output_factor)
        y = self._proj(attn.to_out[0], o2, lora_scale, residual=res2,
                       out_scale=post if (fuse_res or not attn.residual_connection) else 1.0)
        y = attn.to_out[1](y)                                             # dropout (p = 0)
        y = y.view(batch_size, seq_q, -1).to(in_dtype)
        if input_ndim == 4:
            y = y.transpose(-1, -2).reshape(bsz, channel, height, width)
        if attn.residual_connection and not fuse_res:
            y = (y + residual) * post
        return y
