"""Drop-in for the Stage-1 (image UnZipLoRA) attention processor,
unziplora_unet/unzip_attention_processor.py ``AttnProcessor2_0`` (reference :664-759), invoked by
the reference's ``Attention.forward`` (:141-181) with THREE text streams:

    proc(attn, hidden_states, encoder_hidden_states=None, encoder_hidden_states_content=None,
         encoder_hidden_states_style=None, attention_mask=None, temb=None, scale=1.0)

The base projection reads ``encoder_hidden_states`` while the content / style low-rank branches
of to_k / to_v read the content / style prompts (lora_unzip.py:66-75); to_q and to_out alias all
three streams to their own input (:706-708, :745-747) and self-attention aliases them to
``hidden_states`` (:711-714).  Like the reference, the processor does NOT forward ``scale`` to the
projections (they run at their default 1.0).

Device work: every projection is one fused tcgen05 LoRA GEMM (self-attention: q | k | v as ONE
GEMM along N), SDPA is the tcgen05 flash-attention kernel reading column views of the packed
projection, residual add and /rescale ride in the to_out epilogue.  Gradients flow through the
same Functions as the AnimateDiff processor (``_autograd``), so Stage-1 training of A, B and the
mergers runs on this path too.
"""
from typing import Optional

import torch

from .. import ops
from .._host import as_rows_bf16
from ..fused import project_fused


class AttnProcessor2_0:
    def __init__(self):
        if not torch.cuda.is_available():
            raise ImportError("AttnProcessor2_0 (B200, Stage 1) requires a CUDA device; "
                              "there is no CPU fallback.")

    @staticmethod
    def _proj(layer, x2, x_content=None, x_style=None, **kw):
        if not hasattr(layer, "fused_forward"):
            raise TypeError(f"{type(layer).__name__} is not a video_style_transfer_b200 "
                            "LoRACompatibleLinear; wrap the attention projections first")
        return layer.fused_forward(x2, 1.0, x_content, x_style, **kw)

    def __call__(self, attn, hidden_states: torch.Tensor,
                 encoder_hidden_states: Optional[torch.Tensor] = None,
                 encoder_hidden_states_content: Optional[torch.Tensor] = None,
                 encoder_hidden_states_style: Optional[torch.Tensor] = None,
                 attention_mask: Optional[torch.Tensor] = None,
                 temb: Optional[torch.Tensor] = None, scale: float = 1.0) -> torch.Tensor:
        residual = hidden_states
        in_dtype = hidden_states.dtype
        if attn.spatial_norm is not None:                                    # :678-679
            hidden_states = attn.spatial_norm(hidden_states, temb)
        input_ndim = hidden_states.ndim
        if input_ndim == 4:                                                  # :683-685
            bsz, channel, height, width = hidden_states.shape
            hidden_states = hidden_states.view(bsz, channel, height * width).transpose(1, 2)
        batch_size, seq_q, _ = hidden_states.shape
        if attention_mask is not None:                                       # :691-695
            seq_len = seq_q if encoder_hidden_states is None else encoder_hidden_states.shape[1]
            attention_mask = attn.prepare_attention_mask(attention_mask, seq_len, batch_size)
            attention_mask = attention_mask.view(batch_size, attn.heads, -1, attention_mask.shape[-1])
        if attn.group_norm is not None:                                      # :697-698
            hidden_states = attn.group_norm(hidden_states.transpose(1, 2)).transpose(1, 2)

        x2 = as_rows_bf16(hidden_states)
        if encoder_hidden_states is None:
            # :711-714 all three streams are the image tokens -> single-stream q | k | v, one GEMM
            q, k, v = project_fused((attn.to_q, attn.to_k, attn.to_v), x2, 1.0, batch=(batch_size, seq_q))
        else:
            q = self._proj(attn.to_q, x2).view(batch_size, seq_q, -1)        # :706-708
            if encoder_hidden_states_content is None or encoder_hidden_states_style is None:
                raise ValueError("Stage-1 cross-attention needs encoder_hidden_states_content "
                                 "and encoder_hidden_states_style")
            if attn.norm_cross:                                              # :715-718
                encoder_hidden_states = attn.norm_encoder_hidden_states(encoder_hidden_states)
                encoder_hidden_states_content = attn.norm_encoder_hidden_states(encoder_hidden_states_content)
                encoder_hidden_states_style = attn.norm_encoder_hidden_states(encoder_hidden_states_style)
            kv_batch, seq_k = encoder_hidden_states.shape[:2]
            if kv_batch != batch_size:
                raise ValueError("Stage-1 text batch must equal the image batch (reference :729-730)")
            e2 = as_rows_bf16(encoder_hidden_states)
            ec = as_rows_bf16(encoder_hidden_states_content)
            es = as_rows_bf16(encoder_hidden_states_style)
            k = self._proj(attn.to_k, e2, ec, es).view(kv_batch, seq_k, -1)  # :720-725
            v = self._proj(attn.to_v, e2, ec, es).view(kv_batch, seq_k, -1)
        inner = k.shape[-1]
        o = ops.attn_fwd(q, k, v, attn.heads, mask=attention_mask)           # :737-741
        o2 = o.view(batch_size * seq_q, inner)

        drop = attn.to_out[1]        # dropout precedes the residual add (:749-755): fuse only if identity
        drop_id = not (getattr(drop, "training", False) and getattr(drop, "p", 0.0) > 0.0)
        fuse_res = (attn.residual_connection and input_ndim == 3 and residual.dtype == torch.bfloat16
                    and residual.is_contiguous() and drop_id)
        res2 = residual.view(batch_size * seq_q, -1) if fuse_res else None
        post = 1.0 / float(attn.rescale_output_factor)
        y = self._proj(attn.to_out[0], o2, residual=res2,                     # :745-747
                       out_scale=post if (fuse_res or (not attn.residual_connection and drop_id)) else 1.0)
        y = attn.to_out[1](y)                                                # dropout :749
        y = y.view(batch_size, seq_q, -1).to(in_dtype)
        if input_ndim == 4:                                                  # :751-752
            y = y.transpose(-1, -2).reshape(bsz, channel, height, width)
        if attn.residual_connection and not fuse_res:                        # :754-757
            y = (y + residual) * post
        elif not attn.residual_connection and not drop_id:
            y = y * post
        return y
