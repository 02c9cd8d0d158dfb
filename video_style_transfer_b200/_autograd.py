"""torch.autograd glue: each Function's forward and backward are C-ABI kernel calls (ops.*); torch
only records the graph.  Reference: the reference relies on autograd through F.linear / SDPA for
unziplora_linear_layer.py:221-263, temporal_lora.py:29-32 and attention_processor.py:78-80
(train_animatediff.py); here the same gradients come from vst_gemm / vst_colsum /
vst_lora_pack_up_bwd / vst_attn_bwd."""
from typing import Optional

import torch

from . import ops


def _rows_bf16(t: torch.Tensor) -> torch.Tensor:
    t = t.detach()
    if t.dtype != torch.bfloat16:
        t = t.to(torch.bfloat16)
    if t.stride(-1) != 1 or (t.stride(0) % 8) != 0 or (t.data_ptr() % 16) != 0:
        t = t.contiguous()
    return t


class LoraGemmFn(torch.autograd.Function):
    """y = (a1 b1^T + a2 b2^T [grouped] + bias + residual) * out_scale."""

    @staticmethod
    def forward(ctx, a1, b1, a2, b2, bias, residual, out_scale, act, k2_group_n, out_dtype):
        if act != "none":
            raise NotImplementedError("lora_gemm: only act='none' is differentiable (the GELU feed-forward "
                                      "of temporal_transformer.py is inference-only on this path)")
        y = ops.lora_gemm_raw(a1, b1, a2, b2, bias, residual, out_scale, act, k2_group_n, None, out_dtype)
        ctx.save_for_backward(a1, b1, a2, b2)
        ctx.cfg = (out_scale, k2_group_n, None if bias is None else bias.dtype,
                   None if residual is None else residual.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        a1, b1, a2, b2 = ctx.saved_tensors
        out_scale, gn, bias_dtype, res_dtype = ctx.cfg
        need = ctx.needs_input_grad
        g = _rows_bf16(dy)
        m, n = g.shape
        da1 = db1 = da2 = db2 = dbias = dres = None
        if need[0]:     # dx = dy W                      (W stored [N, K1]: trans_b)
            da1 = ops.gemm(g, b1, trans_b=True, alpha=out_scale)
        if need[1]:     # dW = dy^T x                    (contraction over the rows of both)
            db1 = ops.gemm(g, a1, trans_a=True, trans_b=True, alpha=out_scale)
        if a2 is not None and (need[2] or need[3]):
            k2 = b2.shape[1]
            groups = (n // gn) if gn else 1
            if need[2]:
                if groups == 1:
                    da2 = ops.gemm(g, b2, trans_b=True, alpha=out_scale)
                else:
                    da2 = torch.empty((m, groups * k2), dtype=torch.bfloat16, device=g.device)
                    for i in range(groups):
                        ops.gemm(g[:, i * gn:(i + 1) * gn], b2[i * gn:(i + 1) * gn], trans_b=True,
                                 alpha=out_scale, out=da2[:, i * k2:(i + 1) * k2])
            if need[3]:
                if groups == 1:
                    db2 = ops.gemm(g, a2, trans_a=True, trans_b=True, alpha=out_scale)
                else:
                    db2 = torch.empty((n, k2), dtype=torch.bfloat16, device=g.device)
                    for i in range(groups):
                        ops.gemm(g[:, i * gn:(i + 1) * gn], a2[:, i * k2:(i + 1) * k2], trans_a=True,
                                 trans_b=True, alpha=out_scale, out=db2[i * gn:(i + 1) * gn])
        if need[4]:
            dbias = ops.colsum(g, out_scale).to(bias_dtype)
        if need[5]:
            dres = g if out_scale == 1.0 else g * out_scale
            dres = dres.to(res_dtype).reshape(dy.shape)
        return da1, db1, da2, db2, dbias, dres, None, None, None, None


class CastCatFn(torch.autograd.Function):
    """bf16, row-concatenated (optionally row-padded to 8) GEMM operand of one or more parameters;
    the value itself comes from the version-keyed cache of _host.bf16_cached."""

    @staticmethod
    def forward(ctx, pad_rows, value, *tensors):
        ctx.meta = [(t.shape[0], t.dtype) for t in tensors]
        ctx.pad_rows = pad_rows
        return value.detach().view_as(value)

    @staticmethod
    def backward(ctx, g):
        grads, r0 = [], 0
        for rows, dtype in ctx.meta:
            grads.append(g[r0:r0 + rows].to(dtype))
            r0 += (rows + 7) // 8 * 8 if ctx.pad_rows else rows
        return (None, None) + tuple(grads)


class PackUpFn(torch.autograd.Function):
    """B' = scale * [m_c . mask_c . B_c | m_s . mask_s . B_s] (bf16, rank padded to 8)."""

    @staticmethod
    def forward(ctx, value, scale, rank, mask_c, mask_s, b_c, m_c, b_s, m_s):
        ctx.save_for_backward(*[t for t in (b_c, m_c, b_s, m_s) if t is not None])
        ctx.present = [t is not None for t in (b_c, m_c, b_s, m_s)]
        ctx.masks = (mask_c, mask_s)
        ctx.cfg = (scale, rank)
        return value.detach().view_as(value)

    @staticmethod
    def backward(ctx, g):
        it = iter(ctx.saved_tensors)
        b_c, m_c, b_s, m_s = (next(it) if p else None for p in ctx.present)
        scale, rank = ctx.cfg
        cont = lambda t: None if t is None else t.detach().contiguous()
        ref = b_c if b_c is not None else b_s
        same = lambda t: None if t is None else cont(t).to(ref.dtype)
        dbc, dmc, dbs, dms = ops.lora_pack_up_bwd(g.float().contiguous(), rank, cont(b_c), same(m_c),
                                                  ctx.masks[0], cont(b_s), same(m_s), ctx.masks[1], scale)
        like = lambda gr, t: None if (gr is None or t is None) else gr.to(t.dtype)
        return (None, None, None, None, None, like(dbc, b_c), like(dmc, m_c), like(dbs, b_s), like(dms, m_s))


class AttnFn(torch.autograd.Function):
    """softmax(q k^T * scale) v per head (vst_attn_fwd) with the flash backward (vst_attn_bwd)."""

    @staticmethod
    def forward(ctx, q, k, v, heads, scale):
        o, lse = ops.attn_fwd_raw(q, k, v, heads, scale, return_lse=True)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.cfg = (heads, scale)
        return o

    @staticmethod
    def backward(ctx, d_o):
        q, k, v, o, lse = ctx.saved_tensors
        heads, scale = ctx.cfg
        g = d_o.detach()
        if g.dtype != torch.bfloat16 or g.stride(2) != 1 or g.stride(0) != g.shape[1] * g.stride(1) \
                or g.stride(1) % 8:
            g = g.to(torch.bfloat16).contiguous()
        dq, dk, dv = ops.attn_bwd(q.detach(), k.detach(), v.detach(), o, g, lse, heads, scale)
        return dq, dk, dv, None, None
