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


def _transposed(b: torch.Tensor) -> torch.Tensor:
    """(K, N) contiguous copy of a (N, K) GEMM operand, memoised on the operand's version: the
    frozen base weights (the bulk of the stack) are transposed once, so dX = dY W runs on the
    K-major CTA-pair GEMM (vst_lora_gemm) instead of the MN-major path of vst_gemm (r01: 53 -> 45 us
    at the L2 projection shape)."""
    from ._host import cached_build
    return cached_build([b], "transposed", lambda: b.detach().t().contiguous())


class LoraGemmFn(torch.autograd.Function):
    """y = (a1 b1^T + a2 b2^T [grouped] + bias + residual) * out_scale."""

    @staticmethod
    def forward(ctx, a1, b1, a2, b2, bias, residual, out_scale, act, k2_group_n, out_dtype):
        y = ops.lora_gemm_raw(a1, b1, a2, b2, bias, residual, out_scale, act, k2_group_n, None, out_dtype)
        z = None
        if act == "gelu":
            # the pre-activation, kept for gelu'(z): one more GEMM, in training mode only
            z = ops.lora_gemm_raw(a1, b1, a2, b2, bias, None, 1.0, "none", k2_group_n, None, torch.bfloat16)
        ctx.save_for_backward(a1, b1, a2, b2, z)
        ctx.cfg = (out_scale, k2_group_n, None if bias is None else bias.dtype,
                   None if residual is None else residual.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        a1, b1, a2, b2, z = ctx.saved_tensors
        out_scale, gn, bias_dtype, res_dtype = ctx.cfg
        need = ctx.needs_input_grad
        g = _rows_bf16(dy)
        m, n = g.shape
        da1 = db1 = da2 = db2 = dbias = dres = None
        if need[5]:
            dres = g if out_scale == 1.0 else g * out_scale
            dres = dres.to(res_dtype).reshape(dy.shape)
        if z is not None:            # y = (gelu(z) + residual) * out_scale
            g = ops.gelu_bwd(z, g, out_scale)
            out_scale = 1.0
        if need[0]:     # dx = dy W                      (W stored [N, K1])
            if n % 8 == 0 and b1.shape[1] % 8 == 0 and m >= 256:
                da1 = ops.lora_gemm_raw(g, _transposed(b1), out_scale=out_scale)
            else:
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
        return da1, db1, da2, db2, dbias, dres, None, None, None, None


class FusedLoraFn(torch.autograd.Function):
    """y = (x w^T + (x a^T) bp^T [grouped] + bias) * out_scale through the single-launch kernel; u = x a^T comes
    back from the same launch for the backward, which is the arithmetic of LoraGemmFn applied to
    y = x w^T + u bp^T and u = x a^T:  du = dy bp, dx = dy w + du a, dw = dy^T x, da = du^T x, dbp = dy^T u."""

    @staticmethod
    def forward(ctx, x, w, a, bp, bias, out_scale, k2_group_n):
        y, u = ops.lora_gemm_fused_raw(x, w, a, bp, bias, out_scale, k2_group_n, True)
        ctx.save_for_backward(x, w, a, bp, u)
        ctx.cfg = (out_scale, k2_group_n, None if bias is None else bias.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, a, bp, u = ctx.saved_tensors
        out_scale, gn, bias_dtype = ctx.cfg
        return _low_rank_backward(ctx.needs_input_grad, x, w, a, bp, u, _rows_bf16(dy), out_scale, gn,
                                  bias_dtype) + (None, None)


def _low_rank_backward(need, x, w, a, bp, u, g, out_scale, gn, bias_dtype):
    """(dx, dw, da, dbp, dbias) of y = (x w^T + u bp^T [grouped] + bias) * out_scale with u = x a^T, from
    g = dy (2-D bf16 rows); need = needs_input_grad in the order (x, w, a, bp, bias, ...)."""
    m, n = g.shape
    k2 = bp.shape[1]
    groups = (n // gn) if gn else 1
    dx = dw = da = dbp = dbias = None
    du = None
    if need[0] or need[2]:                       # du = dy bp   (per projection group)
        if groups == 1:
            du = ops.gemm(g, bp, trans_b=True, alpha=out_scale)
        else:
            du = torch.empty((m, groups * k2), dtype=torch.bfloat16, device=g.device)
            for i in range(groups):
                ops.gemm(g[:, i * gn:(i + 1) * gn], bp[i * gn:(i + 1) * gn], trans_b=True, alpha=out_scale,
                         out=du[:, i * k2:(i + 1) * k2])
    if need[0]:                                  # dx = dy w + du a   (two operand pairs, one GEMM)
        if (out_scale == 1.0 and m >= 256 and n % 8 == 0 and w.shape[1] % 8 == 0 and du.shape[1] % 8 == 0
                and not need[1]):
            # frozen W (every LoRA site of Stage 1 / Stage 2): W is transposed once (memoised; the rank-sized A
            # again whenever it changed), so the two-pair product runs on the K-major CTA-pair GEMM like
            # LoraGemmFn's dx, not on vst_gemm's single-CTA MN-major path
            dx = ops.lora_gemm_raw(g, _transposed(w), du, _transposed(a))
        elif out_scale == 1.0:
            dx = ops.gemm(g, w, du, a, trans_b=True)
        else:
            dx = ops.gemm(g, w, trans_b=True, alpha=out_scale)
            ops.gemm(du, a, trans_b=True, out=dx, accumulate=True)
    if need[1]:
        dw = ops.gemm(g, x, trans_a=True, trans_b=True, alpha=out_scale)
    if need[2]:
        da = ops.gemm(du, x, trans_a=True, trans_b=True)
    if need[3]:
        if groups == 1:
            dbp = ops.gemm(g, u, trans_a=True, trans_b=True, alpha=out_scale)
        else:
            dbp = torch.empty((n, k2), dtype=torch.bfloat16, device=g.device)
            for i in range(groups):
                ops.gemm(g[:, i * gn:(i + 1) * gn], u[:, i * k2:(i + 1) * k2], trans_a=True, trans_b=True,
                         alpha=out_scale, out=dbp[i * gn:(i + 1) * gn])
    if need[4]:
        dbias = ops.colsum(g, out_scale).to(bias_dtype)
    return dx, dw, da, dbp, dbias


class TwoLaunchLoraFn(torch.autograd.Function):
    """y = (x w^T + (x a^T) bp^T [grouped] + bias + residual) * out_scale as u = x a^T followed by the GEMM with
    extra K blocks -- the form grouped q|k|v and residual-carrying projections take -- recorded as ONE node.
    As two LoraGemmFn nodes (u, then y) autograd summed the two contributions to dx with an elementwise
    kernel, three passes over an (M, in) tensor per projection; here dx = dy w + du a is one GEMM with two
    operand pairs, exactly as in FusedLoraFn."""

    @staticmethod
    def forward(ctx, x, w, a, bp, bias, residual, out_scale, k2_group_n):
        u = ops.lora_gemm_raw(x, a)
        y = ops.lora_gemm_raw(x, w, u, bp, bias, residual, out_scale, "none", k2_group_n)
        ctx.save_for_backward(x, w, a, bp, u)
        ctx.cfg = (out_scale, k2_group_n, None if bias is None else bias.dtype,
                   None if residual is None else residual.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, a, bp, u = ctx.saved_tensors
        out_scale, gn, bias_dtype, res_dtype = ctx.cfg
        need = ctx.needs_input_grad
        g = _rows_bf16(dy)
        dres = None
        if need[5]:
            dres = (g if out_scale == 1.0 else g * out_scale).to(res_dtype).reshape(dy.shape)
        return _low_rank_backward(need, x, w, a, bp, u, g, out_scale, gn, bias_dtype) + (dres, None, None)


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


def _rows3(t2: torch.Tensor, b: int, s_: int) -> torch.Tensor:
    """(b*s, n) column view (row stride ld) -> (b, s, n) with the same row stride."""
    return t2.as_strided((b, s_, t2.shape[1]), (s_ * t2.stride(0), t2.stride(0), 1), t2.storage_offset())


def _packed_grads(ts):
    """If the (B,S,C) tensors `ts` are adjacent column slices of one packed (B*S, len*C) projection
    output, gradient views with the same packing (one fresh buffer); else None.  The gradient of a
    packed q|k|v then reaches the projection's backward GEMM without a single copy."""
    t0 = ts[0]
    b, s_, c = t0.shape
    es = t0.element_size()
    for i, t in enumerate(ts):
        if t.shape != t0.shape or t.stride() != t0.stride() or t.data_ptr() != t0.data_ptr() + i * c * es:
            return None
    if t0.stride(1) != len(ts) * c or t0.stride(0) != s_ * t0.stride(1):
        return None
    buf = torch.empty((b * s_, len(ts) * c), dtype=torch.bfloat16, device=t0.device)
    return [_rows3(buf[:, i * c:(i + 1) * c], b, s_) for i in range(len(ts))]


class SplitColsFn(torch.autograd.Function):
    """y (M, count*n0) -> count column views, 2-D or (b, s, n0).  Backward: when the incoming
    gradients are already the column slices of one packed buffer (AttnFn / TemporalAttnFn write them
    that way) it is returned as is; otherwise they are packed with one copy each -- instead of
    autograd's as_strided / slice backward (a zero fill of the whole packed tensor, a strided copy
    and an add per slice: 27 % of the Stage-2 training step in the r01 launch list)."""

    @staticmethod
    def forward(ctx, y, n0, count, b, s_):
        ctx.cfg = (n0, count, y.shape[0], b, s_)
        ctx.set_materialize_grads(False)
        outs = []
        for i in range(count):
            v = y[:, i * n0:(i + 1) * n0]
            outs.append(v if b is None else _rows3(v, b, s_))
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gs):
        n0, count, m, b, s_ = ctx.cfg
        g0 = next((g for g in gs if g is not None), None)
        if g0 is None:
            return None, None, None, None, None
        es = g0.element_size()
        ld = g0.stride(-2)
        packed = all(g is not None and g.dtype == g0.dtype and g.shape == g0.shape and g.stride() == g0.stride()
                     and g.data_ptr() == g0.data_ptr() + i * n0 * es for i, g in enumerate(gs))
        packed = packed and g0.stride(-1) == 1 and ld == count * n0 and (g0.dim() == 2 or g0.stride(0) == s_ * ld)
        if packed:
            return g0.as_strided((m, count * n0), (ld, 1), g0.storage_offset()), None, None, None, None
        buf = torch.empty((m, count * n0), dtype=g0.dtype, device=g0.device)
        for i, g in enumerate(gs):
            dst = buf[:, i * n0:(i + 1) * n0]
            if g is None:
                dst.zero_()
            else:
                (dst if b is None else _rows3(dst, b, s_)).copy_(g)
        return buf, None, None, None, None


def split_cols(y: torch.Tensor, n0: int, count: int, batch=None):
    """Column views of a packed projection output; differentiable without autograd's slice glue."""
    b, s_ = batch if batch is not None else (None, None)
    if torch.is_grad_enabled() and y.requires_grad:
        return list(SplitColsFn.apply(y, n0, count, b, s_))
    outs = [y[:, i * n0:(i + 1) * n0] for i in range(count)]
    return outs if b is None else [_rows3(v, b, s_) for v in outs]


class AttnFn(torch.autograd.Function):
    """softmax(q k^T * scale) v per head (vst_attn_fwd) with the flash backward (vst_attn_bwd)."""

    @staticmethod
    def forward(ctx, q, k, v, heads, scale, mask=None):
        o, lse = ops.attn_fwd_raw(q, k, v, heads, scale, return_lse=True, mask=mask)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.cfg = (heads, scale)
        ctx.mask = None if mask is None else mask.detach()     # no gradient w.r.t. the mask
        return o

    @staticmethod
    def backward(ctx, d_o):
        q, k, v, o, lse = ctx.saved_tensors
        heads, scale = ctx.cfg
        g = d_o.detach()
        if g.dtype != torch.bfloat16 or g.stride(2) != 1 or g.stride(0) != g.shape[1] * g.stride(1) \
                or g.stride(1) % 8:
            g = g.to(torch.bfloat16).contiguous()
        out = None
        if q.shape == k.shape:
            pk = _packed_grads([q, k, v])                   # self-attention: packed q|k|v
            out = None if pk is None else tuple(pk)
        else:
            pk = _packed_grads([k, v])                      # cross-attention: packed k|v
            if pk is not None:
                out = (torch.empty(q.shape, dtype=torch.bfloat16, device=q.device), pk[0], pk[1])
        dq, dk, dv = ops.attn_bwd(q.detach(), k.detach(), v.detach(), o, g, lse, heads, scale, ctx.mask, out=out)
        return dq, dk, dv, None, None, None


def _dense_bf16(t: torch.Tensor) -> torch.Tensor:
    t = t.detach()
    return t if (t.dtype == torch.bfloat16 and t.is_contiguous()) else t.to(torch.bfloat16).contiguous()


class TemporalAttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, heads, scale):
        ctx.save_for_backward(q, k, v)
        ctx.cfg = (heads, scale)
        return ops.temporal_attn_raw(q, k, v, heads, scale)

    @staticmethod
    def backward(ctx, d_o):
        q, k, v = ctx.saved_tensors
        heads, scale = ctx.cfg
        pk = _packed_grads([q, k, v])
        dq, dk, dv = ops.temporal_attn_bwd(q.detach(), k.detach(), v.detach(), _dense_bf16(d_o), heads, scale,
                                           out=None if pk is None else tuple(pk))
        return dq, dk, dv, None, None


class LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps, pe, frames, out_dtype):
        y = ops.layernorm_raw(x, gamma, beta, eps, pe, frames, out_dtype)
        ctx.save_for_backward(x, gamma)
        ctx.cfg = (eps, None if gamma is None else gamma.dtype, None if beta is None else beta.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma = ctx.saved_tensors
        eps, gdt, bdt = ctx.cfg
        c = x.shape[-1]
        x2 = _dense_bf16(x).reshape(-1, c)
        dy2 = _dense_bf16(dy).reshape(-1, c)
        want = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dx, dg, db = ops.layernorm_bwd(x2, dy2, None if gamma is None else gamma.detach(), eps, want)
        dx = dx.reshape(x.shape).to(x.dtype) if ctx.needs_input_grad[0] else None
        dg = dg.to(gdt) if (ctx.needs_input_grad[1] and dg is not None) else None
        db = db.to(bdt) if (ctx.needs_input_grad[2] and db is not None) else None
        return dx, dg, db, None, None, None, None


class GegluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hg):
        ctx.save_for_backward(hg)
        return ops.geglu_raw(hg)

    @staticmethod
    def backward(ctx, d_out):
        (hg,) = ctx.saved_tensors
        return ops.geglu_bwd(hg.detach(), _dense_bf16(d_out))


def _frames_major_storage(x5: torch.Tensor) -> torch.Tensor:
    """(B,C,F,H,W) view of a contiguous (B,F,C,H,W) tensor -> that (B*F, C, H, W) tensor."""
    b, c, f, h, w = x5.shape
    xf = x5.permute(0, 2, 1, 3, 4)
    if not xf.is_contiguous():
        raise ValueError("the motion-module input must be the (B,C,F,H,W) view of a contiguous "
                         "(B*F,C,H,W) tensor for the backward pass")
    return xf.reshape(b * f, c, h, w)


class GroupNormRearrangeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x5, gamma, beta, groups, eps):
        tok, mean, rstd = ops.groupnorm5d_rearrange_raw(x5, gamma, beta, groups, eps)
        ctx.save_for_backward(x5, gamma, mean, rstd)
        ctx.cfg = (groups, beta.dtype)
        ctx.mark_non_differentiable(mean, rstd)
        return tok, mean, rstd

    @staticmethod
    def backward(ctx, d_tok, _dm, _dr):
        x5, gamma, mean, rstd = ctx.saved_tensors
        groups, bdt = ctx.cfg
        b, c, f, h, w = x5.shape
        x_fm = _frames_major_storage(x5.detach())
        # tokens -> frames-major, then a plain GroupNorm backward in x's own layout
        dy5 = ops.rearrange_bwd_raw(_dense_bf16(d_tok).view(b * h * w, f, c), b, h, w, None, True)
        dy_fm = dy5.permute(0, 2, 1, 3, 4).reshape(b * f, c, h, w)
        dx, dg, db = ops.groupnorm5d_bwd(x_fm, dy_fm, mean, rstd, gamma.detach(), b, f, groups)
        dx5 = dx.view(b, f, c, h, w).permute(0, 2, 1, 3, 4)
        return dx5, dg.to(gamma.dtype), db.to(bdt), None, None


class RearrangeBwdFn(torch.autograd.Function):
    """tokens (B*H*W, F, C) -> (B,C,F,H,W) (+ residual); the adjoint is the forward rearrange."""

    @staticmethod
    def forward(ctx, y, residual5, b, h, w, frames_major):
        ctx.has_res = residual5 is not None
        return ops.rearrange_bwd_raw(y, b, h, w, residual5, frames_major)

    @staticmethod
    def backward(ctx, d_out5):
        g = d_out5.detach()
        if g.dtype not in (torch.bfloat16, torch.float32):
            g = g.to(torch.bfloat16)
        hh, ww = g.shape[3], g.shape[4]
        if hh * ww > 1 and not (g.stride(4) == 1 and (hh == 1 or g.stride(3) == ww)):
            g = g.contiguous()
        dy = ops.rearrange_fwd(g) if ctx.needs_input_grad[0] else None
        dres = d_out5 if (ctx.has_res and ctx.needs_input_grad[1]) else None
        return dy, dres, None, None, None, None


class RearrangeFwdFn(torch.autograd.Function):
    """(B,C,F,H,W) -> tokens (B*H*W, F, C) (+ pe, no gradient); the adjoint is rearrange_bwd."""

    @staticmethod
    def forward(ctx, x5, pe):
        ctx.geom = (x5.shape, x5.dtype)
        with torch.no_grad():
            return ops.rearrange_fwd(x5.detach(), pe)

    @staticmethod
    def backward(ctx, dy):
        (b, c, f, h, w), dtype = ctx.geom
        g = dy.detach().contiguous()
        return ops.rearrange_bwd_raw(g.view(b * h * w, f, c), b, h, w).to(dtype), None
