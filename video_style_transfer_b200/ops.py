"""Tensor-level entry points over the C ABI (include/vst_b200.h).  torch is plumbing here:
device memory, the current stream and dtype checks.  All arithmetic happens in libvst_b200.so.
"""
import math
from typing import Optional, Tuple

import torch

from ._lib import VST_BF16, VST_F32, check, lib

_DT = {torch.float32: VST_F32, torch.bfloat16: VST_BF16}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}: the sm_100a path takes float32 / bfloat16")


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("video_style_transfer_b200 ops need CUDA tensors (no CPU fallback)")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _planar_strides(x5: torch.Tensor) -> Tuple[int, int, int]:
    """(sB, sC, sF) of a (B,C,F,H,W) view whose (H,W) block is contiguous."""
    b, c, f, h, w = x5.shape
    if h * w > 1 and not (x5.stride(4) == 1 and (h == 1 or x5.stride(3) == w)):
        raise ValueError("the (H,W) plane of the (B,C,F,H,W) view must be contiguous")
    return x5.stride(0), x5.stride(1), x5.stride(2)


# ---- a9 -------------------------------------------------------------------------------------------
def rearrange_fwd(x5: torch.Tensor, pe: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(B,C,F,H,W) view -> (B*H*W, F, C) [+ pe[:F]]; temporal_transformer.py:129-133.
    Differentiable w.r.t. x5 (adjoint: rearrange_bwd)."""
    if _needs_grad(x5):
        from ._autograd import RearrangeFwdFn
        return RearrangeFwdFn.apply(x5, pe)
    x5 = x5.detach()
    _cuda(x5, pe)
    b, c, f, h, w = x5.shape
    sB, sC, sF = _planar_strides(x5)
    y = torch.empty((b * h * w, f, c), dtype=x5.dtype, device=x5.device)
    if pe is not None:
        pe = pe.reshape(-1, pe.shape[-1])
        if pe.dtype != torch.float32 or not pe.is_contiguous() or pe.shape[0] < f or pe.shape[1] != c:
            raise ValueError("pe must be a contiguous fp32 (>=F, C) table")
    check(lib.vst_rearrange_fwd(x5.data_ptr(), y.data_ptr(), _ptr(pe), b, c, f, h * w, sB, sC, sF,
                                _dt(x5), _dt(y), _stream()), "vst_rearrange_fwd")
    return y


def rearrange_bwd(y: torch.Tensor, b: int, h: int, w: int, residual5: Optional[torch.Tensor] = None,
                  frames_major: bool = False) -> torch.Tensor:
    """(B*H*W, F, C) -> (B,C,F,H,W) contiguous (temporal_transformer.py:143-144), or, with
    frames_major, the (B,C,F,H,W) view of a fresh (B*F,C,H,W) tensor.  residual5 (same view
    geometry as the output) is added in the store.  Differentiable (adjoint: rearrange_fwd)."""
    if _needs_grad(y, residual5):
        from ._autograd import RearrangeBwdFn
        return RearrangeBwdFn.apply(y, residual5, b, h, w, frames_major)
    return rearrange_bwd_raw(y, b, h, w, residual5, frames_major)


def rearrange_bwd_raw(y: torch.Tensor, b: int, h: int, w: int, residual5: Optional[torch.Tensor] = None,
                      frames_major: bool = False) -> torch.Tensor:
    y = y.detach()
    residual5 = None if residual5 is None else residual5.detach()
    _cuda(y, residual5)
    n, f, c = y.shape
    assert n == b * h * w and y.is_contiguous()
    if frames_major:
        out = torch.empty((b, f, c, h, w), dtype=y.dtype, device=y.device).permute(0, 2, 1, 3, 4)
    else:
        out = torch.empty((b, c, f, h, w), dtype=y.dtype, device=y.device)
    sB, sC, sF = _planar_strides(out)
    if residual5 is not None:
        if residual5.shape != out.shape or residual5.stride() != out.stride() or residual5.dtype != y.dtype:
            raise ValueError("residual must have the output's shape, strides and dtype")
    check(lib.vst_rearrange_bwd(y.data_ptr(), out.data_ptr(), _ptr(residual5), b, c, f, h * w, sB,
                                sC, sF, _dt(y), _dt(out), _stream()), "vst_rearrange_bwd")
    return out


# ---- a10 ------------------------------------------------------------------------------------------
def groupnorm5d_stats(x5: torch.Tensor, groups: int = 32, eps: float = 1e-6):
    _cuda(x5)
    b, c, f, h, w = x5.shape
    sB, sC, sF = _planar_strides(x5)
    nparts = lib.vst_groupnorm5d_partials(c, f, h * w, groups)
    if nparts <= 0:
        raise ValueError("groupnorm5d: C must be divisible by groups")
    partials = torch.empty((b * groups, nparts, 3), dtype=torch.float32, device=x5.device)
    mean = torch.empty(b * groups, dtype=torch.float32, device=x5.device)
    rstd = torch.empty_like(mean)
    check(lib.vst_groupnorm5d_stats(x5.data_ptr(), partials.data_ptr(), b, c, f, h * w, sB, sC, sF,
                                    groups, _dt(x5), _stream()), "vst_groupnorm5d_stats")
    check(lib.vst_groupnorm5d_finalize(partials.data_ptr(), nparts, b, groups, eps, mean.data_ptr(),
                                       rstd.data_ptr(), _stream()), "vst_groupnorm5d_finalize")
    return mean, rstd


def groupnorm5d_partials(x5: torch.Tensor, groups: int = 32) -> torch.Tensor:
    """(count, mean, M2) partials of the local frames, (B*groups, nparts, 3) fp32: what a
    sequence-parallel rank contributes to the GroupNorm statistics."""
    _cuda(x5)
    x5 = x5.detach()
    b, c, f, h, w = x5.shape
    sB, sC, sF = _planar_strides(x5)
    nparts = lib.vst_groupnorm5d_partials(c, f, h * w, groups)
    if nparts <= 0:
        raise ValueError("groupnorm5d: C must be divisible by groups")
    partials = torch.empty((b * groups, nparts, 3), dtype=torch.float32, device=x5.device)
    check(lib.vst_groupnorm5d_stats(x5.data_ptr(), partials.data_ptr(), b, c, f, h * w, sB, sC, sF,
                                    groups, _dt(x5), _stream()), "vst_groupnorm5d_stats")
    return partials


def groupnorm5d_finalize(partials: torch.Tensor, eps: float = 1e-6):
    """Chan-merge (fixed order) of (B*groups, nparts, 3) partials -> (mean, rstd)."""
    _cuda(partials)
    partials = partials.contiguous()
    bg, nparts, _ = partials.shape
    mean = torch.empty(bg, dtype=torch.float32, device=partials.device)
    rstd = torch.empty_like(mean)
    check(lib.vst_groupnorm5d_finalize(partials.data_ptr(), nparts, bg, 1, eps, mean.data_ptr(),
                                       rstd.data_ptr(), _stream()), "vst_groupnorm5d_finalize")
    return mean, rstd


def groupnorm5d_rearrange(x5: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                          groups: int = 32, eps: float = 1e-6):
    """GroupNorm over (C/groups, F, H, W) of the (B,C,F,H,W) view, written straight into the
    (B*H*W, F, C) token layout.  Returns (tokens, mean, rstd).  Differentiable w.r.t. x5, gamma, beta."""
    if _needs_grad(x5, gamma, beta):
        from ._autograd import GroupNormRearrangeFn
        tok, mean, rstd = GroupNormRearrangeFn.apply(x5, gamma, beta, groups, eps)
        return tok, mean, rstd
    return groupnorm5d_rearrange_raw(x5, gamma, beta, groups, eps)


def groupnorm5d_rearrange_raw(x5: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                              groups: int = 32, eps: float = 1e-6):
    x5, gamma, beta = x5.detach(), gamma.detach(), beta.detach()
    _cuda(x5, gamma, beta)
    b, c, f, h, w = x5.shape
    mean, rstd = groupnorm5d_stats(x5, groups, eps)
    sB, sC, sF = _planar_strides(x5)
    if gamma.dtype != beta.dtype:
        raise TypeError("gamma/beta dtype mismatch")
    y = torch.empty((b * h * w, f, c), dtype=x5.dtype, device=x5.device)
    check(lib.vst_groupnorm5d_apply_rearrange(
        x5.data_ptr(), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
        beta.data_ptr(), b, c, f, h * w, sB, sC, sF, groups, _dt(x5), _dt(y), _dt(gamma),
        _stream()), "vst_groupnorm5d_apply_rearrange")
    return y, mean, rstd


# ---- K9 -------------------------------------------------------------------------------------------
def layernorm(x: torch.Tensor, gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor],
              eps: float = 1e-5, pe: Optional[torch.Tensor] = None, frames: int = 1,
              out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """LayerNorm over the last dim (+ pe[row % frames]).  Differentiable w.r.t. x, gamma, beta."""
    if _needs_grad(x, gamma, beta):
        from ._autograd import LayerNormFn
        return LayerNormFn.apply(x, gamma, beta, eps, pe, frames, out_dtype)
    return layernorm_raw(x, gamma, beta, eps, pe, frames, out_dtype)


def layernorm_raw(x: torch.Tensor, gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor],
                  eps: float = 1e-5, pe: Optional[torch.Tensor] = None, frames: int = 1,
                  out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    x = x.detach()
    gamma = None if gamma is None else gamma.detach()
    beta = None if beta is None else beta.detach()
    _cuda(x, gamma, beta, pe)
    c = x.shape[-1]
    x2 = x.reshape(-1, c)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    y = torch.empty(x2.shape, dtype=out_dtype or x.dtype, device=x.device)
    pdt = _dt(gamma) if gamma is not None else (_dt(beta) if beta is not None else VST_F32)
    if pe is not None:
        pe = pe.reshape(-1, pe.shape[-1])
        assert pe.dtype == torch.float32 and pe.is_contiguous() and pe.shape[0] >= frames
    check(lib.vst_layernorm(x2.data_ptr(), y.data_ptr(), _ptr(gamma), _ptr(beta), _ptr(pe),
                            x2.shape[0], c, frames, eps, _dt(x2), _dt(y), pdt, _stream()),
          "vst_layernorm")
    return y.reshape(x.shape)


# ---- a8 -------------------------------------------------------------------------------------------
def temporal_attn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int,
                  scale: Optional[float] = None) -> torch.Tensor:
    """q,k,v (N, F, heads*d) bf16 (last dim contiguous; row stride free) -> o (N, F, heads*d).
    Differentiable: the backward is vst_temporal_attn_bwd."""
    if _needs_grad(q, k, v):
        from ._autograd import TemporalAttnFn
        return TemporalAttnFn.apply(q, k, v, heads, scale)
    return temporal_attn_raw(q, k, v, heads, scale)


def temporal_attn_bwd(q, k, v, d_o, heads: int, scale: Optional[float] = None, out=None):
    """out: optional (dq, dk, dv) bf16 (N,F,C) tensors to write (contiguous C, dense (N,F); e.g.
    column views of one packed buffer)."""
    _cuda(q, k, v, d_o)
    n, f, c = q.shape
    d = c // heads
    for t in (q, k, v, d_o):
        if t.dtype != torch.bfloat16 or t.stride(2) != 1 or t.stride(0) != f * t.stride(1):
            raise ValueError("temporal_attn_bwd: bf16 (N,F,C) tensors with contiguous C and dense (N,F)")
    if out is None:
        dq, dk, dv = (torch.empty((n, f, c), dtype=torch.bfloat16, device=q.device) for _ in range(3))
    else:
        dq, dk, dv = out
        for t in out:
            if t.shape != (n, f, c) or t.dtype != torch.bfloat16 or t.stride(2) != 1 or t.stride(0) != f * t.stride(1):
                raise ValueError("temporal_attn_bwd: out tensors must be bf16 (N,F,C), contiguous C, dense (N,F)")
    check(lib.vst_temporal_attn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), d_o.data_ptr(),
                                    dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), n, f, heads, d,
                                    q.stride(1), k.stride(1), v.stride(1), d_o.stride(1), dq.stride(1),
                                    dk.stride(1), dv.stride(1),
                                    scale if scale is not None else 1.0 / math.sqrt(d), _stream()),
          "vst_temporal_attn_bwd")
    return dq, dk, dv


def temporal_attn_raw(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int,
                      scale: Optional[float] = None) -> torch.Tensor:
    q, k, v = q.detach(), k.detach(), v.detach()
    _cuda(q, k, v)
    n, f, c = q.shape
    d = c // heads
    for t in (q, k, v):
        if t.dtype != torch.bfloat16 or t.stride(2) != 1 or t.stride(0) != f * t.stride(1):
            raise ValueError("temporal_attn: bf16 (N,F,C) tensors with contiguous C and dense (N,F)")
    o = torch.empty((n, f, c), dtype=torch.bfloat16, device=q.device)
    check(lib.vst_temporal_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), n, f,
                                    heads, d, q.stride(1), k.stride(1), v.stride(1), o.stride(1),
                                    scale if scale is not None else 1.0 / math.sqrt(d), _stream()),
          "vst_temporal_attn_fwd")
    return o


# ---- a3/a4/a5/a7 ----------------------------------------------------------------------------------
def _mat(t: torch.Tensor, name: str):
    if t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"lora_gemm: {name} must be a 2-D bf16 matrix with contiguous rows")
    return t


def _needs_grad(*ts) -> bool:
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in ts)


def lora_gemm(a1: torch.Tensor, b1: torch.Tensor, a2: Optional[torch.Tensor] = None,
              b2: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
              residual: Optional[torch.Tensor] = None, out_scale: float = 1.0,
              act: str = "none", k2_group_n: int = 0, out: Optional[torch.Tensor] = None,
              out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """out[M,N] = (act(a1 b1^T + a2 b2^T + bias) + residual) * out_scale; act in {none, gelu}, or
    "geglu" (b1 (2N,K), bias (2N,): hidden * gelu(gate) in the epilogue; inference only).
    Differentiable (act none / gelu): the backward runs vst_gemm / vst_colsum (see _autograd.py)."""
    if _needs_grad(a1, b1, a2, b2, bias, residual):
        if act == "geglu":
            raise RuntimeError("lora_gemm: act='geglu' is the fused inference epilogue; use "
                               "lora_gemm + geglu under autograd")
        if out is not None:
            raise RuntimeError("lora_gemm: out= cannot be combined with autograd")
        from ._autograd import LoraGemmFn
        return LoraGemmFn.apply(a1, b1, a2, b2, bias, residual, float(out_scale), act,
                                int(k2_group_n), out_dtype)
    return lora_gemm_raw(a1, b1, a2, b2, bias, residual, out_scale, act, k2_group_n, out, out_dtype)


def lora_gemm_raw(a1: torch.Tensor, b1: torch.Tensor, a2: Optional[torch.Tensor] = None,
                  b2: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
                  residual: Optional[torch.Tensor] = None, out_scale: float = 1.0,
                  act: str = "none", k2_group_n: int = 0, out: Optional[torch.Tensor] = None,
                  out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    a1, b1, a2, b2 = (None if t is None else t.detach() for t in (a1, b1, a2, b2))
    bias = None if bias is None else bias.detach()
    residual = None if residual is None else residual.detach()
    _cuda(a1, b1, a2, b2, bias, residual, out)
    _mat(a1, "a1"), _mat(b1, "b1")
    m, k1 = a1.shape
    n = b1.shape[0]
    if act == "geglu":           # b1 = [W_hidden ; W_gate] (2N, K), bias (2N,) -> out (M, N)
        assert n % 2 == 0 and a2 is None and residual is None
        n //= 2
    assert b1.shape[1] == k1
    k2 = 0
    if a2 is not None:
        _mat(a2, "a2"), _mat(b2, "b2")
        k2 = b2.shape[1]
        groups = (n // k2_group_n) if k2_group_n else 1
        assert a2.shape == (m, k2 * groups) and b2.shape[0] == n
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a1.device)
    assert out.shape == (m, n) and out.stride(1) == 1
    if residual is not None:
        assert residual.shape == (m, n) and residual.stride(1) == 1 and residual.dtype == out.dtype
    check(lib.vst_lora_gemm(
        a1.data_ptr(), a1.stride(0), b1.data_ptr(), b1.stride(0), k1,
        _ptr(a2), a2.stride(0) if a2 is not None else 0, _ptr(b2),
        b2.stride(0) if b2 is not None else 0, k2, k2_group_n if a2 is not None else 0,
        _ptr(bias), _dt(bias) if bias is not None else VST_F32,
        _ptr(residual), residual.stride(0) if residual is not None else 0, out_scale,
        {"none": 0, "gelu": 1, "geglu": 2}[act], out.data_ptr(), out.stride(0), m, n, _dt(out), _stream()),
        "vst_lora_gemm")
    return out


def lora_linear_two_launch(x: torch.Tensor, w: torch.Tensor, a_down: torch.Tensor, bp: torch.Tensor,
                           bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
                           out_scale: float = 1.0, k2_group_n: int = 0) -> torch.Tensor:
    """y = (x w^T + (x a_down^T) bp^T [grouped] + bias + residual) * out_scale: the u = x a_down^T pre-pass and
    the GEMM with extra K blocks (two launches).  Under autograd it is ONE node (TwoLaunchLoraFn), so dx is a
    single two-pair GEMM instead of two GEMMs and an elementwise sum."""
    if _needs_grad(x, w, a_down, bp, bias, residual):
        from ._autograd import TwoLaunchLoraFn
        return TwoLaunchLoraFn.apply(x, w, a_down, bp, bias, residual, float(out_scale), int(k2_group_n))
    u = lora_gemm_raw(x, a_down)
    return lora_gemm_raw(x, w, u, bp, bias, residual, out_scale, "none", k2_group_n)


def lora_gemm_fused_supported(m: int, n: int, k1: int, k2: int, k2_group_n: int = 0) -> bool:
    return lib.vst_lora_gemm_fused_supported(m, n, k1, k2, k2_group_n) > 0


def lora_gemm_fused(x: torch.Tensor, w: torch.Tensor, a_down: torch.Tensor, bp: torch.Tensor,
                    bias: Optional[torch.Tensor] = None, out_scale: float = 1.0, k2_group_n: int = 0) -> torch.Tensor:
    """y = (x w^T + (x a_down^T) bp^T + bias) * out_scale in ONE launch (vst_lora_gemm_fused): u = x a_down^T is
    computed inside the kernel, per 256-row tile and projection group.  a_down (G*K2, K1), bp (N, K2), K2 in
    {64, 128}.  Differentiable in x, w, a_down, bp, bias (the kernel then also writes u out for the backward)."""
    if _needs_grad(x, w, a_down, bp, bias):
        from ._autograd import FusedLoraFn
        return FusedLoraFn.apply(x, w, a_down, bp, bias, float(out_scale), int(k2_group_n))
    return lora_gemm_fused_raw(x, w, a_down, bp, bias, out_scale, k2_group_n, False)[0]


def lora_gemm_fused_raw(x, w, a_down, bp, bias=None, out_scale: float = 1.0, k2_group_n: int = 0,
                        save_u: bool = False):
    x, w, a_down, bp = (t.detach() for t in (x, w, a_down, bp))
    bias = None if bias is None else bias.detach()
    _cuda(x, w, a_down, bp, bias)
    for t, nm in ((x, "x"), (w, "w"), (a_down, "a_down"), (bp, "bp")):
        _mat(t, nm)
    m, k1 = x.shape
    n, k2 = bp.shape
    groups = (n // k2_group_n) if k2_group_n else 1
    assert w.shape == (n, k1) and a_down.shape == (groups * k2, k1)
    y = torch.empty((m, n), dtype=torch.bfloat16, device=x.device)
    u = torch.empty((m, groups * k2), dtype=torch.bfloat16, device=x.device) if save_u else None
    check(lib.vst_lora_gemm_fused(x.data_ptr(), x.stride(0), w.data_ptr(), w.stride(0), k1, a_down.data_ptr(),
                                  a_down.stride(0), bp.data_ptr(), bp.stride(0), k2, k2_group_n, _ptr(bias),
                                  _dt(bias) if bias is not None else VST_F32, out_scale, 0, _ptr(u),
                                  u.stride(0) if u is not None else 0, y.data_ptr(), y.stride(0), m, n, _stream()),
          "vst_lora_gemm_fused")
    return y, u


def lora_pack_up(b_c: Optional[torch.Tensor], m_c: Optional[torch.Tensor],
                 mask_c: Optional[torch.Tensor], b_s: Optional[torch.Tensor],
                 m_s: Optional[torch.Tensor], mask_s: Optional[torch.Tensor],
                 scale: float) -> torch.Tensor:
    """[scale*m_c*mask_c (.) B_c | scale*m_s*mask_s (.) B_s] as one bf16 (N, nbr*r) matrix."""
    ref = b_c if b_c is not None else b_s
    _cuda(b_c, m_c, mask_c, b_s, m_s, mask_s)
    n, r = ref.shape
    nbr = (b_c is not None) + (b_s is not None)
    for t in (b_c, m_c, b_s, m_s):
        if t is not None and (t.dtype != ref.dtype or not t.is_contiguous()):
            raise ValueError("lora_pack_up: parameters must share a dtype and be contiguous")
    masks = []
    for mk in (mask_c, mask_s):
        masks.append(None if mk is None else mk.to(torch.uint8).contiguous())
    out = torch.empty((n, nbr * r), dtype=torch.bfloat16, device=ref.device)
    check(lib.vst_lora_pack_up(_ptr(b_c), _ptr(m_c), _ptr(masks[0]), _ptr(b_s), _ptr(m_s),
                               _ptr(masks[1]), scale, out.data_ptr(), n, r, _dt(ref), _stream()),
          "vst_lora_pack_up")
    return out


def gemm(a1: torch.Tensor, b1: torch.Tensor, a2: Optional[torch.Tensor] = None,
         b2: Optional[torch.Tensor] = None, trans_a: bool = False, trans_b: bool = False,
         alpha: float = 1.0, out: Optional[torch.Tensor] = None, accumulate: bool = False,
         out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """out[M,N] (+)= alpha * (opA(a1) opB(b1)^T + opA(a2) opB(b2)^T).  opA(a) is [M,K] (a stored
    [K,M] when trans_a); opB(b) is [N,K] (b stored [K,N] when trans_b).  The backward-pass GEMM."""
    _cuda(a1, b1, a2, b2, out)
    _mat(a1, "a1"), _mat(b1, "b1")
    m, k1 = (a1.shape[1], a1.shape[0]) if trans_a else a1.shape
    n = b1.shape[1] if trans_b else b1.shape[0]
    assert (b1.shape[0] if trans_b else b1.shape[1]) == k1, "gemm: contraction extents differ"
    k2 = 0
    if a2 is not None:
        _mat(a2, "a2"), _mat(b2, "b2")
        m2, k2 = (a2.shape[1], a2.shape[0]) if trans_a else a2.shape
        assert m2 == m and (b2.shape[1] if trans_b else b2.shape[0]) == n
        assert (b2.shape[0] if trans_b else b2.shape[1]) == k2
    if out is None:
        assert not accumulate
        out = torch.empty((m, n), dtype=out_dtype, device=a1.device)
    assert out.shape == (m, n) and out.stride(1) == 1
    wsb = lib.vst_gemm_workspace_bytes(m, n, k1, k2)
    ws = torch.empty(wsb, dtype=torch.uint8, device=a1.device) if wsb else None
    check(lib.vst_gemm(a1.data_ptr(), a1.stride(0), b1.data_ptr(), b1.stride(0), k1, _ptr(a2),
                       a2.stride(0) if a2 is not None else 0, _ptr(b2),
                       b2.stride(0) if b2 is not None else 0, k2, int(trans_a), int(trans_b),
                       out.data_ptr(), out.stride(0), m, n, alpha, int(accumulate), _dt(out),
                       _ptr(ws), wsb, _stream()), "vst_gemm")
    return out


def lora_pack_up_bwd(d_bp: torch.Tensor, rank: int, b_c, m_c, mask_c, b_s, m_s, mask_s, scale: float):
    """Adjoint of lora_pack_up: d_bp (N, nbr*pad8(r)) fp32 -> (dB_c, dm_c, dB_s, dm_s) fp32 (None
    where the branch / merger is absent)."""
    ref = b_c if b_c is not None else b_s
    n, r = ref.shape
    assert d_bp.dtype == torch.float32 and d_bp.stride(1) == 1
    rp = (rank + 7) // 8 * 8
    new = lambda like, shape: None if like is None else torch.empty(shape, dtype=torch.float32, device=ref.device)
    dbc, dbs = new(b_c, (n, r)), new(b_s, (n, r))
    dmc, dms = new(m_c if b_c is not None else None, (n,)), new(m_s if b_s is not None else None, (n,))
    masks = [None if mk is None else mk.to(torch.uint8).contiguous() for mk in (mask_c, mask_s)]
    check(lib.vst_lora_pack_up_bwd(d_bp.data_ptr(), d_bp.stride(0), rp, _ptr(b_c), _ptr(m_c),
                                   _ptr(masks[0]), _ptr(b_s), _ptr(m_s), _ptr(masks[1]), scale,
                                   _ptr(dbc), _ptr(dmc), _ptr(dbs), _ptr(dms), n, r, _dt(ref),
                                   _stream()), "vst_lora_pack_up_bwd")
    return dbc, dmc, dbs, dms


def colsum(x: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    """fp32 column sums of a bf16 (M, N) matrix (the bias gradient)."""
    _cuda(x)
    _mat(x, "x")
    m, n = x.shape
    out = torch.empty(n, dtype=torch.float32, device=x.device)
    wsb = lib.vst_colsum_workspace_bytes(m, n)
    ws = torch.empty(wsb, dtype=torch.uint8, device=x.device)
    check(lib.vst_colsum(x.data_ptr(), x.stride(0), m, n, alpha, out.data_ptr(), ws.data_ptr(), wsb,
                         _stream()), "vst_colsum")
    return out


# ---- a2 -------------------------------------------------------------------------------------------
def _mask_args(mask: Optional[torch.Tensor], bt: int, heads: int, sq: int, sk: int):
    """(tensor kept alive, pointer, batch / head / query strides) of an additive attention mask
    broadcastable to (Bt, heads, Sq, Sk); bool masks follow F.scaled_dot_product_attention
    (False = -inf)."""
    if mask is None:
        return None, None, 0, 0, 0
    m = mask.detach()
    if m.dtype == torch.bool:
        m = torch.zeros(m.shape, dtype=torch.float32, device=m.device).masked_fill_(~m, float("-inf"))
    while m.dim() < 4:
        m = m.unsqueeze(0)
    if m.dim() != 4 or m.shape[-1] != sk or any(a not in (1, b) for a, b in zip(m.shape[:3], (bt, heads, sq))):
        raise ValueError(f"attention mask {tuple(mask.shape)} does not broadcast to {(bt, heads, sq, sk)}")
    if m.dtype != torch.float32 or m.stride(3) != 1:
        m = m.float().contiguous()
    _cuda(m)
    st = [0 if m.shape[i] == 1 else m.stride(i) for i in range(3)]
    return m, m.data_ptr(), st[0], st[1], st[2]


def attn_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int,
             scale: Optional[float] = None, return_lse: bool = False,
             mask: Optional[torch.Tensor] = None):
    """q (Bt,Sq,h*64), k/v (Bk,Sk,h*64) bf16, Bt % Bk == 0 (batch b reads kv batch b//(Bt/Bk)).
    mask: optional additive (or bool) mask broadcastable to (Bt, heads, Sq, Sk).
    Differentiable in q, k, v: the backward is vst_attn_bwd (see _autograd.AttnFn)."""
    if _needs_grad(q, k, v) and not return_lse:
        from ._autograd import AttnFn
        return AttnFn.apply(q, k, v, heads, scale, mask)
    return attn_fwd_raw(q, k, v, heads, scale, return_lse, mask)


def attn_fwd_raw(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int,
                 scale: Optional[float] = None, return_lse: bool = False,
                 mask: Optional[torch.Tensor] = None):
    q, k, v = q.detach(), k.detach(), v.detach()
    _cuda(q, k, v)
    bt, sq, c = q.shape
    bk, sk, _ = k.shape
    d = c // heads
    for t in (q, k, v):
        if t.dtype != torch.bfloat16 or t.stride(2) != 1 or t.stride(0) != t.shape[1] * t.stride(1):
            raise ValueError("attn_fwd: bf16 (B,S,C) tensors with contiguous C and dense (B,S)")
    assert v.shape == k.shape and bt % bk == 0
    o = torch.empty((bt, sq, c), dtype=torch.bfloat16, device=q.device)
    lse = torch.empty((bt, heads, sq), dtype=torch.float32, device=q.device) if return_lse else None
    mk, mptr, msb, msh, msq = _mask_args(mask, bt, heads, sq, sk)
    check(lib.vst_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), _ptr(lse), bt,
                           heads, sq, sk, d, q.stride(1), k.stride(1), v.stride(1), o.stride(1),
                           bt // bk, scale if scale is not None else 1.0 / math.sqrt(d), mptr, msb,
                           msh, msq, _stream()),
          "vst_attn_fwd")
    return (o, lse) if return_lse else o


def _bsc(t: torch.Tensor, name: str):
    if t.dtype != torch.bfloat16 or t.stride(2) != 1 or t.stride(0) != t.shape[1] * t.stride(1):
        raise ValueError(f"attn_bwd: {name} must be bf16 (B,S,C) with contiguous C and dense (B,S)")


def attn_bwd(q, k, v, o, d_o, lse, heads: int, scale: Optional[float] = None,
             mask: Optional[torch.Tensor] = None, out=None):
    """(dq, dk, dv) of attn_fwd; dk / dv have the (Bk, Sk, C) shape of k / v.  out: optional
    (dq, dk, dv) tensors to write (e.g. column views of one packed gradient buffer)."""
    _cuda(q, k, v, o, d_o, lse)
    bt, sq, c = q.shape
    bk, sk, _ = k.shape
    d = c // heads
    for t, nm in ((q, "q"), (k, "k"), (v, "v"), (o, "o"), (d_o, "d_o")):
        _bsc(t, nm)
    assert lse.dtype == torch.float32 and lse.is_contiguous() and lse.shape == (bt, heads, sq)
    if out is None:
        dq = torch.empty((bt, sq, c), dtype=torch.bfloat16, device=q.device)
        dk = torch.empty((bk, sk, c), dtype=torch.bfloat16, device=q.device)
        dv = torch.empty((bk, sk, c), dtype=torch.bfloat16, device=q.device)
    else:
        dq, dk, dv = out
        assert dq.shape == (bt, sq, c) and dk.shape == (bk, sk, c) and dv.shape == (bk, sk, c)
        for t, nm in ((dq, "dq"), (dk, "dk"), (dv, "dv")):
            _bsc(t, nm)
    wsb = lib.vst_attn_bwd_workspace_bytes(bt, heads, sq, sk, bt // bk)
    ws = torch.empty(wsb + 256, dtype=torch.uint8, device=q.device)
    off = (-ws.data_ptr()) % 256
    mk, mptr, msb, msh, msq = _mask_args(mask, bt, heads, sq, sk)
    check(lib.vst_attn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), d_o.data_ptr(),
                           lse.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), bt, heads,
                           sq, sk, d, q.stride(1), k.stride(1), v.stride(1), o.stride(1),
                           d_o.stride(1), dq.stride(1), dk.stride(1), dv.stride(1), bt // bk,
                           scale if scale is not None else 1.0 / math.sqrt(d), mptr, msb, msh, msq,
                           ws.data_ptr() + off, wsb, _stream()), "vst_attn_bwd")
    return dq, dk, dv


# ---- §8f-1 ----------------------------------------------------------------------------------------
def geglu(hg: torch.Tensor) -> torch.Tensor:
    """hg (rows, 2*inner) bf16 -> hg[:, :inner] * gelu(hg[:, inner:]).  Differentiable."""
    if _needs_grad(hg):
        from ._autograd import GegluFn
        return GegluFn.apply(hg)
    return geglu_raw(hg)


def geglu_bwd(hg: torch.Tensor, d_out: torch.Tensor) -> torch.Tensor:
    _cuda(hg, d_out)
    rows, two = hg.shape
    inner = two // 2
    assert d_out.shape == (rows, inner) and d_out.dtype == torch.bfloat16 and d_out.stride(1) == 1
    dhg = torch.empty((rows, two), dtype=torch.bfloat16, device=hg.device)
    check(lib.vst_geglu_bwd(hg.data_ptr(), d_out.data_ptr(), dhg.data_ptr(), rows, inner, hg.stride(0),
                            d_out.stride(0), dhg.stride(0), _stream()), "vst_geglu_bwd")
    return dhg


def gelu_bwd(z: torch.Tensor, dy: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    """alpha * dy * gelu'(z) for bf16 (rows, N) matrices."""
    _cuda(z, dy)
    _mat(z, "z"), _mat(dy, "dy")
    rows, n = z.shape
    dz = torch.empty((rows, n), dtype=torch.bfloat16, device=z.device)
    check(lib.vst_gelu_bwd(z.data_ptr(), dy.data_ptr(), dz.data_ptr(), rows, n, z.stride(0), dy.stride(0),
                           dz.stride(0), alpha, _stream()), "vst_gelu_bwd")
    return dz


def layernorm_bwd(x2: torch.Tensor, dy2: torch.Tensor, gamma: Optional[torch.Tensor], eps: float,
                  want_params: bool = True):
    """x2, dy2: contiguous bf16 (rows, C) -> (dx bf16, dgamma fp32 | None, dbeta fp32 | None)."""
    _cuda(x2, dy2, gamma)
    rows, c = x2.shape
    assert x2.dtype == dy2.dtype == torch.bfloat16 and x2.is_contiguous() and dy2.is_contiguous()
    dx = torch.empty_like(x2)
    dg = torch.empty(c, dtype=torch.float32, device=x2.device) if want_params else None
    db = torch.empty(c, dtype=torch.float32, device=x2.device) if want_params else None
    wsb = lib.vst_layernorm_bwd_workspace_bytes(rows, c)
    ws = torch.empty(wsb, dtype=torch.uint8, device=x2.device)
    check(lib.vst_layernorm_bwd(x2.data_ptr(), dy2.data_ptr(), _ptr(gamma), dx.data_ptr(), _ptr(dg),
                                _ptr(db), rows, c, eps, _dt(gamma) if gamma is not None else VST_F32,
                                ws.data_ptr(), wsb, _stream()), "vst_layernorm_bwd")
    return dx, dg, db


def groupnorm5d_bwd(x_fm: torch.Tensor, dy_fm: torch.Tensor, mean, rstd, gamma, b: int, f: int,
                    groups: int, add: Optional[torch.Tensor] = None):
    """x_fm, dy_fm (and add): frames-major contiguous (B*F, C, H, W) bf16 -> (dx, dgamma, dbeta)."""
    _cuda(x_fm, dy_fm, mean, rstd, gamma, add)
    bf, c, h, w = x_fm.shape
    assert bf == b * f and x_fm.is_contiguous() and dy_fm.is_contiguous() and dy_fm.shape == x_fm.shape
    assert x_fm.dtype == dy_fm.dtype == torch.bfloat16
    dx = torch.empty_like(x_fm)
    dg = torch.empty(c, dtype=torch.float32, device=x_fm.device)
    db = torch.empty(c, dtype=torch.float32, device=x_fm.device)
    wsb = lib.vst_groupnorm5d_bwd_workspace_bytes(b, c, f, groups)
    ws = torch.empty(wsb, dtype=torch.uint8, device=x_fm.device)
    check(lib.vst_groupnorm5d_bwd(x_fm.data_ptr(), dy_fm.data_ptr(), _ptr(add), dx.data_ptr(),
                                  mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), dg.data_ptr(),
                                  db.data_ptr(), b, c, f, h * w, groups, _dt(gamma), ws.data_ptr(), wsb,
                                  _stream()), "vst_groupnorm5d_bwd")
    return dx, dg, db


def cfg_euler_step(eps_uncond: torch.Tensor, eps_cond: torch.Tensor, latents: torch.Tensor,
                   guidance: float, sigma: float, sigma_next: float, want_scaled_next: bool = False):
    """eps = eps_u + g (eps_c - eps_u); latents + (sigma_next - sigma) eps (Euler, epsilon
    prediction) and, optionally, the next step's scale_model_input in the eps dtype.
    inference_animatediff.py:122,131."""
    _cuda(eps_uncond, eps_cond, latents)
    assert eps_uncond.shape == eps_cond.shape == latents.shape and eps_uncond.dtype == eps_cond.dtype
    assert eps_uncond.is_contiguous() and eps_cond.is_contiguous() and latents.is_contiguous()
    out = torch.empty_like(latents)
    nxt = torch.empty_like(eps_uncond) if want_scaled_next else None
    check(lib.vst_cfg_euler_step(eps_uncond.data_ptr(), eps_cond.data_ptr(), latents.data_ptr(), out.data_ptr(),
                                 _ptr(nxt), latents.numel(), float(guidance), float(sigma), float(sigma_next),
                                 _dt(eps_uncond), _dt(latents), _stream()), "vst_cfg_euler_step")
    return (out, nxt) if want_scaled_next else out


def geglu_raw(hg: torch.Tensor) -> torch.Tensor:
    hg = hg.detach()
    _cuda(hg)
    assert hg.dtype == torch.bfloat16 and hg.dim() == 2 and hg.stride(1) == 1
    rows, two = hg.shape
    inner = two // 2
    out = torch.empty((rows, inner), dtype=torch.bfloat16, device=hg.device)
    check(lib.vst_geglu(hg.data_ptr(), out.data_ptr(), rows, inner, hg.stride(0), out.stride(0),
                        _stream()), "vst_geglu")
    return out
