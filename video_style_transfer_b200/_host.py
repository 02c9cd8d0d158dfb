"""Host-side glue shared by the drop-in modules: cached bf16 operand copies and the one fused
projection call every linear on the hot path goes through."""
import weakref
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import ops

_cache: Dict[Tuple, Tuple] = {}
_inserts = 0


def _alive(refs):
    return all(r() is not None for r in refs)


def _ident(t: torch.Tensor):
    """Cache identity of an operand.  A tensor DERIVED by cached_build (packed up-projection, folded
    weight ...) carries the key it was built under: when its sources change it is rebuilt as a new
    tensor object, and anything derived from it in turn must overwrite its own entry instead of
    leaving the old one behind (round-1 advisor finding: one folded W_eff leaked per rebuild)."""
    return getattr(t, "_vst_key", None) or id(t)


def _stamp(t: torch.Tensor):
    return getattr(t, "_vst_stamp", None) or (t._version, t.data_ptr())


def _store(key, tensors, stamp, val):
    global _inserts
    _inserts += 1
    if _inserts % 256 == 0:                      # drop entries whose sources are gone or replaced
        for k in [k for k, v in _cache.items() if not _alive(v[0])]:
            del _cache[k]
    _cache[key] = (tuple(weakref.ref(t) for t in tensors), stamp, val)


def invalidate_caches():
    """Forget every cached operand (bf16 copies, packed B', folded W_eff, transposes).  Call after
    writing parameters through ``.data`` on the device (diffusers ``EMAModel.copy_to/restore``,
    ``p.data.copy_()``): such writes do not bump ``tensor._version``, which is what the caches key
    on.  ``load_state_dict``, optimiser steps and in-place ops under ``no_grad`` need no call."""
    _cache.clear()


def cache_stats():
    return {"entries": len(_cache),
            "bytes": sum(v[2].numel() * v[2].element_size() for v in _cache.values()
                         if isinstance(v[2], torch.Tensor))}


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def pad_branch_cols(bp: torch.Tensor, nbr: int, r: int) -> torch.Tensor:
    """(N, nbr*r) packed up-projections -> (N, nbr*pad8(r)), zero columns appended per branch
    (ranks that are not multiples of 8, e.g. the rank-4 default of LoRALinearLayer)."""
    if r % 8 == 0:
        return bp
    n = bp.shape[0]
    return torch.nn.functional.pad(bp.view(n, nbr, r), (0, pad8(r) - r)).reshape(n, nbr * pad8(r))


def bf16_cached(*tensors: torch.Tensor, pad_rows: bool = False) -> torch.Tensor:
    """bf16, contiguous copy of a parameter (or the row-wise concatenation of several), cached
    until any of them is modified in place (tensor._version) or re-allocated (data_ptr).
    pad_rows: zero-pad each tensor's row count to a multiple of 8 first (LoRA down matrices)."""
    pad_rows = pad_rows and any(t.shape[0] % 8 for t in tensors)
    track = torch.is_grad_enabled() and any(t.requires_grad for t in tensors)
    if len(tensors) == 1 and not pad_rows:
        t = tensors[0]
        if t.dtype == torch.bfloat16 and t.is_contiguous():
            # the parameter itself, never a fresh `.detach()` alias: caches of operands derived from it
            # (transposes, folded weights) key on the tensor object's identity, and a new alias per call
            # made them miss every time (nothing is recorded for a frozen tensor or under no_grad anyway)
            return t
    if track:
        from ._autograd import CastCatFn
        with torch.no_grad():
            val = bf16_cached(*tensors, pad_rows=pad_rows)
        return CastCatFn.apply(pad_rows, val, *tensors)
    key = ("bf16",) + tuple(_ident(t) for t in tensors) + (pad_rows,)
    stamp = tuple(_stamp(t) for t in tensors)
    hit = _cache.get(key)
    if hit is not None and hit[1] == stamp and _alive(hit[0]):
        return hit[2]
    with torch.no_grad():
        parts = [t.detach().to(torch.bfloat16) for t in tensors]
        if pad_rows:
            parts = [torch.nn.functional.pad(q, (0, 0, 0, pad8(q.shape[0]) - q.shape[0])) for q in parts]
        val = parts[0].contiguous() if len(parts) == 1 else torch.cat(parts, dim=0).contiguous()
    if len(tensors) > 1 or pad_rows or tensors[0].dtype != torch.bfloat16:
        val._vst_key, val._vst_stamp = key, stamp
    _store(key, tensors, stamp, val)
    return val


def as_rows_bf16(x: torch.Tensor) -> torch.Tensor:
    """(..., K) activation -> contiguous 2-D bf16 (rows, K)."""
    x2 = x.reshape(-1, x.shape[-1])
    if x2.dtype != torch.bfloat16:
        x2 = x2.to(torch.bfloat16)
    if x2.stride(1) != 1 or (x2.stride(0) % 8) != 0:
        x2 = x2.contiguous()
    return x2


def fused_linear(x2: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor],
                 low_rank: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
                 residual: Optional[torch.Tensor] = None, out_scale: float = 1.0,
                 act: str = "none", out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """y = (act(x W^T + u B'^T + bias) + residual) * out_scale on the tcgen05 GEMM.
    low_rank = (u (M,R) bf16, B' (N,R) bf16) or None."""
    w = bf16_cached(weight)
    b = bias
    if b is not None and not (torch.is_grad_enabled() and b.requires_grad):
        b = b.detach()
    if b is not None and b.dtype not in (torch.float32, torch.bfloat16):
        b = b.float()
    u, bp = low_rank if low_rank is not None else (None, None)
    return ops.lora_gemm(x2, w, u, bp, bias=b, residual=residual, out_scale=out_scale, act=act,
                         out_dtype=out_dtype)


def cached_build(tensors: Sequence[torch.Tensor], extra, builder):
    """Memoise ``builder()`` (a derived operand such as the packed, merger-scaled up-projection)
    on the identity + in-place version of ``tensors`` and on ``extra`` (scale, forward type ...)."""
    key = ("build",) + tuple(_ident(t) for t in tensors) + (extra,)
    stamp = tuple(_stamp(t) for t in tensors)
    hit = _cache.get(key)
    if hit is not None and hit[1] == stamp and _alive(hit[0]):
        return hit[2]
    with torch.no_grad():
        val = builder()
    if isinstance(val, torch.Tensor):
        val._vst_key, val._vst_stamp = key, stamp      # identity for operands derived from this one
    _store(key, tensors, stamp, val)
    return val


def packed_up(b_c, m_c, mask_c, b_s, m_s, mask_s, scale: float, rank: int, extra=()) -> torch.Tensor:
    """B' = scale * [m_c . mask_c . B_c | m_s . mask_s . B_s] as a bf16 (N, nbr*pad8(rank)) GEMM operand
    (vst_lora_pack_up), cached on the parameter versions; differentiable w.r.t. B_* and m_* when they
    require grad (vst_lora_pack_up_bwd)."""
    ref = b_c if b_c is not None else b_s
    nbr = (b_c is not None) + (b_s is not None)
    deps = [t for t in (b_c, m_c, mask_c, b_s, m_s, mask_s) if t is not None]
    same = lambda t: None if t is None else t.detach().to(device=ref.device, dtype=ref.dtype).contiguous()
    dev = lambda t: None if t is None else t.to(ref.device)

    def build():
        bp = ops.lora_pack_up(None if b_c is None else b_c.detach().contiguous(), same(m_c), dev(mask_c),
                              None if b_s is None else b_s.detach().contiguous(), same(m_s), dev(mask_s),
                              float(scale))
        return pad_branch_cols(bp, nbr, rank)

    sig = ("packup", float(scale), b_c is not None, m_c is not None, mask_c is not None,
           b_s is not None, m_s is not None, mask_s is not None) + tuple(extra)
    val = cached_build(deps, sig, build)
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in (b_c, m_c, b_s, m_s)):
        from ._autograd import PackUpFn
        return PackUpFn.apply(val, float(scale), rank, dev(mask_c), dev(mask_s), b_c, m_c, b_s, m_s)
    return val
