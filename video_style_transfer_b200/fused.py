"""Projections that read the same activation, fused along N into one tcgen05 GEMM.

to_q | to_k | to_v of a self-attention (or to_k | to_v of a cross-attention, or the q/k/v
TemporalLoRALinear of a motion-module attention) share their input, so
    [q | k | v] = x [W_q; W_k; W_v]^T + u B'^T,   u = x [A_q; A_k; A_v]^T
is two launches instead of nine, x is read twice instead of six times, and the 3C-wide output
tiles the 148 SMs with far less wave-quantisation loss than three C-wide GEMMs.  The low-rank
term is *grouped* (vst_lora_gemm k2_group_n): output columns of projection g only read their own
K2 columns of u, so no block-diagonal zero padding is multiplied.

Inference with frozen adapters (no parameter of the site requires grad) goes one step further:
the low-rank update is folded into the weight once, W_eff = W + B' A (``merged_weight``, built by
vst_gemm and cached on the parameters' versions, i.e. rebuilt after an optimiser step, a merger
clamp, a mask or a forward-type change), and the projection is ONE plain GEMM: no u = x A^T
pre-pass, no extra K blocks.  The reference materialises A^T B^T on every call
(unziplora_linear_layer.py:311-322) and itself ships ``merged_weight`` for the temporal LoRA
(temporal_lora.py:35-41).  Training (any adapter tensor requires grad) and the three-stream
Stage-1 form (lora_unzip.py:66-75, low-rank branches read other activations) keep the fused
low-rank GEMM.  VST_MERGE_LORA=0 (or ``fused.MERGE_LORA = False``) disables the folding.

Every linear taking part implements the "LoRA site" protocol: site_weight(), site_bias(),
site_downs() -> [A_i (r, in)], site_up(scale) -> B' (out, K2) or None.  Concatenated / packed
operands are cached on the parameters' in-place versions (``_host.bf16_cached/cached_build``), so
an optimiser step or a merger clamp invalidates them.
"""
import os
from typing import List, Optional, Sequence

import torch

from . import ops
from ._host import bf16_cached, cached_build, fused_linear, pad8


MERGE_LORA = os.environ.get("VST_MERGE_LORA", "1") != "0"
# The low-rank form in ONE launch (u = x A^T computed inside the main kernel, vst_lora_gemm_fused).
#   1 (default): projections that stand alone (to_q of a cross-attention, to_out without residual): measured
#                103 vs 107 us (L2 to_q, M = 32768) and 135 vs 152 us (L1) against u pre-pass + extra K blocks;
#   2:           also projections fused along N (q|k|v): measured 2-6 % SLOWER there (265 vs 251 us at L2: the
#                u item is recomputed per (256-row tile, group) and a CTA pair's contiguous tile range cuts units);
#   0:           never (two launches everywhere).
FUSED_LORA = int(os.environ.get("VST_FUSED_LORA", "1"))


def _fit_k2(a: torch.Tensor, bp: torch.Tensor, groups: int):
    """Pad the per-group rank of (a (G*k2, in), bp (N, k2)) with zero rows / columns to the 64 or 128 the
    single-launch kernel takes (temporal LoRA: 32 -> 64).  None when the rank does not fit."""
    k2 = bp.shape[1]
    if k2 in (64, 128):
        return a, bp
    k2p = 64 if k2 < 64 else (128 if k2 < 128 else 0)
    if not k2p:
        return None

    def build():
        ap = torch.nn.functional.pad(a.view(groups, k2, a.shape[1]), (0, 0, 0, k2p - k2)).reshape(groups * k2p, a.shape[1])
        return ap.contiguous(), torch.nn.functional.pad(bp, (0, k2p - k2)).contiguous()

    if torch.is_grad_enabled() and (a.requires_grad or bp.requires_grad):
        return build()
    return cached_build([a, bp], ("fit_k2", groups), build)


def _fused_low_rank(x2, weight, bias, a, bp, k2_group_n: int, out_scale: float):
    """The projection in one launch, or None when the shape is not a fused one / the form is switched off."""
    if FUSED_LORA < (2 if k2_group_n else 1):
        return None
    n, groups = bp.shape[0], (bp.shape[0] // k2_group_n if k2_group_n else 1)
    fit = _fit_k2(a, bp, groups)
    if fit is None:
        return None
    a_f, bp_f = fit
    if not ops.lora_gemm_fused_supported(x2.shape[0], n, x2.shape[1], bp_f.shape[1], k2_group_n):
        return None
    b = bias
    if b is not None and not (torch.is_grad_enabled() and b.requires_grad):
        b = b.detach()
    if b is not None and b.dtype not in (torch.float32, torch.bfloat16):
        b = b.float()
    return ops.lora_gemm_fused(x2, bf16_cached(weight), a_f, bp_f, b, out_scale, k2_group_n)


def _gemm_bias(b):
    if b is not None and not (torch.is_grad_enabled() and b.requires_grad):
        b = b.detach()
    if b is not None and b.dtype not in (torch.float32, torch.bfloat16):
        b = b.float()
    return b


def _frozen(tensors) -> bool:
    return not (torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors))


def merged_weight(ws: Sequence[torch.Tensor], downs: Sequence[Sequence[torch.Tensor]],
                  ups: Sequence[Optional[torch.Tensor]]) -> torch.Tensor:
    """bf16 [W_0 + B'_0 A_0 ; W_1 + B'_1 A_1 ; ...] (sum out, in): one vst_gemm per site
    accumulating the rank-K2 product onto a copy of the cached bf16 weight, memoised on every
    tensor involved.  ups[g] is the packed, merger- and scale-folded up-projection of site g."""
    flat = [a for d in downs for a in d]
    deps = list(ws) + flat + [u for u in ups if u is not None]

    def build():
        # W + B'A is accumulated in fp32 from the parameters themselves and rounded to bf16 ONCE
        # (a delta below one bf16 ulp of W still moves the rounding the right way on average;
        # folding onto an already-rounded bf16 W would round twice)
        w32 = torch.cat([w_.detach().float() for w_ in ws], dim=0)
        row = 0
        for wg, dg, ug in zip(ws, downs, ups):
            n = wg.shape[0]
            if ug is not None and dg:
                a = bf16_cached(*dg, pad_rows=True)                     # (K2, in)
                assert a.shape[0] == ug.shape[1]
                ops.gemm(ug.detach(), a, trans_b=True, out=w32[row:row + n], accumulate=True)
            row += n
        return w32.to(torch.bfloat16)

    return cached_build(deps, "merged_weight", build)


def _site(layer):
    if hasattr(layer, "site_weight"):
        return layer
    if isinstance(layer, torch.nn.Linear):
        return _PlainSite(layer)
    raise TypeError(f"{type(layer).__name__} is not a linear the sm_100a path knows how to run")


class _PlainSite:
    """nn.Linear without an adapter."""

    def __init__(self, lin):
        self.lin = lin
        self.out_features = lin.out_features

    def site_weight(self):
        return self.lin.weight

    def site_bias(self):
        return self.lin.bias

    def site_downs(self):
        return []

    def site_up(self, scale=1.0):
        return None


def project_one(layer, x2: torch.Tensor, scale: float = 1.0, residual=None, out_scale: float = 1.0,
                act: str = "none", out_dtype=torch.bfloat16) -> torch.Tensor:
    s = _site(layer)
    downs = s.site_downs()
    low = None
    if downs:
        bp = s.site_up(scale)
        if bp is not None:
            if MERGE_LORA and _frozen([s.site_weight(), bp] + list(downs)):
                w_eff = merged_weight([s.site_weight()], [downs], [bp])
                return fused_linear(x2, w_eff, s.site_bias(), None, residual, out_scale, act, out_dtype)
            a = bf16_cached(*downs, pad_rows=True)
            if residual is None and act == "none" and out_dtype == torch.bfloat16:
                y = _fused_low_rank(x2, s.site_weight(), s.site_bias(), a, bp, 0, out_scale)
                if y is not None:
                    return y
            if act == "none" and out_dtype == torch.bfloat16:
                return ops.lora_linear_two_launch(x2, bf16_cached(s.site_weight()), a, bp, _gemm_bias(s.site_bias()),
                                                  residual, out_scale, 0)
            low = (ops.lora_gemm(x2, a), bp)
    return fused_linear(x2, s.site_weight(), s.site_bias(), low, residual, out_scale, act, out_dtype)


def project_fused(layers: Sequence, x2: torch.Tensor, scale: float = 1.0, batch=None) -> List[torch.Tensor]:
    """[layer(x) for layer in layers] as column views of one (M, sum out) GEMM result when the
    layers are compatible; otherwise one fused GEMM per layer.  batch = (b, s) returns the views as
    (b, s, out) with the packed row stride (what the attention kernels take)."""
    sites = [_site(l) for l in layers]
    ws = [s.site_weight() for s in sites]
    n0 = ws[0].shape[0]
    biases = [s.site_bias() for s in sites]
    downs = [s.site_downs() for s in sites]
    ups = [s.site_up(scale) if d else None for s, d in zip(sites, downs)]
    k2s = {None if u is None else u.shape[1] for u in ups}
    same = (len(sites) > 1 and all(w.shape == ws[0].shape for w in ws) and n0 % 64 == 0
            and len(k2s) == 1 and (all(b is None for b in biases) or all(b is not None for b in biases)))
    if not same:
        ys = [project_one(l, x2, scale) for l in layers]
        return ys if batch is None else [y.view(batch[0], batch[1], y.shape[1]) for y in ys]
    k2_any = next(iter(k2s))
    merge = (k2_any is not None and MERGE_LORA
             and _frozen(list(ws) + [a for d in downs for a in d] + [u for u in ups if u is not None]))
    w = merged_weight(ws, downs, ups) if merge else bf16_cached(*ws)
    bias = None
    if biases[0] is not None:
        if torch.is_grad_enabled() and any(b.requires_grad for b in biases):
            bias = torch.cat([b.float() for b in biases])
        else:
            bias = cached_build(list(biases), "catbias", lambda: torch.cat([b.detach().float() for b in biases]))
    k2 = k2s.pop()
    if k2 is None or merge:
        y = ops.lora_gemm(x2, w, bias=bias)
    else:
        flat = [a for d in downs for a in d]
        a_cat = bf16_cached(*flat, pad_rows=True)                       # (G*K2, in)
        assert a_cat.shape[0] == len(sites) * k2
        if torch.is_grad_enabled() and any(u_.requires_grad for u_ in ups):
            bp = torch.cat(list(ups), dim=0)
        else:
            bp = cached_build(list(ups), "catup", lambda: torch.cat(list(ups), dim=0).contiguous())
        y = _fused_low_rank(x2, w, bias, a_cat, bp, n0, 1.0)            # one launch (VST_FUSED_LORA=2)
        if y is None:                                                   # u = x a_cat^T (M, G*K2), then the GEMM
            y = ops.lora_linear_two_launch(x2, w, a_cat, bp, bias, None, 1.0, n0)
    from ._autograd import split_cols
    return split_cols(y, n0, len(sites), batch)
