"""-m gpu: the streaming-softmax attention kernel (csrc/attn_stream.cu, default for unmasked attention)
against the fp32 oracle, against the exact running-max kernel it falls back to, and on adversarial
inputs where the first key block's maximum does not bound the later scores (the gated exact kernel
must take over)."""
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
bf = torch.bfloat16


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _oracle(q, k, v, heads):
    from oracle import hotpath as hp
    bt, sq, c = q.shape
    bk, sk, _ = k.shape
    d = c // heads
    sp = lambda t, b, s: t.float().reshape(b, s, heads, d).transpose(1, 2)
    kk, vv = sp(k, bk, sk), sp(v, bk, sk)
    if bk != bt:
        kk, vv = kk.repeat_interleave(bt // bk, 0), vv.repeat_interleave(bt // bk, 0)
    return hp.sdpa(sp(q, bt, sq), kk, vv).transpose(1, 2).reshape(bt, sq, c)


@pytest.mark.parametrize("bt,bk,heads,sq,sk", [(2, 2, 2, 128, 128), (2, 2, 3, 256, 384), (3, 3, 2, 200, 136),
                                               (4, 1, 2, 256, 77), (2, 2, 1, 64, 520), (1, 1, 10, 1024, 1024),
                                               (6, 2, 2, 130, 200), (1, 1, 1, 4096, 4096)])
def test_stream_kernel_vs_oracle_and_exact_kernel(bt, bk, heads, sq, sk):
    from video_style_transfer_b200 import ops
    torch.manual_seed(bt * 100 + sq + sk)
    c = heads * 64
    q, k, v = (torch.randn(b, s, c).to(bf) for b, s in ((bt, sq), (bk, sk), (bk, sk)))
    want = _oracle(q, k, v, heads)
    o, lse = ops.attn_fwd_raw(q.cuda(), k.cuda(), v.cuda(), heads, return_lse=True)
    assert rel(o, want) < 1e-2
    # log-sum-exp (saved for the backward) against the oracle's
    d = 64
    sp = lambda t, b, s: t.float().reshape(b, s, heads, d).transpose(1, 2)
    kk = sp(k, bk, sk).repeat_interleave(bt // bk, 0)
    s_ = sp(q, bt, sq) @ kk.transpose(-1, -2) / math.sqrt(d)
    assert (lse.cpu() - torch.logsumexp(s_, dim=-1)).abs().max() < 2e-2
    # the exact kernel (a zero additive mask selects it) agrees to bf16 rounding
    zero_mask = torch.zeros(1, 1, 1, sk, device="cuda")
    o_exact = ops.attn_fwd_raw(q.cuda(), k.cuda(), v.cuda(), heads, mask=zero_mask)
    assert rel(o, o_exact) < 8e-3


def test_stream_kernel_large_later_scores_stay_exact_without_fallback():
    """Scores that grow by ~2^40 after the first key block: p > 1 by design, nothing overflows, the
    result is still the softmax (reference-point invariance)."""
    from video_style_transfer_b200 import ops
    torch.manual_seed(3)
    heads, sq, sk = 2, 256, 512
    c = heads * 64
    q = torch.randn(2, sq, c).to(bf)
    k = torch.randn(2, sk, c) * 0.3
    k[:, 128:] *= 4.0                       # later blocks: scores ~ +-30 (natural log units)
    k, v = k.to(bf), torch.randn(2, sk, c).to(bf)
    want = _oracle(q, k, v, heads)
    o = ops.attn_fwd_raw(q.cuda(), k.cuda(), v.cuda(), heads)
    assert rel(o, want) < 1e-2


def test_stream_kernel_overflow_falls_back_to_exact_kernel():
    """First key block tiny, later keys huge (scores up to ~ +300): 2^(s - m_0) overflows fp32, the row
    sums fail the 2^64 check, the flag gates the exact kernel in, and the output is still correct."""
    from video_style_transfer_b200 import ops
    torch.manual_seed(4)
    heads, sq, sk = 2, 256, 640
    c = heads * 64
    q = torch.randn(2, sq, c).to(bf) * 2
    k = torch.randn(2, sk, c)
    k[:, :128] *= 1e-3
    k[:, 128:] *= 12.0
    k, v = k.to(bf), torch.randn(2, sk, c).to(bf)
    want = _oracle(q, k, v, heads)
    o, lse = ops.attn_fwd_raw(q.cuda(), k.cuda(), v.cuda(), heads, return_lse=True)
    assert torch.isfinite(o.float()).all() and torch.isfinite(lse).all()
    assert rel(o, want) < 1e-2
    # and the launches that follow are unaffected (the flag word is per launch)
    k2 = (torch.randn(2, sk, c) * 0.5).to(bf)
    assert rel(ops.attn_fwd_raw(q.cuda(), k2.cuda(), v.cuda(), heads), _oracle(q, k2, v, heads)) < 1e-2
