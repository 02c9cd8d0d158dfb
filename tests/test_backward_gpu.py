"""-m gpu parity tests of the backward kernels (vst_gemm = lora_gemm_bwd_dx /
_bwd_params, vst_attn_bwd, vst_temporal_attn_bwd, norm / GEGLU gradients), through the C ABI,
against fp32 torch-CPU autograd of the oracle functions on the same seeded inputs.
Tolerance: bf16 operands, fp32 accumulation -> max|a-b| / max|b| <= 1e-2 per tensor."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu
REL_TOL = 1e-2


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def ops():
    from video_style_transfer_b200 import ops as o
    return o


def _bf(t):
    return t.to(torch.bfloat16)


# ---- vst_gemm: every transpose combination, ragged edges, split-K, two operand pairs -------------
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (200, 136, 328), (1000, 264, 72), (136, 128, 16384),
                                   (64, 1288, 4104)])
def test_gemm_transposes(ops, ta, tb, m, n, k):
    if ta and m % 8:
        pytest.skip("MN-major A needs M % 8 == 0")
    if (not ta or not tb) and k % 8:
        pytest.skip("K-major operands need K % 8 == 0")
    torch.manual_seed(m + n + k + 2 * ta + tb)
    a = _bf(torch.randn(m, k))
    b = _bf(torch.randn(n, k))
    want = a.float() @ b.float().T
    ad = (a.T.contiguous() if ta else a).cuda()
    bd = (b.T.contiguous() if tb else b).cuda()
    got = ops.gemm(ad, bd, trans_a=bool(ta), trans_b=bool(tb), out_dtype=torch.float32)
    assert rel_err(got, want) < 2e-3      # fp32 output of bf16 operands: only summation order differs
    got16 = ops.gemm(ad, bd, trans_a=bool(ta), trans_b=bool(tb), alpha=0.5)
    assert rel_err(got16, 0.5 * want) < REL_TOL


def test_gemm_two_pairs_accumulate(ops):
    """dx = dy W + du A (transB = 1, two operand pairs) and D += (accumulate)."""
    torch.manual_seed(3)
    m, n, k1, k2 = 520, 640, 640, 128
    dy, w = _bf(torch.randn(m, k1)), _bf(torch.randn(k1, n))       # W stored [out=k1, in=n]
    du, a = _bf(torch.randn(m, k2)), _bf(torch.randn(k2, n))
    want = dy.float() @ w.float() + du.float() @ a.float()
    got = ops.gemm(dy.cuda(), w.cuda(), du.cuda(), a.cuda(), trans_b=True, out_dtype=torch.float32)
    assert rel_err(got, want) < 2e-3
    base = torch.randn(m, n)
    out = base.clone().cuda()
    ops.gemm(dy.cuda(), w.cuda(), du.cuda(), a.cuda(), trans_b=True, out=out, accumulate=True, alpha=2.0)
    assert rel_err(out, base + 2.0 * want) < 2e-3


def test_gemm_split_k_is_deterministic(ops):
    torch.manual_seed(4)
    tokens, n_out, r2 = 16384, 1280, 128
    dy, u = _bf(torch.randn(tokens, n_out)).cuda(), _bf(torch.randn(tokens, r2)).cuda()
    a = ops.gemm(dy, u, trans_a=True, trans_b=True, out_dtype=torch.float32)
    b = ops.gemm(dy, u, trans_a=True, trans_b=True, out_dtype=torch.float32)
    assert torch.equal(a, b)
    want = dy.float().T @ u.float()
    assert rel_err(a, want) < 2e-3
