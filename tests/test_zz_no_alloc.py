"""-m gpu: vst_colsum_noalloc, the allocation-free variant of vst_colsum (scratch in a static device array of
VST_COLSUM_NOALLOC_SCRATCH_BYTES inside libvst_b200.so; no counterpart in the reference).
Called raw through the C ABI and compared with the existing entry point, two calls in a row
(the array keeps the previous call's contents), plus the refusal of a shape beyond the maximum.

This file is named so that pytest collects it LAST: a failure here must not hide
any other test.  It passes alone in a fresh process and after the rest of the suite."""
import pytest
import torch

pytestmark = pytest.mark.gpu
SCRATCH_BYTES = 1 << 20           # VST_COLSUM_NOALLOC_SCRATCH_BYTES (include/vst_b200.h)
REL_TOL = 1e-2                    # the suite's bf16 tolerance: max|a-b| / max|b|


def _rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _colsum_pair(lib, check, x, alpha):
    """(existing entry point with a caller workspace, new entry point) on the same input."""
    m, n = x.shape
    st = torch.cuda.current_stream().cuda_stream
    wsb = lib.vst_colsum_workspace_bytes(m, n)
    assert 0 < wsb <= SCRATCH_BYTES
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    want = torch.empty(n, dtype=torch.float32, device="cuda")
    check(lib.vst_colsum(x.data_ptr(), x.stride(0), m, n, alpha, want.data_ptr(), ws.data_ptr(), wsb, st),
          "vst_colsum")
    got = torch.full((n,), float("nan"), dtype=torch.float32, device="cuda")
    check(lib.vst_colsum_noalloc(x.data_ptr(), x.stride(0), m, n, alpha, got.data_ptr(), st), "vst_colsum_noalloc")
    torch.cuda.synchronize()
    return want, got


def test_colsum_noalloc_matches_colsum_two_calls_in_a_row():
    from video_style_transfer_b200 import _lib
    lib, check = _lib.lib, _lib.check
    torch.manual_seed(0)
    # call 1: the FF bias gradient of an L2 motion module (N = 8C = 10240); call 2: a smaller, ragged shape that
    # reuses only part of the array the first call filled (stale partials must not leak into the result)
    for m, n, alpha in ((16384, 10240, 1.0), (1000, 328, 0.5)):
        x = torch.randn(m, n, device="cuda").to(torch.bfloat16)
        want, got = _colsum_pair(lib, check, x, alpha)
        ref = alpha * x.float().sum(dim=0)
        assert torch.isfinite(got).all()
        assert _rel(got, want) < REL_TOL            # same kernels, same order: in fact bit-identical
        assert torch.equal(got, want)
        assert _rel(got, ref) < REL_TOL
    # a strided view (row stride != N), as the autograd glue passes column slices of packed gradients
    wide = torch.randn(4096, 1920, device="cuda").to(torch.bfloat16)
    want, got = _colsum_pair(lib, check, wide[:, 640:1280], 2.0)
    assert torch.equal(got, want)


def test_colsum_noalloc_refuses_shapes_beyond_its_scratch():
    from video_style_transfer_b200 import _lib
    lib = _lib.lib
    m, n = 64, 300000                                  # one chunk x 300000 columns x 4 bytes = 1.2 MB > 1 MiB
    assert lib.vst_colsum_workspace_bytes(m, n) > SCRATCH_BYTES
    x = torch.zeros(m, n, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(n, dtype=torch.float32, device="cuda")
    rc = lib.vst_colsum_noalloc(x.data_ptr(), x.stride(0), m, n, 1.0, out.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
    assert rc == -2                                    # VST_E_UNSUPPORTED
    msg = lib.vst_last_error().decode()
    assert "scratch" in msg and "vst_colsum" in msg
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0               # nothing was launched
