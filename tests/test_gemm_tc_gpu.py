"""Tensor-core contraction (csrc/gemm_tc.cu, 3xTF32) against a float64 matmul: the split products must keep fp32
accuracy (north_star tolerance rtol 1e-4 / atol 1e-5 is for the features; here the bar is an order tighter), for the
K / M shapes of the pooled layers (K = 26*Cin, Cin in {64, 82, 130}; skip Linear K = 82 / 130 with row padding), ragged
M, device-side row counts and the split-K path."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(A, W, k, scale, shift, addend, relu):
    r = A[:, :k].double() @ W.double()
    if scale is not None:
        r = r * scale.double()
    if shift is not None:
        r = r + shift.double()
    if addend is not None:
        r = r + addend.double()
    if relu:
        r = r.clamp(min=0)
    return r


@pytest.mark.parametrize("m,k,lda", [(2240, 2132, 2132), (17920, 1664, 1664), (560, 3380, 3380), (35, 1664, 1664),
                                       (1, 64, 64), (4481, 82, 84), (141, 130, 132), (9500, 96, 96), (300, 31, 32)])
def test_gemm64_tc_matches_fp64(m, k, lda):
    from eventad_b200 import ops
    g = torch.Generator().manual_seed(m + k)
    A = torch.randn(m, lda, generator=g)
    A[:, k:] = 7.0                                              # padding columns must not leak into the product
    W = torch.randn(k, 64, generator=g) / k ** 0.5
    scale, shift = torch.rand(64, generator=g) + 0.5, torch.randn(64, generator=g)
    addend = torch.randn(m, 64, generator=g)
    whi, wlo = ops.split_tf32(W.cuda())
    assert (whi + wlo - W.cuda()).abs().max() < 1e-6
    Wd = W.cuda()
    for epi, kw in ((dict(), dict(scale=None, shift=None, addend=None, relu=False)),
                    (dict(scale=scale.cuda(), shift=shift.cuda(), addend=addend.cuda(), relu=True),
                     dict(scale=scale, shift=shift, addend=addend, relu=True))):
        got = ops.gemm64_tc(A.cuda(), Wd, m, None, epi, k=k)
        ref = _ref(A, W, k, **kw)
        err = (got.cpu().double() - ref).abs()
        assert float((err / (2e-6 + 2e-5 * ref.abs())).max()) < 1.0, float(err.max())


def test_gemm64_tc_device_row_count_and_determinism():
    from eventad_b200 import ops
    g = torch.Generator().manual_seed(5)
    m_cap, m_live, k = 2600, 2241, 2132
    A = torch.randn(m_cap, k, generator=g).cuda()
    W = (torch.randn(k, 64, generator=g) / k ** 0.5).cuda()
    m_dev = torch.tensor([m_live], dtype=torch.int32, device="cuda")
    out = torch.full((m_cap, 64), -3.0, device="cuda")
    ops.gemm64_tc(A, W, m_cap, m_dev, dict(), out=out)
    ref = A[:m_live].double() @ W.double()
    assert float((out[:m_live].double() - ref).abs().max()) < 2e-5
    assert torch.all(out[m_live:] == -3.0)                      # rows beyond the device count are not touched
    out2 = torch.empty_like(out)
    ops.gemm64_tc(A, W, m_cap, m_dev, dict(), out=out2)
    assert torch.equal(out[:m_live], out2[:m_live])             # split-K partials are summed in a fixed order
