"""tcgen05 contraction (csrc/gemm_tc5.cu: UMMA kind::tf32, TMEM accumulator, 3xTF32 split) against a float64 matmul, on
the K / M shapes of the pooled layers, ragged M and K, device-side row counts and the split-K path."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,k,lda", [(128, 32, 32), (128, 64, 64), (256, 96, 96), (2240, 2132, 2132), (17920, 1664, 1664),
                                       (560, 3380, 3380), (35, 1664, 1664), (1, 64, 64), (4481, 82, 84), (141, 130, 132),
                                       (300, 31, 32)])
def test_gemm64_tc5_matches_fp64(m, k, lda):
    from eventad_b200 import ops
    g = torch.Generator().manual_seed(m + k)
    A = torch.randn(m, lda, generator=g)
    A[:, k:] = 7.0                                              # padding columns must not leak into the product
    W = torch.randn(k, 64, generator=g) / k ** 0.5
    scale, shift = torch.rand(64, generator=g) + 0.5, torch.randn(64, generator=g)
    addend = torch.randn(m, 64, generator=g)
    wt = ops.transpose_pad_weight(W.cuda())
    for epi, ref_fn in ((dict(), lambda r: r),
                        (dict(scale=scale.cuda(), shift=shift.cuda(), addend=addend.cuda(), relu=True),
                         lambda r: (r * scale.double() + shift.double() + addend.double()).clamp(min=0))):
        got = ops.gemm64_tc5(A.cuda(), wt, m, None, epi, k=k)
        ref = ref_fn(A[:, :k].double() @ W.double())
        err = (got.cpu().double() - ref).abs()
        assert float((err / (2e-6 + 2e-5 * ref.abs())).max()) < 1.0, float(err.max())


def test_gemm64_tc5_device_row_count_and_determinism():
    from eventad_b200 import ops
    g = torch.Generator().manual_seed(5)
    m_cap, m_live, k = 2600, 2241, 2132
    A = torch.randn(m_cap, k, generator=g).cuda()
    W = (torch.randn(k, 64, generator=g) / k ** 0.5).cuda()
    wt = ops.transpose_pad_weight(W)
    m_dev = torch.tensor([m_live], dtype=torch.int32, device="cuda")
    out = torch.full((m_cap, 64), -3.0, device="cuda")
    ops.gemm64_tc5(A, wt, m_cap, m_dev, dict(), out=out)
    ref = A[:m_live].double() @ W.double()
    assert float((out[:m_live].double() - ref).abs().max()) < 2e-5
    assert torch.all(out[m_live:] == -3.0)
    out2 = torch.empty_like(out)
    ops.gemm64_tc5(A, wt, m_cap, m_dev, dict(), out=out2)
    assert torch.equal(out[:m_live], out2[:m_live])
