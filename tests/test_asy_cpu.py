"""CPU: the C restatement of the reference's asy_tools kernels (oracle/asy_oracle.c) against plain torch arithmetic."""
import ctypes

import torch

from oracle import evg


def test_masked_lin_oracle_matches_torch():
    g = torch.Generator().manual_seed(0)
    n, cin, cout, k = 200, 19, 16, 40
    idx = torch.randperm(n, generator=g)[:k].contiguous()
    x, y, w, b = torch.randn(n, cin, generator=g), torch.randn(n, cout, generator=g), torch.randn(cout, cin, generator=g), torch.randn(cout, generator=g)
    for add in (0, 1):
        yc = y.clone()
        evg.asy_lib().asyo_masked_lin(idx.data_ptr(), x.data_ptr(), yc.data_ptr(), w.data_ptr(), b.data_ptr(), k, cin, cout, add)
        ref = y.clone()
        ref[idx] = (y[idx] if add else 0) + x[idx] @ w.t() + b
        assert torch.allclose(yc, ref, rtol=1e-5, atol=1e-5)


def test_masked_isdiff_and_bn_oracle():
    g = torch.Generator().manual_seed(1)
    n, c = 100, 8
    x = torch.randn(n, c, generator=g)
    x2 = x.clone()
    x2[::2] += 0.1
    idx = torch.arange(n)
    evg.asy_lib().asyo_masked_isdiff(idx.data_ptr(), x2.data_ptr(), x.data_ptr(), n, c, ctypes.c_float(1e-8), ctypes.c_float(1e-5))
    # main.cu:36 has no abs() around `other`: an unchanged row with a negative entry has 0 > atol + rtol * other and is
    # reported as different -- reference behaviour, kept
    expect = ((x2 - x).abs() > 1e-8 + 1e-5 * x).any(1).nonzero().view(-1)
    assert idx[idx > -1].tolist() == expect.tolist() and len(expect) > n // 2
    bn = torch.nn.BatchNorm1d(c).eval()
    bn.running_mean.normal_(generator=g)
    bn.running_var.uniform_(0.5, 1.5, generator=g)
    out = torch.zeros(n, c)
    sel = torch.arange(0, n, 3)
    evg.asy_lib().asyo_masked_inplace_bn(sel.data_ptr(), x.data_ptr(), out.data_ptr(), bn.running_mean.data_ptr(),
                                         bn.running_var.data_ptr(), bn.weight.detach().data_ptr(), bn.bias.detach().data_ptr(),
                                         len(sel), c, ctypes.c_float(bn.eps))
    assert torch.allclose(out[sel], bn(x)[sel].detach(), rtol=1e-5, atol=1e-6)
    assert float(out[1].abs().sum()) == 0
