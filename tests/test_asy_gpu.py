"""Masked row-update kernels of the asynchronous layers (csrc/asy.cu) against
  * the UNMODIFIED reference extension oracle/_ref/asy_tools.so (= /root/reference/src/dagr/asynchronous/asy_tools/main.cu
    compiled as-is for sm_100a by oracle/build_ref.py), run on the same B200, and
  * the C restatement oracle/asy_oracle.c (which this also pins to the reference),
bit for bit: same selected rows, same fp32 results."""
import ctypes
import importlib.util
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REF_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "asy_tools.so")


def _ref():
    assert os.path.exists(REF_SO), "oracle/_ref/asy_tools.so missing (python oracle/build_ref.py)"
    spec = importlib.util.spec_from_file_location("asy_tools", REF_SO)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _inputs(n, cin, cout, k, seed):
    g = torch.Generator().manual_seed(seed)
    idx = torch.randperm(n, generator=g)[:k].contiguous()
    return (idx, torch.randn(n, cin, generator=g), torch.randn(n, cout, generator=g), torch.randn(cout, cin, generator=g) * 0.2,
            torch.randn(cout, generator=g))


@pytest.mark.parametrize("n,cin,cout,k", [(1000, 19, 16, 37), (5000, 130, 64, 800), (300, 64, 2, 300), (50, 1, 1, 1)])
@pytest.mark.parametrize("add", [False, True])
@pytest.mark.parametrize("with_bias", [False, True])
def test_masked_lin_bit_exact(n, cin, cout, k, add, with_bias):
    from eventad_b200.asynchronous import asy_tools as ours
    from oracle import evg
    ref = _ref()
    idx, x, y, w, b = _inputs(n, cin, cout, k, seed=cin)
    outs = []
    for mod in (ours, ref):
        yo = y.clone().cuda()
        if with_bias:
            mod.masked_lin(idx.cuda(), x.cuda(), yo, w.cuda(), b.cuda(), add)
        else:
            mod.masked_lin_no_bias(idx.cuda(), x.cuda(), yo, w.cuda(), add)
        outs.append(yo.cpu())
    yc = y.clone()
    evg.asy_lib().asyo_masked_lin(idx.data_ptr(), x.data_ptr(), yc.data_ptr(), w.data_ptr(), b.data_ptr() if with_bias else None,
                                  k, cin, cout, int(add))
    assert torch.equal(outs[0], outs[1]), "kernel vs unmodified reference kernel"
    assert torch.equal(yc, outs[1]), "C oracle vs unmodified reference kernel"
    untouched = torch.ones(n, dtype=torch.bool)
    untouched[idx] = False
    assert torch.equal(outs[0][untouched], y[untouched])


@pytest.mark.parametrize("n,c,k", [(2000, 16, 500), (700, 64, 700), (10, 3, 1)])
def test_masked_bn_and_isdiff_bit_exact(n, c, k):
    from eventad_b200.asynchronous import asy_tools as ours
    from oracle import evg
    ref = _ref()
    g = torch.Generator().manual_seed(c)
    idx = torch.randperm(n, generator=g)[:k].contiguous()
    x = torch.randn(n, c, generator=g)
    mean, var = torch.randn(c, generator=g) * 0.1, torch.rand(c, generator=g) + 0.5
    wt, bs = torch.rand(c, generator=g) + 0.5, torch.randn(c, generator=g)
    outs = []
    for mod in (ours, ref):
        yo = torch.zeros(n, c).cuda()
        mod.masked_inplace_BN(idx.cuda(), x.cuda(), yo, mean.cuda(), var.cuda(), wt.cuda(), bs.cuda(), 1e-5)
        outs.append(yo.cpu())
    yc = torch.zeros(n, c)
    evg.asy_lib().asyo_masked_inplace_bn(idx.data_ptr(), x.data_ptr(), yc.data_ptr(), mean.data_ptr(), var.data_ptr(),
                                         wt.data_ptr(), bs.data_ptr(), k, c, ctypes.c_float(1e-5))
    assert torch.equal(outs[0], outs[1]) and torch.equal(yc, outs[1])
    # isdiff: a third of the rows perturbed beyond / below the tolerance
    x2 = x.clone()
    x2[idx[::3]] += 1e-3
    x2[idx[1::3]] *= (1 + 1e-7)
    res = []
    for mod in (ours, ref):
        res.append(mod.masked_isdiff(idx.clone().cuda(), x2.cuda(), x.cuda(), 1e-8, 1e-5).cpu())
    ic = idx.clone()
    evg.asy_lib().asyo_masked_isdiff(ic.data_ptr(), x2.data_ptr(), x.data_ptr(), k, c, ctypes.c_float(1e-8), ctypes.c_float(1e-5))
    assert torch.equal(res[0], res[1]) and torch.equal(ic[ic > -1], res[1])
    assert set(idx[::3].tolist()) <= set(res[0].tolist())
