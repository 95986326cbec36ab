"""Host half of the tcgen05 layer-1 kernel (csrc/spline_l1_tc5.cu): the K-order map and the block-scaled fp16 x 2 operand
split (ops.l1_pack16), emulated on the CPU with the kernel's arithmetic (fp16 parts, fp32 accumulation) against fp64."""
import numpy as np
import pytest
import torch

from eventad_b200 import ops


@pytest.mark.parametrize("cp", [16, 20])
def test_kmap_is_a_bijection_on_the_used_columns(cp):
    slot, ch = ops.l1_kmap(cp)
    used = ch < cp
    pairs = set(zip(slot[used].tolist(), ch[used].tolist()))
    assert len(pairs) == int(used.sum()) == 16 * cp
    assert pairs == {(s, c) for s in range(16) for c in range(cp)}
    if cp == 16:
        assert bool(used.all())


@pytest.mark.parametrize("cp", [16, 20])
def test_fp16x2_contraction_emulation_keeps_fp32_accuracy(cp):
    g = torch.Generator().manual_seed(cp)
    W = torch.randn(16 * cp, 16, generator=g) * 0.1
    S = torch.zeros((cp + 7) // 8 * 8, 16)
    S[:cp - 1] = torch.randn(cp - 1, 16, generator=g) * 0.2
    P = torch.randn(300, 16 * cp, generator=g) * torch.logspace(-3, 2, 300).view(-1, 1)     # rows of very different scale
    P[:, ::7] *= 1e-3
    pack = ops.l1_pack16(W, S, cp)
    slot, ch = ops.l1_kmap(cp)
    valid = ch < cp
    w16 = pack["w16"].view(torch.float16).float()                          # [2, 16, K]
    Pk = torch.where(valid.view(1, -1), P[:, (slot * cp + ch).clamp(max=16 * cp - 1)], torch.zeros(1))
    m = Pk.abs().amax(1)
    ex = (m.view(torch.int32) >> 23).clamp(20, 240)
    sc = ((267 - ex) << 23).view(torch.float32)
    rs = ((ex - 13) << 23).view(torch.float32)
    ps = Pk * sc.view(-1, 1)
    assert float(ps.abs().max()) < 65504
    p1 = ps.half()
    p2 = ((ps - p1.float()) * 2048.0).half()
    acc0 = p1.float() @ w16[0].t()
    acc1 = p1.float() @ w16[1].t() + p2.float() @ w16[0].t()
    out = (acc0 + acc1 * 2.0 ** -11) * (rs * pack["inv_sw"]).view(-1, 1)
    ref = P.double() @ W.double()
    err = (out.double() - ref).abs()
    bound = 2e-6 * (P.double().abs() @ W.double().abs())                   # relative to the magnitude of the terms
    assert bool((err <= bound + 1e-30).all()), float((err / (bound + 1e-30)).max())
    # skip Linear on the root slot (K chunks 3 and 4)
    s16 = pack["s16"].view(torch.float16).float()                          # [2, 16, 128]
    Pk_pad = torch.cat([Pk, torch.zeros(300, 320 - Pk.shape[1])], 1) if Pk.shape[1] < 320 else Pk
    ps_pad = torch.cat([ps, torch.zeros(300, 320 - ps.shape[1])], 1) if ps.shape[1] < 320 else ps
    q1 = ps_pad[:, 192:320].half()
    q2 = ((ps_pad[:, 192:320] - q1.float()) * 2048.0).half()
    sk = (q1.float() @ s16[0].t() + (q1.float() @ s16[1].t() + q2.float() @ s16[0].t()) * 2.0 ** -11) * \
        (rs * pack["inv_ss"]).view(-1, 1)
    root = P[:, 15 * cp:16 * cp]
    sref = root.double() @ S[:cp].double()
    serr = (sk.double() - sref).abs()
    sbound = 2e-6 * (root.double().abs() @ S[:cp].double().abs())
    assert bool((serr <= sbound + 1e-30).all()), float((serr / (sbound + 1e-30)).max())
