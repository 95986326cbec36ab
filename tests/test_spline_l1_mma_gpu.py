"""evad_spline_conv_l1_mma (tap aggregation on the tensor cores) against an fp64 evaluation of the same sum on the CPU and
against evad_spline_conv_l1 (FFMA aggregation), on synthetic fixed-stride rows: ragged row lengths (0..16), node counts
that are not a multiple of the 32-node tile, both channel widths, with and without the skip Linear, with and without a
processing order.  Semantics: spline_conv.py:39-47,72-76 (LUT message, sum aggregation, root weight) + conv.py:23-28,47-56
(eval BatchNorm folded to scale/shift, ReLU, skip Linear).  Bars: 2e-6 + 2e-5*|ref| + 1e-6*sum|terms| against fp64
(outputs are 320-term sums with cancellation), rtol 1e-4 / atol 1e-5 between the two variants."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(n, cin, with_skip, with_perm, seed, full_rows=False):
    from eventad_b200 import ops
    g = torch.Generator().manual_seed(seed)
    cin_pad = (cin + 3) // 4 * 4
    ncls = 81
    count = torch.full((n,), 16, dtype=torch.int32) if full_rows else torch.randint(0, 17, (n,), generator=g).int()
    nbr = torch.randint(0, n, (n, 16), generator=g).int()
    cls = torch.randint(0, ncls, (n, 16), generator=g).to(torch.uint8)
    # slots beyond count hold garbage in the real layout (torch.empty): make it hostile
    dead = torch.arange(16)[None, :] >= count[:, None]
    nbr[dead] = 2**31 - 1
    cls[dead] = 255
    # degree-1 basis products: four non-zero taps per class out of 15 slots, (1-a, a) x (1-b, b) with arbitrary fractions
    tapw = torch.zeros(ncls, 16)
    for c in range(ncls):
        slots = torch.randperm(15, generator=g)[:4]
        a, b = torch.rand(2, generator=g)
        tapw[c, slots] = torch.stack([(1 - a) * (1 - b), a * (1 - b), (1 - a) * b, a * b])
    x = torch.zeros(n, cin_pad)
    x[:, :cin] = torch.randn(n, cin, generator=g) * 3
    W = torch.zeros(16, cin_pad, 16)
    W[:, :cin] = torch.randn(16, cin, 16, generator=g) * 0.3
    S = torch.zeros((cin_pad + 7) // 8 * 8, 16)
    S[:cin] = torch.randn(cin, 16, generator=g) * 0.3
    scale, shift = torch.rand(16, generator=g) + 0.5, torch.randn(16, generator=g)
    sscale, sshift = torch.rand(16, generator=g) + 0.5, torch.randn(16, generator=g)
    perm = torch.randperm(n, generator=g).int() if with_perm else None

    # fp64 reference: P_i[s] = sum_e tapw[cls_e][s] x_src ; slot 15 = x_i ; out = relu(scale * (P_i . W) + shift)
    xd, Wd = x.double(), W.double()
    P = torch.zeros(n, 16, cin_pad, dtype=torch.float64)
    for e in range(16):
        live = (e < count).double()[:, None, None]
        j = nbr[:, e].long().clamp(0, n - 1)
        P += live * tapw[cls[:, e].long().clamp(0, ncls - 1)].double()[:, :, None] * xd[j][:, None, :]
    P[:, 15] = xd
    ref = torch.relu(torch.einsum("nsc,sco->no", P, Wd) * scale.double() + shift.double())
    mag = torch.einsum("nsc,sco->no", P.abs(), Wd.abs()) * scale.double()   # sum of |terms|: the scale of cancellation errors
    xs = xd[:, :S.shape[0]] if cin_pad >= S.shape[0] else torch.nn.functional.pad(xd, (0, S.shape[0] - cin_pad))
    ref_skip = xs @ S.double() * sscale.double() + sshift.double()
    smag = xs.abs() @ S.double().abs() * sscale.double()

    dev = "cuda"
    whi, wlo = ops.split_tf32(W.view(16 * cin_pad, 16).to(dev))
    pack = dict(tapw=tapw.to(dev), whi=whi, wlo=wlo, shi=None, slo=None, cin_pad=cin_pad)
    if with_skip:
        pack["shi"], pack["slo"] = ops.split_tf32(S.to(dev))
    p16 = ops.l1_pack16(W.view(16 * cin_pad, 16), S if with_skip else None, cin_pad)   # operands of the tcgen05 contraction
    pack.update(w16=p16["w16"].to(dev), inv_sw=p16["inv_sw"], inv_ss=p16["inv_ss"],
                s16=p16["s16"].to(dev) if p16["s16"] is not None else None)
    epi = dict(scale=scale.to(dev), shift=shift.to(dev), relu=True)
    sepi = dict(scale=sscale.to(dev), shift=sshift.to(dev)) if with_skip else None
    args = (count.to(dev), nbr.to(dev).view(-1), cls.to(dev).view(-1), n, x.to(dev), cin, pack, epi, sepi)
    kw = dict(perm=perm.to(dev) if perm is not None else None, ell_stride=16)
    res = {}
    # "h16": both products as fp16-pair mma.sync m16n8k16 (csrc/spline_l1_h16.cu, the default); "h16w": the same with
    # input and output as h2 words (how the two convolutions of the layer are chained); "tc5": the same aggregation +
    # tcgen05 contraction (csrc/spline_l1_tc5.cu); True: both products as 3xTF32 mma.sync; False: FFMA aggregation
    status = ops.new_status(dev)
    for flag, (use_mma, use_tc5, use_h16) in (("h16", (True, False, True)), ("h16w", (True, False, True)),
                                              ("tc5", (True, True, True)), (True, (True, False, False)),
                                              (False, (False, False, False))):
        old = ops.USE_L1_MMA, ops.USE_L1_TC5, ops.USE_L1_H16
        ops.USE_L1_MMA, ops.USE_L1_TC5, ops.USE_L1_H16 = use_mma, use_tc5, use_h16
        try:
            if flag == "h16w":
                a2 = list(args)
                a2[4] = ops.split_h2(args[4], status)
                o, sk = ops.spline_conv_l1(*a2, **kw, x_h2=True, out_h2=True, status=status)
                h = o.view(torch.int32)
                lo16, hi16 = (h & 0xffff).to(torch.int16).view(torch.float16), (h >> 16).to(torch.int16).view(torch.float16)
                res[flag] = (lo16.float() + hi16.float() / 2048.0, sk)
            else:
                res[flag] = ops.spline_conv_l1(*args, **kw, status=status)
        finally:
            ops.USE_L1_MMA, ops.USE_L1_TC5, ops.USE_L1_H16 = old
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    for flag, (out, skip) in res.items():
        err = (out.double().cpu() - ref).abs()
        assert bool((err <= 2e-6 + 2e-5 * ref.abs() + 1e-6 * mag).all()), (flag, float(err.max()), float((err / mag).max()))
        if with_skip:
            serr = (skip.double().cpu() - ref_skip).abs()
            assert bool((serr <= 2e-6 + 2e-5 * ref_skip.abs() + 1e-6 * smag).all()), (flag, float(serr.max()))
        else:
            assert skip is None
    # the two variants sum the same 320 terms in different orders: north_star's rtol 1e-4 / atol 1e-5 plus the
    # cancellation term of the fp64 bar (the synthetic features here are far larger than real layer inputs)
    for other in (False, "tc5", "h16", "h16w"):
        diff = (res[True][0] - res[other][0]).abs().double().cpu()
        assert bool((diff <= 1e-5 + 1e-4 * ref.abs() + 2e-6 * mag).all()), (other, float(diff.max()))


@pytest.mark.parametrize("n,cin,with_skip,with_perm", [(1, 19, True, False), (31, 16, False, True), (33, 19, True, True),
                                                       (4096, 16, False, False), (10_007, 19, True, True),
                                                       (20_000, 16, True, True), (70_001, 20, False, True)])
def test_l1_mma_matches_fp64_and_ffma_variant(n, cin, with_skip, with_perm):
    _case(n, cin, with_skip, with_perm, seed=n % 7)


def test_l1_mma_full_rows():
    _case(5000, 19, True, True, seed=3, full_rows=True)


def test_l1_mma_rejects_what_it_cannot_read():
    """C-ABI argument checks: misaligned rows, unsupported channel width."""
    import ctypes
    from eventad_b200 import _lib
    L = _lib.lib()
    dev = "cuda"
    n = 64
    count = torch.zeros(n, dtype=torch.int32, device=dev)
    nbr = torch.zeros(n * 16 + 4, dtype=torch.int32, device=dev)
    cls = torch.zeros(n * 16 + 16, dtype=torch.uint8, device=dev)
    x = torch.zeros(n, 16, device=dev)
    tapw = torch.zeros(81, 16, device=dev)
    w = torch.zeros(256, 16, device=dev)
    out = torch.empty(n, 16, device=dev)
    e = _lib.EpilogueT()

    def call(nbr_t, cin_pad):
        return L.evad_spline_conv_l1_mma(_lib.ptr(count), _lib.ptr(nbr_t), _lib.ptr(cls), None, n, _lib.ptr(x), 16, cin_pad,
                                         _lib.ptr(tapw), 81, _lib.ptr(w), _lib.ptr(w), None, None, ctypes.byref(e),
                                         _lib.ptr(out), 16, None, None, 0, _lib.stream())
    assert call(nbr, 16) == 0
    assert call(nbr[1:], 16) != 0          # rows not 16-byte aligned
    assert call(nbr, 24) != 0              # channel width without a kernel
    torch.cuda.synchronize()


def test_l1_h16_raises_the_range_bit():
    """|x| >= 2048 does not fit the fp16-pair operand format: evad_split_h2 raises EVAD_STATUS_H16_RANGE."""
    from eventad_b200 import ops
    status = ops.new_status("cuda")
    x = torch.zeros(8, 16, device="cuda")
    ops.split_h2(x, status)
    assert int(status.item()) == 0
    x[3, 5] = -3000.0
    ops.split_h2(x, status)
    assert int(status.item()) == 16
