"""evad_exclusive_scan_static (eventad_b200/csrc/no_alloc.cu): the scan that allocates nothing -- its scratch is a
file-scope `static __device__` array -- through the C ABI, against the existing evad_exclusive_scan (caller workspace)
on the same inputs, bit-exact (integer work), on two calls in a row (the second finds the first call's scratch
contents).  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scan_ws(L, _lib, x):
    n = x.numel()
    out = torch.empty_like(x)
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.empty(L.evad_scan_ws_bytes(n), dtype=torch.uint8, device="cuda")
    _lib.check(L.evad_exclusive_scan(_lib.ptr(x), _lib.ptr(out), n, _lib.ptr(total), _lib.ptr(ws), ws.numel(),
                                     _lib.stream()))
    return out, total


def _scan_static(L, _lib, x, out, total):
    _lib.check(L.evad_exclusive_scan_static(_lib.ptr(x), _lib.ptr(out), x.numel(), _lib.ptr(total), _lib.stream()))


def test_static_scratch_scan_equals_workspace_scan_twice_in_a_row():
    from eventad_b200 import _lib
    L = _lib.lib()
    assert L.evad_exclusive_scan_static_max_n() == 1 << 24
    assert L.evad_exclusive_scan_static_bytes() == 33024
    # first a large scan (fills the scratch with big tile sums), then a smaller multi-tile one over the stale contents,
    # then the single-CTA sizes, then the maximum
    for seed, n in enumerate([3_000_001, 100_003, 8193, 8192, 1, 1 << 24]):
        g = torch.Generator().manual_seed(100 + seed)
        x = torch.randint(0, 17, (n,), generator=g, dtype=torch.int32).cuda()
        want, want_total = _scan_ws(L, _lib, x)
        out = torch.empty_like(x)                      # outputs allocated BEFORE the calls under test
        total = torch.zeros(1, dtype=torch.int32, device="cuda")
        for call in range(2):
            out.fill_(-1)
            total.fill_(-1)
            _scan_static(L, _lib, x, out, total)
            assert torch.equal(out, want), (n, call)
            assert int(total.item()) == int(want_total.item()), (n, call)
    ref = torch.cumsum(x.cpu().long(), 0) - x.cpu().long()        # and against the arithmetic itself at the maximum
    assert torch.equal(out.cpu().long(), ref)


def test_static_scratch_scan_in_place_and_without_total():
    from eventad_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(7)
    x = torch.randint(0, 1000, (1_234_567,), generator=g, dtype=torch.int32).cuda()
    want, _ = _scan_ws(L, _lib, x)
    buf = x.clone()
    _lib.check(L.evad_exclusive_scan_static(_lib.ptr(buf), _lib.ptr(buf), buf.numel(), None, _lib.stream()))
    assert torch.equal(buf, want)


def test_static_scratch_scan_rejects_sizes_beyond_the_maximum():
    from eventad_b200 import _lib
    L = _lib.lib()
    x = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = L.evad_exclusive_scan_static(_lib.ptr(x), _lib.ptr(x), (1 << 24) + 1, None, _lib.stream())
    assert rc == -5                                                # EVAD_ERR_STATIC_LIMIT, before any launch
    assert b"exceeds the static scratch's maximum" in L.evad_no_alloc_last_error()
    with pytest.raises(RuntimeError, match="EVAD_ERR_STATIC_LIMIT"):
        _lib.check(rc)
    assert L.evad_exclusive_scan_static(None, _lib.ptr(x), 16, None, _lib.stream()) == -1
    torch.cuda.synchronize()
