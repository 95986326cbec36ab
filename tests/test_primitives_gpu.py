"""Device-wide scan and stable radix sort (eventad_b200/csrc/primitives.cu) through the C ABI."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scan(x):
    from eventad_b200 import _lib
    L = _lib.lib()
    n = x.numel()
    out = torch.empty_like(x)
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.empty(L.evad_scan_ws_bytes(n), dtype=torch.uint8, device="cuda")
    _lib.check(L.evad_exclusive_scan(_lib.ptr(x), _lib.ptr(out), n, _lib.ptr(total), _lib.ptr(ws), ws.numel(),
                                     _lib.stream()))
    return out, total


@pytest.mark.parametrize("n", [1, 5, 2047, 2048, 2049, 8192, 8193, 100_003, 3_000_001])
def test_exclusive_scan(n):
    g = torch.Generator().manual_seed(n)
    x = torch.randint(0, 17, (n,), generator=g, dtype=torch.int32)
    out, total = _scan(x.cuda())
    ref = torch.cumsum(x.long(), 0) - x.long()
    assert torch.equal(out.cpu().long(), ref)
    assert int(total.item()) == int(x.long().sum())


def _sort(keys, vals, bits):
    from eventad_b200 import _lib
    L = _lib.lib()
    n = keys.numel()
    ko = torch.empty_like(keys)
    vo = torch.empty_like(keys)
    ws = torch.empty(L.evad_sort_ws_bytes(n) + 4 * n + 512, dtype=torch.uint8, device="cuda")
    _lib.check(L.evad_sort_pairs(_lib.ptr(keys), _lib.ptr(vals), _lib.ptr(ko), _lib.ptr(vo), n, bits,
                                 _lib.ptr(ws), ws.numel(), _lib.stream()))
    return ko, vo


@pytest.mark.parametrize("n,bits", [(1, 3), (31, 5), (1000, 8), (1025, 10), (4096, 11), (70_001, 17), (100_000, 20),
                                    (600_000, 23), (1_500_000, 12), (50_000, 32)])
@pytest.mark.parametrize("iota", [True, False])
def test_stable_sort_pairs(n, bits, iota):
    g = torch.Generator().manual_seed(n + bits)
    hi = min(2 ** bits, 2 ** 31 - 1)
    keys = torch.randint(0, hi, (n,), generator=g, dtype=torch.int64)
    if bits == 32:
        keys = keys * 2 + torch.randint(0, 2, (n,), generator=g)
    keys = keys.to(torch.int64)
    vals = torch.arange(n) if iota else torch.randint(0, 2 ** 31 - 1, (n,), generator=g)
    k32 = keys.to(torch.uint32).view(torch.int32) if bits == 32 else keys.to(torch.int32)
    ko, vo = _sort(k32.cuda(), None if iota else vals.to(torch.int32).cuda(), bits)
    order = torch.sort(keys, stable=True).indices
    got_k = ko.cpu().view(torch.uint32).to(torch.int64) if bits == 32 else ko.cpu().long()
    assert torch.equal(got_k, keys[order])
    assert torch.equal(vo.cpu().long(), vals[order].long())
