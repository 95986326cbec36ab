"""Incremental forward of the event branch (eventad_b200/asynchronous/incremental.py): a window is started
with `reset(first events)` and fed micro-batches with `update(new events)`; after every call the outputs must be the
dense fused forward's on ALL events seen so far.  The reference's checker for its asynchronous layers compares dense and
incremental activations within 1e-3 (asynchronous/evaluate_flops.py:141-143); here the two paths run the same kernels on
the same rows, so the bar is bit-exact for node sets, positions and edges, and for the features too whenever both paths
plan their kernels on the same size bounds (capacity >= the 4 480 voxel keys of a window: the pooled layers pick GEMM
variants / split-K plans from upper bounds, and a different plan sums in a different order); north_star's rtol 1e-4 /
atol 1e-5 otherwise."""
import copy

import pytest
import torch

from eventad_b200 import synthetic
from eventad_b200.synthetic import make_backbone, make_data
from tests.util import assert_close

pytestmark = pytest.mark.gpu


def _part(d, a, b):
    from eventad_b200.data import Batch
    p = Batch(x=d.x[a:b].clone(), pos=d.pos[a:b].clone(), batch=d.batch[a:b].clone())
    p.num_graphs, p.width, p.height, p.time_window = d.num_graphs, d.width, d.height, d.time_window
    return p


def _same(got, want, what, exact=True):
    assert len(got) == len(want) == 2
    for o, r in zip(got, want):
        assert torch.equal(o.batch, r.batch), what
        assert torch.equal(o.pos, r.pos), what
        assert torch.equal(o.edge_index, r.edge_index), what
        if exact:
            assert torch.equal(o.x, r.x), (what, float((o.x - r.x).abs().max()))
        else:
            assert_close(o.x, r.x, what=what)


@pytest.mark.parametrize("dist,n0,steps", [("U", 20_000, (1, 1, 10, 100, 1000, 3, 5000)),
                                           ("C", 5_000, (7, 64, 96, 97, 2000, 1)),
                                           ("U", 4_700, (1, 2, 500))])
def test_incremental_updates_equal_dense_forward(dist, n0, steps):
    from eventad_b200.asynchronous.incremental import IncrementalBackbone
    model, args = make_backbone(seed=3, use_image=True)
    model.cache_luts(320, 240, args.radius)
    net = model.cuda().backbone
    dense_net = copy.deepcopy(net)                       # its own graph creator: the dense passes must not touch the stream
    feats = [f.cuda() for f in synthetic.make_image_features(1, seed=6)]
    total = n0 + sum(steps)
    d = make_data(total, 1, seed=11, dist=dist, device="cuda")
    inc = IncrementalBackbone(net, capacity=total, feats=feats)
    with torch.no_grad():
        _same(inc.reset(_part(d, 0, n0)), dense_net(_part(d, 0, n0), image_feat=feats)[0], "reset")
        done = n0
        for k, m in enumerate(steps):
            got = inc.update(_part(d, done, done + m))
            done += m
            _same(got, dense_net(_part(d, 0, done), image_feat=feats)[0], f"update {k} (+{m})")
    assert inc.n == total


def test_incremental_reset_starts_a_new_window_and_capacity_is_checked():
    from eventad_b200.asynchronous.incremental import IncrementalBackbone
    model, args = make_backbone(seed=1, use_image=True)
    model.cache_luts(320, 240, args.radius)
    net = model.cuda().backbone
    dense_net = copy.deepcopy(net)
    feats = [f.cuda() for f in synthetic.make_image_features(1, seed=2)]
    inc = IncrementalBackbone(net, capacity=4000, feats=feats)
    with torch.no_grad():
        for seed in (1, 2):
            d = make_data(3000, 1, seed=seed, dist="C", device="cuda")
            _same(inc.reset(_part(d, 0, 2000)), dense_net(_part(d, 0, 2000), image_feat=feats)[0], f"window {seed} reset", exact=False)
            _same(inc.update(_part(d, 2000, 3000)), dense_net(_part(d, 0, 3000), image_feat=feats)[0], f"window {seed}", exact=False)
        with pytest.raises(RuntimeError, match="capacity"):
            inc.update(_part(d, 0, 1500))


def test_incremental_from_a_single_event():
    """A window that starts with one event (EV_TGN's insert_in_queue_single path) and grows by 1, 2, 500."""
    from eventad_b200.asynchronous.incremental import IncrementalBackbone
    model, args = make_backbone(seed=3, use_image=True)
    model.cache_luts(320, 240, args.radius)
    net = model.cuda().backbone
    dense_net = copy.deepcopy(net)
    feats = [f.cuda() for f in synthetic.make_image_features(1, seed=6)]
    d = make_data(504, 1, seed=11, dist="U", device="cuda")
    inc = IncrementalBackbone(net, capacity=504, feats=feats)
    with torch.no_grad():
        _same(inc.reset(_part(d, 0, 1)), dense_net(_part(d, 0, 1), image_feat=feats)[0], "reset", exact=False)
        done = 1
        for m in (1, 2, 500):
            got = inc.update(_part(d, done, done + m))
            done += m
            _same(got, dense_net(_part(d, 0, done), image_feat=feats)[0], f"+{m}", exact=False)


@pytest.mark.parametrize("name", ["u", "c"])
def test_incremental_update_matches_reference_asynchronous_model(name):
    """tests/golden/ref_async_*.npz hold what the REFERENCE'S OWN asynchronous model (`make_model_asynchronous`, executed
    unmodified on the CPU; generator tests/golden/make_ref_async_golden.py) returns after reset(n0 events) + one update of
    m events.  IncrementalBackbone must give the same voxels (by position) and features within the reference checker's
    1e-3 (evaluate_flops.py:141-143).  The window's last event (t == 1.0) is never streamed: the reference's two modes
    disagree on its voxel (see the generator)."""
    import importlib.util
    import os
    import numpy as np
    from eventad_b200.asynchronous.incremental import IncrementalBackbone
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("make_ref_async_golden", os.path.join(here, "golden", "make_ref_async_golden.py"))
    gold = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gold)
    case = [c for c in gold.CASES if c[0] == name][0]
    _, mseed, dseed, dist, n0, steps = case
    model, args, d, feats = gold.case_inputs(case)
    z = np.load(os.path.join(here, "golden", f"ref_async_{name}.npz"))
    model.cache_luts(gold.W, gold.H, args.radius)
    net = model.cuda().backbone
    for k in ("x", "pos", "batch"):
        setattr(d, k, getattr(d, k).cuda())
    feats = [f.cuda() for f in feats]
    scale = torch.tensor([gold.W, gold.H])
    with torch.no_grad():
        for s, m in enumerate(steps):
            inc = IncrementalBackbone(net, capacity=n0 + max(steps), feats=feats, width=gold.W, height=gold.H)
            inc.reset(_part(d, 0, n0))
            outs = inc.update(_part(d, n0, n0 + m))
            for lvl, o in zip((3, 4), outs):
                rp, rx = torch.from_numpy(z[f"s{s}_out{lvl}_pos"]), torch.from_numpy(z[f"s{s}_out{lvl}_x"])
                ours = {tuple(r): i for i, r in enumerate((o.pos[:, :2].cpu() * scale).round().int().tolist())}
                keys = [tuple(r) for r in (rp[:, :2] * scale).round().int().tolist()]
                assert len(set(keys)) == len(keys) and all(k in ours for k in keys), (name, s, lvl)
                assert len(ours) == len(keys), (name, s, lvl, len(ours), len(keys))
                sel = torch.tensor([ours[k] for k in keys])
                err = float((o.x.cpu()[sel] - rx).abs().max())
                assert err < 1e-3, (name, s, lvl, err)
