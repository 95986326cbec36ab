"""End-to-end parity of the event branch (`Net.forward`, fused and module-by-module) with the torch-CPU oracle
(oracle/dagr_net.py) on seeded synthetic ROL-shaped windows.  Graph structure bit-exact, features within
rtol 1e-4 / atol 1e-5."""
import pytest
import torch

from eventad_b200 import synthetic
from tests.util import assert_close, make_backbone, make_data, net_state_dict

pytestmark = pytest.mark.gpu


def _compare(outs, ref, what):
    assert len(outs) == len(ref) == 2
    for o, r, name in zip(outs, ref, ("out3", "out4")):
        assert o.x.shape == r.x.shape, (what, name, o.x.shape, r.x.shape)
        assert torch.equal(o.batch.cpu(), r.batch), (what, name)
        assert torch.equal(o.edge_index.cpu(), r.edge_index), (what, name)
        assert torch.equal(o.pos[:, :2].cpu(), r.pos[:, :2]), (what, name)
        assert_close(o.pos, r.pos, rtol=1e-6, atol=1e-7, what=f"{what} {name} pos")
        assert_close(o.edge_attr, r.edge_attr, rtol=1e-6, atol=1e-6, what=f"{what} {name} edge_attr")
        assert_close(o.x, r.x, what=f"{what} {name} x")


@pytest.mark.parametrize("n,B,dist,use_image", [(20_000, 1, "U", True), (30_000, 1, "C", True), (12_000, 3, "C", True),
                                                (20_000, 2, "U", False), (50_000, 1, "C", True)])
def test_net_forward_matches_oracle(n, B, dist, use_image):
    from oracle.dagr_net import net_forward
    model, args = make_backbone(seed=n % 5, use_image=use_image)
    model.cache_luts(320, 240, args.radius)
    sd = net_state_dict(model)
    d = make_data(n, B, seed=7, dist=dist)
    feats = synthetic.make_image_features(B, seed=1) if use_image else None
    ref = net_forward(sd, args, d.x, d.pos, d.batch, B, feats, 320, 240)
    model = model.cuda()
    net = model.backbone
    for fused in (True, False):
        net.fused = fused
        dg = make_data(n, B, seed=7, dist=dist, device="cuda")
        with torch.no_grad():
            res = net(dg, image_feat=[f.cuda() for f in feats] if use_image else None)
        outs = res[0] if use_image else res
        _compare(outs, ref, f"fused={fused}")


def test_basis_mode_matches_oracle():
    """Without cache_luts the PyG basis message is used (DAGR training configuration)."""
    from oracle.dagr_net import net_forward
    model, args = make_backbone(seed=3, use_image=False)
    sd = net_state_dict(model)
    d = make_data(15_000, 1, seed=8, dist="C")
    ref = net_forward(sd, args, d.x, d.pos, d.batch, 1, None, 320, 240, use_lut=False)
    net = model.cuda().backbone
    with torch.no_grad():
        outs = net(make_data(15_000, 1, seed=8, dist="C", device="cuda"))
    _compare(outs, ref, "basis")


def test_extract_features_api():
    model, args = make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    model = model.cuda()
    d = make_data(10_000, 2, seed=9, device="cuda")
    d.image = torch.rand((2, 3, 240, 320), device="cuda")
    feats = model.extract_features(d)
    out3, out4 = feats
    assert out4.x.shape[1] == 64 and out3.x.shape[1] == 64
    assert out4.x.shape[0] <= 36 * 2 and out4.pos.shape == (out4.x.shape[0], 3)
    assert torch.isfinite(out4.x).all()


def test_cuda_graph_replay_matches_plain():
    """The fused branch captured in one CUDA graph (engine.GraphedForward) gives the plain path's results on new inputs."""
    from eventad_b200.engine import GraphedForward
    model, args = make_backbone(seed=1, use_image=True)
    model.cache_luts(320, 240, args.radius)
    net = model.cuda().backbone
    feats = [f.cuda() for f in synthetic.make_image_features(2, seed=4)]
    d0 = make_data(6000, 2, seed=1, dist="C", device="cuda")
    gf = GraphedForward(net, d0.pos.shape[0], 2, feats, device="cuda").capture(d0, feats)
    for seed in (2, 3):
        d = make_data(6000, 2, seed=seed, dist="C", device="cuda")
        f2 = [f * (1.0 + 0.1 * seed) for f in feats]
        with torch.no_grad():
            plain = net(make_data(6000, 2, seed=seed, dist="C", device="cuda"), image_feat=f2)[0]
        got = gf(d, f2)
        for o, r in zip(got, plain):
            assert torch.equal(o.batch, r.batch) and torch.equal(o.pos, r.pos)
            assert torch.equal(o.edge_index, r.edge_index)
            assert torch.equal(o.x, r.x)


def test_format_events_matches_format_data():
    """evad_format_events == utils/buffers.py:33-44 on the raw Data layout (int16 pos, int32 t, int8 x, int64 batch),
    bit for bit, including the fp32 division of the converted integers."""
    from eventad_b200 import ops
    g = torch.Generator().manual_seed(11)
    n = 70_001
    pos = torch.stack([torch.randint(0, 320, (n,), generator=g), torch.randint(0, 240, (n,), generator=g)], 1).to(torch.int16)
    t = torch.randint(0, 1_000_001, (n,), generator=g).to(torch.int32)
    x = (torch.randint(0, 2, (n, 1), generator=g) * 2 - 1).to(torch.int8)
    batch = torch.sort(torch.randint(0, 5, (n,), generator=g)).values
    normalizer = torch.tensor([320, 240, 1_000_000])
    ref_pos = torch.cat([pos, t.view(-1, 1)], dim=-1) / normalizer
    gx, gpos, gbatch = ops.format_events(pos.cuda(), t.cuda(), x.view(-1).cuda(), batch.cuda(), 320, 240, 1_000_000)
    assert gpos.dtype == ref_pos.dtype == torch.float32
    assert torch.equal(gpos.cpu(), ref_pos) and torch.equal(gx.cpu(), x.float()) and torch.equal(gbatch.cpu(), batch.int())
    _, _, b0 = ops.format_events(pos.cuda(), t.cuda(), x.view(-1).cuda(), None, 320, 240, 1_000_000)
    assert int(b0.abs().sum()) == 0


def test_streamed_runner_matches_infer():
    """The double-buffered serving loop (copy stream + format kernel + graph replay) returns what the plain end-to-end
    call returns, batch after batch, also when a staging slot is reused."""
    from eventad_b200.engine import EventBranchRunner, StreamedRunner, pin_raw
    model, args = make_backbone(seed=2, use_image=True)
    model.cache_luts(320, 240, args.radius)
    runner = EventBranchRunner(model, "cuda")
    feats = [f.cuda().contiguous(memory_format=torch.channels_last) for f in synthetic.make_image_features(2, seed=6)]
    raws = [pin_raw(synthetic.make_batch(5000, 2, seed=40 + s, dist="C")) for s in range(5)]
    sr = StreamedRunner(runner, 10_000, 2, feats)
    got = list(sr.run(raws, feats))
    assert len(got) == 5
    for raw, (host, nbytes) in zip(raws, got):
        ref, ref_bytes = runner.infer(raw, feats)
        assert nbytes == ref_bytes
        for o, r in zip(host, ref):
            for key in ("x", "pos", "batch"):
                assert torch.equal(o[key], r[key]), key


def test_graphed_image_branch_matches_hooked_forward():
    """engine.GraphedImageBranch (HookModule.forward captured in a CUDA graph, channels-last) against the
    eager hooked forward (net_img.py:122-135) on new images: fp32 within cuDNN's algorithm-choice noise, bf16 autocast
    within 5e-2 of the map's scale; and the whole DAGR forward fed by its static maps equals the eager one's."""
    from eventad_b200.engine import GraphedImageBranch
    model, args = make_backbone(seed=2, use_image=True)
    model.cache_luts(320, 240, args.radius)
    model = model.cuda()
    hook = model.backbone.net
    gib = GraphedImageBranch(hook, 2, 240, 320).capture()
    gib16 = GraphedImageBranch(hook, 2, 240, 320, dtype=torch.bfloat16).capture()
    for seed in (1, 2):
        img = torch.rand((2, 3, 240, 320), device="cuda", generator=torch.Generator("cuda").manual_seed(seed))
        with torch.no_grad():
            want, _ = hook(img)
        got, outs = gib(img)
        assert outs == [] and len(got) == len(want) == 5
        for a, b in zip(got, want):
            assert a.shape == b.shape and a.is_contiguous(memory_format=torch.channels_last)
            scale = float(b.abs().max())
            assert float((a - b).abs().max()) <= 1e-4 * scale + 1e-5
        for a, b in zip(gib16(img)[0], want):
            assert float((a - b).abs().max()) <= 5e-2 * float(b.abs().max())
    d = make_data(8000, 2, seed=3, device="cuda")
    feats, _ = gib(img)
    with torch.no_grad():
        ref = model.backbone(make_data(8000, 2, seed=3, device="cuda"), image_feat=[f.clone() for f in want])[0]
        out = model.backbone(d, image_feat=feats)[0]
    for o, r in zip(out, ref):
        assert torch.equal(o.batch, r.batch) and torch.equal(o.pos, r.pos)
        assert_close(o.x, r.x, what="event branch on graphed CNN maps")
