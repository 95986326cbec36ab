"""world_size-2 gloo tests of the sharding / gather / gradient all-reduce plumbing (eventad_b200/parallel.py)."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from eventad_b200 import parallel


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    clips = parallel.shard_clips(7, world, rank)
    local = {c: [c * 10 + w for (cc, w, n) in parallel.batches_of([c], 5, 2) for w in range(w, w + n)] for c in clips}
    merged = parallel.gather_results(local)
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    unused = torch.nn.Linear(2, 2)
    x = torch.full((2, 4), float(rank + 1))
    lin(x).sum().backward()
    parallel.allreduce_flat_grads(list(lin.parameters()) + list(unused.parameters()))
    q.put((rank, clips, merged, lin.weight.grad.clone(), unused.weight.grad.clone()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_gather_allreduce_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    (r0, c0, m0, g0, u0), (r1, c1, m1, g1, u1) = res
    assert c0 == [0, 2, 4, 6] and c1 == [1, 3, 5]
    assert m1 is None and list(m0.keys()) == list(range(7))
    assert m0[3] == [30, 31, 32, 33, 34]                         # windows of a clip stay in order
    # grads: rank r sees x = r+1 -> d/dW = 2*(r+1) per entry; mean over ranks = 3
    assert torch.allclose(g0, torch.full((3, 4), 3.0)) and torch.equal(g0, g1)
    assert torch.equal(u0, torch.zeros(2, 2)) and torch.equal(u1, torch.zeros(2, 2))


def test_batches_cut_only_inside_a_clip():
    b = parallel.batches_of([4, 9], windows_per_clip=5, batch_size=2)
    assert b == [(4, 0, 2), (4, 2, 2), (4, 4, 1), (9, 0, 2), (9, 2, 2), (9, 4, 1)]


def _train_worker(rank, world, port, q):
    import importlib.util
    import numpy as np
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("g", os.path.join(here, "golden", "make_ref_eventad_golden.py"))
    gold = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gold)
    from eventad_b200 import train
    case = [c for c in gold.CASES if c[0] == "clip_s"][0]
    ead, args, d, feats, labels = gold.case_inputs(case)
    z = np.load(os.path.join(here, "golden", "ref_eventad_clip_s.npz"))
    bf = torch.from_numpy(z["batch_features"]) * (1.0 + rank)          # rank-dependent inputs, same weights
    with torch.enable_grad():
        _, _, _, loss = ead.head(bf, d, labels)
        loss.backward()
    params = train.head_parameters(ead)
    local = [None if p.grad is None else p.grad.clone() for p in params]
    parallel.allreduce_flat_grads(params)
    q.put((rank, [None if g is None else g.numpy() for g in local], [p.grad.numpy().copy() for p in params]))
    dist.barrier()
    dist.destroy_process_group()


def test_head_gradient_allreduce_equals_mean_of_rank_gradients():
    """BASELINE configs[3]: after the flat-bucket all-reduce every rank holds (1/G) * sum of the per-rank head gradients;
    parameters that never get a gradient (fusion_module.attention) come out as zeros on all ranks."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_train_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (_, l0, a0), (_, l1, a1) = res
    import numpy as np
    n_none = 0
    for g0, g1, r0, r1 in zip(l0, l1, a0, a1):
        assert np.array_equal(r0, r1)
        if g0 is None:
            n_none += 1
            assert g1 is None and not r0.any()
        else:
            np.testing.assert_allclose(r0, (g0 + g1) / 2, rtol=1e-6, atol=1e-7)
    assert n_none == 4                                                  # attention.{in_proj_weight,in_proj_bias,out_proj.weight,out_proj.bias}
