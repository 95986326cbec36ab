#!/usr/bin/env python
"""bench.py -- GNN-branch throughput of EventAD's DAGR event branch on B200 (contract: see the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W]            our sm_100a path
  python bench.py --impl reference [--steps K] [--warmup W]      the CPU restatement of the reference path (oracle/)

A "step" is one pass of the hot path (event-graph build -> 5 SplineConv layers -> 4 voxel poolings, CNN-feature
sampling fused) over one batch of `--windows` synthetic ROL-shaped 50 ms windows of `--events` events each
(320x240, T = 1e6 us).  Metric: GNN-branch Mevents/s, whole job (all ranks).  One JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GNN-branch Mevents/s"
UNIT = "Mevents/s"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--events", type=int, default=100_000, help="events per 50 ms window")
    p.add_argument("--windows", type=int, default=64, help="windows per step (batch); BASELINE configs[2]: 64 clips x 100k-event windows")
    p.add_argument("--dist", default="U", choices=["U", "C"])
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--latency-iters", type=int, default=100)
    p.add_argument("--no-extras", action="store_true", help="skip the sweep / streaming / graph-stage / full-model legs")
    return p.parse_args()


def workload_config(a):
    return {"workload": f"dagr-s GNN branch forward (EV_TGN graph r=4px dt=10ms K=16 Q=128 -> Layer(19,16) -> 4x "
                        f"[sample CNN map + voxel pool + Layer(.,64)]), 320x240, {a.events} events/window, "
                        f"{a.windows} windows/step, dist {a.dist}, use_image=True (CNN maps resident, channels-last), LUT mode",
            "events_per_window": a.events, "windows_per_step": a.windows, "distribution": a.dist,
            "l2": "8 rotating input sets (>126 MB with intermediates; per-step traffic ~3 KB/event >> L2)",
            "parallelism": "independent windows per rank, no data-path collective"}


# ------------------------------------------------------------------------------------------------ clocks sampler
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions: NVML polled every 5 ms from a thread (the timed
    region of the default run is ~70 ms, too short for `nvidia-smi -lms`); falls back to one nvidia-smi query."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index=0):
        self.index, self.sm, self.bits, self.max_mhz, self.run, self.thread, self.nvml = index, [], 0, None, False, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None
            return
        self.run = True
        self.thread = threading.Thread(target=self._poll, daemon=True)
        self.thread.start()

    def _poll(self):
        n = self.nvml
        while self.run:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
                self.bits |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        if self.nvml is None:
            try:
                q = "clocks.sm,clocks.max.sm"
                r = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=20).stdout.split(",")
                return {"sm_mhz": float(r[0]), "sm_max_mhz": float(r[1]), "reasons": [], "samples": 1,
                        "source": "nvidia-smi after the timed region (NVML python binding unavailable)"}
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.run = False
        self.thread.join(timeout=1.0)
        n = self.nvml
        masks = [getattr(n, "nvmlClocksThrottleReasonHwSlowdown", 0x8), getattr(n, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 getattr(n, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20), getattr(n, "nvmlClocksThrottleReasonSwPowerCap", 0x4)]
        reasons = [nm for nm, m in zip(self.NAMES, masks) if self.bits & m]
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.sm), "source": "NVML polled every 5 ms across the device-timed and end-to-end timed regions"}


# ------------------------------------------------------------------------------------------------ CPU reference arm
def cpu_reference_run(a, steps, warmup):
    """Times the CPU restatement of the reference path (oracle/: C graph builder with OpenMP + torch-CPU layers with
    all host threads) on one window of `a.events` events per step.  Returns (Mevents/s, ms/step, cores, sample)."""
    from eventad_b200 import synthetic
    from oracle.dagr_net import net_forward
    from eventad_b200.synthetic import make_backbone, make_data, net_state_dict
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model, args = make_backbone(seed=0, use_image=True)
    sd = net_state_dict(model)
    feats = synthetic.make_image_features(1, seed=1)
    times = []
    for s in range(warmup + steps):
        d = make_data(a.events, 1, seed=100 + s, dist=a.dist)
        t0 = time.perf_counter()
        net_forward(sd, args, d.x, d.pos, d.batch, 1, feats, 320, 240)
        if s >= warmup:
            times.append(time.perf_counter() - t0)
    ms = 1e3 * float(np.mean(times))
    sample = f"{steps} x 1 window of {a.events} events (dist {a.dist}), B=1, oracle/ C graph builder (OpenMP) + torch-CPU layers"
    return a.events / (ms * 1e-3) / 1e6, ms, cores, sample


def ref_cuda_graph_stage(a, iters=5, B=1):
    """the reference's OWN CUDA graph builder (src/dagr/graph/ev_graph.cu compiled unmodified into
    oracle/_ref/ev_graph_cuda.so) timed on the GPU for the graph stage of one window, with the host glue of
    graph/utils.py:6-23 / ev_graph.py:45-103 (reset fill, sort, unique, insert, fill_edges, mask); the reference has no
    GPU code for the other stages without PyG.  None when there is no GPU or the module was not built."""
    if not torch.cuda.is_available():
        return None
    from eventad_b200 import synthetic
    from tests.util import RefGraph, load_ref_graph_module
    m = load_ref_graph_module()
    if m is None:
        return None
    raw = synthetic.make_batch(a.events, B, seed=100, dist=a.dist)
    W, H = raw["width"], raw["height"]
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    bt, pt = torch.from_numpy(raw["batch"].astype(np.int32)).cuda(), torch.from_numpy(pos).cuda()
    ts, edges = [], 0
    for it in range(iters + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        g = RefGraph(m, W, H, B, 16, 128, 4, 10000)           # reset(): queue refilled with -1 (ev_graph.py:45-50)
        edges = g.forward(bt, pt).shape[1]
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    return {"ms_per_window": ms / B, "ms_per_call": ms, "Mevents_per_s": a.events * B / ms / 1e3, "edges": int(edges),
            "events": a.events, "windows": B,
            "what": f"reference ev_graph.cu (unmodified, sm_100a) + torch sort/unique/mask glue, graph stage only, B={B}"}


STREAM_SIZES = (1, 10, 100, 1000, 10_000)


def stream_inputs(dist, n0=1_000_000, iters=200):
    """One time-sorted stream: n0 resident events, then `iters` micro-batches of every size in STREAM_SIZES."""
    from eventad_b200 import synthetic
    total = n0 + iters * sum(STREAM_SIZES)
    raw = synthetic.make_batch(total, 1, seed=55, dist=dist)
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    return torch.from_numpy(raw["batch"].astype(np.int32)).cuda(), torch.from_numpy(pos).cuda()


def time_stream(step, batch, pos, n0, iters):
    """step(batch_slice, pos_slice) per micro-batch; CUDA events around every call; -> {size: (p50, p90) ms}."""
    out, off = {}, n0
    for size in STREAM_SIZES:
        ts = []
        for it in range(iters):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step(batch[off:off + size], pos[off:off + size])
            e1.record()
            e1.synchronize()
            off += size
            if it >= 5:
                ts.append(e0.elapsed_time(e1))
        out[str(size)] = {"p50_ms": float(np.median(ts)), "p90_ms": float(np.percentile(ts, 90))}
    return out


def ref_cuda_streaming(a, iters=60):
    """BASELINE configs[4] with the reference's own kernels + glue: 10^6 resident events, sliding window, latency per
    micro-batch (insert_in_queue(_single)_cuda + fill_edges_cuda + sort / unique / mask, ev_graph.py:63-103,121-136)."""
    if not torch.cuda.is_available():
        return None
    from tests.util import RefGraph, load_ref_graph_module
    m = load_ref_graph_module()
    if m is None:
        return None
    n0 = 1_000_000
    batch, pos = stream_inputs(a.dist, n0, iters)
    g = RefGraph(m, 320, 240, 1, 16, 128, 4, 10000)
    g.forward(batch[:n0], pos[:n0])

    def step(b, p):
        g.forward(b, p)
        g.delete_nodes(len(b))
    res = time_stream(step, batch, pos, n0, iters)
    return {"resident_events": n0, "iters": iters - 5, "latency": res,
            "what": "reference ev_graph.cu kernels (unmodified) + its torch glue, append + delete_nodes per micro-batch"}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    REF_STEP_CAP, REF_WARMUP_CAP = 10, 2                         # ~1.6 s of CPU work per step (one 100k-event window)
    steps, warmup = max(1, min(a.steps, REF_STEP_CAP)), max(1, min(a.warmup, REF_WARMUP_CAP))
    v, ms, cores, sample = cpu_reference_run(a, steps, warmup)
    ref_cuda, ref_cuda64, ref_stream = None, None, None
    try:
        ref_cuda = ref_cuda_graph_stage(a)
        ref_cuda64 = ref_cuda_graph_stage(a, iters=3, B=a.windows)
        ref_stream = ref_cuda_streaming(a)
    except Exception as ex:                                       # report, do not hide
        ref_cuda = ref_cuda or {"error": repr(ex)[:200]}
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "ref_cuda_graph_stage": ref_cuda, "ref_cuda_graph_stage_b64": ref_cuda64, "ref_cuda_streaming": ref_stream,
            "note": "the reference has no CPU path for the graph builder and its PyG stack is not installable here; "
                    "this is the CPU restatement in oracle/ (kind=port), timed with all host threads inside one window "
                    f"per step; --steps / --warmup are honoured up to {REF_STEP_CAP} / {REF_WARMUP_CAP} (1.6 s of CPU "
                    "work per step), the `steps` / `warmup` keys say what ran"}
    emit(line)



# ------------------------------------------------------------------------------------------------ extra legs (our arm)
def path_bytes(n_events, edges, n1):
    """SURVEY.md 8(d) module-boundary model of the whole branch."""
    return 1200.0 * n_events + 108.0 * edges + 1100.0 * n1


def leg_sweep(runner, dev, peak, headline):
    """BASELINE north_star: 50k-1M events per window, uniform and clustered.  Windows per step keep ~6M events in flight;
    CUDA-graph replay per step with inputs resident, like `value`."""
    from eventad_b200 import synthetic
    from eventad_b200.engine import GraphedForward, pin_raw
    out = []
    gc = runner.net.events_to_graph.graph_creators
    for n, B in ((50_000, 64), (100_000, 64), (1_000_000, 6)):
        for dist in ("U", "C"):
            if (n, B, dist) == headline[:3]:
                out.append({"events_per_window": n, "windows_per_step": B, "distribution": dist, "Mevents_per_s": headline[3],
                            "ms_per_window": headline[4], "path_roofline_frac": headline[5], "note": "the headline run"})
                continue
            raws = [pin_raw(synthetic.make_batch(n, B, seed=300 + s, dist=dist)) for s in range(2)]
            datas = [runner.to_device(r) for r in raws]
            feats = [f.to(dev).contiguous(memory_format=torch.channels_last) for f in synthetic.make_image_features(B, seed=3)]
            gf = GraphedForward(runner.net, n * B, B, feats, device=dev, static_feats=True)
            gf.capture(datas[0], feats)
            for i in range(3):
                gf(datas[i % 2], feats)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            steps = 6
            e0.record()
            for i in range(steps):
                gf(datas[i % 2], feats)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            denorm = torch.tensor([320, 240, 1_000_000], device=dev)
            gc.reset()
            gc.batch_size = max(gc.batch_size, B)
            E = int(gc.build_csr(datas[0].batch.int(), (denorm.view(1, -1) * datas[0].pos + 1e-3).int())["num_edges"].item())
            ach = path_bytes(n * B, E, 2240.0 * B) / (ms * 1e-3) / 1e9
            out.append({"events_per_window": n, "windows_per_step": B, "distribution": dist,
                        "Mevents_per_s": n * B / ms / 1e3, "ms_per_window": ms / B, "edges_per_event": E / (n * B),
                        "path_roofline_frac": ach / peak})
            del gf, datas, raws, feats
            torch.cuda.empty_cache()
    return out


def leg_graph_stage(runner, a, dev):
    """The repo's graph stage at the SAME B as the reference arm's `ref_cuda_graph_stage` (B = 1 and B = windows): fixed-
    stride rows (what the fused branch consumes) and the full reference output (compacted int64 edge_index)."""
    from eventad_b200 import synthetic
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    res = {}
    for B in (1, a.windows):
        raw = synthetic.make_batch(a.events, B, seed=100, dist=a.dist)
        pos = torch.from_numpy(np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)).to(dev)
        bt = torch.from_numpy(raw["batch"].astype(np.int32)).to(dev)
        g = SlidingWindowGraph(width=320, height=240, batch_size=B, max_num_neighbors=16, max_queue_size=128, radius=4,
                               delta_t_us=10000)
        r = {}
        for name, fn in (("rows", lambda: g.build_rows(bt, pos)), ("edge_index", lambda: g.build_csr(bt, pos, want_edge_index=True))):
            ts = []
            for it in range(7):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                e1.synchronize()
                if it >= 2:
                    ts.append(e0.elapsed_time(e1))
            r[name + "_ms_per_call"] = float(np.median(ts))
            r[name + "_ms_per_window"] = float(np.median(ts)) / B
        res[f"B{B}"] = r
    res["what"] = ("evad_graph_build_rows / evad_graph_build (sort + cell table + spiral search [+ scan + compaction to int64 "
                   "edge_index]) on the inputs of the reference arm's ref_cuda_graph_stage (seed 100)")
    return res


def leg_streaming(a, dev, iters=200):
    """BASELINE configs[4]: 10^6 resident events (320x240), sliding window, latency per micro-batch of 1..10^4 events through
    SlidingWindowGraph.forward(delete_nodes=True) = evad_stream_append (ring insert + neighbour search of the new events)."""
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    n0 = 1_000_000
    batch, pos = stream_inputs(a.dist, n0, iters)
    g = SlidingWindowGraph(width=320, height=240, batch_size=1, max_num_neighbors=16, max_queue_size=128, radius=4,
                           delta_t_us=10000)
    g.forward(batch[:n0], pos[:n0], delete_nodes=False, collect_edges=False)
    res = time_stream(lambda b, p: g.forward(b, p, delete_nodes=True, collect_edges=False), batch, pos, n0, iters)
    out = {"resident_events": n0, "iters": iters - 5, "distribution": a.dist, "latency": res,
           "what": "graph stage of the asynchronous mode (append + search of the new events + delete_nodes), CUDA events "
                   "around every SlidingWindowGraph.forward call incl. its edge-count read-back"}
    del g
    try:
        out["incremental_forward"] = leg_incremental(a, dev)
    except Exception as ex:                                       # report, do not hide
        out["incremental_forward"] = {"error": repr(ex)[:300]}
    try:
        out["ingest"] = leg_ingest(dev)
    except Exception as ex:                                       # report, do not hide
        out["ingest"] = {"error": repr(ex)[:300]}
    return out


def leg_ingest(dev, iters=30):
    """cutting one 50 ms / ~100k-event window out of a device-resident recording (EventStore.load_events =
    _load_events of dsec_utils.py:80-120) and the down-sampling filter (downsample_events.py:91-122) on 1M events."""
    from eventad_b200 import ingest
    rng = np.random.default_rng(5)
    n, span = 4_000_000, 2_000_000
    t = np.sort(rng.integers(0, span, n)).astype(np.int64)
    x, y, p = rng.integers(0, 640, n), rng.integers(0, 480, n), rng.integers(0, 2, n)
    store = ingest.EventStore(x=x, y=y, p=p, t=t, t_offset=0, device=dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    ts, cnt = [], 0
    for i in range(iters + 3):
        e0, e1 = ev(), ev()
        e0.record()
        (xy, tt, pp), tq = store.load_events(100_000 + 50_000 * (i % 30), num_us=50_000, height=430, time_window=50_000)
        e1.record()
        torch.cuda.synchronize()
        cnt = xy.shape[0]
        if i >= 3:
            ts.append(e0.elapsed_time(e1))
    evs = dict(x=store.x[:1_000_000], y=store.y[:1_000_000], t=store.t[:1_000_000], p=2 * store.p[:1_000_000] - 1)
    td = []
    for i in range(8):
        e0, e1 = ev(), ev()
        e0.record()
        out, cm = ingest.downsample_events(evs, 480, 640, 240, 320)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            td.append(e0.elapsed_time(e1))
    return {"resident_events": n, "window_events": cnt, "load_events_p50_ms": float(np.median(ts)),
            "downsample_1M_events_p50_ms": float(np.median(td)), "downsample_survivors": int(out["t"].shape[0]),
            "what": "EventStore.load_events (ms_to_idx lookup on the host, flag + scan + scatter on the device, survivor "
                    "count read back) and downsample_events (stable sort by coarse cell + per-cell walk + compaction) on "
                    "device-resident arrays; CUDA events around the calls"}


def leg_incremental(a, dev, n0=1_000_000, iters=25):
    """the WHOLE event branch per micro-batch on n0 resident events of one window (B = 1): graph append +
    layer 1 on the new nodes + dense tail (asynchronous/incremental.py), against the dense fused forward of the same
    n0 events.  Outputs are bit-identical to the dense forward's (tests/test_incremental_gpu.py)."""
    from eventad_b200 import synthetic
    from eventad_b200.asynchronous.incremental import IncrementalBackbone
    from eventad_b200.data import Batch
    model, args = synthetic.make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    net = model.to(dev).backbone
    feats = [f.to(dev) for f in synthetic.make_image_features(1, seed=0)]
    extra = sum(s * (iters + 3) for s in STREAM_SIZES)
    d = synthetic.make_data(n0 + extra, 1, seed=42, dist=a.dist, device=dev)

    def part(lo, hi):
        p = Batch(x=d.x[lo:hi], pos=d.pos[lo:hi], batch=d.batch[lo:hi])
        p.num_graphs, p.width, p.height, p.time_window = 1, d.width, d.height, d.time_window
        return p
    ev = lambda: torch.cuda.Event(enable_timing=True)
    dense, resets = [], []
    with torch.no_grad():
        for i in range(6):                                   # the plain fused forward of all n0 events (no incremental state)
            e0, e1 = ev(), ev()
            e0.record()
            net(part(0, n0), image_feat=feats)
            e1.record()
            torch.cuda.synchronize()
            if i:
                dense.append(e0.elapsed_time(e1))
        inc = IncrementalBackbone(net, capacity=n0 + extra, feats=feats)
        for i in range(4):
            e0, e1 = ev(), ev()
            e0.record()
            inc.reset(part(0, n0))
            e1.record()
            torch.cuda.synchronize()
            if i:
                resets.append(e0.elapsed_time(e1))
        lat, done = {}, n0
        for sz in STREAM_SIZES:
            ts = []
            for i in range(iters + 3):
                e0, e1 = ev(), ev()
                e0.record()
                inc.update(part(done, done + sz))
                e1.record()
                torch.cuda.synchronize()
                done += sz
                if i >= 3:
                    ts.append(e0.elapsed_time(e1))
            lat[str(sz)] = {"p50_ms": float(np.median(ts)), "p90_ms": float(np.percentile(ts, 90))}
    return {"resident_events": n0, "iters": iters, "dense_forward_ms_p50": float(np.median(dense)),
            "reset_ms_p50": float(np.median(resets)), "latency": lat,
            "what": "IncrementalBackbone.update(micro-batch) -> out3 / out4 of ALL resident events (graph append + rows + "
                    "layer 1 on the new nodes + incremental pool 1 + layers 2-5 as one CUDA-graph replay), CUDA events around "
                    "the call incl. its two host read-backs; dense_forward = the plain fused forward (Net.forward) of the "
                    "same n0 events; reset = the window's first call (dense layer 1 + state initialisation)"}


def leg_full_model(dev, world, rank, windows=100, B=6, events=100_000):
    """BASELINE configs[1]: full EventAD inference (event GNN + ResNet-50 feature branch + anomaly head) on one synthetic
    ROL-shaped clip of `windows` consecutive windows, batches of 6 (config/eventad_config.py:46), 8 boxes per frame."""
    from eventad_b200 import synthetic, train
    from eventad_b200.models.EventAD import EventADModel
    model, args = synthetic.make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    torch.manual_seed(1)
    ead = EventADModel(model).to(dev).eval()
    ead.dagr_model.backbone.net = ead.dagr_model.backbone.net.to(memory_format=torch.channels_last)
    sets = []
    for s in range(3):
        d = synthetic.make_data(events, B, seed=500 + s, dist="U", device=dev)
        for k, v in synthetic.make_boxes(B, seed=600 + s).items():
            setattr(d, k, v.to(dev))
        d.image = torch.rand((B, 3, 240, 320), device=dev).contiguous(memory_format=torch.channels_last)
        sets.append(d)
    nb = (windows + B - 1) // B

    def run(n_batches):
        boxes = 0
        for i in range(n_batches):
            d = sets[i % 3]
            ead(d, d.bbox[:, 4], testing=True)
            boxes += int(d.bbox.shape[0] + d.bbox0.shape[0])
        return boxes
    with torch.no_grad():
        run(3)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        boxes = run(nb)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    res = {"clip_windows": nb * B, "batch": B, "events_per_window": events, "ms_per_batch": ms / nb,
           "windows_per_s": nb * B / (ms * 1e-3), "Mevents_per_s": nb * B * events / ms / 1e3,
           "fps_reference_definition": boxes / (ms * 1e-3),
           "published_fps_other_hardware": 595.48,
           "what": "EventADModel.forward(data, labels, testing=True): ResNet-50 (cuDNN, channels_last, random weights) + fused "
                   "event branch + evad_bbox_pool + batched GRU head; 'FPS' as utils/fps.py:58-59 counts it = boxes of frame t "
                   "and t-1 per second"}
    # ---- the same clip through engine.GraphedEventAD: CNN branch, event branch and pooling + head as three graph replays
    try:
        from eventad_b200.engine import GraphedEventAD
        runner = GraphedEventAD(ead, B, int(sets[0].pos.shape[0]), device=dev)

        def run_g(n_batches):
            for i in range(n_batches):
                d = sets[i % 3]
                runner(d, d.bbox[:, 4])
        with torch.no_grad():
            run_g(3)
            torch.cuda.synchronize()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            run_g(nb)
            g1.record()
            torch.cuda.synchronize()
        gms = g0.elapsed_time(g1)
        res["cuda_graphs"] = {"ms_per_batch": gms / nb, "windows_per_s": nb * B / (gms * 1e-3),
                              "fps_reference_definition": boxes / (gms * 1e-3),
                              "what": "engine.GraphedEventAD on the same clip: the CNN branch, the fused event branch (CNN maps "
                                      "consumed in place) and box pooling + GRU head as three CUDA-graph replays per batch; "
                                      "same kernels and results as the eager forward above (tests/test_eventad_gpu.py)"}
        from eventad_b200.engine import PipelinedEventAD
        pipe = PipelinedEventAD(ead, B, int(sets[0].pos.shape[0]), device=dev, overlap=True)

        def run_p(n_batches):
            for i in range(n_batches):
                d = sets[i % 3]
                pipe.submit(d, d.bbox[:, 4])
            pipe.drain()
        with torch.no_grad():
            run_p(4)
            torch.cuda.synchronize()
            g0.record()
            run_p(nb)
            g1.record()
            torch.cuda.synchronize()
        pms = g0.elapsed_time(g1)
        res["cuda_graphs"]["pipelined"] = {"ms_per_batch": pms / nb, "windows_per_s": nb * B / (pms * 1e-3),
                                           "what": "engine.PipelinedEventAD(overlap=True): two runners on two streams, "
                                                   "batches alternating, the host read of a batch behind the launch of the next; "
                                                   "inside a runner the CNN branch and the event branch are the two arms of one "
                                                   "graph (engine.OverlappedBranches)"}
    except Exception as ex:                                   # noqa: BLE001 - extra key only
        res["cuda_graphs"] = dict(res.get("cuda_graphs") or {}, error=repr(ex)[:200])
    # ---- the CNN branch (HookModule.forward) eager vs captured in a CUDA graph (fp32 / bf16 autocast)
    try:
        from eventad_b200.engine import GraphedImageBranch
        hook = ead.dagr_model.backbone.net
        branch = {}
        for bsz in (1, B):
            img = torch.rand((bsz, 3, 240, 320), device=dev).contiguous(memory_format=torch.channels_last)
            g32 = GraphedImageBranch(hook, bsz, 240, 320, device=dev).capture()
            g16 = GraphedImageBranch(hook, bsz, 240, 320, device=dev, dtype=torch.bfloat16).capture()

            def p50(fn, iters=30):
                ts = []
                with torch.no_grad():
                    for i in range(iters + 5):
                        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        a0.record()
                        fn()
                        a1.record()
                        torch.cuda.synchronize()
                        if i >= 5:
                            ts.append(a0.elapsed_time(a1))
                return float(np.median(ts))
            branch[f"B{bsz}"] = {"eager_ms": p50(lambda: hook(img)), "cuda_graph_ms": p50(lambda: g32(img)),
                                 "cuda_graph_bf16_ms": p50(lambda: g16(img))}
        branch["what"] = ("HookModule.forward (ResNet-50 trunk + 1x1 feature convolutions, cuDNN, channels_last): p50 of "
                          "the eager hooked forward vs engine.GraphedImageBranch (one CUDA-graph replay incl. the image "
                          "copy-in), fp32 and bf16 autocast")
        res["image_branch"] = branch
    except Exception as ex:                                       # report, do not hide
        res["image_branch"] = {"error": repr(ex)[:200]}
    # ---- BASELINE configs[3]: training step, DAGR frozen, flat-bucket gradient all-reduce over NCCL when world > 1
    opt = train.make_optimizer(ead)
    timers = {}
    ead.train()
    ead.dagr_model.eval()
    for i in range(3):
        train.train_step(ead, sets[i % 3], sets[i % 3].bbox[:, 4], opt)
    torch.cuda.synchronize()
    steps = 10
    e0.record()
    for i in range(steps):
        train.train_step(ead, sets[i % 3], sets[i % 3].bbox[:, 4], opt, timers=timers)
    e1.record()
    torch.cuda.synchronize()
    ms_t = e0.elapsed_time(e1) / steps
    ar = float(np.mean([x.elapsed_time(y) for x, y in timers["allreduce"]]))
    n_par = sum(p.numel() for p in train.head_parameters(ead))
    res["train_step"] = {"ms_per_step": ms_t, "allreduce_ms": ar, "allreduce_share": ar / ms_t, "world": world,
                         "trainable_params": n_par, "bucket_bytes": 4 * n_par, "batch": B,
                         "what": "forward (frozen DAGR under no_grad + head) + backward + one flat fp32 gradient all-reduce "
                                 "(NCCL when world > 1; a flatten / unflatten copy at world 1) + clip_grad_norm + AdamW"}
    # the same step with the frozen part (CNN branch, event branch, box pooling) as CUDA-graph replays; single rank only:
    # a rank whose capture failed would otherwise leave the others waiting in the all-reduce
    if world == 1 and isinstance(res.get("cuda_graphs"), dict) and "error" not in res["cuda_graphs"]:
        try:
            for i in range(3):
                train.train_step(ead, sets[i % 3], sets[i % 3].bbox[:, 4], opt, feature_fn=runner.features)
            torch.cuda.synchronize()
            e0.record()
            for i in range(steps):
                train.train_step(ead, sets[i % 3], sets[i % 3].bbox[:, 4], opt, feature_fn=runner.features)
            e1.record()
            torch.cuda.synchronize()
            res["train_step"]["ms_per_step_graphed_features"] = e0.elapsed_time(e1) / steps
        except Exception as ex:                                   # noqa: BLE001 - extra key only
            res["train_step"]["graphed_features_error"] = repr(ex)[:200]
    return res

# ------------------------------------------------------------------------------------------------ our arm
def build_runner(dev):
    from eventad_b200.engine import EventBranchRunner
    from eventad_b200.synthetic import make_backbone
    model, args = make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    return EventBranchRunner(model, dev), args


def run_b200(a):
    import torch.distributed as dist
    from eventad_b200 import _lib, ops, synthetic
    from eventad_b200.engine import pin_raw, raw_bytes
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()
    runner, args = build_runner(dev)
    B, N = a.windows, a.events
    n_sets = 8
    raws = [pin_raw(synthetic.make_batch(N, B, seed=1000 * rank + s, dist=a.dist)) for s in range(n_sets)]
    # CNN maps as a channels-last ResNet (cuDNN's native layout) hands them over: [B,C,h,w] tensors in torch.channels_last
    # memory format, so the sampling kernels read them in place (ops.to_channel_last is a view, not a copy)
    feats = [f.to(dev).contiguous(memory_format=torch.channels_last) for f in synthetic.make_image_features(B, seed=rank)]
    datas = [runner.to_device(r) for r in raws]                      # inputs resident in HBM for `value`
    torch.cuda.synchronize()
    events_per_step = N * B

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- `value`: the fused branch captured once in a CUDA graph (fixed events x windows shape; all data-dependent sizes
    #      live in device counters), inputs resident in HBM: per step 3 device copies into the graph's static inputs,
    #      one replay, one read-back of the 13 counters.  CNN maps are consumed in place.
    from eventad_b200.engine import GraphedForward, StreamedRunner
    gfd = GraphedForward(runner.net, N * B, B, feats, device=dev, static_feats=True)
    c0 = L.evad_launch_count()
    gfd.capture(datas[0], feats)
    launches_per_replay = (L.evad_launch_count() - c0) // 3            # 2 warm-up passes + the captured one

    def step_device(i):
        return gfd(datas[i % n_sets], feats)

    def step_plain(i):
        return runner.forward_device(datas[i % n_sets], feats)

    for i in range(a.warmup):
        step_device(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step_device(a.warmup + i)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = launches_per_replay * a.steps

    # ---- the same steps as plain stream launches, with CUDA events around every C-ABI call: per-stage times for the
    #      roofline block (the graph replay cannot be timed per kernel from the host)
    for i in range(min(a.warmup, 3)):
        step_plain(i)
    barrier()
    timer = ops.KernelTimer()
    ops.TIMER = timer
    e0.record()
    for i in range(a.steps):
        step_plain(a.warmup + i)
    e1.record()
    barrier()
    ops.TIMER = None
    ms_plain = e0.elapsed_time(e1)
    stage = timer.summary()

    # ---- end to end through the public serving call: pinned raw host events -> H2D (copy stream, one step ahead) ->
    #      format_data kernel -> graph replay -> D2H of out3 / out4.  Every step's H2D and D2H are inside the timed region.
    sr = StreamedRunner(runner, N * B, B, feats, static_feats=True)
    for _ in sr.run([raws[i % n_sets] for i in range(max(min(a.warmup, 3), 1))], feats):
        pass
    barrier()
    d2h = 0
    e0.record()
    for _, d2h in sr.run((raws[(a.warmup + i) % n_sets] for i in range(a.steps)), feats):
        pass
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None

    # ---- strong scaling (BASELINE configs[2] as written): the SAME `windows` windows in total, windows / world per rank
    ms_strong = None
    if world > 1 and B % world == 0:
        Bs = B // world
        fs = [f[:Bs].contiguous(memory_format=torch.channels_last) for f in feats]
        ds = [runner.to_device(pin_raw(synthetic.make_batch(N, Bs, seed=2000 + 10 * rank + s, dist=a.dist))) for s in range(4)]
        gfs = GraphedForward(runner.net, N * Bs, Bs, fs, device=dev, static_feats=True)
        gfs.capture(ds[0], fs)
        for i in range(a.warmup):
            gfs(ds[i % 4], fs)
        barrier()
        e0.record()
        for i in range(a.steps):
            gfs(ds[i % 4], fs)
        e1.record()
        barrier()
        ms_strong = e0.elapsed_time(e1)
        del gfs, ds

    t = torch.tensor([ms_total, ms_e2e, ms_strong or 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, ms_strong = float(t[0]), float(t[1]), float(t[2])
    full_model = None
    if not a.no_extras:
        try:
            full_model = leg_full_model(dev, world, rank)         # every rank takes part (NCCL all-reduce of the train step)
        except Exception as ex:                                   # report, do not hide
            full_model = {"error": repr(ex)[:300]}
    value = world * events_per_step * a.steps / (ms_total * 1e-3) / 1e6
    e2e_value = world * events_per_step * a.steps / (ms_e2e * 1e-3) / 1e6
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (CUDA events around every C-ABI call inside the timed region)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_kind = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    gc = runner.net.events_to_graph.graph_creators
    e_sets, n1_sets = [], []
    for d in datas:
        denorm = torch.tensor([320, 240, 1_000_000], device=dev)
        pos_int = (denorm.view(1, -1) * d.pos + 1e-3).int()
        gc.reset()
        e_sets.append(int(gc.build_csr(d.batch.int(), pos_int)["num_edges"].item()))
    E = float(np.mean(e_sets))
    Ntot = float(events_per_step)
    stage_ms = {k: v[1] / a.steps for k, v in stage.items()}
    ms_plain_step = ms_plain / a.steps
    dom = "spline_conv_l1_c19"
    dom_ms = stage_ms.get(dom, None)
    # algorithmic bytes of one conv1a launch: per node the x0 row (80 B), rowptr (4 B) and pixel-order
    # index (4 B) read, conv output 64 B + skip output 64 B written; per edge the source index (4 B) and the offset-class
    # byte (1 B); gathered rows counted once
    dom_bytes = Ntot * (80 + 4 + 4 + 64 + 64) + 5 * E
    # DRAM bytes of the same launch from the committed `ncu --set full` capture of THIS kernel (profiles/
    # r02n_l1_h16_full_raw.csv, round 2: dram__bytes_read.sum 134.54 MB + dram__bytes_write.sum 75.21 MB for
    # spline_l1_h16_kernel<20> at 8 x 100k uniform events), scaled per event; null for other distributions / kernels
    from eventad_b200 import ops as _ops
    h16 = _ops.USE_L1_MMA and _ops.USE_L1_H16 and not _ops.USE_L1_TC5
    ncu_traffic = 209.75e6 / 800000.0 * Ntot if (a.dist == "U" and h16) else None
    roofline = None
    if dom_ms:
        ach = dom_bytes / (dom_ms * 1e-3) / 1e9
        kname = ("spline_l1_h16_kernel<20> (conv_block1.conv_block1: 19->16 LUT conv + BN + ReLU + skip Linear/BN; per-node tap "
                 "aggregation and the weight contraction both as fp16-pair mma.sync m16n8k16 products, one launch)") if h16 else \
                ("layer-1 kernel selected by EVAD_L1_H16 / EVAD_L1_MMA / EVAD_L1_TC5 (not the default)")
        roofline = {"bound": "hbm", "kernel": kname,
                    "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": ncu_traffic,
                    "traffic_source": "profiles/r02n_l1_h16_full_raw.csv (ncu --set full, round 2), scaled per event",
                    "peak_source": peak_kind, "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": dom_ms,
                    "share_of_step": dom_ms / ms_plain_step,
                    "note": "ncu: DRAM traffic = the algorithmic bytes (6 % of the HBM peak); the kernel is bound on chip: L1 "
                            "data-pipe LSU wavefronts 83 % (shared-memory fragments + row gathers), issue slots 50 %, 16 warps "
                            "per SM at 128 registers, top stall long_scoreboard; see profiles/r02_summary.md"}
    n1 = 2240.0 * B                                              # pool-1 nodes (upper bound without the t==T extras)
    path_bytes = 1200 * Ntot + 108 * E + 1100 * n1                # SURVEY.md 8(d) module-boundary model
    path_ach = path_bytes / (ms_total / a.steps * 1e-3) / 1e9
    path_roofline = {"bound": "hbm", "model": "SURVEY 8(d): 1200*N + 108*E + 1100*N1 bytes per step", "achieved": path_ach,
                     "peak": peak, "unit": "GB/s", "frac": path_ach / peak, "edges_per_step": E}

    # ---- p50 latency of one window (B = 1), device-resident inputs, CUDA events, single stream:
    #      plain launches (host-enqueue bound) and the same pipeline replayed from one CUDA graph
    lat = None
    if a.latency_iters > 0:
        from eventad_b200.engine import GraphedForward
        raw1 = [pin_raw(synthetic.make_batch(N, 1, seed=77 + s, dist=a.dist)) for s in range(4)]
        d1 = [runner.to_device(r) for r in raw1]
        f1 = [f[:1].contiguous(memory_format=torch.channels_last) for f in feats]

        def p50(fn):
            for i in range(10):
                fn(i)
            torch.cuda.synchronize()
            ts = []
            for i in range(a.latency_iters):
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s0.record()
                fn(i)
                s1.record()
                s1.synchronize()
                ts.append(s0.elapsed_time(s1))
            return float(np.median(ts)), float(np.percentile(ts, 90))

        plain = p50(lambda i: runner.forward_device(d1[i % 4], f1))
        lat = {"events": N, "iters": a.latency_iters, "plain_p50_ms_per_window": plain[0], "plain_p90_ms_per_window": plain[1],
               "note": "B=1, events resident in HBM (the 1.7 MB H2D of a window's events is NOT inside), CNN maps consumed in "
                       "place (as in `value`); includes the copies of the events into the graph's static buffers and the "
                       "final count read-back; the < 1 ms figure needs the CUDA graph captured for this shape"}
        try:
            gf = GraphedForward(runner.net, N, 1, f1, device=dev, static_feats=True)   # CNN maps consumed in place, as in `value`
            gf.capture(d1[0], f1)
            gr = p50(lambda i: gf(d1[i % 4], f1))
            lat.update({"p50_ms_per_window": gr[0], "p90_ms_per_window": gr[1], "mode": "cuda_graph"})
        except Exception as ex:                                   # report, do not hide
            lat.update({"p50_ms_per_window": plain[0], "p90_ms_per_window": plain[1], "mode": "plain",
                        "cuda_graph_error": repr(ex)[:300]})

    cpu = None
    if not a.no_cpu_baseline and world == 1:
        v, ms, cores, sample = cpu_reference_run(a, 8, 1)        # ~13 s of CPU work at 100k events per window
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "ms_per_window": ms}

    strong = None
    if world > 1 and ms_strong > 0:
        strong = {"windows_total": B, "windows_per_rank": B // world, "ms_per_step": ms_strong / a.steps,
                  "value": B * N * a.steps / (ms_strong * 1e-3) / 1e6, "unit": UNIT,
                  "what": "strong scaling: the same 64 windows per step in total, sharded over the ranks (max over ranks)"}
    sweep = graph_stage = streaming = None
    if not a.no_extras:
        try:
            graph_stage = leg_graph_stage(runner, a, dev)
            streaming = leg_streaming(a, dev)
            sweep = leg_sweep(runner, dev, peak, (N, B, a.dist, value / world, ms_total / a.steps / B, path_ach / peak))
        except Exception as ex:                                   # report, do not hide
            sweep = sweep or {"error": repr(ex)[:300]}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": raw_bytes(raws[0]), "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / a.steps},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "path_roofline": path_roofline,
            "cpu_baseline": cpu, "latency": lat, "stage_ms_per_step": stage_ms, "plain_launch_ms_per_step": ms_plain_step,
            "mode": "value: CUDA-graph replay per step (inputs resident); e2e: double-buffered H2D on a copy stream + "
                    "format_data kernel + graph replay + D2H; stage_ms/roofline: plain stream launches with CUDA events",
            "ms_per_window": ms_total / a.steps / B, "strong_scaling": strong, "sweep": sweep, "graph_stage": graph_stage,
            "streaming": streaming, "full_model": full_model}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def _json_only_stdout():
    """Everything a library writes to fd 1 during the run (NCCL's version banner, warnings of C extensions) goes to
    stderr; the one JSON line is written to the real stdout by `emit`."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


_REAL_STDOUT = None


def main():
    a = parse()
    _json_only_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the sm_100a path has no CPU fallback); use --impl reference")
        run_b200(a)


if __name__ == "__main__":
    main()
