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


def ref_cuda_graph_stage(a, iters=5):
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
    raw = synthetic.make_batch(a.events, 1, seed=100, dist=a.dist)
    W, H = raw["width"], raw["height"]
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    bt, pt = torch.from_numpy(raw["batch"].astype(np.int32)).cuda(), torch.from_numpy(pos).cuda()
    ts, edges = [], 0
    for it in range(iters + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        g = RefGraph(m, W, H, 1, 16, 128, 4, 10000)           # reset(): queue refilled with -1 (ev_graph.py:45-50)
        edges = g.forward(bt, pt).shape[1]
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    return {"ms_per_window": ms, "Mevents_per_s": a.events / ms / 1e3, "edges": int(edges), "events": a.events,
            "what": "reference ev_graph.cu (unmodified, sm_100a) + torch sort/unique/mask glue, graph stage only, B=1"}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(a.steps, 5)), max(1, min(a.warmup, 1))
    v, ms, cores, sample = cpu_reference_run(a, steps, warmup)
    try:
        ref_cuda = ref_cuda_graph_stage(a)
    except Exception as ex:                                       # report, do not hide
        ref_cuda = {"error": repr(ex)[:200]}
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "ref_cuda_graph_stage": ref_cuda,
            "note": "the reference has no CPU path for the graph builder and its PyG stack is not installable here; "
                    "this is the CPU restatement in oracle/ (kind=port), timed with all host threads"}
    print(json.dumps(line), flush=True)


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

    t = torch.tensor([ms_total, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])
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
    # DRAM bytes of the same launch from the committed `ncu --set full` capture (profiles/r01g_spline_l1_mma_full_raw.csv:
    # dram__bytes_read.sum 134.5 MB + dram__bytes_write.sum 74.4 MB at 8 x 100k uniform events), scaled per event
    ncu_traffic = 208.9e6 / 800000.0 * Ntot if a.dist == "U" else None
    roofline = None
    if dom_ms:
        ach = dom_bytes / (dom_ms * 1e-3) / 1e9
        from eventad_b200 import ops as _ops
        kname = ("spline_l1_mma_kernel<20> (conv_block1.conv_block1: 19->16 LUT conv + BN + ReLU + skip Linear/BN; per-node "
                 "tap aggregation and the weight contraction both as 3xTF32 mma.sync products, one launch)") if _ops.USE_L1_MMA else \
                ("spline_l1_kernel<20> (conv_block1.conv_block1: 19->16 LUT conv + BN + ReLU + skip Linear/BN; "
                 "register tap aggregation + 3xTF32 mma contraction, one launch)")
        roofline = {"bound": "hbm", "kernel": kname,
                    "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": ncu_traffic,
                    "peak_source": peak_kind, "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": dom_ms,
                    "share_of_step": dom_ms / ms_plain_step,
                    "note": "ncu: DRAM 4.3 % of peak, L1 data pipe (LSU wavefronts: row gathers + shared-memory fragment "
                            "traffic) 75 %, tensor pipe 42 %, issue slots 53 % -- bound by the L1 data pipe and by latency at "
                            "16 warps per SM, not by HBM; see profiles/r01g_summary.md"}
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
               "note": "B=1, events resident, CNN maps consumed in place (as in `value`); includes the copies of the events into the graph's static buffers and the final count read-back"}
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

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": raw_bytes(raws[0]), "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / a.steps},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "path_roofline": path_roofline,
            "cpu_baseline": cpu, "latency": lat, "stage_ms_per_step": stage_ms, "plain_launch_ms_per_step": ms_plain_step,
            "mode": "value: CUDA-graph replay per step (inputs resident); e2e: double-buffered H2D on a copy stream + "
                    "format_data kernel + graph replay + D2H; stage_ms/roofline: plain stream launches with CUDA events",
            "ms_per_window": ms_total / a.steps / B}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the sm_100a path has no CPU fallback); use --impl reference")
        run_b200(a)


if __name__ == "__main__":
    main()
