"""Generates tests/golden/ref_ingest_*.npz by EXECUTING THE REFERENCE'S OWN FUNCTION BODIES (oracle/ingest.py:
ref_functions -- create_ms_to_idx, downsample_events / _filter_events_resize of scripts/downsample_events.py and
_load_events of src/dagr/data/dsec_utils.py, AST-extracted; h5py.File is a dict stand-in, @numba.jit is dropped).

    python tests/golden/make_ref_ingest_golden.py        (build container only: needs /root/reference; seconds)

Inputs are regenerated from seeds by `recording(seed)`; the fixtures hold only outputs."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

LOADS = [dict(dt=120_000, num_us=50_000, time_window=50_000, height=430),
         dict(dt=7_000, num_events=3000, time_window=50_000, height=None),
         dict(dt=250_000, num_us=20_000, time_window=50_000, height=215),
         dict(dt=299_000, num_us=0, time_window=50_000, height=None)]          # empty window


def recording(seed, n=60_000, width=640, height=480, span_us=300_000):
    """A synthetic recording: ascending t (with ties), blobs of same-polarity events so that the down-sampling filter fires."""
    rng = np.random.default_rng(seed)
    t = np.sort(rng.integers(0, span_us, n)).astype("uint32")
    cx, cy = rng.integers(0, width, 40), rng.integers(0, height, 40)
    k = rng.integers(0, 40, n)
    x = np.clip(cx[k] + rng.integers(-6, 7, n), 0, width - 1).astype("uint16")
    y = np.clip(cy[k] + rng.integers(-6, 7, n), 0, height - 1).astype("uint16")
    p = ((k + (t // 50_000)) % 2).astype("uint8")                                 # blobs flip polarity every 50 ms
    return dict(x=x, y=y, p=p, t=t, t_offset=1_000_000 + 17 * seed)


def main():
    from oracle import ingest
    R = ingest.ref_functions()
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for seed in (0, 1):
        rec = recording(seed)
        z = {}
        z["ms_to_idx"] = R["create_ms_to_idx"](rec["t"]).astype(np.int64)
        ev = dict(x=rec["x"].copy(), y=rec["y"].copy(), t=rec["t"].copy(), p=(2 * rec["p"].astype("int8") - 1))
        cm = None
        for c, (a, b) in enumerate(((0, 35_000), (35_000, 60_000))):          # two chunks: change_map is threaded through
            e, cm = R["downsample_events"]({k: v[a:b].copy() for k, v in ev.items()}, 480, 640, 240, 320, cm)
            for k in "xytp":
                z[f"ds{c}_{k}"] = e[k]
            z[f"ds{c}_map"] = cm.copy()
        f = {"t_offset": rec["t_offset"], "ms_to_idx": z["ms_to_idx"], "events/x": rec["x"], "events/y": rec["y"],
             "events/t": rec["t"], "events/p": rec["p"]}
        for i, kw in enumerate(LOADS):
            kw = dict(kw)
            dt = kw.pop("dt")
            with np.errstate(over="ignore"):
                (xy, t, p), tq = R["_load_events"](f, rec["t_offset"] + dt, **kw)
            z[f"load{i}_xy"], z[f"load{i}_t"], z[f"load{i}_p"], z[f"load{i}_tq"] = xy, t, p, np.int64(tq)
        np.savez_compressed(os.path.join(out_dir, f"ref_ingest_{seed}.npz"), **z)
        print(seed, {k: v.shape for k, v in z.items() if k.endswith(("_t", "idx"))})


if __name__ == "__main__":
    main()
