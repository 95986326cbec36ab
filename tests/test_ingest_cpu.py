"""Event ingest: the numpy restatement (oracle/ingest.py) against outputs of the reference's own function
bodies (tests/golden/ref_ingest_*.npz, generator tests/golden/make_ref_ingest_golden.py), and against those bodies live
where /root/reference is mounted."""
import importlib.util
import os

import numpy as np
import pytest

from oracle import ingest

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_ref_ingest_golden", os.path.join(HERE, "golden", "make_ref_ingest_golden.py"))
gold = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gold)


def run_oracle(rec):
    """Every golden quantity from the restatement."""
    z = {"ms_to_idx": ingest.create_ms_to_idx(rec["t"]).astype(np.int64)}
    ev = dict(x=rec["x"].copy(), y=rec["y"].copy(), t=rec["t"].copy(), p=(2 * rec["p"].astype("int8") - 1))
    cm = None
    for c, (a, b) in enumerate(((0, 35_000), (35_000, 60_000))):
        e, cm = ingest.downsample_events({k: v[a:b].copy() for k, v in ev.items()}, 480, 640, 240, 320, cm)
        for k in "xytp":
            z[f"ds{c}_{k}"] = e[k]
        z[f"ds{c}_map"] = cm.copy()
    f = {"t_offset": rec["t_offset"], "ms_to_idx": z["ms_to_idx"], "events/x": rec["x"], "events/y": rec["y"],
         "events/t": rec["t"], "events/p": rec["p"]}
    for i, kw in enumerate(gold.LOADS):
        kw = dict(kw)
        dt = kw.pop("dt")
        (xy, t, p), tq = ingest.load_events(f, rec["t_offset"] + dt, **kw)
        z[f"load{i}_xy"], z[f"load{i}_t"], z[f"load{i}_p"], z[f"load{i}_tq"] = xy, t, p, np.int64(tq)
    return z


@pytest.mark.parametrize("seed", [0, 1])
def test_ingest_oracle_matches_reference_golden(seed):
    g = np.load(os.path.join(HERE, "golden", f"ref_ingest_{seed}.npz"))
    z = run_oracle(gold.recording(seed))
    assert set(z) == set(g.files)
    for k in g.files:
        assert z[k].shape == g[k].shape and np.array_equal(z[k], g[k]), k
    assert g["ds0_t"].shape[0] > 1000 and g["load0_t"].shape[0] > 1000 and g["load3_t"].shape[0] == 0


@pytest.mark.skipif(not ingest.available(), reason="/root/reference is only mounted in the build container")
def test_reference_bodies_live_reproduce_golden():
    R = ingest.ref_functions()
    rec = gold.recording(1)
    g = np.load(os.path.join(HERE, "golden", "ref_ingest_1.npz"))
    assert np.array_equal(R["create_ms_to_idx"](rec["t"]).astype(np.int64), g["ms_to_idx"])
    ev = dict(x=rec["x"][:35_000].copy(), y=rec["y"][:35_000].copy(), t=rec["t"][:35_000].copy(),
              p=(2 * rec["p"][:35_000].astype("int8") - 1))
    e, cm = R["downsample_events"](ev, 480, 640, 240, 320)
    assert np.array_equal(e["t"], g["ds0_t"]) and np.array_equal(cm, g["ds0_map"])
