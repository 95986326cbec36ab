"""oracle/ingest.py -- CPU restatement (numpy) of the reference's event ingest for SURVEY 8 f-4.

TEST INFRASTRUCTURE ONLY (imported by tests/ and tests/golden/make_ref_ingest_golden.py; never by eventad_b200/).

Follows, function by function:
  create_ms_to_idx        /root/reference/scripts/downsample_events.py:36-42
  filter_events_resize    /root/reference/scripts/downsample_events.py:108-122   (numba loop, restated as a Python loop)
  downsample_events       /root/reference/scripts/downsample_events.py:91-106
  load_events             /root/reference/src/dagr/data/dsec_utils.py:80-120    (the h5py.File becomes a dict of arrays)
  to_data                 /root/reference/src/dagr/data/utils.py:6-20           (arrays only, no torch_geometric)

Pinned: the reference's own function bodies are executed (AST-extracted, h5py / numba / hdf5plugin replaced by
stand-ins: `ref_functions()` below) by tests/golden/make_ref_ingest_golden.py, which writes tests/golden/ref_ingest_*.npz;
tests/test_ingest_cpu.py checks this restatement against those files, and against the live functions where
/root/reference is mounted.
"""
import ast
import os

import numpy as np

REF_ROOT = "/root/reference"


def create_ms_to_idx(t_us):
    t_ms = np.asarray(t_us) // 1000
    x, counts = np.unique(t_ms, return_counts=True)
    out = np.zeros(shape=(int(t_ms[-1]) + 2,), dtype="uint64")
    out[x + 1] = counts
    return out[:-1].cumsum()


def filter_events_resize(x, y, p, mask, change_map, fx, fy):
    """One pass in event order: every coarse cell integrates p / (fx * fy) in its float32 cell; an event whose cell reaches
    |v| >= 1 passes and takes p back out.  numba evaluates `cell + p * 1.0 / (fx * fy)` in float64 and stores float32."""
    for i in range(len(x)):
        xl, yl = int(x[i]) // fx, int(y[i]) // fy
        change_map[yl, xl] = np.float32(np.float64(change_map[yl, xl]) + np.float64(p[i]) * 1.0 / (fx * fy))
        if np.abs(change_map[yl, xl]) >= 1:
            mask[i] = True
            change_map[yl, xl] = np.float32(np.float64(change_map[yl, xl]) - np.float64(p[i]))
    return mask, change_map


def downsample_events(events, input_height, input_width, output_height, output_width, change_map=None):
    if change_map is None:
        change_map = np.zeros((output_height, output_width), dtype="float32")
    fx, fy = int(input_width / output_width), int(input_height / output_height)
    mask = np.zeros(shape=(len(events["t"]),), dtype="bool")
    mask, change_map = filter_events_resize(events["x"], events["y"], events["p"], mask, change_map, fx, fy)
    events = {k: v[mask] for k, v in events.items()}
    events["x"] = (events["x"] / fx).astype("uint16")
    events["y"] = (events["y"] / fy).astype("uint16")
    return events, change_map


def load_events(f, t0, num_events=None, num_us=None, height=None, time_window=None):
    """f: dict with 't_offset' (int), 'ms_to_idx' (array) and 'events/{x,y,t,p}' arrays (the HDF5 file's datasets)."""
    ms = int((t0 - f["t_offset"]) / 1e3)
    idx0 = int(f["ms_to_idx"][ms])
    if num_events is not None:
        idx1 = idx0 + num_events
    if num_us is not None:
        idx1 = int(f["ms_to_idx"][ms + int(num_us / 1e3)])
    idx0, idx1 = sorted([idx0, idx1])
    idx0, idx1 = max(idx0, 0), max(idx1, 0)
    ev = {k: np.asarray(f[f"events/{k}"][idx0:idx1]) for k in "xytp"}
    tq = ev["t"][-1] if idx1 > idx0 else f["events/t"][max([idx1 - 1, idx0])]
    p = 2 * ev["p"][..., None].astype("int8") - 1
    xy = np.stack([ev["x"], ev["y"]], axis=-1).astype("int16")
    t = (np.int64(time_window) - np.int64(tq) + ev["t"][..., None].astype(np.int64)).astype("int32")
    tq_abs = np.int64(tq) + np.int64(f["t_offset"])
    mask = t[:, 0] > 0
    if height is not None:
        mask &= xy[:, 1] < height
    return (xy[mask], t[mask], p[mask]), tq_abs


def to_data(x, y, t, p):
    """data/utils.py:6-20 on arrays: x <- polarity [n, 1], pos <- int16 [n, 2], t <- int32."""
    return dict(x=np.asarray(p).reshape((-1, 1)), pos=np.stack([x, y], axis=-1).astype("int16"), t=np.asarray(t).astype("int32"))


# ------------------------------------------------------------------------------------------------ the reference's own bodies
def _extract(path, names):
    """Source text of the named top-level functions of a reference file, decorators (@numba.jit) dropped."""
    tree = ast.parse(open(path).read())
    out = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            node.decorator_list = []
            out.append(ast.unparse(node))
    return "\n\n".join(out)


class _FakeH5:
    """h5py.File stand-in over a dict: `with h5py.File(file) as f`, f['name'][()] / f['name'][a:b] / f[f'events/{k}']."""

    def __init__(self, d):
        self.d = d

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def __getitem__(self, k):
        return np.asarray(self.d[k])


def ref_functions():
    """dict of the reference's create_ms_to_idx, downsample_events, _filter_events_resize, _load_events, executed from their
    own source text (no h5py / numba needed).  Only where /root/reference is mounted."""
    ns = {"np": np}
    code = _extract(os.path.join(REF_ROOT, "scripts", "downsample_events.py"),
                    {"create_ms_to_idx", "downsample_events", "_filter_events_resize"})
    exec(compile(code, "ref:downsample_events.py", "exec"), ns)

    class _H5Mod:
        File = staticmethod(lambda file, mode="r": _FakeH5(file))
    ns2 = {"np": np, "h5py": _H5Mod}
    code2 = _extract(os.path.join(REF_ROOT, "src", "dagr", "data", "dsec_utils.py"), {"_load_events"})
    exec(compile(code2, "ref:dsec_utils.py", "exec"), ns2)
    ns["_load_events"] = ns2["_load_events"]
    return ns


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "scripts", "downsample_events.py"))
