"""ctypes front-end of oracle/evg_oracle.c -- TEST INFRASTRUCTURE ONLY (see that file's header).

`OracleSlidingWindowGraph` mirrors the host-side bookkeeping of the reference's
`AsyncGraph` / `SlidingWindowGraph` (/root/reference/src/dagr/graph/ev_graph.py:18-166):
accumulated `edges`, `all_timestamps`, `max_index`, `min_index`, `reset()`, `delete_nodes()`,
`forward(batch, pos, return_node_counts, return_total_edges, delete_nodes, collect_edges)`.
Arrays are numpy; the product path never imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libevg_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "evg_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.check_call(["make", "-s", "-C", _HERE], env=env)
    return _LIB_PATH


_asy = None


def asy_lib():
    """oracle/asy_oracle.c (CPU restatement of the reference's asy_tools kernels)."""
    global _asy
    if _asy is None:
        build()
        path = os.path.join(_HERE, "_build", "libasy_oracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "asy_oracle.c")):
            env = dict(os.environ)
            env.pop("CC", None)
            subprocess.check_call(["make", "-s", "-C", _HERE], env=env)
        A = ctypes.CDLL(path)
        p, i64, i, f = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float
        A.asyo_masked_isdiff.argtypes = [p, p, p, i64, i, f, f]
        A.asyo_masked_inplace_bn.argtypes = [p, p, p, p, p, p, p, i64, i, f]
        A.asyo_masked_lin.argtypes = [p, p, p, p, p, i64, i, i, i]
        _asy = A
    return _asy


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        L.evgo_create.restype = ctypes.c_void_p
        L.evgo_create.argtypes = [ctypes.c_int] * 5 + [ctypes.c_float, ctypes.c_float]
        L.evgo_destroy.argtypes = [ctypes.c_void_p]
        L.evgo_reset.argtypes = [ctypes.c_void_p]
        L.evgo_delete_nodes.argtypes = [ctypes.c_void_p, ctypes.c_int64]
        for f in ("evgo_max_index", "evgo_min_index", "evgo_num_timestamps"):
            getattr(L, f).restype = ctypes.c_int64
            getattr(L, f).argtypes = [ctypes.c_void_p]
        L.evgo_queue.restype = ctypes.POINTER(ctypes.c_int32)
        L.evgo_queue.argtypes = [ctypes.c_void_p]
        L.evgo_forward.restype = ctypes.c_int64
        L.evgo_forward.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                   ctypes.c_void_p, ctypes.c_void_p]
        L.evgo_spiral.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


def spiral(count: int) -> np.ndarray:
    dx = np.zeros(count, np.int32)
    dy = np.zeros(count, np.int32)
    lib().evgo_spiral(count, dx.ctypes.data, dy.ctypes.data)
    return np.stack([dx, dy], 1)


class OracleSlidingWindowGraph:
    """CPU twin of SlidingWindowGraph(width, height, batch_size, max_num_neighbors, max_queue_size, radius, delta_t_us)."""

    def __init__(self, width=640, height=480, batch_size=1, max_num_neighbors=16, max_queue_size=1024,
                 radius=7, delta_t_us=600000):
        self.width, self.height, self.batch_size = int(width), int(height), int(batch_size)
        self.max_num_neighbors, self.max_queue_size = int(max_num_neighbors), int(max_queue_size)
        self.radius, self.delta_t_us = radius, delta_t_us
        self._h = lib().evgo_create(self.width, self.height, self.batch_size, self.max_num_neighbors,
                                    self.max_queue_size, float(radius), float(delta_t_us))
        self.edges = np.zeros((2, 0), np.int64)
        self.all_timestamps = np.zeros((0,), np.int32)

    def __del__(self):
        try:
            if self._h:
                lib().evgo_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def max_index(self):
        return int(lib().evgo_max_index(self._h))

    @property
    def min_index(self):
        return int(lib().evgo_min_index(self._h))

    @property
    def init(self):
        return len(self.all_timestamps) > 0

    def queue(self) -> np.ndarray:
        n = self.batch_size * self.max_queue_size * self.height * self.width
        arr = np.ctypeslib.as_array(lib().evgo_queue(self._h), shape=(n,))
        return arr.reshape(self.batch_size, self.max_queue_size, self.height, self.width).copy()

    def reset(self):
        lib().evgo_reset(self._h)
        self.edges = np.zeros((2, 0), np.int64)
        self.all_timestamps = np.zeros((0,), np.int32)

    def _async_forward(self, batch, pos, collect_edges=True):
        batch = np.ascontiguousarray(batch, np.int32)
        pos = np.ascontiguousarray(pos, np.int32).reshape(-1, 3)
        n = len(batch)
        if n == 0:
            return np.zeros((2, 0), np.int32)  # ev_graph.py:71 (int32 for the empty case)
        cap = self.max_num_neighbors * n
        src = np.empty(cap, np.int64)
        dst = np.empty(cap, np.int64)
        e = lib().evgo_forward(self._h, batch.ctypes.data, pos.ctypes.data, n, src.ctypes.data, dst.ctypes.data)
        self.all_timestamps = np.concatenate([self.all_timestamps, pos[:, 2]])
        edges = np.stack([src[:e], dst[:e]])
        if collect_edges:
            self.edges = np.concatenate([self.edges, edges], axis=1)
        return edges

    def delete_nodes(self, n_delete, delete_edges=True, return_edges=True):
        self.all_timestamps = self.all_timestamps[n_delete:]
        lib().evgo_delete_nodes(self._h, int(n_delete))
        deleted = None
        if delete_edges:
            mask = (self.edges[0] < n_delete) | (self.edges[1] < n_delete)
            deleted = self.edges[:, mask].copy()
            self.edges = self.edges[:, ~mask]
        self.edges = self.edges - n_delete
        if delete_edges and return_edges:
            return deleted

    def forward(self, batch, pos, return_node_counts=False, return_total_edges=False, delete_nodes=True,
                collect_edges=True):
        n_delete = len(batch) if self.init else 0
        edges = self._async_forward(batch, pos, collect_edges=collect_edges)
        ret = [edges]
        if return_total_edges:
            total_edges = self.edges.copy()
        if return_node_counts:
            tot_nodes = len(self.all_timestamps)
        if delete_nodes:
            ret.append(self.delete_nodes(n_delete))
        if return_total_edges:
            ret.append(total_edges)
        if return_node_counts:
            ret.append([n_delete, len(batch), tot_nodes])
        return ret[0] if len(ret) == 1 else ret
