"""AsyncGraph / SlidingWindowGraph with the reference's constructor, attributes and call signatures
(/root/reference/src/dagr/graph/ev_graph.py:18-166), running on the sm_100a kernels of libeventad_b200.so.

The reference keeps a `[B, Q, H, W]` int32 FIFO tensor (39 MB per 320x240 sample) that it clears and re-fills per
call; here a call after `reset()` is served by `evad_graph_build` (pixel-bucket sort + neighbour scan, graph.cu) and
produces the identical `edge_index`.  Besides the reference's return value the dst-major CSR of the last call is
kept in `self.csr = (rowptr[N+1] int32, src int32, num_edges int32[1])` for the fused layers.
"""
import torch

from .. import _lib


class AsyncGraph:
    def __init__(self, width=640, height=480, batch_size=1, max_num_neighbors=16, max_queue_size=512, radius=7,
                 delta_t_us=600000):
        self.radius = radius
        self.delta_t_us = delta_t_us
        self.max_index = 0
        self.min_index = 0
        self.max_queue_size = max_queue_size
        self.max_num_neighbors = max_num_neighbors
        self.width = width
        self.height = height
        self.batch_size = batch_size
        self.device = None
        self.edges = torch.zeros((2, 0), dtype=torch.long)
        self._ts_buf, self._ts_lo, self._ts_hi = torch.zeros((0,), dtype=torch.int32), 0, 0
        self._ts_pending, self._ts_npend = [], 0
        self.csr = None
        self._ws = _lib.Workspace()
        self._status = None
        self._meta = None
        self._stream_state = None
        self._history = []          # (batch, pos, base_index) of reset-mode calls, replayed if streaming starts

    # `all_timestamps` (ev_graph.py:75: torch.cat on every call, an O(resident) copy) is a window [lo, hi) of a buffer
    # that grows geometrically: appending a micro-batch costs O(micro-batch), deleting the oldest nodes moves `lo`
    # The kernels never read it (the ring carries the timestamps), so appended chunks are only queued and written into
    # the buffer when somebody looks at the attribute.
    @property
    def all_timestamps(self):
        self._ts_flush()
        return self._ts_buf[self._ts_lo:self._ts_hi]

    @all_timestamps.setter
    def all_timestamps(self, t):
        self._ts_buf, self._ts_lo, self._ts_hi, self._ts_pending, self._ts_npend = t, 0, t.shape[0], [], 0

    @property
    def num_resident(self):
        """len(all_timestamps) without materialising it."""
        return self._ts_hi - self._ts_lo + self._ts_npend

    def _ts_append(self, t):
        self._ts_pending.append(t)
        self._ts_npend += t.shape[0]
        if len(self._ts_pending) > 4096:
            self._ts_flush()

    def _ts_flush(self):
        if not self._ts_pending:
            return
        new = torch.cat(self._ts_pending) if len(self._ts_pending) > 1 else self._ts_pending[0]
        n = new.shape[0]
        if self._ts_hi + n > self._ts_buf.shape[0]:
            live = self._ts_hi - self._ts_lo
            buf = torch.empty(max(2 * (live + n), 1 << 16), dtype=torch.int32, device=new.device)
            buf[:live] = self._ts_buf[self._ts_lo:self._ts_hi]
            self._ts_buf, self._ts_lo, self._ts_hi = buf, 0, live
        self._ts_buf[self._ts_hi:self._ts_hi + n] = new
        self._ts_hi += n
        self._ts_pending, self._ts_npend = [], 0

    def _ts_delete(self, n):
        if self._ts_npend and self._ts_lo + n > self._ts_hi:
            self._ts_flush()
        self._ts_lo = min(self._ts_lo + n, self._ts_hi)

    def initialize(self, n_ev, device):
        self.edges = torch.zeros((2, 0), dtype=torch.long, device=device)
        self.all_timestamps = torch.zeros((0,), dtype=torch.int32, device=device)
        self._meta = torch.zeros(2, dtype=torch.int32, device=device)   # [edge count of an append, status word]
        self._status = self._meta[1:2]

    def reset(self):
        self.edges = torch.zeros((2, 0), dtype=torch.long, device=self.device)
        self.all_timestamps = torch.zeros((0,), dtype=torch.int32, device=self.device)
        self.max_index = 0
        self.min_index = 0
        self.csr = None
        self._stream_state = None
        self._history = []

    # ------------------------------------------------------------------ device-resident build (no host sync)
    def build_csr(self, batch, pos, want_edge_index=False, want_cells=False, want_cls=False):
        """Enqueues evad_graph_build for a fresh graph.  Returns dict(rowptr, src, num_edges, edge_index, cell_perm,
        cell_start); all sizes are upper bounds, `num_edges` lives on the device."""
        if not (batch.is_cuda and pos.is_cuda):
            raise RuntimeError("batch must be a CUDA tensor")            # CHECK_CUDA, ev_graph.cu:9
        if batch.device != pos.device:
            raise RuntimeError("batch and pos must be in same CUDA device")
        assert batch.dtype == torch.int32, [type(batch), batch.dtype]   # ev_graph.py:73
        batch = batch.contiguous()
        pos = pos.contiguous()
        if pos.dtype != torch.int32:
            pos = pos.int()
        dev = batch.device
        n = batch.shape[0]
        if self.device is None:
            self.device = dev
            self.initialize(n, dev)
        K = int(self.max_num_neighbors)
        L = _lib.lib()
        rowptr = torch.empty(n + 1, dtype=torch.int32, device=dev)
        src = torch.empty(max(n * K, 1), dtype=torch.int32, device=dev)
        num_edges = torch.empty(1, dtype=torch.int32, device=dev)
        edge_index = torch.empty((2, max(n * K, 1)), dtype=torch.int64, device=dev) if want_edge_index else None
        ncells = self.batch_size * self.height * self.width
        cell_perm = torch.empty(max(n, 1), dtype=torch.int32, device=dev) if want_cells else None
        cell_start = torch.empty(ncells + 2, dtype=torch.int32, device=dev) if want_cells else None
        edge_cls = torch.empty(max(n * K, 1), dtype=torch.uint8, device=dev) if want_cls else None
        nbytes = L.evad_graph_ws_bytes(n, self.batch_size, self.height, self.width, K)
        ws = self._ws.get(nbytes, dev)
        from .. import ops
        with ops._timed("graph_build"):
          rc = L.evad_graph_build(_lib.ptr(batch), _lib.ptr(pos), n, int(self.batch_size), int(self.height),
                                int(self.width), int(self.radius), int(self.delta_t_us), int(self.max_queue_size), K,
                                _lib.ptr(rowptr), _lib.ptr(src), _lib.ptr(num_edges), _lib.ptr(edge_index),
                                max(n * K, 1), _lib.ptr(cell_perm), _lib.ptr(cell_start), _lib.ptr(edge_cls),
                                _lib.ptr(self._status),
                                _lib.ptr(ws), ws.numel(), _lib.stream())
        _lib.check(rc, "evad_graph_build")
        return dict(rowptr=rowptr, src=src, num_edges=num_edges, edge_index=edge_index, cell_perm=cell_perm,
                    cell_start=cell_start, edge_cls=edge_cls)

    def build_rows(self, batch, pos, want_cls=True):
        """Enqueues evad_graph_build_rows for a fresh graph: fixed-stride rows nbr [n, K] / count [n] / nbr_cls [n, K]
        (+ cell_perm), no scan and no compaction.  Same edges and order as build_csr."""
        if not (batch.is_cuda and pos.is_cuda):
            raise RuntimeError("batch must be a CUDA tensor")
        assert batch.dtype == torch.int32, [type(batch), batch.dtype]
        batch, pos = batch.contiguous(), pos.contiguous()
        if pos.dtype != torch.int32:
            pos = pos.int()
        dev, n = batch.device, batch.shape[0]
        if self.device is None:
            self.device = dev
            self.initialize(n, dev)
        K = int(self.max_num_neighbors)
        L = _lib.lib()
        nbr = torch.empty((max(n, 1), K), dtype=torch.int32, device=dev)
        count = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        nbr_cls = torch.empty((max(n, 1), K), dtype=torch.uint8, device=dev) if want_cls else None
        cell_perm = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        ws = self._ws.get(L.evad_graph_ws_bytes(n, self.batch_size, self.height, self.width, K), dev)
        from .. import ops
        with ops._timed("graph_build"):
            rc = L.evad_graph_build_rows(_lib.ptr(batch), _lib.ptr(pos), n, int(self.batch_size), int(self.height),
                                         int(self.width), int(self.radius), int(self.delta_t_us), int(self.max_queue_size),
                                         K, _lib.ptr(nbr), _lib.ptr(count), _lib.ptr(nbr_cls), _lib.ptr(cell_perm), None,
                                         _lib.ptr(self._status), _lib.ptr(ws), ws.numel(), _lib.stream())
        _lib.check(rc, "evad_graph_build_rows")
        return dict(nbr=nbr, count=count, nbr_cls=nbr_cls, cell_perm=cell_perm, K=K)

    def forward(self, batch, pos, collect_edges=True):
        n_ev = len(batch)
        if self.device is None:
            if not batch.is_cuda:
                raise RuntimeError("batch must be a CUDA tensor")
            self.device = batch.device
            self.initialize(n_ev, self.device)
        if n_ev == 0:
            return torch.zeros((2, 0), device=self.device, dtype=torch.int32)   # ev_graph.py:71
        assert type(batch) is torch.Tensor and batch.dtype == torch.int32, [type(batch), batch.dtype]
        pos = pos.int() if pos.dtype != torch.int32 else pos

        if self.max_index == 0 and self.min_index == 0 and self._stream_state is None:
            out = self.build_csr(batch, pos, want_edge_index=True)
            self._history = [(batch, pos, 0)]
            self._ts_append(pos[:, 2])
            self.max_index += n_ev
            e = int(out["num_edges"].item())                    # the reference syncs here too (mask compaction)
            _lib.check_status(self._status, "graph build")
            self.csr = (out["rowptr"], out["src"][:e], out["num_edges"])
            edge_indices = out["edge_index"][:, :e]
        else:
            from .stream import append_events
            edge_indices = append_events(self, batch, pos)
        if collect_edges:
            self.edges = torch.cat([self.edges, edge_indices], dim=-1)
        return edge_indices


class SlidingWindowGraph(AsyncGraph):
    def __init__(self, width=640, height=480, batch_size=1, max_num_neighbors=16, max_queue_size=1024, radius=7,
                 delta_t_us=600000):
        AsyncGraph.__init__(self, width, height, batch_size, max_num_neighbors, max_queue_size, radius, delta_t_us)

    @property
    def init(self):
        return self.num_resident > 0

    def delete_nodes(self, n_delete, delete_edges=True, return_edges=True):
        self._ts_delete(n_delete)
        self.min_index += n_delete
        if self._stream_state is not None:
            self._stream_state.delete(n_delete)
        deleted_edges = None
        if self.edges.shape[1] == 0:                          # nothing collected (collect_edges=False): no masks to evaluate
            deleted_edges = self.edges if delete_edges else None
        else:
            if delete_edges:
                mask = (self.edges[0] < n_delete) | (self.edges[1] < n_delete)
                deleted_edges = self.edges[:, mask].clone()
                self.edges = self.edges[:, ~mask]
            self.edges.add_(-n_delete)
        if delete_edges and return_edges:
            return deleted_edges

    def forward(self, batch, pos, return_node_counts=False, return_total_edges=False, delete_nodes=True,
                collect_edges=True):
        n_delete = len(batch) if self.init else 0
        edges = AsyncGraph.forward(self, batch, pos, collect_edges=collect_edges)
        if return_total_edges:
            total_edges = self.edges.clone()
        if return_node_counts:
            tot_nodes = self.num_resident
        ret = [edges]
        if delete_nodes:
            ret.append(self.delete_nodes(n_delete))
        if return_total_edges:
            ret.append(total_edges)
        if return_node_counts:
            ret.append([n_delete, len(batch), tot_nodes])
        if len(ret) == 1:
            ret = ret[0]
        return ret
