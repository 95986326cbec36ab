"""Streaming (append) mode of AsyncGraph / SlidingWindowGraph on the per-pixel ring of csrc/stream.cu.

The first call after `reset()` is served by the sort-based builder (graph.cu); if a later call arrives without a
reset (EV_TGN.forward(reset=False), /root/reference/src/dagr/model/layers/ev_tgn.py:45-56) the events seen so far are
replayed into the ring once (insert only) and from then on every call is one `evad_stream_append`."""
import torch

from .. import _lib


class StreamState:
    def __init__(self, graph, device):
        L = _lib.lib()
        self.B, self.H, self.W, self.Q = int(graph.batch_size), int(graph.height), int(graph.width), int(graph.max_queue_size)
        nbytes = L.evad_stream_state_bytes(self.B, self.H, self.W, self.Q)
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _lib.check(L.evad_stream_reset(_lib.ptr(self.buf), self.B, self.H, self.W, self.Q, _lib.stream()), "evad_stream_reset")

    def delete(self, n):
        pass  # node deletion is min_index bookkeeping only; ring entries < min_index terminate the walks

    def append(self, graph, batch, pos, base_index, do_search=True):
        L = _lib.lib()
        dev = batch.device
        n = batch.shape[0]
        K = int(graph.max_num_neighbors)
        batch = batch.contiguous()
        pos = pos.contiguous()
        rowptr = torch.empty(n + 1, dtype=torch.int32, device=dev) if do_search else None
        src = torch.empty(max(n * K, 1), dtype=torch.int32, device=dev) if do_search else None
        num_edges = graph._meta[0:1] if do_search else None      # next to the status word: one read-back for both
        ei = torch.empty((2, max(n * K, 1)), dtype=torch.int64, device=dev) if do_search else None
        ws = graph._ws.get(L.evad_stream_ws_bytes(n, K), dev)
        rc = L.evad_stream_append(_lib.ptr(self.buf), self.B, self.H, self.W, self.Q, _lib.ptr(batch), _lib.ptr(pos), n,
                                  int(base_index), int(graph.min_index), int(graph.radius), int(graph.delta_t_us), K,
                                  int(do_search), _lib.ptr(rowptr), _lib.ptr(src), _lib.ptr(num_edges), _lib.ptr(ei),
                                  max(n * K, 1), _lib.ptr(graph._status), _lib.ptr(ws), ws.numel(), _lib.stream())
        _lib.check(rc, "evad_stream_append")
        return rowptr, src, num_edges, ei


def append_events(graph, batch, pos):
    """AsyncGraph.forward for a call that finds resident events (ev_graph.py:75-103)."""
    dev = batch.device
    if graph._stream_state is None:
        graph._stream_state = StreamState(graph, dev)
        for b0, p0, base in graph._history:                 # replay earlier reset-mode calls into the ring
            graph._stream_state.append(graph, b0, p0, base, do_search=False)
        graph._history = []
    n_ev = batch.shape[0]
    base = graph.max_index
    rowptr, src, num_edges, ei = graph._stream_state.append(graph, batch, pos, base, do_search=True)
    graph._ts_append(pos[:, 2])
    graph.max_index += n_ev
    e, st = graph._meta.tolist()                              # the one host synchronisation of the call
    if st:
        _lib.check_status(graph._status, "graph append")
    graph.csr = (rowptr, src[:e], num_edges.clone())
    return ei[:, :e]
