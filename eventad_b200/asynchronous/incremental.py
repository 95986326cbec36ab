"""Incremental ("asynchronous") forward of the event branch: SURVEY 8 f-1, BASELINE configs[4].

Reference behaviour: `make_model_asynchronous` (/root/reference/src/dagr/asynchronous/__init__.py:39-110) turns every
layer into a stateful one that, on `forward(new_events, reset=False)`, recomputes only what the new events change
(asynchronous/conv.py:94-227: new nodes from their in-edges, old nodes by message differences; max_pool.py:123-243:
voxels that received nodes) and must agree with the dense forward on all events (evaluate_flops.py:82-165).

What the DAGR graph gives us for free: edges run from OLDER to NEWER events only (ev_graph.cu:57-59, the causal cut), so
appending events never changes an existing node of the full-resolution graph -- not its in-edges, not its input row,
hence not its layer-1 output.  The incremental update is therefore exact, not a difference scheme:

  graph     evad_stream_append   the new events' in-edges from the per-pixel ring (csrc/stream.cu)
  rows      evad_stream_rows     ... written into the resident fixed-stride rows nbr / count / nbr_cls at the new ids
  x0        evad_sample_concat_h2 on the new events only
  layer 1   evad_spline_conv_l1_h16 over perm = the new node ids (two launches: conv + skip, conv + residual)
  pool 1    evad_pool_inc_update folds the new nodes and their in-edges into per-voxel state (running maxima of
            [layer-1 output | sampled CNN map], serial position sums, counts, one bitmap row of coarse in-edges per
            voxel); evad_pool_inc_emit writes the pooled graph of everything seen so far, bit-identical to
            evad_voxel_pool over all resident nodes (max is order independent; appended nodes have the highest ids, so
            the position sums are continued in the dense kernel's order)
  layer2 .. the dense tail (Net._enqueue_tail) on the 2 240 voxels per window, captured once in a CUDA graph together
  layer5    with the emit step: its cost is independent of the micro-batch and of the number of resident events.

Results equal the dense fused forward on the accumulated events BIT FOR BIT (same kernels, same rows, same per-node
arithmetic; tests/test_incremental_gpu.py), which is stronger than the reference checker's 1e-3.  Not built: node deletion
inside a window (SlidingWindowGraph.delete_nodes; it would need the arg-max bookkeeping of max_pool.py): `update` is
append-only until the next `reset`.  EVAD_INCR_POOL=0 keeps pool 1 in the dense tail (evad_voxel_pool over all nodes).
"""
import ctypes
import os

import torch

from .. import _lib, ops
from ..graph import stream as gstream
from ..model.layers.conv import _fold


USE_TAIL_GRAPH = os.environ.get("EVAD_INCR_GRAPH", "1") != "0"   # 0: the dense tail as plain launches (debugging)
USE_INC_POOL = os.environ.get("EVAD_INCR_POOL", "1") != "0"      # 0: pool 1 recomputed over all resident nodes by the tail


class IncrementalBackbone:
    def __init__(self, net, capacity, feats=None, width=320, height=240, time_window=1_000_000, num_graphs=1,
                 device="cuda"):
        """net: eventad_b200 Net (eval, LUTs cached, on `device`); capacity: most events resident between two resets;
        feats: the CNN maps [B, C, h, w] of the window (None without image branch -- not supported by layer 1's kernel)."""
        self.net, self.cap, self.dev = net, int(capacity), torch.device(device)
        self.meta = (int(width), int(height), int(time_window), int(num_graphs))
        self.feats = [ops.to_channel_last(f.detach().float()) for f in feats] if feats is not None else None
        if not (net.use_image and self.feats is not None and net.conv_block1.in_channel == 19):
            raise RuntimeError("IncrementalBackbone: the layer-1 kernel needs the 19-channel input of the image configuration")
        c, dev = self.cap, self.dev
        self.n = 0
        self.x = torch.zeros((c, 1), dtype=torch.float32, device=dev)
        self.pos = torch.zeros((c, 3), dtype=torch.float32, device=dev)
        self.batch = torch.zeros(c, dtype=torch.int32, device=dev)
        self.pos_int = torch.zeros((c, 3), dtype=torch.int32, device=dev)
        self.count = torch.zeros(c, dtype=torch.int32, device=dev)
        self.nbr = torch.zeros((c, 16), dtype=torch.int32, device=dev)
        self.cls = torch.zeros((c, 16), dtype=torch.uint8, device=dev)
        self.x0 = torch.zeros((c, 20), dtype=torch.float32, device=dev)      # h2 words
        self.y1 = torch.zeros((c, 16), dtype=torch.float32, device=dev)      # h2 words: output of conv_block1.conv_block1
        self.skip = torch.zeros((c, 16), dtype=torch.float32, device=dev)
        self.xk = torch.zeros((c, 16), dtype=torch.float32, device=dev)      # layer-1 output
        self._packs = None
        # the dense tail is captured once in a CUDA graph over the full-capacity buffers; the number of resident events
        # is a device cell (every kernel of the tail takes its row count from device memory)
        self._n_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self._scratch = _lib.Scratch()
        self._tail_graph, self._tail_res = None, None
        self._pool = None
        if USE_INC_POOL:
            p = net.pool1
            P = ops.pool_params(p._vs, p._nv, self.meta[3], p._inv_wh, p.aggr, self.meta[0], self.meta[1], self.feats[1],
                                append_pos=True, ell_stride=16)
            self._pool = ops.PoolIncState(P, 16, c, dev)

    # ------------------------------------------------------------------------------------------------ pieces
    def _graph_creator(self):
        from ..data import Data
        w, h, tw, b = self.meta
        g = self.net.events_to_graph
        if g.graph_creators is None:
            g.init_graph_creator(Data(pos=self.pos[:1], width=[w], height=[h], time_window=[tw], _num_graphs=b))
        gc = g.graph_creators
        if gc.device is None:
            gc.device = self.dev
            gc.initialize(self.cap, self.dev)
        if int(gc.max_num_neighbors) != 16:
            raise RuntimeError("IncrementalBackbone: rows of 16 neighbours")
        return gc

    def _layer1(self, perm, n_compute):
        """conv_block1 (two SplineConvs + skip) for the nodes perm[0:n_compute] of the resident graph."""
        L = _lib.lib()
        layer = self.net.conv_block1
        b1, b2 = layer.conv_block1, layer.conv_block2
        if self._packs is None:
            relu = b1.activation is torch.nn.functional.relu
            if not relu or not layer.h16_chain_ok(self.nbr.view(-1), self.cls.view(-1), 16):
                raise RuntimeError("IncrementalBackbone: layer 1 does not fit evad_spline_conv_l1_h16")
            s1, h1 = _fold(b1.norm, b1.conv.bias)
            s2, h2 = _fold(b2.norm, b2.conv.bias)
            ss, hs = _fold(b2.norm_skip, b2.lin.mlp.bias)
            self._packs = dict(p1=b1.conv.l1_pack(b2.lin.mlp.weight), p2=b2.conv.l1_pack(),
                               e1=ops._epi_struct(scale=s1, shift=h1, relu=True), es=ops._epi_struct(scale=ss, shift=hs),
                               e2=ops._epi_struct(scale=s2, shift=h2, addend=self.skip, relu=True), keep=(s1, h1, s2, h2, ss, hs))
        pk, N = self._packs, self.n
        p1, p2 = pk["p1"], pk["p2"]
        ptr, status = _lib.ptr, self._gc._status
        with ops._timed('incr_layer1'):
            _lib.check(L.evad_spline_conv_l1_h16(
                ptr(self.count), ptr(self.nbr), ptr(self.cls), ptr(perm), n_compute, N, ptr(self.x0), 20, 20,
                ptr(p1["tapw"]), p1["tapw"].shape[0], ptr(p1["whi"]), ptr(p1["wlo"]), ptr(p1["shi"]), ptr(p1["slo"]),
                ctypes.byref(pk["e1"]), ptr(self.y1), 16, 1, ctypes.byref(pk["es"]), ptr(self.skip), 16, ptr(status),
                _lib.stream()), "evad_spline_conv_l1_h16 (incremental, conv 1)")
            _lib.check(L.evad_spline_conv_l1_h16(
                ptr(self.count), ptr(self.nbr), ptr(self.cls), ptr(perm), n_compute, N, ptr(self.y1), 16, 16,
                ptr(p2["tapw"]), p2["tapw"].shape[0], ptr(p2["whi"]), ptr(p2["wlo"]), None, None,
                ctypes.byref(pk["e2"]), ptr(self.xk), 16, 0, None, None, 0, ptr(status), _lib.stream()),
                "evad_spline_conv_l1_h16 (incremental, conv 2)")

    def _enqueue_tail(self):
        w, h, tw, b = self.meta
        cur = dict(x=self.xk, pos=self.pos, batch=self.batch, rowptr=self.count, src=self.nbr.view(-1), n_cap=self.cap,
                   n_dev=self._n_dev, e_cap=self.cap * 16, ell=16)
        saved = ops._SCRATCH
        ops._SCRATCH = self._scratch
        try:
            status = self._scratch.status(self.pos.device)
            status.zero_()
            return self.net._enqueue_tail(cur, self.feats, b, w, h, status,
                                          pool1=self._pool.emit if self._pool is not None else None)
        finally:
            ops._SCRATCH = saved

    def _tail(self):
        w, h, tw, b = self.meta
        self._n_dev.fill_(self.n)
        if not USE_TAIL_GRAPH:
            res = self._enqueue_tail()
            return self._finish(self.net.finalize_fused(res))
        if self._tail_graph is None:
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):                           # warm-up: workspaces, folded BN, function attributes
                    self._enqueue_tail()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            self._tail_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._tail_graph):
                self._tail_res = self._enqueue_tail()
        self._tail_graph.replay()
        return self._finish(self.net.finalize_fused(self._tail_res))

    def _finish(self, outs):
        w, h, tw, b = self.meta
        for o in outs:
            o.width, o.height = torch.tensor([w] * b), torch.tensor([h] * b)
        return outs

    def _take(self, x, pos, batch):
        n0, n = self.n, pos.shape[0]
        if n0 + n > self.cap:
            raise RuntimeError(f"IncrementalBackbone: {n0 + n} events exceed the capacity {self.cap}")
        w, h, tw, b = self.meta
        pos = pos.float().contiguous()
        self.x[n0:n0 + n] = x.float().view(-1, 1)
        self.pos[n0:n0 + n] = pos
        self.batch[n0:n0 + n] = batch.int()
        pi = ops.denormalize_pos(pos, w, h, tw)
        self.pos_int[n0:n0 + n] = pi
        x0 = ops.sample_concat_h2(self.x[n0:n0 + n], self.pos[n0:n0 + n], self.batch[n0:n0 + n], self.feats[0], w, h,
                                  self._gc._status)
        if x0 is None:
            raise RuntimeError("IncrementalBackbone: layer-1 rows must have the [polarity | 16 CNN channels | xy] shape")
        self.x0[n0:n0 + n] = x0
        return pi

    # ------------------------------------------------------------------------------------------------ API
    @torch.no_grad()
    def reset(self, data):
        """Dense pass on the first events of a window (EV_TGN.forward(reset=True), ev_tgn.py:39-44); keeps the state."""
        gc = self._gc = self._graph_creator()
        gc._status.zero_()
        gc.batch_size = max(gc.batch_size, self.meta[3])
        gc.reset()
        self.n = 0
        n = data.pos.shape[0]
        pi = self._take(data.x, data.pos, data.batch)
        self.n = n
        bi = self.batch[:n]
        graph = gc.build_rows(bi, pi, want_cls=True)
        self.count[:n] = graph["count"]
        self.nbr[:n] = graph["nbr"].view(n, 16)
        self.cls[:n] = graph["nbr_cls"].view(n, 16)
        gc.max_index, gc.all_timestamps, gc._history = n, pi[:, 2], [(bi.clone(), pi, 0)]
        self._layer1(graph["cell_perm"], n)
        if self._pool is not None:
            self._pool.reset()
            self._pool.update(0, n, self.pos, self.batch, self.xk, self.count, self.nbr, gc._status)
        return self._tail()

    @torch.no_grad()
    def update(self, data):
        """The new events of one micro-batch (EV_TGN.forward(reset=False), ev_tgn.py:45-56): outputs for ALL resident events."""
        gc, n0, n = self._gc, self.n, data.pos.shape[0]
        if n == 0:
            return self._tail()
        gc._status.zero_()
        pi = self._take(data.x, data.pos, data.batch)
        gstream.append_events(gc, self.batch[n0:n0 + n], pi)            # one host sync: edge count + status
        rowptr, src, _ = gc.csr
        self.n = n0 + n
        _lib.check(_lib.lib().evad_stream_rows(_lib.ptr(rowptr), _lib.ptr(src), n, n0, int(gc.min_index), _lib.ptr(self.pos_int),
                                               int(gc.radius), 16, _lib.ptr(self.nbr), _lib.ptr(self.count), _lib.ptr(self.cls),
                                               _lib.ptr(gc._status), _lib.stream()), "evad_stream_rows")
        perm = torch.arange(n0, n0 + n, dtype=torch.int32, device=self.dev)
        self._layer1(perm, n)
        if self._pool is not None:
            self._pool.update(n0, n, self.pos, self.batch, self.xk, self.count, self.nbr, gc._status)
        return self._tail()
