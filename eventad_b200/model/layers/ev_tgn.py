"""EV_TGN: events -> spatiotemporal graph, reference API (/root/reference/src/dagr/model/layers/ev_tgn.py:7-58)."""
import torch

from ...data import Batch
from ...graph.ev_graph import SlidingWindowGraph


def _get_value_as_int(obj, key):
    val = getattr(obj, key)
    return val if type(val) is int else val[0]


def denormalize_pos(events):
    if hasattr(events, "pos_denorm"):
        return events.pos_denorm
    denorm = torch.tensor([int(events.width[0]), int(events.height[0]), int(events.time_window[0])],
                          device=events.pos.device)
    return (denorm.view(1, -1) * events.pos + 1e-3).int()


class EV_TGN(torch.nn.Module):
    def __init__(self, args):
        torch.nn.Module.__init__(self)
        self.radius = args.radius
        self.max_neighbors = args.max_neighbors
        self.max_queue_size = 128
        self.graph_creators = None

    def init_graph_creator(self, data):
        delta_t_us = int(self.radius * _get_value_as_int(data, "time_window"))
        radius = int(self.radius * _get_value_as_int(data, "width") + 1)
        batch_size = data.num_graphs
        width = int(_get_value_as_int(data, "width"))
        height = int(_get_value_as_int(data, "height"))
        self.graph_creators = SlidingWindowGraph(width=width, height=height, max_num_neighbors=self.max_neighbors,
                                                 max_queue_size=self.max_queue_size, batch_size=batch_size,
                                                 radius=radius, delta_t_us=delta_t_us)

    def forward(self, events, reset=True):
        if events.batch is None:
            events = Batch.from_data_list([events])
        if self.graph_creators is None:
            self.init_graph_creator(events)
        else:
            if reset:
                self.graph_creators.reset()
            # the reference sizes its FIFO tensor by the first call's batch; grow instead of faulting
            if events.num_graphs > self.graph_creators.batch_size:
                self.graph_creators.batch_size = events.num_graphs
        pos = denormalize_pos(events)
        events.edge_index = self.graph_creators.forward(events.batch.int(), pos, delete_nodes=False, collect_edges=reset)
        events.edge_index = events.edge_index.long()
        g = self.graph_creators
        if g.csr is not None and reset:
            events._csr = dict(rowptr=g.csr[0], src=g.csr[1], eid=None,
                               edge_index_ptr=(events.edge_index.data_ptr(), tuple(events.edge_index.shape)))
        return events
