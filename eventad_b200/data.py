"""Minimal PyG-style `Data` / `Batch` attribute bags.

torch_geometric is not a dependency of this package; the reference's modules only use `Data`/`Batch` as mutable
attribute bags (`data.x = ...`, `hasattr(data, "pos_denorm")`, `data.num_graphs`, `Batch.from_data_list`;
/root/reference/src/dagr/model/layers/ev_tgn.py:39-58, pooling.py:79-83, model/utils.py:158-166).  Objects of the
real torch_geometric classes are accepted everywhere, because only attribute access is used.
"""
import torch


class Data:
    def __init__(self, x=None, edge_index=None, edge_attr=None, pos=None, batch=None, **kwargs):
        self.x = x
        self.edge_index = edge_index
        self.edge_attr = edge_attr
        self.pos = pos
        self.batch = batch
        for k, v in kwargs.items():
            setattr(self, k, v)

    def keys(self):
        return [k for k, v in self.__dict__.items() if not k.startswith("_") and v is not None]

    def __contains__(self, key):
        return key in self.__dict__ and self.__dict__[key] is not None

    @property
    def num_nodes(self):
        if self.x is not None:
            return self.x.shape[0]
        return self.pos.shape[0] if self.pos is not None else 0

    @property
    def num_edges(self):
        return 0 if self.edge_index is None else self.edge_index.shape[1]

    @property
    def num_graphs(self):
        if "_num_graphs" in self.__dict__ and self.__dict__["_num_graphs"] is not None:
            return self.__dict__["_num_graphs"]
        if self.batch is None:
            return 1
        return int(self.batch.max().item()) + 1 if self.batch.numel() > 0 else 0

    @num_graphs.setter
    def num_graphs(self, v):
        self.__dict__["_num_graphs"] = v

    def to(self, device, non_blocking=False):
        for k, v in list(self.__dict__.items()):
            if torch.is_tensor(v):
                self.__dict__[k] = v.to(device, non_blocking=non_blocking)
            elif isinstance(v, (list, tuple)) and len(v) > 0 and all(torch.is_tensor(t) for t in v):
                self.__dict__[k] = type(v)(t.to(device, non_blocking=non_blocking) for t in v)
        return self

    def clone(self):
        out = self.__class__.__new__(self.__class__)
        for k, v in self.__dict__.items():
            out.__dict__[k] = v.clone() if torch.is_tensor(v) else v
        return out

    def __repr__(self):
        parts = []
        for k in self.keys():
            v = getattr(self, k)
            parts.append(f"{k}={list(v.shape)}" if torch.is_tensor(v) else f"{k}={v!r}")
        return f"{self.__class__.__name__}({', '.join(parts)})"


class Batch(Data):
    @classmethod
    def from_data_list(cls, data_list):
        """Concatenates node-level tensors, offsets edge_index, builds `batch` (PyG semantics for the keys used on
        the path: x, pos, t, edge_index, edge_attr; per-sample scalars such as width/height/time_window are stacked)."""
        out = cls()
        n_off = 0
        cat = {}
        batch = []
        for b, d in enumerate(data_list):
            n = d.num_nodes
            for k, v in d.__dict__.items():
                if k.startswith("_") or v is None or k == "batch":
                    continue
                if k == "edge_index":
                    cat.setdefault(k, []).append(v + n_off)
                elif torch.is_tensor(v):
                    cat.setdefault(k, []).append(v if v.dim() > 0 else v.view(1))
                else:
                    cat.setdefault(k, []).append(torch.as_tensor([v]) if isinstance(v, (int, float)) else v)
            ref = d.x if d.x is not None else d.pos
            batch.append(torch.full((n,), b, dtype=torch.long, device=ref.device if ref is not None else None))
            n_off += n
        for k, vs in cat.items():
            if all(torch.is_tensor(v) for v in vs):
                setattr(out, k, torch.cat(vs, dim=-1 if k == "edge_index" else 0))
            else:
                setattr(out, k, vs)
        out.batch = torch.cat(batch) if batch else None
        out.num_graphs = len(data_list)
        return out

    def to_data_list(self):
        res = []
        for b in range(self.num_graphs):
            m = self.batch == b
            d = Data(x=None if self.x is None else self.x[m], pos=None if self.pos is None else self.pos[m])
            res.append(d)
        return res
