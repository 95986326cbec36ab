"""EventADModel: frozen DAGR event branch -> per-box graph features -> per-track GRUs + soft attention -> 2-class logits,
with the reference's constructor, module tree and state-dict keys (/root/reference/models/EventAD.py:11-176), so
`best_rol.pth['model']` loads unchanged:
    dagr_model.*, gru_net_event.gru.*, gru_net_cor.gru.*, fusion_module.{event_proj,coord_proj,attention,fusion.0,fusion.3}.*,
    soft_attention.weight, soft_attention_cor.weight.

What runs underneath is re-designed:
  * `extract_features` (EventAD.py:335-404): the python loops over samples and boxes with four `.item()` syncs per box
    become two launches of `evad_bbox_pool` (csrc/bbox_pool.cu), one per frame list; no host synchronisation.
  * `forward` (EventAD.py:178-333): the reference walks track ids 1..30 per sample and calls both GRUs, the fusion MLP and
    the loss once per (sample, track) on [1, 1, C] tensors (hundreds of tiny launches and `.item()` calls).  Here a sample
    ("frame" of the clip) is ONE batched step over all 30 track slots: a validity mask replaces the `continue`s, the
    hidden-state dictionaries become two state tensors [layers, 30, H] that only valid tracks update, and the soft
    attention is a masked softmax over the valid tracks.  The recurrence over the samples of a batch stays sequential,
    as in the reference (hidden states flow from sample b to sample b + 1).
  The arithmetic per valid (sample, track) is the reference's; rows of a batched GEMM round like the reference's
  single-row products to ~1e-6 (parity bar of north_star: logits rtol 1e-4 / atol 1e-5).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _lib, ops


class Fuse_GRUNet(nn.Module):
    """EventAD.py:11-59: two projections to 256, concat, MLP to 2 logits.  `attention` is constructed and never called
    by the reference (its keys exist in the checkpoint)."""

    def __init__(self, event_dim, coord_dim):
        super().__init__()
        self.event_proj = nn.Linear(event_dim, 256)
        self.coord_proj = nn.Linear(coord_dim, 256)
        self.attention = nn.MultiheadAttention(256, 4)
        self.fusion = nn.Sequential(nn.Linear(256 * 2, 256), nn.ReLU(), nn.Dropout(0.3), nn.Linear(256, 2))

    def forward(self, event_feature, coord_feature):
        return self.fusion(torch.cat([self.event_proj(event_feature), self.coord_proj(coord_feature)], dim=1))


def _init_gru(gru):
    for name, param in gru.named_parameters():
        if "bias" in name:
            nn.init.constant_(param, 0.0)
        elif "weight_ih" in name:
            nn.init.kaiming_normal_(param)
        elif "weight_hh" in name:
            nn.init.orthogonal_(param)


class Cor_GRUNet(nn.Module):
    def __init__(self, input_dim, hidden_dim, n_layers):
        super().__init__()
        self.hidden_dim, self.n_layers = hidden_dim, n_layers
        self.gru = nn.GRU(input_dim, hidden_dim, n_layers, batch_first=True)
        _init_gru(self.gru)

    def forward(self, x, h):
        return self.gru(x, h)


class Event_GRUNet(nn.Module):
    def __init__(self, input_dim, hidden_dim, n_layers):
        super().__init__()
        self.hidden_dim, self.n_layers = hidden_dim, n_layers
        self.gru = nn.GRU(input_dim, hidden_dim, n_layers, batch_first=True, dropout=0.3)
        _init_gru(self.gru)

    def forward(self, x, h):
        return self.gru(x, h)


class SpatialAttention(nn.Module):
    """EventAD.py:98-138: alpha = softmax over the objects of a frame of tanh(h) @ weight; h <- h * alpha."""

    def __init__(self, h_dim):
        super().__init__()
        self.weight = nn.Parameter(torch.Tensor(h_dim, 1))
        nn.init.kaiming_normal_(self.weight, a=math.sqrt(5))

    def masked(self, h, valid):
        """h [layers, T, H], valid bool [T] -> h * alpha with the softmax taken over the valid objects only."""
        score = torch.matmul(torch.tanh(h), self.weight)                    # [layers, T, 1]
        score = score.masked_fill(~valid.view(1, -1, 1), float("-inf"))
        alpha = torch.softmax(score, 1)
        return h * torch.nan_to_num(alpha)                                  # no valid object: all -inf -> zeros (unused)

    def forward(self, h_all_in):
        """Dictionary form of the reference (track id -> [layers, 1, H])."""
        keys = list(h_all_in)
        if not keys:
            return h_all_in
        h = torch.cat([h_all_in[k] for k in keys], dim=1)
        out = self.masked(h, torch.ones(h.shape[1], dtype=torch.bool, device=h.device))
        return {k: out[:, i:i + 1].contiguous() for i, k in enumerate(keys)}


class EventADModel(nn.Module):
    def __init__(self, dagr_model, x_dim=64, h_dim=256, n_frames=2, fps=30):
        super().__init__()
        self.dagr_model = dagr_model
        for p in self.dagr_model.parameters():                 # EventAD.py:149-150: DAGR stays frozen
            p.requires_grad = False
        self.gru_net_event = Event_GRUNet(x_dim, h_dim, n_layers=2)
        self.gru_net_cor = Cor_GRUNet(4, 32, n_layers=1)
        self.fusion_module = Fuse_GRUNet(h_dim, 32)
        self.soft_attention = SpatialAttention(h_dim)
        self.soft_attention_cor = SpatialAttention(32)
        self.ce_loss = nn.CrossEntropyLoss()
        self.batch_feature_cache = None
        self.batch_input_hash = None
        self.emulate_feature_cache = True      # EventAD.py:349-354: features are reused while hash(str(x.sum())) repeats
        self.n_frames = n_frames
        self.fps = fps
        self.max_boxes = 30

    # ------------------------------------------------------------------------------------------ features
    def pool_boxes(self, out4, data, batch_size):
        """batch_features [B, 2, max_boxes + 1, C] from the out4 graph (EventAD.py:376-397) on the device."""
        L = _lib.lib()
        feat = out4.x.float().contiguous()
        pos = out4.pos.float().contiguous()
        nb = out4.batch.int().contiguous()
        dev = feat.device
        C = feat.shape[1]
        out = torch.zeros((batch_size, 2, self.max_boxes + 1, C), dtype=torch.float32, device=dev)
        width = torch.as_tensor(data.width, device=dev).float().view(-1).contiguous()
        height = torch.as_tensor(data.height, device=dev).float().view(-1).contiguous()
        if feat.shape[0] == 0:
            return out
        for frame, bkey, bbkey in ((1, "bbox", "bbox_batch"), (0, "bbox0", "bbox0_batch")):
            bb, bbb = getattr(data, bkey, None), getattr(data, bbkey, None)
            if bb is None or bbb is None or bb.shape[0] == 0:
                continue
            bb = bb.float().contiguous()
            bbb = bbb.long().contiguous()
            _lib.check(L.evad_bbox_pool(_lib.ptr(feat), feat.stride(0), C, _lib.ptr(pos), pos.stride(0), _lib.ptr(nb),
                                        feat.shape[0], None, _lib.ptr(bb), bb.stride(0), _lib.ptr(bbb), bb.shape[0], frame,
                                        _lib.ptr(width), _lib.ptr(height), 2, self.max_boxes, _lib.ptr(out), _lib.stream()),
                       "evad_bbox_pool")
        return out

    def extract_features(self, data):
        """EventAD.py:335-404 -> [B, n_frames = 2, max_boxes + 1, C]; frame 1 = current boxes, frame 0 = previous boxes."""
        if self.emulate_feature_cache:
            x_sum = data.x.sum().item() if getattr(data, "x", None) is not None else 0
            input_hash = hash(str(x_sum))
            if self.batch_feature_cache is not None and self.batch_input_hash == input_hash:
                return self.batch_feature_cache
        with torch.no_grad():
            self.dagr_model.eval()
            out4 = self.dagr_model.extract_features(data)[1]
            batch_features = self.pool_boxes(out4, data, data.num_graphs)
        if self.emulate_feature_cache:
            self.batch_feature_cache, self.batch_input_hash = batch_features, input_hash
        return batch_features

    # ------------------------------------------------------------------------------------------ head
    def box_table(self, data, labels, batch_size, device):
        """Per (sample, track slot): has a current-frame box, its normalised [x, y, w, h] and its label -- the FIRST box
        of the sample with that track id, as `torch.where(...)[0][0]` picks it (EventAD.py:239-256).  No host sync."""
        T = self.max_boxes
        has = torch.zeros((batch_size, T + 1), dtype=torch.bool, device=device)
        coord = torch.zeros((batch_size, T + 1, 4), dtype=torch.float32, device=device)
        target = torch.zeros((batch_size, T + 1), dtype=torch.long, device=device)
        bb = getattr(data, "bbox", None)
        if bb is None or getattr(data, "bbox_batch", None) is None or bb.shape[0] == 0:
            return has, coord, target
        bbb = data.bbox_batch.long()
        # `data.bbox[:, 5] == track_id` compares the fp32 id with the integer: only exact integers 1..T match
        tid_f = bb[:, 5]
        tid = tid_f.long().clamp(0, T)
        ok = (tid_f == tid.to(tid_f.dtype)) & (tid >= 1)
        m = bb.shape[0]
        order = torch.arange(m, device=device)
        # rows that do not count go to a spare slot behind the table: fixed shapes, no host sync (CUDA-graph capturable)
        key = torch.where(ok, bbb * (T + 1) + tid, torch.full_like(bbb, batch_size * (T + 1)))
        first = torch.full((batch_size * (T + 1) + 1,), m, dtype=torch.long, device=device)
        first.scatter_reduce_(0, key, order, reduce="amin", include_self=True)
        first = first[:-1].view(batch_size, T + 1)
        has = first < m
        idx = first.clamp(max=m - 1)
        width = torch.as_tensor(data.width, device=device).double().view(-1, 1)
        height = torch.as_tensor(data.height, device=device).double().view(-1, 1)
        b4 = bb[idx][..., :4].double()                          # python-float arithmetic of the reference, then fp32
        coord = torch.stack([b4[..., 0] / width, b4[..., 1] / height, b4[..., 2] / width, b4[..., 3] / height], -1).float()
        if labels is not None:
            target = labels[idx].long()
        return has, coord, target

    def head(self, batch_features, data, labels=None):
        """Batched form of EventAD.py:199-333.  Returns (logits [B, T + 1, 2], valid [B, T + 1], target [B, T + 1],
        loss): slot t of sample b is valid where the reference produces an output for track id t."""
        B = batch_features.shape[0]
        dev = batch_features.device
        T = self.max_boxes
        has, coord, target = self.box_table(data, labels, B, dev)
        curr = batch_features[:, 1]                                             # [B, T + 1, C]
        valid = has & (curr.sum(-1) != 0)                                       # EventAD.py:229 and :239-260
        valid[:, 0] = False
        he = torch.zeros((self.gru_net_event.n_layers, T + 1, self.gru_net_event.hidden_dim), device=dev)
        hc = torch.zeros((self.gru_net_cor.n_layers, T + 1, self.gru_net_cor.hidden_dim), device=dev)
        logits = []
        # fp32 GRUs: cuDNN's RNN kernels would otherwise use TF32 tensor cores (10-bit mantissa) and miss the score
        # tolerance of north_star (rtol 1e-4 / atol 1e-5) by a factor of three
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            for b in range(B):                                                  # the recurrence over the samples of a batch
                v = valid[b]
                eo, he_out = self.gru_net_event(curr[b].unsqueeze(1), he)       # [T+1, 1, H], [layers, T+1, H]
                co, hc_out = self.gru_net_cor(coord[b].unsqueeze(1), hc)
                logits.append(self.fusion_module(eo[:, -1, :], co[:, -1, :]))
                vm = v.view(1, -1, 1)
                he = torch.where(vm, self.soft_attention.masked(he_out, v), he)     # only tracks seen in this sample move on
                hc = torch.where(vm, self.soft_attention_cor.masked(hc_out, v), hc)
        logits = torch.stack(logits)                                            # [B, T + 1, 2]
        per = F.cross_entropy(logits.view(-1, 2), target.view(-1), reduction="none").view(B, T + 1)
        loss = (per * valid.to(per.dtype)).sum()
        return logits, valid, target, loss

    def forward(self, data, labels=None, testing=False):
        """Reference signature and return structure (EventAD.py:178-333): (losses, outputs, labels) with `outputs` a list
        over the samples that produced at least one prediction, each a list of [1, 2] logits in track-id order."""
        batch_features = self.extract_features(data)
        logits, valid, target, loss = self.head(batch_features, data, labels)
        return self.collect(logits, valid, target, loss)

    @staticmethod
    def collect(logits, valid, target, loss):
        """The reference's return structure from the batched head's tensors (one host read)."""
        losses = {"cross_entropy": loss}
        v = valid.cpu()                                                         # the one host read of the head
        all_outputs, all_labels = [], []
        for b in range(v.shape[0]):
            ts = v[b].nonzero().view(-1).tolist()
            if ts:
                all_outputs.append([logits[b, t].unsqueeze(0) for t in ts])
                all_labels.append([target[b, t].view(1) for t in ts])
        if not bool(v.any()):
            losses["cross_entropy"] = torch.tensor(0.0, device=logits.device, requires_grad=True)
        return losses, all_outputs, all_labels
