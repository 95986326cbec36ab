"""Data-parallel training step of EventAD (BASELINE configs[3]; reference loop: /root/reference/utils/train.py:100-152,
optimiser :38-43): DAGR frozen (models/EventAD.py:149-150), gradients only for the GRU / fusion / attention head
(1 115 234 fp32 parameters), one flat-bucket all-reduce (sum, / world) between backward and the clip + AdamW step.

One process per GPU; every rank steps its own clip batch.  The forward of the frozen event branch runs under no_grad on
the sm_100a kernels; only the head builds an autograd graph."""
import torch

from . import parallel


def head_parameters(model):
    return [p for n, p in model.named_parameters() if not n.startswith("dagr_model.") and p.requires_grad]


def make_optimizer(model, lr=1e-3, weight_decay=1e-5):
    """AdamW over the trainable head (utils/train.py:38-43 passes model.parameters(); frozen ones never get a gradient)."""
    return torch.optim.AdamW(head_parameters(model), lr=lr, weight_decay=weight_decay)


def train_step(model, data, labels, optimizer, grad_clip=1.0, timers=None, feature_fn=None):
    """zero_grad -> forward -> backward -> all-reduce of the flat head gradient -> clip -> step.  Returns the loss tensor
    (no host sync).  timers: optional dict filled with CUDA events around the all-reduce.  feature_fn: replaces
    model.extract_features (e.g. engine.GraphedEventAD(model, ...).features: the frozen part as CUDA-graph replays)."""
    optimizer.zero_grad(set_to_none=False)
    feats = (feature_fn or model.extract_features)(data)      # frozen DAGR + box pooling (no autograd graph)
    with torch.enable_grad():
        logits, valid, target, loss = model.head(feats, data, labels)
        loss.backward()
    params = head_parameters(model)
    if timers is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    parallel.allreduce_flat_grads(params)
    if timers is not None:
        e1.record()
        timers.setdefault("allreduce", []).append((e0, e1))
    torch.nn.utils.clip_grad_norm_(params, grad_clip)
    optimizer.step()
    return loss.detach()
