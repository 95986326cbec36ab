"""Multi-GPU plumbing: one process per GPU, clips/windows sharded with no data-path collective.

A window never interacts with another window (FIFO, edges and clusters are keyed by the batch index); the only
cross-sample state is per-clip (the EventAD head's GRU states and the t == 1.0 pooling quirk inside one batch), so the
unit of sharding is the clip: clip c -> rank c mod G, windows of a clip stay in order on one rank.  Results are
gathered to rank 0 with one `gather_object` at the end (a few KB of per-frame scores)."""
import torch.distributed as dist


def shard_clips(n_clips: int, world_size: int, rank: int):
    """Clip ids owned by `rank` (round-robin)."""
    return list(range(rank, n_clips, world_size))


def batches_of(clip_ids, windows_per_clip: int, batch_size: int):
    """(clip, first_window, n_windows) chunks: consecutive windows of one clip per batch, cut only at batch boundaries."""
    out = []
    for c in clip_ids:
        for w in range(0, windows_per_clip, batch_size):
            out.append((c, w, min(batch_size, windows_per_clip - w)))
    return out


def gather_results(local: dict, dst: int = 0):
    """local: {clip_id: result}.  Returns the merged dict on rank `dst` (None elsewhere); works on gloo and nccl."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(local)
    world, rank = dist.get_world_size(), dist.get_rank()
    bucket = [None] * world if rank == dst else None
    dist.gather_object(local, bucket, dst=dst)
    if rank != dst:
        return None
    merged = {}
    for part in bucket:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError(f"clip {k} produced on two ranks")
            merged[k] = v
    return dict(sorted(merged.items()))


def allreduce_flat_grads(params, world_size=None):
    """Training step plumbing (BASELINE config 4): one NCCL/gloo all-reduce (sum, then / G) over ONE flat fp32 bucket
    of the trainable head gradients; parameters without a gradient contribute zeros so that all ranks agree
    (the never-called `fusion_module.attention`)."""
    import torch
    params = [p for p in params if p.requires_grad]
    if not params:
        return
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float() for p in params])
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat /= (world_size or dist.get_world_size())
    off = 0
    for p in params:
        n = p.numel()
        g = flat[off:off + n].view_as(p).to(p.dtype)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += n
