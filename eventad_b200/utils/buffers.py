"""format_data: the input contract of the path (/root/reference/src/dagr/utils/buffers.py:33-44)."""
import torch


def format_data(data, normalizer=None):
    if normalizer is None:
        normalizer = torch.stack([torch.as_tensor(data.width[0]), torch.as_tensor(data.height[0]),
                                  torch.as_tensor(data.time_window[0])], dim=-1).to(data.pos.device)
    if hasattr(data, "image") and data.image is not None:
        data.image = data.image.float() / 255.0
    data.pos = torch.cat([data.pos, data.t.view((-1, 1))], dim=-1)
    data.t = None
    data.x = data.x.float()
    data.pos = data.pos / normalizer
    return data
