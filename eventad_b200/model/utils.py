"""shallow_copy and voxel_size_to_params of /root/reference/src/dagr/model/utils.py:112-116,158-166."""
import numpy as np

from ..data import Data


def voxel_size_to_params(pooling_layer, height, width):
    rx = int(np.ceil(2 * pooling_layer.voxel_size[0].cpu().numpy() * width))
    ry = int(np.ceil(2 * pooling_layer.voxel_size[1].cpu().numpy() * height))
    M = pooling_layer.transform.max
    return rx, ry, M


def shallow_copy(data):
    out = Data(x=data.x.clone(), edge_index=data.edge_index, edge_attr=data.edge_attr, pos=data.pos, batch=data.batch)
    for key in ["active_clusters", "_changed_attr", "_changed_attr_indices", "diff_idx", "diff_pos_idx", "pooling",
                "num_image_channels", "skipped", "pooled", "height", "width", "_csr"]:
        if hasattr(data, key):
            setattr(out, key, getattr(data, key))
    for key in ["diff_idx", "diff_pos_idx"]:
        if hasattr(data, key):
            setattr(out, key, getattr(data, key).clone())
    return out
