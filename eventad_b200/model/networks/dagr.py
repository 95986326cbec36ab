"""DAGR wrapper as EventAD uses it (/root/reference/src/dagr/model/networks/dagr.py:14-130): `backbone` (Net),
`cache_luts`, `extract_features`.  The YOLOX detection head (`head.*`, needs the un-vendored yolox submodule) is
never executed on the EventAD path (models/EventAD.py:357-362); its checkpoint tensors are kept in an opaque
container so `load_state_dict(checkpoint['ema'])` stays strict for `backbone.*` and loss-free for `head.*`."""
import torch

from ..utils import voxel_size_to_params
from .net import Net


class OpaqueHead(torch.nn.Module):
    """Holds `head.*` checkpoint tensors verbatim (buffers named after the checkpoint keys)."""

    def __init__(self):
        super().__init__()
        self._store = {}

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        for k in [k for k in state_dict if k.startswith(prefix)]:
            self._store[k[len(prefix):]] = state_dict[k]

    def state_dict(self, *args, destination=None, prefix="", keep_vars=False):
        destination = destination if destination is not None else {}
        for k, v in self._store.items():
            destination[prefix + k] = v
        return destination

    def forward(self, *a, **k):
        raise NotImplementedError("the YOLOX/GNN detection head is outside the EventAD hot path (SURVEY section 2, #5)")


class DAGR(torch.nn.Module):
    def __init__(self, args, height, width):
        super().__init__()
        self.conf_threshold = 0.001
        self.nms_threshold = 0.65
        self.height = height
        self.width = width
        self.batch_size = args.batch_size
        self.backbone = Net(args, height=height, width=width)
        self.head = OpaqueHead()

    def cache_luts(self, width, height, radius):
        """dagr.py:36-71 for the backbone layers (the head's convolutions are not instantiated)."""
        M = 2 * float(int(radius * width + 2) / width)
        r = int(radius * width + 1)
        b = self.backbone
        b.conv_block1.conv_block1.conv.init_lut(height=height, width=width, Mx=M, rx=r)
        b.conv_block1.conv_block2.conv.init_lut(height=height, width=width, Mx=M, rx=r)
        for pool, layer in ((b.pool1, b.layer2), (b.pool2, b.layer3), (b.pool3, b.layer4), (b.pool4, b.layer5)):
            rx, ry, M = voxel_size_to_params(pool, height, width)
            layer.conv_block1.conv.init_lut(height=height, width=width, Mx=M, rx=rx, ry=ry)
            layer.conv_block2.conv.init_lut(height=height, width=width, Mx=M, rx=rx, ry=ry)

    def forward(self, x, reset=True, return_targets=True, filtering=True):
        raise NotImplementedError("detection forward (YOLOX head, NMS) is outside the EventAD hot path")

    def extract_features(self, x):
        """dagr.py:108-130."""
        self.eval()
        with torch.no_grad():
            features = self.backbone(x)
            if isinstance(features, (list, tuple)):
                return features[0]
            return features
