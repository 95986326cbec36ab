"""Default hyper-parameters of the dagr-s configuration EventAD runs.  `config/dagr-s-dsec.yaml` is
not in the reference tree; every value stays a plain attribute that the layer constructors read exactly like the
reference's argparse namespace (net.py:34-97, ev_tgn.py:22-23, spline_conv.py:13-14, conv.py:13-14)."""
from types import SimpleNamespace


def default_args(**overrides):
    a = SimpleNamespace(
        radius=0.01, max_neighbors=16, edge_attr_dim=2, kernel_size=5, aggr="sum", activation="relu",
        pooling_aggr="max", pooling_dim_at_output="5x7", base_width=0.5, after_pool_width=1, net_stem_width=0.5,
        yolo_stem_width=0.5, num_scales=2, use_image=True, img_net="resnet50", batch_size=6,
        keep_temporal_ordering=False, no_events=False, dataset="dsec", img_net_pretrained=False)
    for k, v in overrides.items():
        setattr(a, k, v)
    return a
