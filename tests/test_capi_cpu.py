"""The C-ABI library loads without a GPU and exports every symbol include/eventad_b200.h declares; the Python mirror
of the reference interface keeps the reference's names, constructor arguments and state-dict keys."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "eventad_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(evad_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from eventad_b200 import build
    lib = ctypes.CDLL(build.build())
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/eventad_b200.h but not exported"
    lib.evad_compiled_arch.restype = ctypes.c_int
    assert lib.evad_compiled_arch() == 100


def test_ctypes_signatures_cover_the_header():
    from eventad_b200 import _lib
    L = _lib.lib()
    for n in _declared():
        assert getattr(L, n).argtypes is not None or n in ("evad_version", "evad_compiled_arch", "evad_launch_count"), n


def test_no_cpu_fallback():
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    g = SlidingWindowGraph(width=8, height=8, batch_size=1, max_queue_size=4, radius=1, delta_t_us=10)
    with pytest.raises(RuntimeError):
        g.forward(torch.zeros(3, dtype=torch.int32), torch.zeros((3, 3), dtype=torch.int32), delete_nodes=False)


def test_state_dict_keys_match_reference_layout():
    """keys/shapes of dagr-s so that dagr_s_50.pth['ema'] loads unchanged."""
    from eventad_b200.config import default_args
    from eventad_b200.model.networks.dagr import DAGR
    m = DAGR(default_args(), 240, 320)
    sd = m.state_dict()
    exp = {"backbone.conv_block1.conv_block1.conv.weight": (25, 19, 16), "backbone.conv_block1.conv_block1.conv.lin.weight": (16, 19),
           "backbone.conv_block1.conv_block2.lin.mlp.weight": (16, 19), "backbone.layer2.conv_block1.conv.weight": (25, 82, 64),
           "backbone.layer3.conv_block1.conv.weight": (25, 130, 64), "backbone.layer5.conv_block2.conv.weight": (25, 64, 64),
           "backbone.layer5.conv_block2.norm_skip.module.running_var": (64,), "backbone.layer2.conv_block1.conv.kernel_size": (2,),
           "backbone.layer2.conv_block1.conv.is_open_spline": (2,), "backbone.net.feature_dconv.0.weight": (16, 64, 1, 1),
           "backbone.net.feature_dconv.4.weight": (64, 2048, 1, 1), "backbone.net.output_dconv.1.weight": (256, 2048, 1, 1),
           "backbone.net.module.layer4.2.conv3.weight": (512 * 4, 512, 1, 1)}
    for k, shp in exp.items():
        assert k in sd and tuple(sd[k].shape) == shp, k
    assert not any(k.startswith("backbone.pool") or k.startswith("backbone.edge_attrs") or
                   k.startswith("backbone.events_to_graph") for k in sd)
    assert sd["backbone.layer2.conv_block1.conv.kernel_size"].dtype == torch.int64
    assert sd["backbone.layer2.conv_block1.conv.is_open_spline"].dtype == torch.uint8
    # head.* tensors of a checkpoint are kept verbatim and round-trip
    sd2 = dict(sd)
    sd2["head.stem1.conv.weight"] = torch.zeros(25, 64, 64)
    m.load_state_dict(sd2, strict=True)
    assert "head.stem1.conv.weight" in m.state_dict()
    assert m.backbone.pool1._nv == [56, 40, 1] and m.backbone.pool4._nv == [7, 5, 1]
    assert int(m.backbone.pool1.num_grid_cells) == 2240 and m.backbone.strides == [24, 48]
