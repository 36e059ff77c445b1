"""CPU: the C-ABI library builds, loads, and exports every symbol include/can_b200.h declares.
No compute is launched here (there is no GPU in the build container)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    import can_b200
    return can_b200


def header_symbols():
    src = open(os.path.join(ROOT, "include", "can_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(can_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    lib = built.lib()
    syms = header_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/can_b200.h but not exported"
    # and the Python binding covers the same set
    assert sorted(built.exported_symbols()) == syms


def test_argument_validation_needs_no_gpu(built):
    from can_b200 import _lib
    lib = built.lib()
    assert lib.can_version() >= 100
    assert lib.can_conv_tc(None, None) < 0
    assert b"null" in lib.can_last_error_string()
    a = _lib.ConvArgs()
    a.N, a.H, a.W, a.Cin, a.Cout, a.taps, a.dil, a.nsplit = 1, 8, 8, 60, 64, 9, 1, 1
    assert lib.can_conv_tc(a, None) < 0  # Cin not a multiple of 64
    assert b"Cin" in lib.can_last_error_string()
    assert lib.can_ctx_pool_workspace_bytes(2, 96) == 2 * 96 * 12 * 512 * 4
    assert lib.can_sgd_momentum_step(None, None, None, 0, 0.0, 0.0, 1.0, 0, None) < 0


def test_struct_layout_matches_header(built):
    """ctypes mirror of can_conv_args: same field order as the header."""
    from can_b200 import _lib
    src = open(os.path.join(ROOT, "include", "can_b200.h")).read()
    body = src[src.index("typedef struct can_conv_args"):src.index("} can_conv_args;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split("{", 1)[1].split(";"):
        decl = decl.strip()
        if not decl:
            continue
        parts = decl.replace("*", " ").split()
        rest = decl[decl.index(parts[1] if parts[0] == "const" else parts[0]) + len(parts[1] if parts[0] == "const" else parts[0]):]
        names += [n.strip(" *") for n in rest.split(",")]
    assert names == [f[0] for f in _lib.ConvArgs._fields_]


def test_module_is_a_drop_in_without_gpu():
    import torch
    from can_b200.model import CANNet, make_layers
    from oracle import cannet_oracle as O, harness
    m = CANNet(load_weights=True)
    sd = m.state_dict()
    assert list(sd.keys()) == O.state_dict_keys()
    shapes = O.state_dict_shapes()
    assert all(tuple(sd[k].shape) == shapes[k] for k in sd)
    m.load_state_dict(harness.make_state_dict("he"), strict=True)
    assert sum(p.numel() for p in m.parameters()) == 20_719_937
    assert [k for k, _ in m.named_parameters()] == O.state_dict_keys()
    seq = make_layers([8, "M", 16], in_channels=4, dilation=True)
    assert [type(x).__name__ for x in seq] == ["Conv2d", "ReLU", "MaxPool2d", "Conv2d", "ReLU"]
    assert seq[0].dilation == (2, 2) and seq[0].padding == (2, 2)
    m._initialize_weights()
    assert float(m.output_layer.bias.abs().sum()) == 0.0
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 3, 64, 64))  # no CPU fallback
    with pytest.raises(ValueError):
        m(torch.zeros(1, 1, 64, 64))


def test_wgrad_workspace_query_covers_every_patch_shape(built, monkeypatch):
    """can_conv_wgrad_workspace_bytes does not know the dilation, while the patch shape of the weight-gradient tiles (16 x 2
    pixels, or 8 x 4 with the halo X tile for dil <= 2) and hence the split-K factor can depend on it: the query must
    cover both plans, on ragged sizes too (host arithmetic only, no GPU)."""
    lib = built.lib()
    shapes = [(4, 96, 128, 512, 512), (1, 37, 41, 128, 128), (2, 23, 24, 256, 512), (1, 9, 21, 512, 512), (3, 50, 70, 256, 256)]
    for N, H, W, Cin, Cout in shapes:
        monkeypatch.delenv("CAN_B200_WGHX", raising=False)
        q = lib.can_conv_wgrad_workspace_bytes(N, H, W, Cin, Cout, 9, 1)
        monkeypatch.setenv("CAN_B200_WGHX", "0")
        q_plain = lib.can_conv_wgrad_workspace_bytes(N, H, W, Cin, Cout, 9, 1)
        per_split = 9 * Cout * Cin * 4
        assert q > 0 and q % per_split == 0 and q_plain % per_split == 0
        assert q >= q_plain, (N, H, W, Cin, Cout, q, q_plain)
    assert lib.can_conv_wgrad_workspace_bytes(1, 8, 8, 60, 64, 9, 1) < 0   # Cin not a multiple of 64
