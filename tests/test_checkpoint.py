"""Checkpoint compatibility helpers -- CPU only."""
import collections
import io

import pytest
import torch

from oracle import harness


@pytest.fixture()
def model():
    from can_b200.model import CANNet
    m = CANNet(load_weights=True)
    m.load_state_dict(harness.make_state_dict("default"))
    return m


def test_expected_keys_are_the_reference_layout(model):
    from can_b200.utils import checkpoint as ck
    assert ck.expected_keys() == list(model.state_dict().keys()) == harness_keys()


def harness_keys():
    from oracle import cannet_oracle as O
    return list(O.state_dict_keys())


def test_roundtrip_plain_and_ddp_prefixed(model):
    from can_b200.model import CANNet
    from can_b200.utils import checkpoint as ck
    for prefixed in (False, True):
        buf = io.BytesIO()
        ck.save_checkpoint(model, buf, ddp_prefix=prefixed)
        buf.seek(0)
        raw = torch.load(io.BytesIO(buf.getvalue()))
        assert all(k.startswith("module.") == prefixed for k in raw)
        fresh = CANNet(load_weights=True)
        missing, unexpected = ck.load_checkpoint(fresh, buf)
        assert not missing and not unexpected
        for (k, a), (_, b) in zip(model.state_dict().items(), fresh.state_dict().items()):
            assert torch.equal(a, b), k


def test_strict_load_names_the_problem(model):
    from can_b200.utils import checkpoint as ck
    sd = collections.OrderedDict(model.state_dict())
    sd.pop("conv6_2.weight")
    with pytest.raises(KeyError, match="conv6_2.weight"):
        ck.load_checkpoint(model, sd)
    missing, unexpected = ck.load_checkpoint(model, sd, strict=False)
    assert missing == ["conv6_2.weight"] and unexpected == []
    sd = collections.OrderedDict(model.state_dict())
    sd["frontend.0.weight"] = torch.zeros(64, 3, 5, 5)
    with pytest.raises(ValueError, match="frontend.0.weight"):
        ck.load_checkpoint(model, sd)
    mixed = collections.OrderedDict(model.state_dict())
    mixed["module.extra"] = torch.zeros(1)
    with pytest.raises(ValueError, match="mixes"):
        ck.load_checkpoint(model, mixed)


def test_load_is_in_place(model):
    """Parameters keep their storage (a FusedTrainer arena or DDP bucket view must survive a load)."""
    from can_b200.utils import checkpoint as ck
    ptrs = [p.data_ptr() for p in model.parameters()]
    ck.load_checkpoint(model, {"state_dict": harness.make_state_dict("he")})
    assert ptrs == [p.data_ptr() for p in model.parameters()]
    assert torch.equal(model.backend[0].weight, harness.make_state_dict("he")["backend.0.weight"])


def test_vgg16_positional_graft(model):
    """model/CANNet.py:30-35 with a local torchvision-style state_dict (13 convs + classifier; first 20 tensors used)."""
    from can_b200.utils import checkpoint as ck
    g = torch.Generator().manual_seed(3)
    cfg = [(64, 3), (64, 64), (128, 64), (128, 128), (256, 128), (256, 256), (256, 256), (512, 256), (512, 512), (512, 512),
           (512, 512), (512, 512), (512, 512)]
    idx = [0, 2, 5, 7, 10, 12, 14, 17, 19, 21, 24, 26, 28]
    vgg = collections.OrderedDict()
    for i, (co, ci) in zip(idx, cfg):
        vgg[f"features.{i}.weight"] = torch.randn(co, ci, 3, 3, generator=g)
        vgg[f"features.{i}.bias"] = torch.randn(co, generator=g)
    vgg["classifier.0.weight"] = torch.zeros(8, 8)
    before = model.backend[0].weight.clone()
    ck.graft_vgg16(model, vgg)
    fr = list(model.frontend.state_dict().values())
    for t, (k, v) in zip(fr, list(vgg.items())[:20]):
        assert torch.equal(t, v), k
    assert torch.equal(model.backend[0].weight, before)
    with pytest.raises(ValueError):
        ck.graft_vgg16(model, collections.OrderedDict(list(vgg.items())[:10]))
