"""CPU: convnets_b200.convert() applied to the REFERENCE's own classes (the north_star API surface:
"a model switches over by swapping its Conv2d/BatchNorm2d/pooling layers"), imported read-only from
/root/reference through oracle/refimport.py.  Structural only — the converted layers have no CPU
path — and skipped where the reference is absent (the GPU box).

For every family: state_dict keys / shapes / Parameter objects are unchanged, every nn.Conv2d /
BatchNorm2d / pooling layer is re-classed to its library-backed subclass, BatchNorm -> ReLU pairs
are fused without moving any child index, the composite blocks get their fused forwards, exactly
one Conv2d is the network entry, and the reference's strict load_state_dict (src/basemodel.py:940)
round-trips between a converted and an un-converted instance.
"""
import pytest
import torch
import torch.nn as nn

from oracle import refimport

pytestmark = pytest.mark.skipif(not refimport.available(), reason="reference checkout not present")

CASES = [("VGGNet", 11), ("InceptionNetV1", ""), ("ResNet", 18), ("ResNet", 50), ("ResNeXt", 50),
         ("SqueezeNet", "1.1"), ("DenseNet", "121"), ("MobileNetV1", ""), ("ShuffleNetV1", "g3"),
         ("SEResNet", 50), ("SENet", 26), ("SKResNet", 26), ("SKNet", 50)]

_PATCHED = {"ResBlock", "ResBottleneck", "SEResBlock", "SEResBottleneck", "SEBlock", "SEBottleneck",
            "SKResBottleneck", "SKBottleneck", "SEUnit", "SKConv", "ShuffleUnit", "Fire",
            "InceptionBlock", "DenseLayer"}


@pytest.mark.parametrize("family,kind", CASES, ids=lambda v: str(v))
def test_convert_on_reference_classes(family, kind):
    import convnets_b200 as cnb
    from convnets_b200 import layers as L
    model = refimport.make_model(family, kind, hw=32)
    twin = refimport.make_model(family, kind, hw=32)
    spec = [(k, tuple(v.shape)) for k, v in model.state_dict().items()]
    ids = [id(p) for p in model.parameters()]
    n_plain = {t: sum(isinstance(m, t) for m in model.modules())
               for t in (nn.Conv2d, nn.BatchNorm2d, nn.MaxPool2d, nn.AvgPool2d, nn.AdaptiveAvgPool2d)}
    cnb.convert(model)
    assert [(k, tuple(v.shape)) for k, v in model.state_dict().items()] == spec
    assert [id(p) for p in model.parameters()] == ids
    swapped = {nn.Conv2d: L.Conv2d, nn.BatchNorm2d: L.BatchNorm2d, nn.MaxPool2d: L.MaxPool2d,
               nn.AvgPool2d: L.AvgPool2d, nn.AdaptiveAvgPool2d: L.AdaptiveAvgPool2d}
    for plain, ours in swapped.items():
        assert sum(isinstance(m, ours) for m in model.modules()) == n_plain[plain], plain.__name__
        assert not any(type(m) is plain for m in model.modules()), plain.__name__
    convs = [m for m in model.modules() if isinstance(m, L.Conv2d)]
    assert sum(m.cast_input for m in convs) == 1 and convs[0].cast_input
    assert all(m.weight.data.is_contiguous(memory_format=torch.channels_last) for m in convs)
    # composite blocks carry the fused forwards (instance attribute shadows the class method)
    for m in model.modules():
        if type(m).__name__ in _PATCHED:
            assert "forward" in m.__dict__, type(m).__name__
    # conv -> BatchNorm pairs: statistics from the conv epilogue; BatchNorm -> ReLU pairs fused in place
    for seq in (m for m in model.modules() if isinstance(m, nn.Sequential)):
        ch = list(seq.children())
        for a, b in zip(ch, ch[1:]):
            if isinstance(a, L.Conv2d) and isinstance(b, L.BatchNorm2d):
                assert a.emit_bn_stats
            if isinstance(a, L.BatchNorm2d) and a.fuse_relu:
                assert isinstance(b, L.Identity)
    assert not any(type(m) is nn.ReLU for seq in model.modules() if isinstance(seq, nn.Sequential)
                   for a, m in zip(list(seq.children()), list(seq.children())[1:])
                   if isinstance(a, L.BatchNorm2d))
    # the reference's strict checkpoint load works in both directions
    twin.load_state_dict(model.state_dict(), strict=True)
    model.load_state_dict(twin.state_dict(), strict=True)
    # the reference's own isinstance-based init still recognises the layers (src/basemodel.py:92-104)
    model.init_params()
