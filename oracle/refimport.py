"""Import the UNMODIFIED reference (milan-pavlovic-ai/ConvNets) from /root/reference/src.

TEST INFRASTRUCTURE ONLY, and only usable in the build container: /root/reference does not exist
on the GPU box, so nothing under `-m gpu`, smoke() or bench.py may call this.  It is used by
oracle/make_golden.py (to write tests/golden/*) and by the CPU tests that are skipped when the
reference is absent.

The reference needs three modules that are not installed here (torchsummary, matplotlib, seaborn:
imported at src/basemodel.py:15, src/mngrdata.py:6-8, src/mngrplot.py:4-5) — stubbed from
tests/_stubs — and one keyword torch 2.x removed (`ReduceLROnPlateau(verbose=...)`,
src/basemodel.py:73-78) — dropped by a wrapper.  The reference source is never edited or copied.
"""
from __future__ import annotations

import importlib
import os
import sys

REF_ROOT = os.environ.get("CONVNETS_REFERENCE", "/root/reference")
REF_SRC = os.path.join(REF_ROOT, "src")
_STUBS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "_stubs")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_SRC, "basemodel.py"))


_ready = False


def _prepare() -> None:
    global _ready
    if _ready:
        return
    if not available():
        raise RuntimeError(f"reference not found under {REF_SRC}")
    for p in (REF_SRC, _STUBS):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch

    base = torch.optim.lr_scheduler.ReduceLROnPlateau
    if not getattr(base, "_cnb_patched", False):
        class _Plateau(base):
            _cnb_patched = True

            def __init__(self, *a, verbose=None, **k):
                super().__init__(*a, **k)

        torch.optim.lr_scheduler.ReduceLROnPlateau = _Plateau
    _ready = True


# family -> (module, class)
FAMILIES = {
    "VGGNet": ("vggnet", "VGGNet"),
    "InceptionNetV1": ("inceptionnet_v1", "InceptionNetV1"),
    "ResNet": ("resnet", "ResNet"),
    "ResNeXt": ("resnext", "ResNeXt"),
    "SqueezeNet": ("squeezenet", "SqueezeNet"),
    "DenseNet": ("densenet", "DenseNet"),
    "MobileNetV1": ("mobilenet_v1", "MobileNetV1"),
    "ShuffleNetV1": ("shufflenet_v1", "ShuffleNetV1"),
    "SEResNet": ("se_resnet", "SEResNet"),
    "SENet": ("senet", "SENet"),
    "SKResNet": ("sk_resnet", "SKResNet"),
    "SKNet": ("sknet", "SKNet"),
}


def module(name: str):
    _prepare()
    return importlib.import_module(name)


def make_model(family: str, kind, hw: int = 32, num_classes: int = 10, seed: int = 21, **settings):
    """Construct the reference model on CPU exactly as its process_fit() does (Settings ->
    model ctor with init_params), after the reference's own reproducibility helper."""
    _prepare()
    import torch
    st = module("settings")
    mod, cls = FAMILIES[family]
    util = module("mngrutility")
    util.UtilityMngr.set_reproducible_mode(seed=seed)
    kw = dict(batch_size=8, dropout_rate=0.0, mixed_precision=False, num_workers=0, seed=seed)
    kw.update(settings)
    setting = st.Settings(kind=kind, input_size=(3, hw, hw), num_classes=num_classes, **kw)
    setting.device.set_device("cpu")
    model = getattr(module(mod), cls)(setting)
    torch.use_deterministic_algorithms(False)
    return model
