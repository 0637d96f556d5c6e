"""Model zoo: host-side mirrors of the reference's nn.Module classes (src/*.py), built from the
same layer factories so module names, parameter order and state_dict keys are identical."""
from .settings import Settings, DeviceMngr  # noqa: F401
from .base import MultiClassBaseModel, Conv2dBlock  # noqa: F401
from .resnet import ResNet, ResBlock, ResBottleneck  # noqa: F401
