"""Model zoo: host-side mirrors of the reference's nn.Module classes (src/*.py), built from the
same layer factories so module names, parameter order and state_dict keys are identical."""
from .settings import Settings, DeviceMngr  # noqa: F401
from .base import MultiClassBaseModel, Conv2dBlock  # noqa: F401
from .resnet import ResNet, ResBlock, ResBottleneck  # noqa: F401
from .vggnet import VGGNet  # noqa: F401
from .inceptionnet_v1 import InceptionNetV1, InceptionBlock  # noqa: F401
from .resnext import ResNeXt  # noqa: F401
from .squeezenet import SqueezeNet, Fire  # noqa: F401
from .densenet import DenseNet, DenseLayer, DenseBlock, TransitionBlock  # noqa: F401
from .mobilenet_v1 import MobileNetV1, Conv2dBlockDW  # noqa: F401
from .shufflenet_v1 import ShuffleNetV1, ShuffleUnit  # noqa: F401
from .se import SEResNet, SENet, SEUnit  # noqa: F401
from .sk import SKResNet, SKNet, SKConv  # noqa: F401

FAMILIES = {c.__name__: c for c in (VGGNet, InceptionNetV1, ResNet, ResNeXt, SqueezeNet, DenseNet,
                                    MobileNetV1, ShuffleNetV1, SEResNet, SENet, SKResNet, SKNet)}
