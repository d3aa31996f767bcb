from .convnet import ConvNet  # noqa: F401
from .resnet import ResNet, resnet18, resnet34, resnet50  # noqa: F401
