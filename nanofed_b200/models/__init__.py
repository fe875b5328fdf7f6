"""Bundled architectures (reference ``nanofed/models``) + the B200 cohort engines."""

from .mnist import MNISTModel
from .resnet import ResNet18

__all__ = ["MNISTModel", "ResNet18"]
