"""B200-native (sm_100a) hot path of thnak/image_super_resolution: the conv network's forward pass
and pixel-loss training step behind the reference's own Python surfaces.  See DESIGN.md."""
from .modelpt import ModelPT, load, RF_RADIUS, FLOP_PER_PX  # noqa: F401

from .models import Model, ResNet, EResNet, SRGAN, fuse_conv_and_bn  # noqa: F401,E402

__all__ = ["ModelPT", "load", "RF_RADIUS", "FLOP_PER_PX", "Model", "ResNet", "EResNet", "SRGAN", "fuse_conv_and_bn"]
