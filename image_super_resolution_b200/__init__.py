"""B200-native (sm_100a) hot path of thnak/image_super_resolution: the conv network's forward pass
and pixel-loss training step behind the reference's own Python surfaces.  See DESIGN.md."""
from .modelpt import ModelPT, load, RF_RADIUS, FLOP_PER_PX  # noqa: F401

from .models import Model, ResNet, EResNet, SRGAN, Denoise, fuse_conv_and_bn  # noqa: F401,E402

from .training import L1Loss, MSELoss  # noqa: F401,E402
from .train import ModelEMA, init_distributed, shard_of, broadcast_parameters  # noqa: F401,E402
from .optim import FlatAdam, GradScaler, clip_grad_norm_  # noqa: F401,E402
# (`train()` itself lives in the submodule of the same name: `from image_super_resolution_b200.train import train`)

__all__ = ["ModelPT", "load", "RF_RADIUS", "FLOP_PER_PX", "Model", "ResNet", "EResNet", "SRGAN", "Denoise", "fuse_conv_and_bn",
           "L1Loss", "MSELoss", "ModelEMA", "init_distributed", "shard_of", "broadcast_parameters", "FlatAdam", "GradScaler", "clip_grad_norm_"]
