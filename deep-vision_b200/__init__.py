"""deep_vision_b200 — B200-native (sm_100a) implementation of deep-vision's conv-BN-ReLU hot path.

    import deep_vision_b200 as dv
    net = dv.models.ResNet50().cuda()          # same class / state_dict as ResNet/pytorch/models/resnet50.py
    dv.set_compute_dtype(torch.bfloat16)       # tcgen05 path (default); torch.float32 = 1e-4 parity path
    logits = net(images_nchw_fp32)             # fp32 [B,1000]

The package directory is ``deep-vision_b200/`` (not an identifier), so the repo root carries a one-file import
alias ``deep_vision_b200.py``.
"""
from . import checkpoint, config, evaluate, functional, inference, lib, models, nn  # noqa: F401
from . import data, epochs  # noqa: F401,E402
from .config import bump_weight_epoch, compute_dtype, set_compute_dtype, set_conv_algo, set_deterministic, set_fold_bn, set_fuse_bn_bwd, set_fuse_bn_stats, set_fuse_shortcut_grad, set_fuse_stem_pool  # noqa: F401

__version__ = "0.1.0"
