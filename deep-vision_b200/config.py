"""Process-wide settings of the B200 path."""
import os

import torch

_state = {"compute_dtype": torch.bfloat16, "weight_epoch": 0, "conv_algo": 0,
          "fuse_bn_stats": os.environ.get("DV_FUSE_STATS", "1") != "0",
          "deterministic": os.environ.get("DV_DETERMINISTIC", "0") == "1",
          "atomic_wgrad": os.environ.get("DV_ATOMIC_WGRAD", "1") == "1",
          "atomic_bn_bwd": os.environ.get("DV_ATOMIC_BN", "0") == "1",
          "fold_bn": os.environ.get("DV_FOLD_BN", "1") != "0",
          "side_wgrad": os.environ.get("DV_SIDE_WGRAD", "1") == "1",
          "fuse_bn_bwd": os.environ.get("DV_FUSE_BN_BWD", "1") == "1",
          "stem_direct": os.environ.get("DV_STEM_DIRECT", "1") == "1",
          "stem_raw_input": os.environ.get("DV_STEM_RAW", "1") == "1",
          "fuse_stem_pool": os.environ.get("DV_FUSE_STEM_POOL", "1") == "1",
          "fuse_shortcut_grad": os.environ.get("DV_FUSE_SHORTCUT", "0") == "1"}


def compute_dtype():
    return _state["compute_dtype"]


def set_compute_dtype(dtype):
    """torch.bfloat16 (tensor-core path, the speed configuration) or torch.float32 (CUDA-core path that meets the
    1e-4 parity bound of BASELINE.json)."""
    if dtype not in (torch.bfloat16, torch.float32):
        raise TypeError("compute dtype must be torch.bfloat16 or torch.float32")
    _state["compute_dtype"] = dtype


def weight_epoch():
    return _state["weight_epoch"]


def bump_weight_epoch():
    """Invalidate every cached compute-dtype filter copy.  Call after parameters were changed behind torch's back
    (the fused optimiser kernels write through raw pointers and do not move Tensor._version)."""
    _state["weight_epoch"] += 1


def conv_algo():
    return _state["conv_algo"]


def set_conv_algo(algo):
    """0 = auto (tcgen05 where the shape qualifies), 1 = CUDA-core kernels only, 2 = tensor cores or error."""
    _state["conv_algo"] = int(algo)


def fuse_bn_stats():
    return _state["fuse_bn_stats"] and not _state["deterministic"]


def set_fuse_bn_stats(flag):
    """True (default): tensor-core convolutions accumulate the per-channel sums BatchNorm needs in their epilogue, so the
    statistics pass never re-reads the activation.  False: BatchNorm runs its own reduction pass."""
    _state["fuse_bn_stats"] = bool(flag)


def deterministic():
    return _state["deterministic"]


def set_deterministic(flag):
    """False (default): split reductions (wgrad tiles, BN backward sums, fused BN statistics) are combined with fp32
    atomics — fewer launches and less traffic, summation order varies run to run.  True: fixed-order second-stage
    reductions, bit-reproducible results."""
    _state["deterministic"] = bool(flag)   # fuse_bn_stats() answers False while this is set (the fused sums use atomics)


def atomic_wgrad():
    """Add wgrad tiles straight into dW with fp32 vector reductions instead of split partials + a reduce pass."""
    return _state["atomic_wgrad"] and not _state["deterministic"]


def atomic_bn_bwd():
    """Add BN-backward block sums straight into dgamma/dbeta with fp32 atomics instead of partials + a finalize pass."""
    return _state["atomic_bn_bwd"] and not _state["deterministic"]


def fold_bn():
    return _state["fold_bn"]


def set_fold_bn(flag):
    """True (default): under net.eval() + torch.no_grad() in bf16, each conv -> BatchNorm (-> +identity) (-> ReLU) group of
    the models runs as ONE kernel: the convolution with the BN folded into its epilogue (inference.py).  False: eval-mode
    BN stays a separate normalise pass."""
    _state["fold_bn"] = bool(flag)


def side_wgrad():
    """Trainer: launch dense weight-gradient kernels on a second stream beside the dgrad -> BN-backward chain."""
    return _state["side_wgrad"] and not _state["deterministic"]


def fuse_bn_bwd():
    """The dgrad of a dense convolution whose input is relu(bn(x0)) also produces that BatchNorm's backward sums (dgamma, dbeta)
    in its epilogue (dv_conv_params.bnb_*), so the BatchNorm's own backward reduce pass is not launched.  The sums are combined
    with fp32 atomics, hence off in the fixed-order mode."""
    return _state["fuse_bn_bwd"] and not _state["deterministic"]


def set_fuse_bn_bwd(flag):
    _state["fuse_bn_bwd"] = bool(flag)


def stem_direct():
    """C <= 4 stride-2 stems run straight from a padded NHWC4 image (dv_stem_*) instead of through a patch matrix."""
    return _state["stem_direct"]


def set_stem_direct(flag):
    _state["stem_direct"] = bool(flag)


def fuse_stem_pool():
    """Train-mode bn1 -> relu -> maxpool of the ResNet stem as one forward pass and two backward passes (dv_bn_relu_maxpool_*)."""
    return _state["fuse_stem_pool"]


def set_fuse_stem_pool(flag):
    _state["fuse_stem_pool"] = bool(flag)


def fuse_shortcut_grad():
    """Identity shortcuts: the gradient `dout * (out > 0)` is added by the first convolution's dgrad epilogue
    (dv_conv_params.ep_residual / ep_residual_mask) instead of being written by BatchNorm's backward and accumulated onto.
    Off by default: it saves 4 bytes per element of HBM traffic per identity block, but the epilogue's extra arithmetic per
    element costs as much on the K = 64..512 point-wise dgrads it rides on (ResNet-50 step 16.98 vs 16.86 ms, round 2)."""
    return _state["fuse_shortcut_grad"]


def set_fuse_shortcut_grad(flag):
    _state["fuse_shortcut_grad"] = bool(flag)


def stem_raw_input():
    """The direct stem takes the loader's fp32 NCHW batch itself (dv_stem_pad_input_nchw) instead of a converted NHWC copy."""
    return _state["stem_raw_input"]
