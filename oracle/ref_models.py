"""torch-functional restatement of the reference's four model files, driven by a state_dict.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The product's model classes are not used here: the forward passes
below are written against `torch.nn.functional` directly from the reference's `forward` methods, so that a parity
test compares two independent implementations.  Runs on CPU (fp32 or fp64); it is also the "port" that bench.py times
as the CPU baseline, because /root/reference does not exist on the GPU box.

`sd` is a {name: tensor} mapping with the reference's state_dict keys; `bn_train` selects batch statistics (the
reference trains with nn.BatchNorm2d defaults: eps 1e-5, momentum 0.1, running stats updated in place).
"""
import torch
import torch.nn.functional as F


def _bn(x, sd, prefix, bn_train, momentum=0.1, eps=1e-5):
    # nn.BatchNorm2d(C) — resnet50.py:30,108,119,129,144; mobilenet_v1.py:31,126,149
    rm, rv = sd.get(prefix + "running_mean"), sd.get(prefix + "running_var")
    y = F.batch_norm(x, rm, rv, sd[prefix + "weight"], sd[prefix + "bias"], training=bn_train, momentum=momentum, eps=eps)
    if bn_train and (prefix + "num_batches_tracked") in sd:
        sd[prefix + "num_batches_tracked"] += 1
    return y


def _conv(x, sd, prefix, stride=1, pad=0, groups=1):
    # nn.Conv2d(..., bias=False)
    return F.conv2d(x, sd[prefix + "weight"], None, stride=stride, padding=pad, groups=groups)


def bottleneck_block(x, sd, prefix, stride, downsample, bn_train=True):
    """BottleneckBlock.forward — resnet50.py:147-165 (stride on conv1, projection = 1x1(stride)+BN)."""
    identity = x
    if downsample:
        identity = _bn(_conv(x, sd, prefix + "projection.0.", stride=stride), sd, prefix + "projection.1.", bn_train)
    x = F.relu(_bn(_conv(x, sd, prefix + "conv1.", stride=stride), sd, prefix + "bn1.", bn_train))
    x = F.relu(_bn(_conv(x, sd, prefix + "conv2.", pad=1), sd, prefix + "bn2.", bn_train))
    x = _bn(_conv(x, sd, prefix + "conv3."), sd, prefix + "bn3.", bn_train)
    return F.relu(x + identity)


def basic_block(x, sd, prefix, stride, downsample, bn_train=True):
    """BasicBlock.forward — resnet34.py:130-143."""
    identity = x
    if downsample:
        identity = _bn(_conv(x, sd, prefix + "projection.0.", stride=stride), sd, prefix + "projection.1.", bn_train)
    x = F.relu(_bn(_conv(x, sd, prefix + "conv1.", stride=stride, pad=1), sd, prefix + "bn1.", bn_train))
    x = _bn(_conv(x, sd, prefix + "conv2.", pad=1), sd, prefix + "bn2.", bn_train)
    return F.relu(x + identity)


def dwsep_block(x, sd, prefix, dw_stride, bn_train=True):
    """DepthwiseSeparableConv.forward — mobilenet_v1.py:103-106,129-133,152-156."""
    c = x.shape[1]
    x = F.relu(_bn(_conv(x, sd, prefix + "dw.conv.", stride=dw_stride, pad=1, groups=c), sd, prefix + "dw.bn.", bn_train))
    return F.relu(_bn(_conv(x, sd, prefix + "pw.conv."), sd, prefix + "pw.bn.", bn_train))


_RESNET_STAGES = {
    "resnet34": ("basic", [2, 2, 2, 2]),      # resnet34.py:38-41 — [2,2,2,2] as coded
    "resnet50": ("bottleneck", [3, 4, 6, 3]),   # resnet50.py:37-40
    "resnet152": ("bottleneck", [3, 8, 36, 3]),  # resnet152.py:37-40
}


def resnet_forward(name, x, sd, bn_train=True):
    """ResNet34/50/152.forward — resnet50.py:47-62, resnet34.py:48-63."""
    kind, counts = _RESNET_STAGES[name]
    block = basic_block if kind == "basic" else bottleneck_block
    x = F.relu(_bn(_conv(x, sd, "conv1.", stride=2, pad=3), sd, "bn1.", bn_train))
    x = F.max_pool2d(x, kernel_size=3, stride=2, padding=1)
    for si, (stage, n) in enumerate(zip(("conv2x", "conv3x", "conv4x", "conv5x"), counts)):
        for b in range(n):
            stride = (1 if si == 0 else 2) if b == 0 else 1
            x = block(x, sd, f"{stage}.{b}.", stride, downsample=(b == 0), bn_train=bn_train)
    x = F.adaptive_avg_pool2d(x, (1, 1))
    x = x.view(x.size(0), -1)
    return F.linear(x, sd["linear.weight"], sd["linear.bias"])


_MOBILENET_DW_STRIDES = [1, 2, 1, 2, 1, 2, 1, 1, 1, 1, 1, 2, 1]  # mobilenet_v1.py:35-81


def mobilenet_forward(x, sd, bn_train=True):
    """MobileNetV1.forward — mobilenet_v1.py:89-93."""
    x = F.relu(_bn(_conv(x, sd, "features.0.", stride=2, pad=1), sd, "features.1.", bn_train))
    for i, s in enumerate(_MOBILENET_DW_STRIDES):
        x = dwsep_block(x, sd, f"features.{i + 3}.", s, bn_train)
    x = F.adaptive_avg_pool2d(x, (1, 1))
    x = x.view(x.size(0), -1)
    return F.linear(x, sd["linear.weight"], sd["linear.bias"])


def model_forward(name, x, sd, bn_train=True):
    if name == "mobilenet1":
        return mobilenet_forward(x, sd, bn_train)
    return resnet_forward(name, x, sd, bn_train)


def clone_state(sd, dtype=None, requires_grad=False):
    """Detached copy of a state_dict; floating tensors optionally cast (fp64 = arbiter) and made leaves."""
    out = {}
    for k, v in sd.items():
        t = torch.as_tensor(v).detach().clone()
        if t.is_floating_point():
            if dtype is not None:
                t = t.to(dtype)
            if requires_grad and not k.endswith(("running_mean", "running_var")):
                t.requires_grad_(True)
        out[k] = t
    return out


def train_step(name, sd, images, labels, opt_state, optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=1e-4,
               alpha=0.9, eps=1.0):
    """One iteration of train() — ResNet/pytorch/train.py:431-485: forward, mean cross-entropy, backward,
    SGD(momentum, wd on every parameter) or RMSprop step.  `sd` holds leaf tensors (clone_state(requires_grad=True));
    updated in place.  Returns the loss."""
    logits = model_forward(name, images, sd, bn_train=True)
    loss = F.cross_entropy(logits, labels)
    params = [(k, v) for k, v in sd.items() if v.requires_grad]
    grads = torch.autograd.grad(loss, [v for _, v in params])
    with torch.no_grad():
        for (k, p), g in zip(params, grads):
            g = g + weight_decay * p
            if optimizer == "sgd":
                buf = opt_state.get(k)
                buf = g.clone() if buf is None else buf.mul_(momentum).add_(g)
                opt_state[k] = buf
                p.sub_(lr * buf)
            else:
                sq = opt_state.get(k)
                sq = torch.zeros_like(p) if sq is None else sq
                sq.mul_(alpha).addcmul_(g, g, value=1 - alpha)
                opt_state[k] = sq
                p.sub_(lr * g / (sq.sqrt() + eps))
    return float(loss.detach())
