"""Evaluation pass — the restatement of `validate()` and `accuracy()` in ResNet/pytorch/train.py:488-538.

`validate` runs the model in eval mode (BatchNorm normalises with its running statistics through dv_bn_apply's
eval path, no statistics pass, no gradient state) and reports mean cross-entropy, top-1 and top-5 accuracy in percent,
as the reference prints them (train.py:509-518)."""
import torch

from . import functional as F


def accuracy(output, target, topk=(1,)):
    """train.py:524-538 — precision@k in percent for each k of `topk`.  output: [B,classes] logits, target: [B]."""
    with torch.no_grad():
        maxk = max(topk)
        batch_size = target.size(0)
        _, pred = output.topk(maxk, 1, True, True)
        pred = pred.t()
        correct = pred.eq(target.view(1, -1).expand_as(pred))
        res = []
        for k in topk:
            correct_k = correct[:k].reshape(-1).float().sum(0, keepdim=True)
            res.append(correct_k.mul_(100.0 / batch_size))
        return res


def validate(net, loader, device="cuda"):
    """train.py:488-520.  `loader` yields dicts {'image': [B,3,H,W] float, 'annotation': [B] int} like the reference's
    dataset (data_load.py) or (image, label) tuples.  Returns (mean loss, top-1 %, top-5 %)."""
    net.eval()
    total_loss, top1, top5, batches = 0.0, 0.0, 0.0, 0
    with torch.no_grad():
        for batch in loader:
            if isinstance(batch, dict):
                image, label = batch["image"], batch["annotation"]
            else:
                image, label = batch
            image = image.to(device=device, dtype=torch.float32, non_blocking=True)
            label = label.to(device=device, dtype=torch.long, non_blocking=True)
            output = net(image)
            total_loss += float(F.cross_entropy(output, label))
            p1, p5 = accuracy(output, label, topk=(1, 5))
            top1 += float(p1)
            top5 += float(p5)
            batches += 1
    n = max(batches, 1)
    return total_loss / n, top1 / n, top5 / n
