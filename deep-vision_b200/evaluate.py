"""Evaluation pass — the restatement of `validate()` and `accuracy()` in ResNet/pytorch/train.py:488-538.

`validate` runs the model in eval mode (BatchNorm normalises with its running statistics through dv_bn_apply's
eval path, no statistics pass, no gradient state) and reports mean cross-entropy, top-1 and top-5 accuracy in percent,
as the reference prints them (train.py:509-518)."""
import torch

from . import functional as F


def accuracy(output, target, topk=(1,)):
    """Precision@k in percent for each k of `topk` (train.py:524-538, same signature and return type: a list of 1-element
    tensors).  output: [B,classes] logits, target: [B].  A sample counts for k when its label is among its k largest logits."""
    with torch.no_grad():
        ranked = output.topk(max(topk), dim=1).indices                 # [B, maxk], best first
        hit_rank = (ranked == target.view(-1, 1)).float().cumsum(dim=1)  # [B, maxk]: 1 from the column where the label sits
        scale = 100.0 / target.size(0)
        return [hit_rank[:, k - 1].sum().reshape(1) * scale for k in topk]


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
