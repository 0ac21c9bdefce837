"""Checkpoint files in the reference's format (ResNet/pytorch/train.py:293-307,417-428):

    {'epoch': int, 'model': state_dict, 'optimizer': ..., 'scheduler': ..., 'loggers': dict of lists}

`model` holds exactly the reference's keys and OIHW fp32 tensors, so files move in both directions; a reference
checkpoint written under nn.DataParallel carries a 'module.' prefix on every key (ResNet/pytorch/README.md:87-96), which
`load_checkpoint` strips."""
import torch

from . import config


def _strip_module_prefix(state):
    if all(k.startswith("module.") for k in state):
        return {k[len("module."):]: v for k, v in state.items()}
    return state


def save_checkpoint(path, net, epoch, trainer=None, scheduler=None, loggers=None):
    model_state = {k: (v.detach().contiguous().cpu() if torch.is_tensor(v) else v) for k, v in net.state_dict().items()}
    torch.save({"epoch": epoch, "model": model_state, "optimizer": trainer.state_dict() if trainer is not None else None,
                "scheduler": scheduler.state_dict() if scheduler is not None else None, "loggers": loggers or {}}, path)


def load_checkpoint(path, net, trainer=None, map_location="cpu"):
    """Loads a checkpoint written by save_checkpoint OR by the reference's train.py.  Returns (next epoch, loggers)."""
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    net.load_state_dict(_strip_module_prefix(ckpt["model"]))
    config.bump_weight_epoch()  # cached compute-dtype filter copies are stale now
    if trainer is not None:
        trainer.refresh_from_parameters()
        opt = ckpt.get("optimizer")
        if isinstance(opt, dict) and "state" in opt and opt.get("optimizer") == trainer.kind:
            trainer.flat_s.copy_(opt["state"].to(trainer.flat_s.device))
            trainer.steps_done = int(opt.get("steps", 0))
    return int(ckpt.get("epoch", 0)) + 1, ckpt.get("loggers", {})
