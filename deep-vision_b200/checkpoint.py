"""Checkpoint files in the reference's format (ResNet/pytorch/train.py:293-307,417-428):

    {'epoch': int, 'model': state_dict, 'optimizer': ..., 'scheduler': ..., 'loggers': dict of lists}

`model` holds exactly the reference's keys and OIHW fp32 tensors and `optimizer` is torch.optim's own state_dict layout
(per-parameter momentum_buffer / square_avg in OIHW + one param_group, Trainer.state_dict), so files move in both
directions: the reference's load_checkpoint (train.py:293-307) accepts a file written here and vice versa.  A reference
checkpoint written under nn.DataParallel carries a 'module.' prefix on every key (ResNet/pytorch/README.md:87-96), which
`load_checkpoint` strips."""
import pickle

import torch

from . import config
from .train import flat_view


def _strip_module_prefix(state):
    if all(k.startswith("module.") for k in state):
        return {k[len("module."):]: v for k, v in state.items()}
    return state


def _load_torch_optimizer_state(trainer, opt):
    """torch.optim.SGD keeps 'momentum_buffer', RMSprop 'square_avg', keyed by the parameter's index in
    net.parameters() order — the order FlatParams lays the flat buffers out in."""
    key = {"sgd": "momentum_buffer", "rmsprop": "square_avg"}[trainer.kind]
    params = [p for p in trainer.model.parameters()]
    steps = 0
    for idx, p in enumerate(params):
        st = opt["state"].get(idx)
        if not st or st.get(key) is None:
            continue
        buf = st[key].to(device=trainer.flat_s.device, dtype=torch.float32)
        off = (p.data_ptr() - trainer.flat_p.data_ptr()) // 4
        # the flat state pairs with flat_p element by element, and 4-D filters sit there as KRSC: a plain .view(p.shape)
        # would scramble every 3x3 / 7x7 buffer
        flat_view(trainer.flat_s, off, p.shape).copy_(buf.reshape(p.shape))
        steps = max(steps, int(st.get("step", 1)) if not torch.is_tensor(st.get("step", 1)) else int(st["step"].item()))
    trainer.steps_done = max(trainer.steps_done, steps, 1 if opt["state"] else 0)
    groups = opt.get("param_groups") or []
    if groups and "dv_steps" in groups[0]:
        trainer.steps_done = int(groups[0]["dv_steps"])
    if groups and "lr" in groups[0]:
        trainer.set_lr(groups[0]["lr"])


def save_checkpoint(path, net, epoch, trainer=None, scheduler=None, loggers=None):
    model_state = {k: (v.detach().contiguous().cpu() if torch.is_tensor(v) else v) for k, v in net.state_dict().items()}
    torch.save({"epoch": epoch, "model": model_state, "optimizer": trainer.state_dict() if trainer is not None else None,
                "scheduler": scheduler.state_dict() if scheduler is not None else None, "loggers": loggers or {}}, path)


def load_checkpoint(path, net, trainer=None, map_location="cpu", scheduler=None):
    """Loads a checkpoint written by save_checkpoint OR by the reference's train.py.  Returns (next epoch, loggers).
    A reference checkpoint's 'optimizer' entry is torch.optim's per-parameter state_dict (train.py:295-302): its momentum
    / squared-average buffers are copied into the trainer's flat state in parameter order."""
    try:    # tensors, numbers, lists and dicts only: what both writers put into the file
        ckpt = torch.load(path, map_location=map_location, weights_only=True)
    except (pickle.UnpicklingError, RuntimeError):
        ckpt = torch.load(path, map_location=map_location, weights_only=False)   # e.g. a scheduler state holding a function
    net.load_state_dict(_strip_module_prefix(ckpt["model"]))
    config.bump_weight_epoch()  # cached compute-dtype filter copies are stale now
    if trainer is not None:
        trainer.refresh_from_parameters()
        opt = ckpt.get("optimizer")
        if isinstance(opt, dict) and "state" in opt and opt.get("optimizer") == trainer.kind:
            trainer.flat_s.copy_(opt["state"].to(trainer.flat_s.device))
            trainer.steps_done = int(opt.get("steps", 0))
            if "lr" in opt:
                trainer.set_lr(opt["lr"])
        elif isinstance(opt, dict) and "param_groups" in opt and isinstance(opt.get("state"), dict):
            _load_torch_optimizer_state(trainer, opt)
    if scheduler is not None and ckpt.get("scheduler") is not None:
        scheduler.load_state_dict(ckpt["scheduler"])
    return int(ckpt.get("epoch", 0)) + 1, ckpt.get("loggers", {})
