"""Epoch orchestration in the reference's shape: `training_config`, learning-rate schedulers,
loggers, checkpoints and the `run_epochs` / `train` / `validate` loop of ResNet/pytorch/train.py:26-215,218-307,310-428,
431-520 (MobileNet/pytorch/train.py likewise), driving the per-rank `Trainer` instead of nn.DataParallel + torch.optim.

Data loading stays the caller's: `run_epochs` takes the two loaders (any iterable of {'image','annotation'} dicts or
(image, label) tuples, as data_load.py:14-297 yields) where the reference builds them from ../dataset.
"""
import os
import time

import torch
from torch import optim

from . import evaluate, models
from .checkpoint import load_checkpoint, save_checkpoint
from .train import Trainer

# The hot-path entries of the reference's dict (train.py:141-214): optimiser and scheduler classes are kept as the torch
# classes the reference names, so a config written for it reads the same here.
training_config = {
    "resnet34": {
        "name": "resnet34", "batch_size": 256, "num_workers": 16, "model": models.ResNet34,
        "optimizer": optim.SGD, "optimizer_params": {"lr": 0.1, "momentum": 0.9, "weight_decay": 0.0001},
        "scheduler": optim.lr_scheduler.ReduceLROnPlateau, "scheduler_params": {"factor": 0.1, "mode": "max"},
        "total_epochs": 200,
    },
    "resnet50": {
        "name": "resnet50", "batch_size": 256, "num_workers": 16, "model": models.ResNet50,
        "optimizer": optim.SGD, "optimizer_params": {"lr": 0.1, "momentum": 0.9, "weight_decay": 0.0001},
        "scheduler": optim.lr_scheduler.ReduceLROnPlateau, "scheduler_params": {"factor": 0.1, "mode": "max"},
        "total_epochs": 200,
    },
    "resnet152": {   # same recipe; the reference's README trains it with the resnet50 entry
        "name": "resnet152", "batch_size": 128, "num_workers": 16, "model": models.ResNet152,
        "optimizer": optim.SGD, "optimizer_params": {"lr": 0.1, "momentum": 0.9, "weight_decay": 0.0001},
        "scheduler": optim.lr_scheduler.ReduceLROnPlateau, "scheduler_params": {"factor": 0.1, "mode": "max"},
        "total_epochs": 200,
    },
    "mobilenet1": {
        "name": "mobilenet1", "batch_size": 128, "num_workers": 16, "model": models.MobileNetV1, "model_params": {"alpha": 1},
        "optimizer": optim.RMSprop, "optimizer_params": {"lr": 0.045, "alpha": 0.9, "eps": 1.0},
        "scheduler": optim.lr_scheduler.StepLR, "scheduler_params": {"step_size": 2, "gamma": 0.94},
        "total_epochs": 200,
    },
}


class SchedulerTarget(optim.Optimizer):
    """What torch's lr_scheduler classes need from an optimiser — `param_groups` with an 'lr' — backed by the fused
    optimiser of a `Trainer`.  `step()` is a no-op (the Trainer steps inside its own launch sequence); after a
    scheduler moved the rate, `sync()` hands it to the trainer."""

    def __init__(self, trainer):
        self.trainer = trainer
        anchor = torch.nn.Parameter(torch.zeros(1), requires_grad=False)   # torch insists on a non-empty parameter list
        super().__init__([anchor], {"lr": float(trainer.lr)})

    def step(self, closure=None):
        self._opt_called = True
        return None

    def sync(self):
        self.trainer.set_lr(self.param_groups[0]["lr"])
        return self.trainer.lr


def initialize_loggers():
    """train.py:260-279."""
    return {k: {"epochs": [], "value": []} for k in ("train_loss", "val_loss", "val_top1_acc", "val_top5_acc")}


def log_metrics(loggers, name, value, epoch):
    """train.py:282-285."""
    logger = loggers.get(name)
    logger.get("epochs").append(epoch)
    logger.get("value").append(value)


def get_lr(optimizer):
    """train.py:288-290."""
    for param_group in optimizer.param_groups:
        return param_group["lr"]


def make_trainer(net, config, **trainer_kwargs):
    """Trainer with the optimiser the config names (torch class or "sgd"/"rmsprop") and its parameters."""
    opt = config.get("optimizer")
    kind = {optim.SGD: "sgd", optim.RMSprop: "rmsprop"}.get(opt, opt)
    if kind not in ("sgd", "rmsprop"):
        raise NotImplementedError(f"optimizer {opt!r}: the hot path has fused SGD (train.py:166-184) and RMSprop (train.py:199-205)")
    # What the config leaves out takes TORCH's default for that optimiser class, as it does in the reference (train.py:361-365):
    # the mobilenet1 entry names no weight_decay, so RMSprop runs with none ("very little or no weight decay", train.py:199-205).
    defaults = {"sgd": dict(momentum=0.0, weight_decay=0.0), "rmsprop": dict(lr=0.01, alpha=0.99, eps=1e-8, weight_decay=0.0)}[kind]
    params = dict(defaults)
    params.update(config.get("optimizer_params", {}))
    return Trainer(net, optimizer=kind, **params, **trainer_kwargs)


def _unpack(batch, device):
    if isinstance(batch, dict):
        image, label = batch["image"], batch["annotation"]
    else:
        image, label = batch
    return (image.to(device=device, dtype=torch.float32, non_blocking=True),
            label.to(device=device, dtype=torch.long, non_blocking=True))


def train(train_loader, trainer, epoch, loggers, optimizer=None, device="cuda", log_every=10, verbose=True):
    """One epoch — train.py:431-485.  The loss is read back (`.item()`, train.py:470) only every `log_every` batches,
    when the running mean is printed and logged; in between it is accumulated on the device."""
    if verbose:
        print("Start training epoch {}".format(epoch))
    acc = None
    for batch_i, data in enumerate(train_loader):
        image, annotation = _unpack(data, device)
        loss = trainer.step(image, annotation)
        acc = loss.clone() if acc is None else acc + loss
        if batch_i % log_every == log_every - 1:
            batches_loss = float(acc.item()) / log_every
            if verbose:
                print("Time, {}, Epoch: {}, Batch: {}, Training Loss: {}, LR: {}".format(
                    time.strftime("%Y-%m-%dT%H:%M:%S", time.localtime()), epoch, batch_i + 1, batches_loss,
                    get_lr(optimizer) if optimizer is not None else trainer.lr))
            log_metrics(loggers, "train_loss", batches_loss, epoch)
            acc = None


def validate(val_loader, net, epoch, loggers, device="cuda", verbose=True):
    """train.py:488-520."""
    val_loss, top1_acc, top5_acc = evaluate.validate(net, val_loader, device)
    if verbose:
        print("Epoch: {}, Validation Top 1 acc: {}".format(epoch, top1_acc))
        print("Epoch: {}, Validation Top 5 acc: {}".format(epoch, top5_acc))
        print("Epoch: {}, Validation Set Loss: {}".format(epoch, val_loss))
    log_metrics(loggers, "val_top1_acc", top1_acc, epoch)
    log_metrics(loggers, "val_top5_acc", top5_acc, epoch)
    log_metrics(loggers, "val_loss", val_loss, epoch)
    return val_loss, top1_acc, top5_acc


def run_epochs(config, checkpoint_path=None, train_loader=None, val_loader=None, model_dir="./saved_models/", device="cuda",
               use_cuda_graph=True, total_epochs=None, verbose=True):
    """train.py:310-428 with the loaders passed in.  Returns (net, trainer, scheduler, loggers)."""
    if train_loader is None or val_loader is None:
        raise ValueError("run_epochs: pass train_loader and val_loader (the reference builds them from ../dataset, train.py:218-257)")
    Model = config.get("model")
    model_params = config.get("model_params")
    net = Model(**model_params) if model_params is not None else Model()
    net.to(device=device)
    trainer = make_trainer(net, config, use_cuda_graph=use_cuda_graph)
    optimizer = SchedulerTarget(trainer)
    Sched = config.get("scheduler")
    scheduler = Sched(optimizer, **config.get("scheduler_params"))
    loggers = initialize_loggers()
    model_id = time.strftime("%Y-%m-%dT%H:%M:%S", time.localtime())
    model_name = config.get("name")
    start_epoch = 1
    if checkpoint_path is not None:
        start_epoch, saved = load_checkpoint(checkpoint_path, net, trainer, scheduler=scheduler)
        loggers = saved or loggers
        optimizer.param_groups[0]["lr"] = trainer.lr   # what optimizer.load_state_dict restores in the reference (train.py:296)
    os.makedirs(model_dir, exist_ok=True)
    validate(val_loader, net, 0, loggers, device, verbose)
    last = (total_epochs if total_epochs is not None else config.get("total_epochs"))
    for epoch in range(start_epoch, last + 1):
        train(train_loader, trainer, epoch, loggers, optimizer, device, verbose=verbose)
        val_loss, top1_acc, top5_acc = validate(val_loader, net, epoch, loggers, device, verbose)
        optimizer.step()   # keeps torch's "scheduler before optimizer" warning quiet; the real step ran inside Trainer.step
        if isinstance(scheduler, optim.lr_scheduler.ReduceLROnPlateau):
            scheduler.step(top1_acc)      # accuracy, hence mode='max' (train.py:190,408-411)
        else:
            scheduler.step()
        optimizer.sync()
        save_checkpoint(os.path.join(model_dir, "{}-{}-epoch-{}.pt".format(model_name, model_id, epoch)), net, epoch, trainer,
                        scheduler, loggers)
    return net, trainer, scheduler, loggers
