"""GPU input pipeline after decode + rescale.

The reference transforms every image on a DataLoader worker (ResNet/pytorch/data_load.py:102-297 composed at
train.py:315-331) and was input-bound at 99 images/s/GPU.  Here the host only decodes and rescales (cv2, unchanged) and
ships uint8 batches; flip, crop, colour jitter, the float conversion (no /255, as the reference) and the normalisation run
as `dv_augment` on the device and write the model's input layout directly.

`GpuTransform.draw` makes the random draws exactly as the reference's classes do, in the same order and from the same
generators (`random`, `numpy.random`), so a seeded run picks the same crops and factors image for image.
"""
import ctypes as C
import random

import numpy as np
import torch

from . import config
from .functional import _CL, _dv_dtype, _require_cuda, _stream
from .lib import AugmentParams, RescaleParams, check, lib

ORDERS = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]   # 0 brightness, 1 contrast, 2 saturation
IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)   # train.py:324


class GpuTransform:
    """train=True: RandomHorizontalFlip(0.5) -> RandomCrop(crop) -> ColorJitter(brightness, contrast, saturation) -> ToTensor
    -> Normalize (train.py:315-325).  train=False: CenterCrop(crop) -> ToTensor -> Normalize (train.py:327-331)."""

    def __init__(self, crop=224, train=True, brightness=0.2, contrast=0.2, saturation=0.2, flip_p=0.5, mean=IMAGENET_MEAN,
                 std=IMAGENET_STD, dtype=None, nchw=False):
        self.crop, self.train = int(crop), bool(train)
        self.brightness, self.contrast, self.saturation, self.flip_p = brightness, contrast, saturation, flip_p
        self.mean, self.std = tuple(mean), tuple(std)
        self.dtype, self.nchw = dtype, bool(nchw)

    def _to_device(self, arr, dev):
        """Small per-batch tables go through a ring of PINNED staging buffers: a copy from pageable memory makes the host wait
        for everything already queued on the stream (the previous training step), which serialises host and GPU."""
        ring = self.__dict__.setdefault("_ring", {})
        key = (arr.shape, arr.dtype.str)
        slots = ring.setdefault(key, {"i": 0, "bufs": [], "events": []})
        if len(slots["bufs"]) < 4:
            slots["bufs"].append(torch.from_numpy(np.empty_like(arr)).pin_memory())
            slots["events"].append(None)
        i = slots["i"] % len(slots["bufs"])
        slots["i"] += 1
        if slots["events"][i] is not None:
            slots["events"][i].synchronize()       # the copy that last used this slot has run
        slots["bufs"][i].numpy()[...] = arr
        out = slots["bufs"][i].to(dev, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(dev))
        slots["events"][i] = ev
        return out

    def draw(self, sizes):
        """sizes: [(h, w)] of the rescaled images.  Returns (geom int32 [B,4], factors float32 [B,3]) — one image after the
        other, each consuming the generators in the order flip, crop top/left, brightness, contrast, saturation, shuffle
        (data_load.py:110,137-138,248-270)."""
        geom, fac = [], []
        for h, w in sizes:
            if not self.train:
                geom.append([(h - self.crop) // 2, (w - self.crop) // 2, 0, -1])      # data_load.py:165-166
                fac.append([1.0, 1.0, 1.0])
                continue
            flip = int(random.random() < self.flip_p)
            top = int(np.random.randint(0, h - self.crop))
            left = int(np.random.randint(0, w - self.crop))
            f = [1.0, 1.0, 1.0]
            ops = []
            for k, amount in enumerate((self.brightness, self.contrast, self.saturation)):
                if amount > 0:
                    f[k] = random.uniform(max(0, 1 - amount), 1 + amount)
                    ops.append(k)
            random.shuffle(ops)
            order = -1
            if ops:
                rest = [k for k in (0, 1, 2) if k not in ops]   # switched-off operations run as identity blends
                order = ORDERS.index(tuple(ops + rest))
            geom.append([top, left, flip, order])
            fac.append(f)
        return np.asarray(geom, np.int32), np.asarray(fac, np.float32)

    def __call__(self, images_u8, sizes=None, params=None):
        """images_u8: uint8 CUDA tensor [B,Hs,Ws,3] (RGB, each image in the top-left corner of its slot); sizes: [(h,w)] or
        None when every image fills its slot; params: (geom, factors) to replay given draws instead of drawing.
        Returns [B,3,crop,crop] — channels_last strides in the compute dtype (what the models consume without a further
        layout pass), or contiguous NCHW fp32-like when nchw=True."""
        _require_cuda(images_u8, "GpuTransform")
        if images_u8.dtype != torch.uint8 or images_u8.dim() != 4 or images_u8.shape[-1] != 3 or not images_u8.is_contiguous():
            raise TypeError("GpuTransform: expected a contiguous uint8 [B,H,W,3] tensor")
        B, Hs, Ws, _ = images_u8.shape
        if sizes is None:
            sizes = [(Hs, Ws)] * B
        geom, fac = params if params is not None else self.draw(sizes)
        dev = images_u8.device
        geom_d = self._to_device(np.ascontiguousarray(geom, np.int32), dev)
        fac_d = self._to_device(np.ascontiguousarray(fac, np.float32), dev)
        widths_d = self._to_device(np.asarray([w for _, w in sizes], np.int32), dev)
        sums = torch.empty(B, dtype=torch.int32, device=dev)
        dtype = self.dtype or config.compute_dtype()
        if self.nchw:
            out = torch.empty((B, 3, self.crop, self.crop), dtype=dtype, device=dev)
        else:
            out = torch.empty((B, 3, self.crop, self.crop), dtype=dtype, device=dev, memory_format=_CL)
        p = AugmentParams()
        p.src, p.dst, p.geom, p.factors = images_u8.data_ptr(), out.data_ptr(), geom_d.data_ptr(), fac_d.data_ptr()
        p.widths, p.gray_sums = widths_d.data_ptr(), sums.data_ptr()
        p.B, p.Hs, p.Ws, p.crop, p.dtype, p.nchw = B, Hs, Ws, self.crop, _dv_dtype(dtype), int(self.nchw)
        p.mean = (C.c_float * 3)(*self.mean)
        p.std = (C.c_float * 3)(*self.std)
        check(lib.dv_augment(C.byref(p), _stream()), "dv_augment")
        return out


class GpuRescale:
    """Rescale(output_size) of data_load.py:72-99 for a batch of decoded images of different sizes: the host hands over the
    uint8 HWC arrays as decoded (numpy or torch, any mix of sizes), `dv_rescale` resizes them all in one launch with OpenCV's
    INTER_LINEAR arithmetic bit for bit, and the result is the slot tensor + size list that `GpuTransform` consumes."""

    def __init__(self, output_size=256):
        self.output_size = int(output_size)

    def sizes(self, h, w):
        s = self.output_size                      # data_load.py:88-95, float arithmetic and int() as there
        if h > w:
            new_h, new_w = s * h / w, s
        else:
            new_h, new_w = s, s * w / h
        return int(new_h), int(new_w)

    def __call__(self, images, device="cuda"):
        imgs = [torch.as_tensor(im) for im in images]
        for im in imgs:
            if im.dtype != torch.uint8 or im.dim() != 3 or im.shape[2] != 3:
                raise TypeError("GpuRescale: expected uint8 [H,W,3] images")
        src_hw = [(int(im.shape[0]), int(im.shape[1])) for im in imgs]
        dst_hw = [self.sizes(h, w) for h, w in src_hw]
        offs, total = [], 0
        for h, w in src_hw:
            offs.append(total)
            total += (h * w * 3 + 15) // 16 * 16
        staged = torch.empty(total, dtype=torch.uint8).pin_memory() if torch.cuda.is_available() else torch.empty(total, dtype=torch.uint8)
        for im, o, (h, w) in zip(imgs, offs, src_hw):
            staged[o:o + h * w * 3] = im.contiguous().reshape(-1)
        src = staged.to(device, non_blocking=True)
        B = len(imgs)
        Hs, Ws = max(h for h, _ in dst_hw), max(w for _, w in dst_hw)
        out = torch.empty((B, Hs, Ws, 3), dtype=torch.uint8, device=device)
        off_d = torch.tensor(offs, dtype=torch.int64).to(device, non_blocking=True)
        shw_d = torch.tensor(src_hw, dtype=torch.int32).to(device, non_blocking=True)
        dhw_d = torch.tensor(dst_hw, dtype=torch.int32).to(device, non_blocking=True)
        p = RescaleParams()
        p.src, p.src_off, p.src_hw, p.dst, p.dst_hw = src.data_ptr(), off_d.data_ptr(), shw_d.data_ptr(), out.data_ptr(), dhw_d.data_ptr()
        p.B, p.Hs, p.Ws = B, Hs, Ws
        check(lib.dv_rescale(C.byref(p), _stream()), "dv_rescale")
        return out, dst_hw
