"""Freeze outputs of the reference's own transform classes (ResNet/pytorch/data_load.py, imported unmodified from
/root/reference) for tests/golden/augment_ref.npz, together with the parameters its RNG draws produced — recovered by
replaying the same `random` / `numpy.random` call sequence (data_load.py:110,137-138,248-270) from the same seeds.
Run in the build container only:  python oracle/make_golden_augment.py"""
import os
import random
import sys

import numpy as np

REF = "/root/reference/ResNet/pytorch"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "augment_ref.npz")
ORDERS = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]


def main():
    sys.path.insert(0, REF)
    import data_load as dl
    from torchvision import transforms
    crop = 32
    tf = transforms.Compose([dl.RandomHorizontalFlip(0.5), dl.RandomCrop(crop), dl.ColorJitter(brightness=0.2, contrast=0.2, saturation=0.2, hue=0),
                             dl.ToTensor(), dl.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    vf = transforms.Compose([dl.CenterCrop(crop), dl.ToTensor(), dl.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    rng = np.random.default_rng(7)
    rec = {"crop": crop}
    n = 12
    imgs, outs, vals, params = [], [], [], []
    for i in range(n):
        h, w = 40 + 3 * (i % 4), 44 + 5 * (i % 3)
        kind = i % 3
        if kind == 0:
            img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        elif kind == 1:   # smooth, saturated: exercises the clipping branch of blend
            yy, xx = np.mgrid[0:h, 0:w]
            img = np.stack([(yy * 6) % 256, (xx * 5 + 128) % 256, np.full_like(yy, 250 if i % 2 else 3)], -1).astype(np.uint8)
        else:
            img = np.clip(rng.normal(200, 60, (h, w, 3)), 0, 255).astype(np.uint8)
        seed = 100 + i
        random.seed(seed)
        np.random.seed(seed)
        out = tf({"image": img, "annotation": i})["image"].numpy()
        # replay the draws
        random.seed(seed)
        np.random.seed(seed)
        flip = int(random.random() < 0.5)
        top = int(np.random.randint(0, h - crop))
        left = int(np.random.randint(0, w - crop))
        bf = random.uniform(0.8, 1.2)
        cf = random.uniform(0.8, 1.2)
        sf = random.uniform(0.8, 1.2)
        perm = [0, 1, 2]
        random.shuffle(perm)
        order = ORDERS.index(tuple(perm))
        pad = np.zeros((52, 56, 3), np.uint8)
        pad[:h, :w] = img
        imgs.append(pad)
        outs.append(out)
        vals.append(vf({"image": img, "annotation": i})["image"].numpy())
        params.append([h, w, top, left, flip, order, bf, cf, sf])
    rec.update(images=np.stack(imgs), train_out=np.stack(outs), val_out=np.stack(vals), params=np.asarray(params, np.float64))
    # Rescale (data_load.py:72-99, cv2.resize): down- and up-scaling, both orientations, a degenerate 1-pixel-high image
    rs_in, rs_out, rs_hw = [], [], []
    target = 40
    for j, (h, w) in enumerate([(75, 100), (100, 66), (40, 40), (30, 55), (17, 23), (1, 7), (90, 41), (64, 200)]):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        out = dl.Rescale(target)({"image": img, "annotation": j})["image"]
        pi = np.zeros((100, 200, 3), np.uint8)
        pi[:h, :w] = img
        po = np.zeros((300, 300, 3), np.uint8)
        po[:out.shape[0], :out.shape[1]] = out
        rs_in.append(pi)
        rs_out.append(po)
        rs_hw.append([h, w, out.shape[0], out.shape[1]])
    rec.update(rescale_target=target, rescale_in=np.stack(rs_in), rescale_out=np.stack(rs_out), rescale_hw=np.asarray(rs_hw, np.int64))
    import PIL
    import torchvision
    import cv2
    rec["versions"] = np.asarray([PIL.__version__, torchvision.__version__, cv2.__version__])
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
