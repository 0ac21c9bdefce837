"""CPU restatement (test infrastructure only) of the reference's training / validation image transforms after decoding
— ResNet/pytorch/data_load.py: Rescale :72-99 (cv2.resize, INTER_LINEAR), RandomHorizontalFlip :102-112, RandomCrop :115-143, CenterCrop :146-172,
ColorJitter :213-297 (torchvision's PIL path: brightness, contrast, saturation in a shuffled order; hue = 0),
ToTensor :175-193 (float, NOT divided by 255), Normalize :196-210 — as composed at ResNet/pytorch/train.py:315-331.

The arithmetic lives in two third-party dependencies absent from /root/reference: Pillow (ImageEnhance / Image.blend /
convert("L"); installed 12.2.0) and torchvision (functional.normalize; installed 0.26.0).  Their published algorithms are
restated here on uint8 / float32 numpy arrays with explicit parameters instead of RNG draws; `oracle/make_golden_augment.py`
pins this file against the reference's own transform classes (tests/golden/augment_ref.npz).

    Image.blend(deg, img, a):  v = float32(deg) + a * float32(img - deg);  0 <= a <= 1: trunc(v)  else clip to [0,255], trunc
    convert("L"):              (19595 R + 38470 G + 7471 B + 0x8000) >> 16
    brightness: deg = 0        contrast: deg = int(mean(L) + 0.5)        saturation: deg = L
"""
import numpy as np


def _resize_coefs(d, s, vertical):
    """Source indices and 11-bit fixed-point weights of OpenCV's INTER_LINEAR for uint8 (resize.cpp, the HResizeLinear /
    VResizeLinear pair): pixel centres aligned, the fraction taken in float32, weights = cvRound(w * 2048).  Along x the
    fraction is zeroed where the window leaves the image; along y it is kept and the ROW INDICES are clamped instead."""
    scale = 1.0 / (np.float64(d) / np.float64(s))
    f = ((np.arange(d) + 0.5) * scale - 0.5).astype(np.float32)
    s0 = np.floor(f).astype(np.int64)
    fr = (f - s0.astype(np.float32)).astype(np.float32)
    if not vertical:
        lo = s0 < 0
        fr[lo] = 0
        s0[lo] = 0
        hi = s0 >= s - 1
        fr[hi] = 0
        s0[hi] = s - 1
    w1 = np.rint(fr * np.float32(2048)).astype(np.int64)
    w0 = np.rint((np.float32(1) - fr) * np.float32(2048)).astype(np.int64)
    return np.clip(s0, 0, s - 1), np.clip(s0 + 1, 0, s - 1), w0, w1


def resize_bilinear_u8(img, new_h, new_w):
    """cv2.resize(img, (new_w, new_h)) for a uint8 HWC image, bit for bit (third-party dependency: opencv-python, installed
    4.13.0; pinned against cv2 itself and against the reference's Rescale in tests/test_augment_oracle_cpu.py)."""
    h, w = img.shape[:2]
    x0, x1, ax0, ax1 = _resize_coefs(new_w, w, False)
    y0, y1, ay0, ay1 = _resize_coefs(new_h, h, True)
    src = img.astype(np.int64)
    rows = src[:, x0] * ax0[None, :, None] + src[:, x1] * ax1[None, :, None]
    r0, r1 = rows[y0], rows[y1]
    out = (((ay0[:, None, None] * (r0 >> 4)) >> 16) + ((ay1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2
    return out.astype(np.uint8)


def rescale_size(h, w, output_size):
    """Rescale.__call__ (data_load.py:86-97): the smaller edge becomes output_size, the other keeps the aspect ratio (int())."""
    if h > w:
        new_h, new_w = output_size * h / w, output_size
    else:
        new_h, new_w = output_size, output_size * w / h
    return int(new_h), int(new_w)


def rescale(img, output_size):
    """data_load.py:72-99 for an int output_size."""
    nh, nw = rescale_size(img.shape[0], img.shape[1], output_size)
    return resize_bilinear_u8(img, nh, nw)


ORDERS = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]   # 0 brightness, 1 contrast, 2 saturation


def _gray(img):
    r, g, b = (img[..., i].astype(np.int64) for i in range(3))
    return ((r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16).astype(np.uint8)


def _blend(deg, img, alpha):
    a = np.float32(alpha)
    v = deg.astype(np.float32) + a * (img.astype(np.int32) - deg.astype(np.int32)).astype(np.float32)
    if not (0.0 <= float(a) <= 1.0):
        v = np.clip(v, np.float32(0), np.float32(255))
    return v.astype(np.uint8)   # truncation; v is within [0, 255]


def color_jitter(img, order, brightness, contrast, saturation):
    """img: [H,W,3] uint8 RGB.  order: index into ORDERS.  Factors as drawn by ColorJitter.get_params (data_load.py:246-270)."""
    for op in ORDERS[order]:
        if op == 0:
            img = _blend(np.zeros_like(img), img, brightness)
        elif op == 1:
            g = _gray(img)
            mean = int(g.astype(np.int64).sum() / g.size + 0.5)
            img = _blend(np.full_like(img, mean), img, contrast)
        else:
            img = _blend(np.repeat(_gray(img)[..., None], 3, axis=2), img, saturation)
    return img


def train_transform(img, top, left, flip, order, brightness, contrast, saturation, crop=224,
                    mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)):
    """[Hs,Ws,3] uint8 (already Rescale'd) -> [3,crop,crop] float32, train.py:315-325."""
    if flip:
        img = img[:, ::-1]
    img = img[top:top + crop, left:left + crop]
    img = color_jitter(np.ascontiguousarray(img), order, brightness, contrast, saturation)
    return normalize(img, mean, std)


def val_transform(img, crop=224, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)):
    """Rescale'd uint8 image -> CenterCrop -> ToTensor -> Normalize, train.py:327-331."""
    h, w = img.shape[:2]
    top, left = (h - crop) // 2, (w - crop) // 2
    return normalize(img[top:top + crop, left:left + crop], mean, std)


def normalize(img_u8, mean, std):
    x = img_u8.astype(np.float32).transpose(2, 0, 1)
    m = np.asarray(mean, np.float32)[:, None, None]
    s = np.asarray(std, np.float32)[:, None, None]
    return (x - m) / s
