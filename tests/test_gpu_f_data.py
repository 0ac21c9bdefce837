"""GPU parity of the input-pipeline kernel (dv_augment) against outputs of the reference's own transform
classes frozen in tests/golden/augment_ref.npz, and against the CPU restatement on full-size images."""
import random

import numpy as np
import pytest
import torch

import deep_vision_b200 as dv
from deep_vision_b200.data import GpuTransform
from oracle import ref_augment

pytestmark = pytest.mark.gpu


def _golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "augment_ref.npz"))


@pytest.mark.parametrize("nchw", [True, False], ids=["nchw", "nhwc"])
def test_augment_matches_reference_transforms_bit_for_bit(nchw):
    d = _golden()
    crop = int(d["crop"])
    P = d["params"]
    sizes = [(int(p[0]), int(p[1])) for p in P]
    geom = np.asarray([[int(p[2]), int(p[3]), int(p[4]), int(p[5])] for p in P], np.int32)
    fac = np.asarray([[p[6], p[7], p[8]] for p in P], np.float32)
    imgs = torch.from_numpy(d["images"]).cuda()
    out = GpuTransform(crop, train=True, dtype=torch.float32, nchw=nchw)(imgs, sizes, (geom, fac))
    torch.cuda.synchronize()
    assert tuple(out.shape) == (len(P), 3, crop, crop)
    assert out.is_contiguous() if nchw else out.is_contiguous(memory_format=torch.channels_last)
    assert np.array_equal(out.cpu().numpy(), d["train_out"])              # fp32: exact
    val = GpuTransform(crop, train=False, dtype=torch.float32, nchw=nchw)(imgs, sizes)
    assert np.array_equal(val.cpu().numpy(), d["val_out"])
    bf = GpuTransform(crop, train=True, dtype=torch.bfloat16, nchw=nchw)(imgs, sizes, (geom, fac))
    assert torch.equal(bf.float().cpu(), torch.from_numpy(d["train_out"]).to(torch.bfloat16).float())


def test_augment_draws_like_the_reference_and_full_size_matches_oracle():
    """Seeded draws reproduce the fixture's parameters (same generators, same order), and a full-size batch (256 x 341 -> 224)
    equals the CPU restatement exactly."""
    d = _golden()
    crop = int(d["crop"])
    t = GpuTransform(crop, train=True)
    for i, p in enumerate(d["params"]):
        random.seed(100 + i)
        np.random.seed(100 + i)
        geom, fac = t.draw([(int(p[0]), int(p[1]))])
        assert list(geom[0]) == [int(p[2]), int(p[3]), int(p[4]), int(p[5])]
        assert np.allclose(fac[0], p[6:9], rtol=1e-7)
    rng = np.random.default_rng(3)
    B, Hs, Ws = 6, 256, 341
    imgs = rng.integers(0, 256, (B, Hs, Ws, 3), dtype=np.uint8)
    imgs[1] = 255
    imgs[2, :, :, 1] = 0
    sizes = [(256, 341), (256, 300), (256, 341), (240, 341), (256, 256), (256, 341)]
    random.seed(11)
    np.random.seed(11)
    t = GpuTransform(224, train=True, dtype=torch.float32, nchw=True)
    geom, fac = t.draw(sizes)
    out = t(torch.from_numpy(imgs).cuda(), sizes, (geom, fac)).cpu().numpy()
    for b in range(B):
        h, w = sizes[b]
        ref = ref_augment.train_transform(imgs[b, :h, :w], int(geom[b, 0]), int(geom[b, 1]), int(geom[b, 2]), int(geom[b, 3]),
                                          fac[b, 0], fac[b, 1], fac[b, 2], 224)
        assert np.array_equal(out[b], ref), b


def test_augmented_batch_feeds_the_model_without_a_layout_pass():
    dv.set_compute_dtype(torch.bfloat16)
    imgs = torch.randint(0, 256, (4, 72, 80, 3), dtype=torch.uint8).cuda()
    x = GpuTransform(64, train=True)(imgs)
    assert x.dtype == torch.bfloat16 and x.is_contiguous(memory_format=torch.channels_last)
    assert dv.functional.to_compute_layout(x) is x
    torch.manual_seed(0)
    net = dv.models.MobileNetV1().cuda().eval()
    with torch.no_grad():
        y = net(x * 0.02)   # the reference's un-scaled 0..255 normalisation gives inputs of magnitude ~500
    assert tuple(y.shape) == (4, 1000) and torch.isfinite(y).all()


def test_rescale_matches_reference_rescale_bit_for_bit():
    """dv_rescale on a ragged batch == outputs of the reference's Rescale (cv2.resize) frozen in the fixture, and the composed
    pipeline decode -> GpuRescale -> GpuTransform(validation) == the oracle's rescale + centre crop + normalise."""
    from deep_vision_b200.data import GpuRescale
    d = _golden()
    target = int(d["rescale_target"])
    imgs = [img[:h, :w].copy() for img, (h, w, _, _) in zip(d["rescale_in"], d["rescale_hw"])]
    slots, sizes = GpuRescale(target)(imgs)
    torch.cuda.synchronize()
    got = slots.cpu().numpy()
    for b, (out, (h, w, nh, nw)) in enumerate(zip(d["rescale_out"], d["rescale_hw"])):
        assert sizes[b] == (int(nh), int(nw))
        assert np.array_equal(got[b, :nh, :nw], out[:nh, :nw]), b
        assert not got[b, nh:].any() and not got[b, :, nw:].any()
    # full-size: 375x500 and 500x333 photographs -> 256 on the short side -> centre crop 224
    rng = np.random.default_rng(9)
    big = [rng.integers(0, 256, (375, 500, 3), dtype=np.uint8), rng.integers(0, 256, (500, 333, 3), dtype=np.uint8),
           rng.integers(0, 256, (120, 90, 3), dtype=np.uint8)]
    slots, sizes = GpuRescale(256)(big)
    val = GpuTransform(224, train=False, dtype=torch.float32, nchw=True)(slots, sizes).cpu().numpy()
    for b, im in enumerate(big):
        ref = ref_augment.val_transform(ref_augment.rescale(im, 256), 224)
        assert np.array_equal(val[b], ref), b
