"""The CPU restatement of the reference's image transforms (oracle/ref_augment.py) against outputs of the reference's own
classes frozen by oracle/make_golden_augment.py — exact equality (the uint8 stages are integer / single-rounded float)."""
import os

import numpy as np

from oracle import ref_augment


def test_augment_oracle_equals_reference_outputs():
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "augment_ref.npz"))
    crop = int(d["crop"])
    orders = set()
    for img, out, val, p in zip(d["images"], d["train_out"], d["val_out"], d["params"]):
        h, w, top, left, flip, order = (int(v) for v in p[:6])
        orders.add(order)
        got = ref_augment.train_transform(img[:h, :w], top, left, flip, order, p[6], p[7], p[8], crop)
        assert got.dtype == np.float32 and np.array_equal(got, out)
        assert np.array_equal(ref_augment.val_transform(img[:h, :w], crop), val)
    assert len(orders) >= 4   # the fixture exercises several operation orders


def test_blend_branches():
    img = np.array([[[0, 128, 255]]], np.uint8)
    assert ref_augment._blend(np.zeros_like(img), img, 0.5).tolist() == [[[0, 64, 127]]]        # interpolation truncates
    assert ref_augment._blend(np.zeros_like(img), img, 1.2).tolist() == [[[0, 153, 255]]]       # extrapolation clips
    g = ref_augment._gray(np.array([[[255, 255, 255], [255, 0, 0]]], np.uint8))
    assert g.tolist() == [[255, 76]]


def test_rescale_oracle_equals_reference_rescale():
    """oracle.ref_augment.rescale == the reference's Rescale (cv2.resize) on frozen outputs: exact, up- and down-scaling."""
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "augment_ref.npz"))
    target = int(d["rescale_target"])
    for img, out, (h, w, nh, nw) in zip(d["rescale_in"], d["rescale_out"], d["rescale_hw"]):
        assert ref_augment.rescale_size(int(h), int(w), target) == (int(nh), int(nw))
        got = ref_augment.rescale(img[:h, :w], target)
        assert got.shape == (nh, nw, 3) and np.array_equal(got, out[:nh, :nw])
