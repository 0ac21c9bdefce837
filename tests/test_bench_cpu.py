"""CPU checks of the measurement harness: the reference arm of bench.py and the state it is driven by."""
import json
import os
import subprocess
import sys

import pytest
import torch

from helpers import GOLDEN

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_init, ref_models  # noqa: E402

NAMES = {"resnet34": "ResNet34", "resnet50": "ResNet50", "resnet152": "ResNet152", "mobilenet1": "MobileNetV1"}


@pytest.mark.parametrize("name", sorted(NAMES))
def test_ref_init_has_the_reference_state_dict_layout(name):
    """oracle/ref_init.py lays out the state_dict key by key and shape by shape as the reference classes do
    (tests/golden/reference_interface.json was frozen from them), with the reference's initial distributions."""
    iface = json.load(open(os.path.join(GOLDEN, "reference_interface.json")))
    want = [(k, tuple(s)) for k, s in iface["state_dict"][NAMES[name]]]
    sd = ref_init.init_state(name)
    assert [(k, tuple(v.shape)) for k, v in sd.items()] == want
    n_params = sum(v.numel() for k, v in sd.items() if not k.endswith(("running_mean", "running_var", "num_batches_tracked")))
    assert n_params == iface["param_count"][NAMES[name]]
    if name != "mobilenet1":   # Kaiming-normal fan_out: std = sqrt(2 / (K*R*S))   (resnet50.py:84-93)
        w = sd["conv2x.0.conv2.weight"]
        assert abs(float(w.std()) / (2.0 / (w.shape[0] * 9)) ** 0.5 - 1.0) < 0.05


def test_ref_init_drives_the_oracle_forward():
    sd = ref_models.clone_state(ref_init.init_state("mobilenet1"), dtype=torch.float32)
    y = ref_models.model_forward("mobilenet1", torch.randn(2, 3, 64, 64), sd, bn_train=True)
    assert tuple(y.shape) == (2, 1000) and bool(torch.isfinite(y).all())


def test_reference_arm_prints_one_line_and_never_loads_the_product():
    """`bench.py --impl reference` must time the oracle port only: the process asserts itself that neither the package nor
    libdv_b200.so was mapped (the round-1 arm imported the product to obtain its initial weights)."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--model", "resnet34", "--steps", "1",
                          "--cpu-batch", "2"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and rec["cpu_baseline"]["kind"] == "port" and rec["value"] > 0
    assert rec["e2e"]["h2d_bytes_per_step"] == 0 and rec["cpu_baseline"]["cores"] >= 1 and rec["cpu_baseline"]["cpu"]
