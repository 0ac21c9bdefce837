"""GPU parity of the whole models against logits frozen from the reference classes (tests/golden/model_*.npz).

fp32: eval-mode logits <= 1e-4 of the reference's fp64 logits; train-mode logits are gated at the larger of 1e-4 and
10x the reference's own fp32-vs-fp64 distance on that fixture (random-init BN networks amplify rounding; SURVEY.md
App. B).  bf16: eval-mode logits <= 2e-2 with identical top-1 (north_star) for all four models; train-mode is reported against torch
autocast's own distance on the same fixture (stored in it) and gated at max(3e-1, 2x that).
"""
import numpy as np
import pytest
import torch

import deep_vision_b200 as dv
from helpers import load_golden, rel_l2

pytestmark = pytest.mark.gpu

CTORS = {"resnet34": lambda: dv.models.ResNet34(), "resnet50": lambda: dv.models.ResNet50(),
         "resnet152": lambda: dv.models.ResNet152(), "mobilenet1": lambda: dv.models.MobileNetV1()}
FIXTURES = {
    "model_resnet34_b4_224_eval": "resnet34",     # BASELINE.json configs[0]
    "model_resnet34_b4_128_train": "resnet34",
    "model_resnet50_b4_128_train": "resnet50",
    "model_resnet50_b2_64_eval": "resnet50",
    "model_resnet152_b4_128_train": "resnet152",
    "model_resnet152_b2_64_eval": "resnet152",
    "model_mobilenet1_b4_128_train": "mobilenet1",
    "model_mobilenet1_b2_64_eval": "mobilenet1",
}


@pytest.fixture(autouse=True)
def _restore_dtype():
    yield
    dv.set_compute_dtype(torch.bfloat16)


def _logits(name, dtype):
    rec = load_golden(name)
    torch.manual_seed(int(rec["seed"]))
    m = CTORS[FIXTURES[name]]()
    psum = float(sum(p.detach().double().sum() for p in m.parameters()))
    if abs(psum - float(rec["param_sum"])) > 1e-9 * float(rec["param_abs_sum"]):
        pytest.skip("torch RNG differs from the build that froze the fixture")
    g = torch.Generator().manual_seed(int(rec["seed"]) + 1)
    x = torch.randn(tuple(int(v) for v in rec["in_shape"]), generator=g)
    dv.set_compute_dtype(dtype)
    m = m.cuda()
    m.train(bool(rec["bn_train"]))
    with torch.no_grad():
        y = m(x.cuda())
    torch.cuda.synchronize()
    assert y.dtype == torch.float32 and tuple(y.shape) == (x.shape[0], 1000)
    return y.cpu().numpy(), rec


@pytest.mark.parametrize("name", sorted(FIXTURES))
def test_model_fp32_logits(name):
    y, rec = _logits(name, torch.float32)
    err = rel_l2(y, rec["logits_fp64"])
    floor = rel_l2(rec["logits_fp32"], rec["logits_fp64"])
    bar = 1e-4 if not bool(rec["bn_train"]) else max(1e-4, 10 * floor)
    assert err < bar, (err, floor)


@pytest.mark.parametrize("name", sorted(FIXTURES))
def test_model_bf16_logits(name):
    y, rec = _logits(name, torch.bfloat16)
    err = rel_l2(y, rec["logits_fp64"])
    if bool(rec["bn_train"]):
        # chaotic at random init: torch's own bf16 autocast of the reference module sits at `floor` from fp64
        floor = float(rec["torch_bf16_autocast_rel"])
        print(f"{name}: ours {err:.3e}  torch bf16 autocast {floor:.3e}")
        assert err < max(3e-1, 2.0 * floor), (err, floor)
    else:
        floor = float(rec["torch_bf16_autocast_rel"])
        print(f"{name}: ours {err:.3e}  torch bf16 autocast {floor:.3e}")
        # north_star bar for every model, ResNet-152 included: 2e-2 and identical top-1.  (The default eval path folds
        # BatchNorm into the convolution's epilogue: one bf16 rounding per conv->BN(+identity)(+ReLU) group; measured
        # 5.6e-3 on the 152-layer stack, 5.8e-3 on ResNet-50.)
        assert err < 2e-2, (err, floor)
        assert np.array_equal(y.argmax(1), rec["logits_fp64"].argmax(1))


UNFUSED_EVAL = ["model_resnet34_b4_224_eval", "model_resnet50_b2_64_eval", "model_mobilenet1_b2_64_eval",
                pytest.param("model_resnet152_b2_64_eval", marks=pytest.mark.xfail(
                    strict=True, reason="set_fold_bn(False) rounds the conv output to bf16 BEFORE the eval-mode BN (two roundings per "
                                        "group): 3.08e-2 on the 152-layer stack; a CPU emulation of exactly that arithmetic on the "
                                        "reference weights gives 3.08e-2 as well (tools/diag_eval_stream.py).  Not the default path."))]


@pytest.mark.parametrize("name", UNFUSED_EVAL)
def test_model_bf16_logits_unfused_eval_path(name):
    """The debugging switch set_fold_bn(False): eval-mode BN as its own pass after a bf16 convolution output.  Held to the
    same 2e-2 bar; the one case that misses it is marked, not hidden behind a wider gate."""
    dv.set_fold_bn(False)
    try:
        y, rec = _logits(name, torch.bfloat16)
    finally:
        dv.set_fold_bn(True)
    err = rel_l2(y, rec["logits_fp64"])
    print(f"{name}: unfused eval path {err:.3e}")
    assert err < 2e-2, err


@pytest.mark.parametrize("dtype,bar", [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)], ids=["fp32", "bf16"])
def test_resnet50_eval_logits_at_the_bench_configuration(dtype, bar):
    """BASELINE.json configs[1]'s shape — 256 x 3 x 224 x 224 — through the whole ResNet-50 in eval mode, so every kernel runs
    with the grid, tile walk and tensor maps of the benchmarked configuration.  Eval-mode images are independent, so the oracle
    port (oracle.ref_models, fp64 on the host) is evaluated on twelve of the 256 images: the first, middle and last four.
    Measured: fp32 2.0e-6, bf16 4.7e-3."""
    from oracle import ref_models
    torch.manual_seed(0)
    m = dv.models.ResNet50()
    sd = ref_models.clone_state(m.state_dict(), dtype=torch.float64)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(256, 3, 224, 224, generator=g)
    sel = [0, 1, 2, 3, 126, 127, 128, 129, 252, 253, 254, 255]
    xs = x[sel]
    if dtype == torch.bfloat16:
        xs = xs.to(torch.bfloat16).float()      # the product rounds the image once at the entrance; not a model error
    with torch.no_grad():
        ref = ref_models.model_forward("resnet50", xs.double(), sd, bn_train=False).numpy()
    dv.set_compute_dtype(dtype)
    m = m.cuda().eval()
    with torch.no_grad():
        y = m(x.cuda())
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    assert y.shape == (256, 1000) and np.isfinite(y).all()
    err = rel_l2(y[sel], ref)
    print(f"resnet50 eval B=256 224x224 {dtype}: {err:.3e}")
    assert err < bar, err
    # identical top-1 (north_star) wherever the reference's own decision is not a near-tie of its two best logits
    top = np.sort(ref, axis=1)
    decided = (top[:, -1] - top[:, -2]) > 4 * np.abs(y[sel] - ref).max()
    print(f"top-1 decided on {int(decided.sum())} of {len(sel)} images")
    assert np.array_equal(y[sel].argmax(1)[decided], ref.argmax(1)[decided])


def test_resnet50_train_step_fp32_matches_oracle_gradients():
    """One forward+backward of the whole ResNet-50 in fp32 against the torch-functional oracle in fp64 on CPU
    (B=2, 64x64).  Whole-model gradients are chaotic at random init (torch NHWC-vs-NCHW fp32 kernels already differ by
    2.2e-2), so the gate is 5e-2 on the flattened gradient; per-block gradients are gated at 1e-4 in
    test_gpu_c_blocks.py."""
    from oracle import ref_models
    torch.manual_seed(0)
    m = dv.models.ResNet50()
    sd = ref_models.clone_state(m.state_dict(), dtype=torch.float64, requires_grad=True)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 3, 64, 64, generator=g)
    labels = torch.tensor([3, 977])
    loss_ref = torch.nn.functional.cross_entropy(ref_models.model_forward("resnet50", x.double(), sd, True), labels)
    loss_ref.backward()
    dv.set_compute_dtype(torch.float32)
    m = m.cuda().train()
    loss = torch.nn.functional.cross_entropy(m(x.cuda()), labels.cuda())
    loss.backward()
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_ref)) < 1e-4 * abs(float(loss_ref))
    gr = torch.cat([sd[k].grad.flatten() for k, _ in m.named_parameters()]).numpy()
    gg = torch.cat([p.grad.flatten().double().cpu() for _, p in m.named_parameters()]).numpy()
    assert rel_l2(gg, gr) < 5e-2


@pytest.mark.parametrize("name", ["model_resnet50_b2_64_eval", "model_resnet34_b4_224_eval", "model_mobilenet1_b2_64_eval"])
def test_folded_batchnorm_inference_matches_unfused_and_golden(name):
    """net.eval() + no_grad in bf16 runs every conv->BN(->add)(->ReLU) group as one kernel (inference.py).  Same bar as
    the unfused path against the frozen reference logits, one launch fewer per BatchNorm2d, and the two paths agree."""
    from deep_vision_b200.lib import lib
    rec = load_golden(name)
    torch.manual_seed(int(rec["seed"]))
    m = CTORS[FIXTURES[name]]()
    psum = float(sum(p.detach().double().sum() for p in m.parameters()))
    if abs(psum - float(rec["param_sum"])) > 1e-9 * float(rec["param_abs_sum"]):
        pytest.skip("torch RNG differs from the build that froze the fixture")
    g = torch.Generator().manual_seed(int(rec["seed"]) + 1)
    x = torch.randn(tuple(int(v) for v in rec["in_shape"]), generator=g).cuda()
    dv.set_compute_dtype(torch.bfloat16)
    m = m.cuda().eval()

    def run():
        n0 = lib.dv_launch_count()
        with torch.no_grad():
            y = m(x)
        torch.cuda.synchronize()
        return y.cpu().numpy(), lib.dv_launch_count() - n0

    run()                                   # first call also launches one dv_bn_fold per BatchNorm2d
    y_fold, n_fold = run()
    dv.set_fold_bn(False)
    try:
        run()                               # (first unfused call packs the stem's filters once)
        y_plain, n_plain = run()
    finally:
        dv.set_fold_bn(True)
    n_bn = sum(isinstance(mod, torch.nn.BatchNorm2d) for mod in m.modules())
    assert n_plain - n_fold == n_bn, (n_plain, n_fold, n_bn)
    err = rel_l2(y_fold, rec["logits_fp64"])
    print(f"{name}: folded {err:.3e}  unfused {rel_l2(y_plain, rec['logits_fp64']):.3e}  launches {n_fold} vs {n_plain}")
    assert err < 2e-2
    assert np.array_equal(y_fold.argmax(1), rec["logits_fp64"].argmax(1))
    assert rel_l2(y_fold, y_plain) < 2e-2
    # a changed running statistic invalidates the cached scale/shift
    bn = next(mod for mod in m.modules() if isinstance(mod, torch.nn.BatchNorm2d))
    bn.running_var.mul_(4.0)
    y2, _ = run()
    dv.set_fold_bn(False)
    try:
        y2_plain, _ = run()
    finally:
        dv.set_fold_bn(True)
    assert rel_l2(y2, y2_plain) < 2e-2
    if "mobilenet" not in name:   # (a fresh MobileNetV1 in eval mode has logits ~ the classifier bias: nothing to see there)
        assert rel_l2(y2, y_fold) > 1e-3


def test_captured_forward_replays_the_eager_inference_pass():
    """inference.CapturedForward: the eval-mode forward recorded into a CUDA graph gives the eager pass's logits bit for bit,
    for every new batch copied into its static input, and refuses another shape."""
    torch.manual_seed(0)
    dv.set_compute_dtype(torch.bfloat16)
    net = dv.models.ResNet34().cuda().eval()
    g = torch.Generator().manual_seed(3)
    xs = [torch.randn(8, 3, 64, 64, generator=g).cuda() for _ in range(3)]
    fwd = dv.inference.CapturedForward(net, xs[0])
    for x in xs:
        with torch.no_grad():
            want = net(x).clone()
        got = fwd(x)
        torch.cuda.synchronize()
        assert torch.equal(got, want)
    with pytest.raises(RuntimeError):
        fwd(xs[0][:4])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
def test_resnet_stem_fused_bn_relu_maxpool_equals_separate_modules(dtype):
    """set_fuse_stem_pool(True) (default): bn1 -> relu -> maxpool of the stem (resnet34.py:48-51) run as one forward pass
    and a reduce + finalize + apply backward instead of two forward and four backward launches.  With a fixed-order forward the
    logits are bit-identical; every gradient agrees to the summation order of dgamma/dbeta of bn1."""
    from deep_vision_b200.lib import lib
    dv.set_compute_dtype(dtype)
    dv.set_fuse_bn_stats(False)      # fixed-order statistics: both runs see the same activations
    dv.set_fuse_bn_bwd(False)        # ... and fixed-order dgamma / dbeta, so the layers behind the stem see the same gradients
    torch.manual_seed(3)
    ref = dv.models.ResNet34()
    state = {k: v.clone() for k, v in ref.state_dict().items()}
    x = torch.randn(4, 3, 96, 80, generator=torch.Generator().manual_seed(4)).cuda()
    labels = torch.tensor([1, 500, 999, 42]).cuda()
    out, launches = {}, {}
    try:
        for fuse in (False, True):
            dv.set_fuse_stem_pool(fuse)
            m = dv.models.ResNet34()
            m.load_state_dict(state)
            m = m.cuda().train()
            n0 = lib.dv_launch_count()
            logits = m(x)
            loss = torch.nn.functional.cross_entropy(logits.float(), labels)
            loss.backward()
            torch.cuda.synchronize()
            launches[fuse] = lib.dv_launch_count() - n0
            out[fuse] = {"logits": logits.detach().double().cpu().numpy(), "rm": m.bn1.running_mean.double().cpu().numpy(),
                         "rv": m.bn1.running_var.double().cpu().numpy(), "nbt": int(m.bn1.num_batches_tracked)}
            out[fuse].update({"grad/" + k: p.grad.double().cpu().numpy() for k, p in m.named_parameters()})
    finally:
        dv.set_fuse_stem_pool(True)
        dv.set_fuse_bn_stats(True)
        dv.set_fuse_bn_bwd(True)
    assert launches[False] - launches[True] == 2, launches     # apply + pool -> 1 forward; pool + reduce + finalize + apply -> 3 backward
    assert np.array_equal(out[True]["logits"], out[False]["logits"])
    assert np.array_equal(out[True]["rm"], out[False]["rm"]) and np.array_equal(out[True]["rv"], out[False]["rv"])
    assert out[True]["nbt"] == out[False]["nbt"] == 1
    for k in out[True]:
        if k.startswith("grad/"):
            bar = 1e-5 if (dtype == torch.float32 or not k.startswith("grad/conv1")) else 2e-3
            assert rel_l2(out[True][k], out[False][k]) < bar, k
