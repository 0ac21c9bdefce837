"""CPU tests (no GPU): the oracle against torch's CPU kernels and against the golden fixtures frozen from the
UNMODIFIED reference model files by oracle/make_golden.py.  This is what pins the oracle."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import golden_state, load_golden, rel_l2
from oracle import ref_models, ref_ops

BLOCKS = {
    "block_bottleneck_256_64_256": ("bottleneck", dict(stride=1, downsample=False)),
    "block_bottleneck_128_64_256_s2_proj": ("bottleneck", dict(stride=2, downsample=True)),
    "block_bottleneck_64_64_256_s1_proj": ("bottleneck", dict(stride=1, downsample=True)),
    "block_basic_64_64": ("basic", dict(stride=1, downsample=False)),
    "block_basic_64_128_s2_proj": ("basic", dict(stride=2, downsample=True)),
    "block_dwsep_64_128_s2": ("dwsep", dict(dw_stride=2)),
    "block_dwsep_32_64_s1": ("dwsep", dict(dw_stride=1)),
}


@pytest.mark.parametrize("shape", [
    # N, C, H, W, K, R, stride, pad, groups
    (2, 3, 12, 12, 8, 7, 2, 3, 1),     # stem-like
    (2, 8, 9, 9, 16, 3, 1, 1, 1),
    (2, 8, 9, 9, 16, 3, 2, 1, 1),
    (3, 16, 7, 7, 8, 1, 2, 0, 1),      # strided 1x1 (Bottleneck conv1 / projection)
    (2, 8, 8, 8, 8, 3, 2, 1, 8),       # depth-wise
    (1, 4, 5, 6, 4, 3, 1, 1, 4),       # ragged H != W
])
def test_numpy_conv_matches_torch(shape):
    N, C, H, W, K, R, stride, pad, groups = shape
    g = torch.Generator().manual_seed(0)
    x = torch.randn(N, C, H, W, generator=g, dtype=torch.float64, requires_grad=True)
    w = torch.randn(K, C // groups, R, R, generator=g, dtype=torch.float64, requires_grad=True)
    y = F.conv2d(x, w, None, stride, pad, 1, groups)
    dy = torch.randn(y.shape, generator=g, dtype=torch.float64)
    y.backward(dy)
    yo = ref_ops.conv2d(x.detach().numpy(), w.detach().numpy(), stride, pad, groups)
    dxo, dwo = ref_ops.conv2d_backward(x.detach().numpy(), w.detach().numpy(), dy.numpy(), stride, pad, groups)
    assert rel_l2(yo, y.detach().numpy()) < 1e-12
    assert rel_l2(dxo, x.grad.numpy()) < 1e-12
    assert rel_l2(dwo, w.grad.numpy()) < 1e-12


def test_numpy_bn_matches_torch():
    g = torch.Generator().manual_seed(1)
    x = torch.randn(4, 6, 5, 5, generator=g, dtype=torch.float64, requires_grad=True) * 3 + 2
    x.retain_grad()
    gamma = torch.rand(6, generator=g, dtype=torch.float64, requires_grad=True)
    beta = torch.randn(6, generator=g, dtype=torch.float64, requires_grad=True)
    rm, rv = torch.zeros(6, dtype=torch.float64), torch.ones(6, dtype=torch.float64)
    y = F.batch_norm(x, rm, rv, gamma, beta, True, 0.1, 1e-5)
    dy = torch.randn(y.shape, generator=g, dtype=torch.float64)
    y.backward(dy)
    yo, rmo, rvo, mean, invstd = ref_ops.batch_norm_train(x.detach().numpy(), gamma.detach().numpy(), beta.detach().numpy(),
                                                         np.zeros(6), np.ones(6))
    dxo, dgo, dbo = ref_ops.batch_norm_backward(x.detach().numpy(), gamma.detach().numpy(), mean, invstd, dy.numpy())
    assert rel_l2(yo, y.detach().numpy()) < 1e-12
    assert rel_l2(rmo, rm.numpy()) < 1e-12 and rel_l2(rvo, rv.numpy()) < 1e-12
    assert rel_l2(dxo, x.grad.numpy()) < 1e-10
    assert rel_l2(dgo, gamma.grad.numpy()) < 1e-10 and rel_l2(dbo, beta.grad.numpy()) < 1e-10
    ye = ref_ops.batch_norm_eval(x.detach().numpy(), gamma.detach().numpy(), beta.detach().numpy(), rm.numpy(), rv.numpy())
    yt = F.batch_norm(x.detach(), rm, rv, gamma.detach(), beta.detach(), False, 0.1, 1e-5)
    assert rel_l2(ye, yt.numpy()) < 1e-12


def test_numpy_pool_linear_ce_optim_match_torch():
    g = torch.Generator().manual_seed(2)
    x = torch.relu(torch.randn(2, 4, 9, 11, generator=g, dtype=torch.float64)).requires_grad_(True)  # many ties at 0
    y, idx = F.max_pool2d(x, 3, 2, 1, return_indices=True)
    dy = torch.randn(y.shape, generator=g, dtype=torch.float64)
    y.backward(dy)
    yo, tap = ref_ops.max_pool_3x3_s2_p1(x.detach().numpy())
    assert np.array_equal(yo, y.detach().numpy())
    assert rel_l2(ref_ops.max_pool_backward(dy.numpy(), tap, 9, 11), x.grad.numpy()) < 1e-12
    assert rel_l2(ref_ops.global_avg_pool(x.detach().numpy()), F.adaptive_avg_pool2d(x.detach(), (1, 1)).numpy()) < 1e-12
    w = torch.randn(5, 7, generator=g, dtype=torch.float64)
    b = torch.randn(5, generator=g, dtype=torch.float64)
    a = torch.randn(3, 7, generator=g, dtype=torch.float64)
    assert rel_l2(ref_ops.linear(a.numpy(), w.numpy(), b.numpy()), F.linear(a, w, b).numpy()) < 1e-12
    logits = torch.randn(6, 10, generator=g, dtype=torch.float64, requires_grad=True)
    labels = torch.randint(0, 10, (6,), generator=g)
    loss = F.cross_entropy(logits, labels)
    loss.backward()
    lo, dl = ref_ops.cross_entropy(logits.detach().numpy(), labels.numpy())
    assert abs(lo - float(loss)) < 1e-12 and rel_l2(dl, logits.grad.numpy()) < 1e-12
    # optimisers, two steps each (train.py:166-205)
    p = torch.randn(50, generator=g, dtype=torch.float64).requires_grad_(True)
    opt = torch.optim.SGD([p], lr=0.1, momentum=0.9, weight_decay=1e-4)
    pn, buf = p.detach().numpy().copy(), None
    for step in range(2):
        gr = torch.randn(50, generator=g, dtype=torch.float64)
        p.grad = gr.clone()
        opt.step()
        pn, buf = ref_ops.sgd_step(pn, gr.numpy(), buf, 0.1, 0.9, 1e-4, first=(step == 0))
        assert rel_l2(pn, p.detach().numpy()) < 1e-12
    p = torch.randn(50, generator=g, dtype=torch.float64).requires_grad_(True)
    opt = torch.optim.RMSprop([p], lr=0.045, alpha=0.9, eps=1.0)
    pn, sq = p.detach().numpy().copy(), np.zeros(50)
    for step in range(2):
        gr = torch.randn(50, generator=g, dtype=torch.float64)
        p.grad = gr.clone()
        opt.step()
        pn, sq = ref_ops.rmsprop_step(pn, gr.numpy(), sq, 0.045, 0.9, 1.0)
        assert rel_l2(pn, p.detach().numpy()) < 1e-12


def _run_block(kind, kw, sd, x):
    if kind == "bottleneck":
        return ref_models.bottleneck_block(x, sd, "", bn_train=True, **kw)
    if kind == "basic":
        return ref_models.basic_block(x, sd, "", bn_train=True, **kw)
    return ref_models.dwsep_block(x, sd, "", bn_train=True, **kw)


@pytest.mark.parametrize("name", sorted(BLOCKS))
def test_functional_oracle_matches_reference_blocks(name):
    """ref_models (fp64, CPU) reproduces the frozen outputs/gradients of the reference's own block classes."""
    kind, kw = BLOCKS[name]
    rec = load_golden(name)
    sd = ref_models.clone_state(golden_state(rec), dtype=torch.float64, requires_grad=True)
    x = torch.from_numpy(rec["x"]).double().requires_grad_(True)
    y = _run_block(kind, kw, sd, x)
    y.backward(torch.from_numpy(rec["dy"]).double())
    assert rel_l2(y.detach().numpy(), rec["y"]) < 1e-6      # fixtures are stored in fp32
    assert rel_l2(x.grad.numpy(), rec["dx"]) < 1e-6
    for k, v in rec.items():
        if k.startswith("grad/"):
            assert rel_l2(sd[k[5:]].grad.numpy(), v) < 1e-5, k
        if k.startswith("after/"):
            assert rel_l2(sd[k[6:]].detach().numpy(), v) < 1e-6, k


def test_numpy_oracle_matches_reference_block():
    """ref_ops (numpy) chained by hand over one Bottleneck reproduces the reference fixture (forward)."""
    rec = load_golden("block_bottleneck_128_64_256_s2_proj")
    s = {k[6:]: v.astype(np.float64) for k, v in rec.items() if k.startswith("state/")}
    x = rec["x"].astype(np.float64)

    def bn(t, p):
        return ref_ops.batch_norm_train(t, s[p + "weight"], s[p + "bias"], s[p + "running_mean"], s[p + "running_var"])[0]

    ident = bn(ref_ops.conv2d(x, s["projection.0.weight"], 2, 0), "projection.1.")
    t = ref_ops.relu(bn(ref_ops.conv2d(x, s["conv1.weight"], 2, 0), "bn1."))
    t = ref_ops.relu(bn(ref_ops.conv2d(t, s["conv2.weight"], 1, 1), "bn2."))
    t = bn(ref_ops.conv2d(t, s["conv3.weight"], 1, 0), "bn3.")
    y = ref_ops.relu(t + ident)
    assert rel_l2(y, rec["y"]) < 1e-6


MODELS = {
    "model_resnet34_b4_128_train": "resnet34",
    "model_resnet50_b2_64_eval": "resnet50",
    "model_mobilenet1_b4_128_train": "mobilenet1",
    "model_mobilenet1_b2_64_eval": "mobilenet1",
}


@pytest.mark.parametrize("name", sorted(MODELS))
def test_functional_oracle_matches_reference_models(name):
    """Whole-model forward of ref_models == frozen logits of the reference class (weights rebuilt from the seed through
    the product's constructors, whose RNG consumption mirrors the reference's; checked by the parameter checksum)."""
    import deep_vision_b200 as dv
    rec = load_golden(name)
    ctor = {"resnet34": dv.models.ResNet34, "resnet50": dv.models.ResNet50, "mobilenet1": dv.models.MobileNetV1}[MODELS[name]]
    torch.manual_seed(int(rec["seed"]))
    m = ctor()
    psum = float(sum(p.detach().double().sum() for p in m.parameters()))
    if str(rec["torch_version"]) != torch.__version__ and abs(psum - float(rec["param_sum"])) > 1e-6 * float(rec["param_abs_sum"]):
        pytest.skip("torch RNG differs from the build that froze the fixture")
    assert abs(psum - float(rec["param_sum"])) <= 1e-9 * float(rec["param_abs_sum"]), "constructor RNG order differs from the reference"
    g = torch.Generator().manual_seed(int(rec["seed"]) + 1)
    x = torch.randn(tuple(int(v) for v in rec["in_shape"]), generator=g)
    sd = ref_models.clone_state(m.state_dict(), dtype=torch.float64)
    with torch.no_grad():
        y = ref_models.model_forward(MODELS[name], x.double(), sd, bn_train=bool(rec["bn_train"]))
    assert rel_l2(y.numpy(), rec["logits_fp64"]) < 1e-9
