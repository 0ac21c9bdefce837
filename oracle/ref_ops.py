"""numpy restatement of the primitives on the hot path, in the arithmetic torch defines for them.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference calls these through torch.nn; torch is an
un-vendored, unpinned dependency (ResNet/pytorch/requirements.in:2), so what is restated here is torch's documented
semantics at the reference's call sites.  float64 throughout so the functions can arbitrate between two fp32 results.
Layouts are the reference's: activations NCHW, filters OIHW.
"""
import numpy as np


def conv2d(x, w, stride=1, pad=0, groups=1):
    """nn.Conv2d(bias=False): cross-correlation with zero padding, out = floor((H+2p-R)/s)+1.
    Call sites: resnet50.py:20-28,101-143; resnet34.py:20-28,96-126; mobilenet_v1.py:30,112-122,139-145."""
    x = np.asarray(x, np.float64)
    w = np.asarray(w, np.float64)
    N, C, H, W = x.shape
    K, Cg, R, S = w.shape
    assert C == Cg * groups and K % groups == 0
    P = (H + 2 * pad - R) // stride + 1
    Q = (W + 2 * pad - S) // stride + 1
    xp = np.zeros((N, C, H + 2 * pad, W + 2 * pad), np.float64)
    xp[:, :, pad:pad + H, pad:pad + W] = x
    y = np.zeros((N, K, P, Q), np.float64)
    kg = K // groups
    for g in range(groups):
        xs = xp[:, g * Cg:(g + 1) * Cg]
        ws = w[g * kg:(g + 1) * kg]
        for r in range(R):
            for s in range(S):
                patch = xs[:, :, r:r + stride * (P - 1) + 1:stride, s:s + stride * (Q - 1) + 1:stride]
                y[:, g * kg:(g + 1) * kg] += np.einsum("ncpq,kc->nkpq", patch, ws[:, :, r, s])
    return y


def conv2d_backward(x, w, dy, stride=1, pad=0, groups=1):
    """autograd of conv2d (train.py:460): returns (dx, dw)."""
    x = np.asarray(x, np.float64)
    w = np.asarray(w, np.float64)
    dy = np.asarray(dy, np.float64)
    N, C, H, W = x.shape
    K, Cg, R, S = w.shape
    _, _, P, Q = dy.shape
    xp = np.zeros((N, C, H + 2 * pad, W + 2 * pad), np.float64)
    xp[:, :, pad:pad + H, pad:pad + W] = x
    dxp = np.zeros_like(xp)
    dw = np.zeros_like(w)
    kg = K // groups
    for g in range(groups):
        cs = slice(g * Cg, (g + 1) * Cg)
        ks = slice(g * kg, (g + 1) * kg)
        for r in range(R):
            for s in range(S):
                hs = slice(r, r + stride * (P - 1) + 1, stride)
                wsl = slice(s, s + stride * (Q - 1) + 1, stride)
                dw[ks, :, r, s] = np.einsum("nkpq,ncpq->kc", dy[:, ks], xp[:, cs, hs, wsl])
                dxp[:, cs, hs, wsl] += np.einsum("nkpq,kc->ncpq", dy[:, ks], w[ks, :, r, s])
    return dxp[:, :, pad:pad + H, pad:pad + W], dw


def batch_norm_train(x, gamma, beta, running_mean, running_var, momentum=0.1, eps=1e-5):
    """nn.BatchNorm2d in training mode (resnet50.py:30,108,...): normalise with the batch mean and BIASED variance over
    (N,H,W); running stats move by `momentum` towards the batch mean and the UNBIASED variance.
    Returns (y, new_running_mean, new_running_var, mean, invstd)."""
    x = np.asarray(x, np.float64)
    n = x.shape[0] * x.shape[2] * x.shape[3]
    mean = x.mean(axis=(0, 2, 3))
    var = x.var(axis=(0, 2, 3))
    invstd = 1.0 / np.sqrt(var + eps)
    y = (x - mean[None, :, None, None]) * invstd[None, :, None, None] * np.asarray(gamma, np.float64)[None, :, None, None] \
        + np.asarray(beta, np.float64)[None, :, None, None]
    unbiased = var * n / max(n - 1, 1)
    rm = (1 - momentum) * np.asarray(running_mean, np.float64) + momentum * mean
    rv = (1 - momentum) * np.asarray(running_var, np.float64) + momentum * unbiased
    return y, rm, rv, mean, invstd


def batch_norm_eval(x, gamma, beta, running_mean, running_var, eps=1e-5):
    x = np.asarray(x, np.float64)
    sc = np.asarray(gamma, np.float64) / np.sqrt(np.asarray(running_var, np.float64) + eps)
    return (x - np.asarray(running_mean, np.float64)[None, :, None, None]) * sc[None, :, None, None] \
        + np.asarray(beta, np.float64)[None, :, None, None]


def batch_norm_backward(x, gamma, mean, invstd, dy):
    """Training-mode BN backward: dgamma = sum(dy*xhat), dbeta = sum(dy),
    dx = gamma*invstd*(dy - dbeta/n - xhat*dgamma/n)."""
    x = np.asarray(x, np.float64)
    dy = np.asarray(dy, np.float64)
    n = x.shape[0] * x.shape[2] * x.shape[3]
    xhat = (x - mean[None, :, None, None]) * invstd[None, :, None, None]
    dbeta = dy.sum(axis=(0, 2, 3))
    dgamma = (dy * xhat).sum(axis=(0, 2, 3))
    g = np.asarray(gamma, np.float64) * invstd
    dx = g[None, :, None, None] * (dy - dbeta[None, :, None, None] / n - xhat * dgamma[None, :, None, None] / n)
    return dx, dgamma, dbeta


def relu(x):
    """nn.ReLU(inplace=True) — resnet50.py:31,131."""
    return np.maximum(np.asarray(x, np.float64), 0.0)


def max_pool_3x3_s2_p1(x):
    """nn.MaxPool2d(3, 2, 1) — resnet50.py:34: -inf padding, floor mode.  Returns (y, argmax tap 0..8, first max wins)."""
    x = np.asarray(x, np.float64)
    N, C, H, W = x.shape
    P, Q = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    xp = np.full((N, C, H + 2, W + 2), -np.inf)
    xp[:, :, 1:H + 1, 1:W + 1] = x
    y = np.full((N, C, P, Q), -np.inf)
    idx = np.zeros((N, C, P, Q), np.int64)
    for r in range(3):
        for s in range(3):
            patch = xp[:, :, r:r + 2 * (P - 1) + 1:2, s:s + 2 * (Q - 1) + 1:2]
            better = patch > y
            y = np.where(better, patch, y)
            idx = np.where(better, r * 3 + s, idx)
    return y, idx


def max_pool_backward(dy, idx, H, W):
    dy = np.asarray(dy, np.float64)
    N, C, P, Q = dy.shape
    dxp = np.zeros((N, C, H + 2, W + 2))
    for r in range(3):
        for s in range(3):
            dxp[:, :, r:r + 2 * (P - 1) + 1:2, s:s + 2 * (Q - 1) + 1:2] += np.where(idx == r * 3 + s, dy, 0.0)
    return dxp[:, :, 1:H + 1, 1:W + 1]


def global_avg_pool(x):
    """nn.AdaptiveAvgPool2d((1,1)) — resnet50.py:43."""
    return np.asarray(x, np.float64).mean(axis=(2, 3), keepdims=True)


def linear(x, w, b):
    """nn.Linear — resnet50.py:44."""
    return np.asarray(x, np.float64) @ np.asarray(w, np.float64).T + np.asarray(b, np.float64)


def cross_entropy(logits, labels):
    """nn.CrossEntropyLoss() — train.py:358,452: mean over the batch of -log_softmax[label].  Returns (loss, dlogits)."""
    z = np.asarray(logits, np.float64)
    z = z - z.max(axis=1, keepdims=True)
    logp = z - np.log(np.exp(z).sum(axis=1, keepdims=True))
    n = z.shape[0]
    loss = -logp[np.arange(n), labels].mean()
    d = np.exp(logp)
    d[np.arange(n), labels] -= 1.0
    return loss, d / n


def sgd_step(p, g, buf, lr, momentum, weight_decay, first):
    """optim.SGD(lr, momentum, weight_decay) — train.py:166-184,467 (no dampening, no Nesterov)."""
    g = g + weight_decay * p
    buf = g.copy() if first else momentum * buf + g
    return p - lr * buf, buf


def rmsprop_step(p, g, sq, lr, alpha, eps, weight_decay=0.0):
    """optim.RMSprop(lr, alpha, eps) — train.py:199-205 (no momentum, not centred)."""
    g = g + weight_decay * p
    sq = alpha * sq + (1 - alpha) * g * g
    return p - lr * g / (np.sqrt(sq) + eps), sq
