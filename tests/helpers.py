"""Shared test helpers: error metrics, golden loading, direct C-ABI calls on torch tensors."""
import ctypes as C
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel_l2(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = np.linalg.norm(a - b)
    n = np.linalg.norm(b)
    return float(d / n) if n > 0 else float(d)


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def golden_state(rec, prefix="state/"):
    out = {}
    for k, v in rec.items():
        if k.startswith(prefix):
            t = torch.from_numpy(v)
            out[k[len(prefix):]] = t
    return out


def nchw_to_nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


def nhwc_to_nchw(t):
    return t.permute(0, 3, 1, 2).contiguous()


def bf16_round(a):
    """numpy fp32/fp64 array -> values exactly representable in bf16 (as float64)."""
    return torch.from_numpy(np.asarray(a, np.float32)).to(torch.bfloat16).to(torch.float64).numpy()


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
