"""CPU tests of the drop-in boundary: the shared library loads without a GPU, exports exactly what include/dv.h
declares, and rejects bad arguments with an error code + message instead of crashing."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "dv.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dv_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from deep_vision_b200 import lib
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib.lib, n), f"{n} declared in include/dv.h but not exported"
    # and the Python binding knows every one of them (argtypes set)
    missing = [n for n in names if n not in lib.EXPORTED_SYMBOLS]
    assert not missing, missing


def test_version_and_error_reporting_without_gpu():
    from deep_vision_b200 import lib
    assert lib.lib.dv_version() == 100
    # null params: rejected before any CUDA call
    assert lib.lib.dv_conv_fprop(None, None) == 1
    assert b"null" in lib.lib.dv_last_error()
    p = lib.ConvParams()
    p.x = p.w = p.y = 16  # non-null dummies; dims are zero
    assert lib.lib.dv_conv_fprop(C.byref(p), None) == 1
    assert b"dimension" in lib.lib.dv_last_error()
    assert lib.lib.dv_bn_stats(None, None) == 1
    with pytest.raises(lib.DvError):
        lib.check(lib.lib.dv_bn_apply(None, None), "dv_bn_apply")


def test_algo_selection_is_host_logic():
    from deep_vision_b200 import lib
    p = lib.ConvParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = 256, 56, 56, 64, 64, 3, 3, 1, 1
    p.dtype = lib.DV_BF16
    for ps in (lib.DV_CONV_FPROP, lib.DV_CONV_DGRAD, lib.DV_CONV_WGRAD):
        assert lib.lib.dv_conv_algo(C.byref(p), ps) == lib.DV_ALGO_TC
    p.dtype = lib.DV_F32                       # fp32 parity path is CUDA-core
    assert lib.lib.dv_conv_algo(C.byref(p), lib.DV_CONV_FPROP) == lib.DV_ALGO_SIMT
    p.dtype, p.C, p.R, p.S, p.stride, p.pad, p.H, p.W = lib.DV_BF16, 3, 7, 7, 2, 3, 224, 224   # the stem
    assert lib.lib.dv_conv_algo(C.byref(p), lib.DV_CONV_FPROP) == lib.DV_ALGO_SIMT
    p.C, p.K, p.R, p.S, p.stride, p.pad, p.H, p.W = 64, 128, 3, 3, 2, 1, 56, 56                # ResNet-34 3x3 stride 2
    assert lib.lib.dv_conv_algo(C.byref(p), lib.DV_CONV_FPROP) == lib.DV_ALGO_TC
    assert lib.lib.dv_conv_algo(C.byref(p), lib.DV_CONV_DGRAD) == lib.DV_ALGO_TC      # four parity-class sub-convolutions
    assert lib.lib.dv_conv_algo(C.byref(p), lib.DV_CONV_WGRAD) == lib.DV_ALGO_TC
    p.R, p.S, p.pad = 5, 5, 2                                                          # any other strided RxS dgrad: CUDA cores
    assert lib.lib.dv_conv_algo(C.byref(p), lib.DV_CONV_DGRAD) == lib.DV_ALGO_SIMT
    p.R, p.S, p.pad = 3, 3, 1
    # wgrad scratch is sized by host logic and is a whole number of [K,R,S,C] fp32 slabs
    need = lib.lib.dv_conv_workspace_bytes(C.byref(p), lib.DV_CONV_WGRAD)
    assert need > 0 and need % (128 * 9 * 64 * 4) == 0
    assert lib.lib.dv_bn_workspace_bytes(256 * 56 * 56, 64) > 0


def test_no_cpu_path():
    import torch
    import deep_vision_b200 as dv
    m = dv.models.MobileNetV1()
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 3, 32, 32))


def test_stem_geometry_is_host_logic():
    """dv_stem_supported: which C <= 4 convolutions the direct (no patch matrix) path takes, and the padded image extents."""
    from deep_vision_b200 import lib
    p = lib.StemParams()
    p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.stride, p.pad = 256, 224, 224, 3, 64, 7, 7, 2, 3      # resnet50.py:20-28
    hp, wp = C.c_int(0), C.c_int(0)
    assert lib.lib.dv_stem_supported(C.byref(p), C.byref(hp), C.byref(wp)) == 1
    assert (hp.value, wp.value) == (229, 230)          # 3 + 224 rows (+2 for the last window), row pitch 230 * 8 B = a multiple of 16
    p.K, p.R, p.S, p.pad = 32, 3, 3, 1                 # mobilenet_v1.py:30
    assert lib.lib.dv_stem_supported(C.byref(p), C.byref(hp), C.byref(wp)) == 1 and wp.value % 2 == 0
    p.stride = 1
    assert lib.lib.dv_stem_supported(C.byref(p), None, None) == 0
    assert lib.lib.dv_stem_fprop(None, None) == 1 and b"null" in lib.lib.dv_last_error()
