import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deep_vision_b200.lib import DV_BF16, check, lib
x = torch.randn(256, 224, 224, 3, device="cuda").to(torch.bfloat16)
out = torch.empty(256 * 112 * 112, 192, device="cuda", dtype=torch.bfloat16)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    check(lib.dv_im2col(x.data_ptr(), out.data_ptr(), 256, 224, 224, 3, 7, 7, 2, 3, 192, DV_BF16, st), "im2col")
torch.cuda.synchronize()
print("ok")
