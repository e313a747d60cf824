"""Run the shifted-descriptor probe (csrc/probe.cu) for several row shifts and both base_offset conventions."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200.ops import _native
lib = _native.require()
vp, i = ctypes.c_void_p, ctypes.c_int
lib.pdt_probe_shifted_desc.argtypes = [vp, vp, vp, i, i, vp]
lib.pdt_probe_shifted_desc.restype = i
dev = torch.device("cuda", 0)
torch.manual_seed(0)
X = torch.randn(256, 64, device=dev).bfloat16()
B = torch.randn(64, 64, device=dev).bfloat16()
for mode in (0, 1):
    for shift in (0, 1, 2, 3, 7, 8, 9, 57, 58, 59, 60, 116, 117, 118):
        out = torch.zeros(128, 64, device=dev)
        rc = lib.pdt_probe_shifted_desc(X.data_ptr(), B.data_ptr(), out.data_ptr(), shift, mode, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ref = X[shift:shift + 128].float() @ B.float().t()
        err = (out - ref).abs().max().item() / ref.abs().max().item()
        print("base_offset_mode", mode, "shift", shift, "rc", rc, "rel err", round(err, 5), flush=True)
