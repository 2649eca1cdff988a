import ctypes, os, sys, subprocess
sys.path.insert(0, "/root/repo/implicitglobalgrid.jl_b200")
import torch
torch.cuda.init(); torch.zeros(1, device="cuda")
rt = ctypes.CDLL("libcudart.so.12")
val = ctypes.c_size_t(0)
LIM = 0x05  # cudaLimitMaxL2FetchGranularity
print("get rc", rt.cudaDeviceGetLimit(ctypes.byref(val), LIM), "default granularity", val.value)
g = int(sys.argv[1])
print("set rc", rt.cudaDeviceSetLimit(LIM, ctypes.c_size_t(g)))
print("get rc", rt.cudaDeviceGetLimit(ctypes.byref(val), LIM), "now", val.value)
sys.argv = ["dims_sweep.py"]
exec(open("/root/repo/profiles/dims_sweep.py").read())
