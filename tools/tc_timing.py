import ctypes, os, sys, subprocess
os.environ["RELAYRL_B200_SO"] = "/root/repo/tools/_timing/librelayrl_b200_timing.so"
sys.path.insert(0, "/root/repo/relayrl-prototype_b200"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
from relayrl_b200 import _lib
lib = _lib.dev_lib()
f = lib.rrl_debug_tc_timing
f.argtypes = [ctypes.c_void_p, ctypes.c_int32]; f.restype = ctypes.c_int32
import bench
sys.argv = ["bench.py", "--transitions", "1048576", "--steps", "1", "--warmup", "1", "--cpu-sample", "2048", "--no-extras"]
out = (ctypes.c_ulonglong * 16)()
f(None, 1)
bench.main()
torch.cuda.synchronize()
f(out, 0)
v = list(out)
names = ["mma:wait go23", "mma:G3 issue", "mma:G2 issue", "mma:wait go1", "mma:G1 issue", "wrk:wait bar1", "wrk:epi1+H1T->go23", "wrk:wait bar3+drain", "wrk:stage1(next)", "wrk:(unused)", "wrk:wait bar2+epi2", "wrk:spill"]
m = sum(v[:5]); w = sum(v[5:12])
for i, (n, x) in enumerate(zip(names, v)):
    tot = m if i < 5 else w
    print(f"{n:20s} {100.0*x/tot:6.1f} %   ~{14700.0*x/tot:7.0f} cycles/tile")
print("mma total", m, "worker total/16", w / 16)
