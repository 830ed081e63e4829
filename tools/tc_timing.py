import ctypes, os, sys, subprocess
VAR = sys.argv[1] if len(sys.argv) > 1 else "tctiming"   # make tc_variants: tctiming, tctiming_skipg1, tctiming_skipg2
os.environ["RELAYRL_B200_SO"] = "/root/repo/tools/_timing/librelayrl_b200_tc_%s.so" % VAR
sys.path.insert(0, "/root/repo/relayrl-prototype_b200"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
from relayrl_b200 import _lib
lib = _lib.lib()
f = lib.rrl_debug_tc_timing
f.argtypes = [ctypes.c_void_p, ctypes.c_int32]; f.restype = ctypes.c_int32
import bench
sys.argv = ["bench.py", "--transitions", "1048576", "--steps", "1", "--warmup", "1", "--cpu-sample", "2048", "--no-extras", "--no-check"]
out = (ctypes.c_ulonglong * 16)()
f(None, 1)
bench.main()
torch.cuda.synchronize()
f(out, 0)
v = list(out)
print("variant", VAR)
names = ["mma:wait go23", "mma:G3 issue", "mma:G2 issue", "mma:wait go1", "mma:G1 issue", "wrk:wait bar1", "wrk:e1 dz2+wait G2+go23", "wrk:wait bar3+drain", "wrk:stage1(next)", "wrk:(unused)", "wrk:epi2 math", "wrk:spill", "wrk:e1 D1+head reduce", "wrk:e1 barrier", "wrk:e1 H1T->TMEM", "wrk:e1 row loss"]
m = sum(v[:5]); w = sum(v[5:16])
TILES = 2 * 81 * 111   # launches of the bench run (2 epochs x 81 passes) x tiles of CTA 0 at 1 Mi transitions (approx.)
for i, (n, x) in enumerate(zip(names, v)):
    tot = m if i < 5 else w
    print(f"{n:20s} {100.0*x/tot:6.1f} %   {x/TILES/(1 if i < 5 else 16):8.0f} cycles/tile")
print("mma total", m, "worker total/16", w / 16)
