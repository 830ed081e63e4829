"""Developer probe: cycles per tcgen05.mma (M=128, K=16) with the A operand in tensor memory, B in shared memory."""
import ctypes, sys, torch
sys.path.insert(0, "/root/repo/relayrl-prototype_b200")
from relayrl_b200 import _lib
lib = _lib.dev_lib()
f = lib.rrl_debug_tc_rate
f.argtypes = [ctypes.c_int32] * 5 + [ctypes.c_void_p, ctypes.c_void_p]
f.restype = ctypes.c_int32
out = torch.zeros(2, dtype=torch.int64, device="cuda")
for amn in (0, 2):
    for N in (32, 64, 128):
        for count in (48, 480):
            rc = f(N, amn, 0, count, 1, out.data_ptr(), None)
            torch.cuda.synchronize()
            o = out.tolist()
            print(f"A={'tmem' if amn == 2 else 'smem'} N={N:3d} count={count:3d} rc={rc} issue={o[0]/count:7.1f} total={o[1]/count:7.1f} cyc/mma floor={128*N/256:.0f}", flush=True)
