import ctypes, sys, torch
sys.path.insert(0, "/root/repo/relayrl-prototype_b200")
from relayrl_b200 import _lib
lib = _lib.dev_lib()
f = lib.rrl_debug_tc_rate
f.argtypes = [ctypes.c_int32] * 5 + [ctypes.c_void_p, ctypes.c_void_p]
f.restype = ctypes.c_int32
out = torch.zeros(2, dtype=torch.int64, device="cuda")
for uni in (0, 1):
    for (amn, bmn) in [(0, 0), (1, 0), (1, 1)]:
        for N in (32, 64, 128, 256):
            if amn and bmn and N == 256:
                continue
            for count in (48, 480):
                rc = f(N, amn, bmn, count, uni, out.data_ptr(), None)
                torch.cuda.synchronize()
                o = out.tolist()
                print(f"uni={uni} a_mn={amn} b_mn={bmn} N={N:3d} count={count:3d} rc={rc} issue={o[0]/count:7.1f} total={o[1]/count:7.1f} cyc/mma floor={128*N/256:.0f}", flush=True)
