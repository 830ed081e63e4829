"""Developer probe: how the tensor core rounds its fp32 accumulation.  Operands are exact bf16 values (only the first of
the three planes is non-zero and every product is exact in fp32), K = 128 -> what differs from the exact sum is the
accumulator's rounding alone.  Prints the mean signed error for positive sums, for the negated problem, and for sums
that cancel, in units of 2^-24 of the mean |sum| / of the mean sum of |terms|."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(ROOT, "relayrl-prototype_b200"))
from relayrl_b200 import _lib  # noqa: E402

L = _lib.dev_lib()
fn = L.rrl_debug_tc_gemm
fn.restype, fn.argtypes = C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
dev = torch.device("cuda:0")


def bf16(x):
    return torch.from_numpy(x.astype(np.float32)).bfloat16().float().numpy()


def run(A, B):
    dA, dB = torch.from_numpy(A).to(dev), torch.from_numpy(B).to(dev)
    dD = torch.zeros((128, 64), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    assert fn(0, dA.data_ptr(), dB.data_ptr(), dD.data_ptr(), None) == 0
    torch.cuda.synchronize()
    return dD.cpu().numpy().astype(np.float64)


rng = np.random.default_rng(0)
for name, mk in (("positive terms", lambda s: np.abs(rng.standard_normal(s)) + 0.5),
                 ("cancelling terms", lambda s: rng.standard_normal(s))):
    A, B = bf16(mk((128, 128))), bf16(mk((64, 128)))
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    mag = np.abs(A).astype(np.float64) @ np.abs(B).astype(np.float64).T
    for sign in (1.0, -1.0):
        got = run((sign * A).astype(np.float32), B)
        err = got - sign * ref
        u = 2.0 ** -24
        print("%-17s x %+d: mean err %+.3f, mean |err| %.3f (units of 2^-24 mean|sum| = %.3g); mean err / (2^-24 mean sum|terms|) %+.3f; "
              "fraction exact %.3f" % (name, int(sign), err.mean() / (u * np.abs(ref).mean()), np.abs(err).mean() / (u * np.abs(ref).mean()),
                                       np.abs(ref).mean(), err.mean() / (u * mag.mean()), (err == 0).mean()))
    f32 = (A.astype(np.float32) @ B.astype(np.float32).T).astype(np.float64)
    print("   numpy float32 matmul: mean err %+.3f, mean |err| %.3f (same units)" % ((f32 - ref).mean() / (u * np.abs(ref).mean()),
                                                                                    np.abs(f32 - ref).mean() / (u * np.abs(ref).mean())))
