"""tcgen05 building blocks (csrc/tc_common.cuh): one bf16x3 GEMM per operand orientation used by the
fused learner kernel, against a float64 product.  bf16x3 keeps the six largest partial products,
so the result is fp32-accurate: error <= ~1e-6 of sum |a||b|."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_tc_gemm_bf16x3(mode):
    import torch
    from relayrl_b200 import _lib
    L = _lib.dev_lib()   # the self-test kernels are not part of the product library
    fn = L.rrl_debug_tc_gemm
    fn.restype, fn.argtypes = C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(mode)
    ra, rb = (64 if mode in (2, 4) else 128), (128 if mode == 4 else 64)
    A = (rng.standard_normal((ra, 128)) * np.exp(rng.uniform(-3, 3, (ra, 128)))).astype(np.float32)
    B = (rng.standard_normal((rb, 128)) * np.exp(rng.uniform(-3, 3, (rb, 128)))).astype(np.float32)
    A64, B64 = A.astype(np.float64), B.astype(np.float64)
    def prod(a, b):
        if mode == 3:  # only the first 64 features of A take part
            return a[:, :64] @ b
        if mode == 4:  # A^T from tensor memory; only the first 64 features of B take part
            return a.T @ b[:, :64].T
        return a @ b.T if mode == 0 else a.T @ b.T if mode == 1 else a.T @ b
    ref, mag = prod(A64, B64), prod(np.abs(A64), np.abs(B64))
    dev = torch.device("cuda:0")
    dA, dB = torch.from_numpy(A).to(dev), torch.from_numpy(B).to(dev)
    dD = torch.zeros(ref.shape, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    assert fn(mode, dA.data_ptr(), dB.data_ptr(), dD.data_ptr(), None) == 0
    torch.cuda.synchronize()
    got = dD.cpu().numpy().astype(np.float64)
    err = np.abs(got - ref) / mag
    assert err.max() < 2e-6, "max scaled error %.3e" % err.max()
