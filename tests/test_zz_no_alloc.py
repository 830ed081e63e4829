"""rrl_scan_device_noalloc -- the scan for hosts that are not allowed to allocate device memory.

Some embedders get one slab of GPU memory at start-up and may never call cudaMalloc again: a
trajectory server with a locked memory pool, a process under a quota with no head-room, a plug-in
living inside another application's CUDA context.  They can hand the library their inputs and
outputs but not a workspace, so this entry point keeps its scratch in a fixed array inside the
library (which caps the size at RRL_SCAN_NOALLOC_MAX transitions).

The test drives the entry point through the C ABI twice in a row -- with different sizes and both
baseline modes, so the second call runs over whatever the first one left in the shared scratch --
and compares every output with the ordinary entry point (rrl_scan_device + caller workspace) AND
with the CPU oracle (oracle/scan_ref.c, the sequential float64 recurrence) at the tolerance of
tests/test_gpu_scan.py.  The second call goes out on a stream of the caller's own.  Around each
call the free device memory reported by the driver must not move: the point of the entry point is
that it allocates nothing.  It lives in a file that sorts last so that a run stopping at the first
failure has already executed every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-6, 1e-6   # as tests/test_gpu_scan.py


def _case(rng, n, dev):
    import torch
    lens, tot = [], 0
    while tot < n:
        k = int(rng.integers(1, 3000)) if rng.random() < 0.97 else int(rng.integers(3000, 300000))
        k = min(k, n - tot)
        lens.append(k)
        tot += k
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    E = len(lens)
    h = dict(rew=rng.uniform(-1, 1, n).astype(np.float32), val=rng.standard_normal(n).astype(np.float32),
             offs=offs, lvs=rng.uniform(-1, 1, E).astype(np.float32))
    d = {k: torch.from_numpy(v).to(dev) for k, v in h.items()}
    return d, E


def test_noalloc_matches_workspace_entry_point_twice():
    import torch
    from relayrl_b200 import _lib, scan_workspace_bytes
    from oracle import scan_ref as sr
    L = _lib.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    rng = np.random.default_rng(2024)
    g, l = float(np.float32(0.98)), float(np.float32(0.97))
    p = lambda t: C.c_void_p(t.data_ptr())
    own_stream = torch.cuda.Stream(device=dev)
    # The first call on a device is the documented set-up call (it resolves the scratch array's device address, and with
    # CUDA's lazy module loading the driver may bring the kernel image in at that moment): make it a tiny one, outside
    # the free-memory bracket.
    z, zo = torch.zeros(16, device=dev), torch.tensor([0, 16], dtype=torch.int64, device=dev)
    assert L.rrl_scan_device_noalloc(p(z), p(z), p(zo), p(z), 16, 1, g, l, 1, p(z.clone()), p(z.clone()), None, None) == 0
    assert L.rrl_scan_device_noalloc(p(z), p(z), p(zo), p(z), 16, 1, g, l, 0, p(z.clone()), p(z.clone()), None, None) == 0
    torch.cuda.synchronize()
    # two consecutive calls: a large with-baseline batch on the default stream, then a smaller without-baseline one on
    # the caller's own stream
    for (n, wb), stream in zip(((3_000_001, 1), (777_777, 0)), (None, own_stream)):
        d, E = _case(rng, n, dev)
        outs = []
        for which in ("noalloc", "workspace"):
            adv, ret = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
            stats = torch.zeros(8, dtype=torch.float64, device=dev)
            torch.cuda.synchronize()
            if which == "noalloc":
                free0, _ = torch.cuda.mem_get_info(dev)
                st = L.rrl_scan_device_noalloc(p(d["rew"]), p(d["val"]), p(d["offs"]), p(d["lvs"]), n, E, g, l, wb,
                                               p(adv), p(ret), p(stats),
                                               C.c_void_p(stream.cuda_stream) if stream is not None else None)
                torch.cuda.synchronize()
                free1, _ = torch.cuda.mem_get_info(dev)
                assert free1 == free0, "the no-allocation entry point changed free device memory: %d -> %d" % (free0, free1)
            else:
                ws = torch.empty(scan_workspace_bytes(n), dtype=torch.uint8, device=dev)
                st = L.rrl_scan_device(p(d["rew"]), p(d["val"]), p(d["offs"]), p(d["lvs"]), n, E, g, l, wb, p(adv),
                                       p(ret), p(stats), p(ws), ws.numel(), None)
            assert st == 0, L.rrl_last_error(None)
            torch.cuda.synchronize()
            outs.append((adv.cpu().numpy(), ret.cpu().numpy(), stats.cpu().numpy()))
        (a0, r0, s0), (a1, r1, s1) = outs
        np.testing.assert_allclose(a0, a1, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(r0, r1, rtol=RTOL, atol=ATOL)
        assert s0[2] == n == s1[2]
        np.testing.assert_allclose(s0[:2], s1[:2], rtol=1e-9)          # sum, sum of squares (f64)
        np.testing.assert_allclose(s0[3:5], s1[3:5], rtol=1e-6, atol=1e-7)   # mean, std (f32 values)
        assert np.abs(a0).max() > 0 and np.isfinite(a0).all() and np.isfinite(r0).all()
        # the checker of record: the sequential float64 recurrence on the host
        a_ref, r_ref = sr.finish_paths(d["rew"].cpu().numpy(), d["val"].cpu().numpy(), d["offs"].cpu().numpy(),
                                       d["lvs"].cpu().numpy(), g, l, bool(wb))
        np.testing.assert_allclose(a0, a_ref, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(r0, r_ref, rtol=RTOL, atol=ATOL)
        assert np.mean(a0 != a_ref) < 1e-3


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from relayrl_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    x = torch.zeros(16, device=dev)
    offs = torch.zeros(2, dtype=torch.int64, device=dev)
    p = lambda t: C.c_void_p(t.data_ptr())
    too_big = 16 * 1024 * 1024 + 1
    st = L.rrl_scan_device_noalloc(p(x), p(x), p(offs), p(x), too_big, 1, 0.98, 0.97, 1, p(x), p(x), None, None)
    assert st == -1
    msg = L.rrl_last_error(None).decode()
    assert "exceeds the static scratch limit" in msg and str(16 * 1024 * 1024) in msg
