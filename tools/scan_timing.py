"""Developer probe: phase breakdown of the scan kernel (the middle CTA) from a -DSCAN_TIMING build.
Usage: RELAYRL_B200_SO=tools/_timing/librelayrl_b200_timing.so python tools/scan_timing.py [N]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200")):
    sys.path.insert(0, p)
import bench  # noqa: E402
from relayrl_b200 import _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 26
dev = torch.device("cuda", 0)
rng = np.random.default_rng(64)
lens = bench.episode_lengths(rng, n)
offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
cols = {"rew": torch.empty(n, device=dev).uniform_(-1, 1), "val": torch.randn(n, device=dev)}
last_vals = rng.uniform(-1, 1, len(lens)).astype(np.float32)
L = _lib.lib()
L.rrl_debug_scan_timing.argtypes = [C.c_void_p, C.c_int32]
ms = bench.measure_scan(cols, offsets, last_vals, n, dev, reps=5)
out = (C.c_uint64 * 16)()
L.rrl_debug_scan_timing(None, 1)
ms = bench.measure_scan(cols, offsets, last_vals, n, dev, reps=5)
L.rrl_debug_scan_timing(out, 0)
names = {0: "consumer: loads + pass 1", 1: "consumer: warp scan", 2: "consumer: barrier 1", 3: "consumer: combine + look-back wait",
         6: "consumer: pass 2 arithmetic", 7: "consumer: pass 2 stores", 4: "consumer: pass 2 statistics reduce", 5: "consumer: wait for producer", 8: "producer: issue copies", 9: "producer: search + mask",
         10: "producer: wait for data", 11: "producer: wait for consumers"}
tiles = max(out[13], 1)
print("scan %.3f ms, %d tiles by CTA 0 (over 8 launches)" % (ms, tiles))
for k, v in names.items():
    print("  %-40s %8.0f cycles per tile" % (v, out[k] / tiles))
