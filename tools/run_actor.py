"""Developer driver for profiles: a few batched actor steps at BASELINE config C4 (65 536 envs, d = 128, 2 x 256, 18
actions, with baseline), device-resident inputs.  Usage: python tools/run_actor.py [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200")):
    sys.path.insert(0, p)
from relayrl_b200 import Actor  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    B, d, A, H = 65536, 128, 18, 256
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(0)
    n_pi, n_v = H * d + H + H * H + H + A * H + A, H * d + H + H * H + H + H + 1
    a = Actor(d, A, hidden=H, with_baseline=True, max_batch=B)
    a.set_params(0, (rng.standard_normal(n_pi) * 0.05).astype(np.float32))
    a.set_params(1, (rng.standard_normal(n_v) * 0.05).astype(np.float32))
    obs = torch.randn(B, d, device=dev)
    mask = torch.ones(B, A, device=dev)
    act = torch.empty(B, dtype=torch.int64, device=dev)
    lp, v = torch.empty(B, device=dev), torch.empty(B, device=dev)
    st = torch.cuda.ExternalStream(a.stream, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(reps):
        if i == reps - 1:
            e0.record(st)
        a.step_device(obs, mask, act, lp, v, seed=1, offset=i)
    e1.record(st)
    torch.cuda.synchronize()
    print("last call %.4f ms" % e0.elapsed_time(e1))
    a.close()


if __name__ == "__main__":
    main()
