"""Developer probe: where the value/policy gradient of the CUDA path departs from a float64 evaluation as N grows
(per parameter block: W1,b1,W2,b2,W3,b3).  Usage: python tools/diag_grad_scale.py N [N ...]; RRL_DISABLE_TC=1 selects
the CUDA-core kernel."""
import os

# RRL_DISABLE_TC is honoured by the developer library only
os.environ.setdefault("RELAYRL_B200_SO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "relayrl-prototype_b200", "relayrl_b200", "librelayrl_b200_dev.so"))
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import f64_eval  # noqa: E402
from relayrl_b200 import Learner  # noqa: E402
from test_gpu_scale import A, D, H, _c3_lengths, _linear_init  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    for n in [int(x) for x in sys.argv[1:]] or [1 << 22]:
        rng = np.random.default_rng(n % 1000)
        lens = _c3_lengths(rng, n)
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        last_vals = rng.uniform(-1, 1, len(lens)).astype(np.float32)
        g = torch.Generator(device=dev).manual_seed(n % 1000)
        cols = {"obs": torch.randn(n, D, device=dev, generator=g),
                "act": torch.randint(0, A, (n,), device=dev, generator=g).float(),
                "mask": torch.ones(n, A, device=dev),
                "rew": torch.empty(n, device=dev).uniform_(-1, 1, generator=g),
                "val": torch.randn(n, device=dev, generator=g), "logp": torch.randn(n, device=dev, generator=g)}
        pi0, v0 = _linear_init(rng, D, H, A), _linear_init(rng, D, H, 1)
        L = Learner(D, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=1)
        L.set_params(0, pi0), L.set_params(1, v0)
        torch.cuda.synchronize()
        L.ingest_device(cols["obs"], cols["act"], cols["mask"], cols["rew"], cols["val"], cols["logp"], n, offsets, last_vals)
        L.finish()
        adv = torch.from_numpy(L.advantages(n, True).copy()).to(dev)
        ret = torch.from_numpy(L.returns(n).copy()).to(dev)
        gp, lp = L.eval_grad(0)
        gv, lv = L.eval_grad(1)
        L.close()
        r64 = f64_eval.evaluate(pi0, v0, D, H, A, cols, adv, ret, dtype=torch.float64)
        r32 = f64_eval.evaluate(pi0, v0, D, H, A, cols, adv, ret, dtype=torch.float32)
        print("n=%d TC=%s loss_v %.9g (f64 %.9g) loss_pi %.9g (f64 %.9g)" % (n, "off" if os.environ.get("RRL_DISABLE_TC") else "on",
                                                                       lv[0], r64["loss_v"], lp[0], r64["loss_pi"]))
        for name, got, key, out in (("policy", gp, "pi_grad", A), ("value", gv, "v_grad", 1)):
            want, w32 = r64[key], r32[key]
            o = 0
            for seg, cnt in (("W1", H * D), ("b1", H), ("W2", H * H), ("b2", H), ("W3", out * H), ("b3", out)):
                a, b, c = got[o:o + cnt], want[o:o + cnt], w32[o:o + cnt]
                e = np.abs(a - b)
                i = int(e.argmax())
                print("  %s %s: max|g| %.3g  max err %.3g (rel to max %.2e; at entry %.3g)  torch-f32 err %.3g  mean signed err/|g| %.2e"
                      % (name, seg, np.abs(b).max(), e.max(), e.max() / np.abs(b).max(), b[i], np.abs(c - b).max(),
                         np.mean((a - b) / np.maximum(np.abs(b), 1e-30) * (np.abs(b) > 0.1 * np.abs(b).max()))))
                o += cnt


if __name__ == "__main__":
    main()
