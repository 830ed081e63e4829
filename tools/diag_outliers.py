"""Developer probe: which parameters miss rel 1e-5 + abs 5e-7 after the golden epochs, and what their gradients and
Adam moments look like (basis for the bounds in tests/test_gpu_learner.py::assert_weights_close)."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from conftest import load_golden  # noqa: E402
from relayrl_b200 import Learner  # noqa: E402


def main():
    for name in ("epoch_vf_h128", "epoch_vf_h64", "epoch_novf_h128"):
        g = load_golden(name)
        d, A, H, wb = int(g["d"]), int(g["A"]), int(g["H"]), bool(g["with_baseline"])
        L = Learner(d, A, 8192, hidden=H, with_baseline=wb, gamma=float(g["gamma"]), lam=float(g["lam"]),
                    pi_lr=float(g["pi_lr"]), vf_lr=float(g["vf_lr"]), train_vf_iters=int(g["iters"]))
        L.set_params(0, g["pi0"])
        if wb:
            L.set_params(1, g["v0"])
        for ep in range(int(g["epochs"])):
            p = "e%d_" % ep
            ing = lambda: L.ingest(g[p + "in_obs"], g[p + "in_act"], g[p + "in_mask"], g[p + "in_rew"], g[p + "in_val"],
                                   g[p + "in_logp"], g[p + "in_done"])
            ing()
            L.train_epoch()
            for which, key, lr in ((0, "pi", float(g["pi_lr"])), (1, "v", float(g["vf_lr"]))):
                if which == 1 and not wb:
                    continue
                got, want = L.get_params(which), g[p + key]
                err = np.abs(got - want)
                out = np.flatnonzero(err > 5e-7 + 1e-5 * np.abs(want))
                m, v, step = L.get_opt_state(which)
                ing()
                grad, _ = L.eval_grad(which)
                L.reset_buffer()
                sv = np.sqrt(v / (1 - 0.999 ** step))
                print("%s epoch %d %s: %d outliers of %d (step %d); max err/lr %.3g" %
                      (name, ep, key, len(out), len(got), step, (err / lr).max()))
                for i in out[:12]:
                    print("   i=%d err/lr=%.3g |g|/max=%.3g sqrt(vhat)=%.3g (max %.3g) |m|=%.3g w=%.4g" %
                          (i, err[i] / lr, abs(grad[i]) / np.abs(grad).max(), sv[i], sv.max(), abs(m[i]), want[i]))
        L.close()


if __name__ == "__main__":
    main()
