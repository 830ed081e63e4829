"""Developer probe: value-gradient error of the CUDA path on a C1-sized batch after N epochs of C1 training (large value
targets, ~2500 transitions), per parameter block, against float64 and next to a torch float32 evaluation.
RELAYRL_B200_SO=<dev lib> RRL_DISABLE_TC=1 selects the CUDA-core kernel."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import f64_eval  # noqa: E402
import pipe_driver as pd  # noqa: E402
from oracle import learner_oracle as lo  # noqa: E402
from relayrl_b200 import Learner  # noqa: E402

d, A, H, epochs = 4, 2, 128, int(sys.argv[1]) if len(sys.argv) > 1 else 49
dev = torch.device("cuda", 0)
rng = np.random.default_rng(1)
L = Learner(d, A, 8 * 500 + 8, hidden=H, with_baseline=True, train_vf_iters=80)
L.set_params(0, lo.flatten(lo.init_params(d, H, A, 1)).numpy()), L.set_params(1, lo.flatten(lo.init_params(d, H, 1, 2)).numpy())
for ep in range(epochs + 1):
    def policy(obs):
        with torch.no_grad():
            act, out = lo.actor_step(lo.unflatten(L.get_params(0), d, H, A), lo.unflatten(L.get_params(1), d, H, 1),
                                     torch.from_numpy(obs), torch.ones(len(obs), A))
        return act.reshape(-1).numpy(), out["logp_a"].numpy(), out["v"].numpy()
    acts = []
    for _ in range(8):
        acts += pd.c1_trajectory(rng, policy=policy)
    cols = pd.actions_to_columns(acts, d, A)
    L.ingest(*cols)
    if ep < epochs:
        L.train_epoch()
n, _ = L.finish()
keep = cols[6] == 0
dc = {"obs": torch.from_numpy(cols[0][keep].astype(np.float32)).to(dev), "act": torch.from_numpy(cols[1][keep]).to(dev),
      "mask": torch.from_numpy(cols[2][keep].astype(np.float32)).to(dev), "logp": torch.from_numpy(cols[5][keep]).to(dev)}
adv = torch.from_numpy(L.advantages(n, True).copy()).to(dev)
ret = torch.from_numpy(L.returns(n).copy()).to(dev)
pi, vf = L.get_params(0), L.get_params(1)
gv, lv = L.eval_grad(1)
r64 = f64_eval.evaluate(pi, vf, d, H, A, dc, adv, ret, dtype=torch.float64)
r32 = f64_eval.evaluate(pi, vf, d, H, A, dc, adv, ret, dtype=torch.float32)
print("n=%d loss_v %.9g (f64 %.9g) TC=%s" % (n, lv[0], r64["loss_v"], "off" if os.environ.get("RRL_DISABLE_TC") else "on"))
o = 0
for seg, cnt in (("W1", H * d), ("b1", H), ("W2", H * H), ("b2", H), ("W3", H), ("b3", 1)):
    a, b, c = gv[o:o + cnt], r64["v_grad"][o:o + cnt], r32["v_grad"][o:o + cnt]
    e, e32 = np.abs(a - b), np.abs(c - b)
    print("  %s: max|g| %.3g; cuda: max err %.3g, mean signed err %+.3g, rms %.3g; torch-f32: max err %.3g, rms %.3g"
          % (seg, np.abs(b).max(), e.max(), (a - b).mean(), np.sqrt((e ** 2).mean()), e32.max(), np.sqrt((e32 ** 2).mean())))
    o += cnt
W2e = (gv - r64["v_grad"])[H * d + H:H * d + H + H * H].reshape(H, H)
print("  W2 error by row (out unit) rms: max %.3g median %.3g; by column (in unit): max %.3g median %.3g"
      % (np.sqrt((W2e ** 2).mean(1)).max(), np.median(np.sqrt((W2e ** 2).mean(1))), np.sqrt((W2e ** 2).mean(0)).max(),
         np.median(np.sqrt((W2e ** 2).mean(0)))))
