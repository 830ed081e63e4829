"""Debug probe: first-iteration losses of the tensor-core path against the CUDA-core path (RRL_DISABLE_TC=1)."""
import os, sys, subprocess, json
sys.path.insert(0, "/root/repo/relayrl-prototype_b200"); sys.path.insert(0, "/root/repo")
import numpy as np
if len(sys.argv) > 1 and sys.argv[1] == "child":
    from relayrl_b200 import Learner
    from oracle import learner_oracle as lo
    out = {}
    for n in [64 * 148, 64 * 148 + 64, 64 * 148 * 2, 64 * 148 * 3, 64 * 148 * 5 + 17]:
        rng = np.random.default_rng(1)
        d, A, H = 4, 2, 128
        done = np.zeros(n + 1, np.uint8); done[-1] = 1
        obs = rng.standard_normal((n + 1, d)).astype(np.float32)
        act = rng.integers(0, A, n + 1).astype(np.float32)
        mask = np.ones((n + 1, A), np.float32)
        rew = rng.uniform(-1, 1, n + 1).astype(np.float32); val = rng.standard_normal(n + 1).astype(np.float32)
        logp = rng.standard_normal(n + 1).astype(np.float32)
        L = Learner(d, A, n + 2, hidden=H, with_baseline=True, train_vf_iters=1)
        L.set_params(0, lo.flatten(lo.init_params(d, H, A, 1)).numpy()); L.set_params(1, lo.flatten(lo.init_params(d, H, 1, 2)).numpy())
        L.ingest(obs, act, mask, rew, val, logp, done)
        st = L.train_epoch()
        out[n] = [st["loss_pi"], st["loss_v_old"], st["entropy"], float(np.abs(L.get_params(1)).sum())]
        L.close()
    print(json.dumps(out))
else:
    res = {}
    for tc in (1, 0):
        env = dict(os.environ)
        if not tc: env["RRL_DISABLE_TC"] = "1"
        r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
        res[tc] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else r.stderr[-2000:]
    for n in res[1]:
        print(n, "TC ", res[1][n]); print(n, "SIMT", res[0][n])
