"""Generates tests/golden/*.npz by running the UNMODIFIED reference learner
(/root/reference/relayrl_framework/src/native/python/algorithms/REINFORCE/) in place.

Run in the build container only:   python tests/golden/make_golden.py
The .npz files are committed; the GPU box never needs /root/reference.

What is recorded per case: the action stream fed to REINFORCE.receive_trajectory, the model's
initial state (flat, torch Linear order), and -- per epoch -- the successive path_start_idx values
(episode offsets), the raw adv/ret buffers, the dict ReplayBuffer.get() returned, the scalars
train_model() stored in the logger and the flat parameters after the epoch.
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
from oracle import ref_bridge as rb  # noqa: E402


def flat(module_seq):
    return torch.cat([p.detach().reshape(-1) for p in module_seq.parameters()]).numpy().copy()


def make_stream(rng, lens, d, A, mask_p=0.0, tail=0, double_done_at=None):
    """A flat action stream: for each length L, L transitions then one terminal marker."""
    rows = []
    for k, L in enumerate(lens):
        for _ in range(int(L)):
            rows.append(0)
        rows.append(1)
        if double_done_at is not None and k == double_done_at:
            rows.append(1)  # an empty path: two terminal markers in a row
    rows += [0] * tail      # unfinished tail
    done = np.array(rows, np.uint8)
    n = len(done)
    obs = rng.standard_normal((n, d))                       # f64 on the wire
    act = rng.integers(0, A, size=n).astype(np.float32)
    mask = np.ones((n, A), np.float64)
    if mask_p > 0:
        for i in range(n):
            if rng.random() < mask_p:
                j = int(rng.integers(0, A))
                if j != int(act[i]):
                    mask[i, j] = 0.0
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    logp = rng.standard_normal(n).astype(np.float32)
    return dict(done=done, obs=obs, act=act, mask=mask, rew=rew, val=val, logp=logp)


def to_actions(st, lo, hi, with_baseline):
    acts = []
    for i in range(lo, hi):
        if st["done"][i]:
            acts.append(rb.DuckAction(rew=float(st["rew"][i]), data=None, done=True))
        else:
            data = {"logp_a": np.array([st["logp"][i]], np.float32)}
            if with_baseline:
                data["v"] = np.float32(st["val"][i])
            acts.append(rb.DuckAction(st["obs"][i], np.array([st["act"][i]], np.float32),
                                      st["mask"][i], float(st["rew"][i]), data, False))
    return acts


def run_case(name, d, A, H, with_baseline, epochs, lens_per_epoch, iters, seed, mask_p=0.0,
             tail_epoch=None, double_done=None):
    hp = dict(with_vf_baseline=with_baseline, traj_per_epoch=1, train_vf_iters=iters, seed=seed)
    tmp = tempfile.mkdtemp(prefix="rrl_golden_")
    R, K, B = rb.load(hp, os.path.join(tmp, "server_model.pt"))
    torch.manual_seed(seed)
    algo = R.REINFORCE(env_dir=tmp, config_path="unused", obs_dim=d, act_dim=A, buf_size=200000)
    torch.manual_seed(1000 + seed)
    if H != 128:  # hidden sizes are hard-coded [128,128] in REINFORCE.py:46,49
        if with_baseline:
            algo._model = K.PolicyWithBaseline(d, A, True, [H, H], torch.nn.ReLU)
            algo._vf_optimizer = torch.optim.Adam(algo._model.baseline.parameters(), lr=algo._vf_lr)
        else:
            algo._model = K.PolicyWithoutBaseline(d, A, True, [H, H])
        algo._pi_optimizer = torch.optim.Adam(algo._model.parameters(), lr=algo._pi_lr)

    out = dict(d=d, A=A, H=H, with_baseline=int(with_baseline), iters=iters, epochs=epochs,
               gamma=algo._gamma, lam=algo._lam, pi_lr=algo._pi_lr, vf_lr=algo._vf_lr,
               pi0=flat(algo._model.policy.pi_network))
    if with_baseline:
        out["v0"] = flat(algo._model.baseline.v_network)

    buf = algo._replay_buffer
    rec = {}
    orig_finish, orig_get, orig_store = buf.finish_path, buf.get, algo.logger.store

    def finish(last_val=0):
        orig_finish(last_val)
        rec["offsets"].append(buf.path_start_idx)

    def get():
        n = buf.ptr
        rec["adv_raw"] = buf.adv_buf[:n].copy()
        rec["ret_raw"] = buf.ret_buf[:n].copy()
        data = orig_get()
        rec["data"] = {k: v.numpy().copy() for k, v in data.items()}
        return data

    def store(**kw):
        for k, v in kw.items():
            if k in ("LossPi", "DeltaLossPi", "KL", "Entropy", "LossV", "DeltaLossV"):
                rec[k] = float(v)
        orig_store(**kw)

    buf.finish_path, buf.get, algo.logger.store = finish, get, store

    rng = np.random.default_rng(seed)
    for ep in range(epochs):
        tail = 7 if tail_epoch == ep else 0
        st = make_stream(rng, lens_per_epoch[ep], d, A, mask_p, tail,
                         double_done_at=double_done if ep == 0 else None)
        rec.clear()
        rec["offsets"] = [0]
        updated = algo.receive_trajectory(rb.DuckTrajectory(to_actions(st, 0, len(st["done"]), with_baseline)))
        assert updated
        p = "e%d_" % ep
        for k, v in st.items():
            out[p + "in_" + k] = v
        out[p + "offsets"] = np.array(rec["offsets"], np.int64)
        out[p + "adv_raw"], out[p + "ret_raw"] = rec["adv_raw"], rec["ret_raw"]
        for k, v in rec["data"].items():
            out[p + "get_" + k] = v
        for k in ("LossPi", "DeltaLossPi", "KL", "Entropy", "LossV", "DeltaLossV"):
            if k in rec:
                out[p + k] = np.float64(rec[k])
        out[p + "pi"] = flat(algo._model.policy.pi_network)
        if with_baseline:
            out[p + "v"] = flat(algo._model.baseline.v_network)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print("wrote", name, {k: getattr(v, "shape", v) for k, v in out.items() if k.startswith("e0_")})


def scan_case(name, with_baseline, seed):
    """ReplayBuffer alone (replay_buffer.py): ragged lengths incl. 1, an empty path and a tail."""
    R, K, B = rb.load()
    rng = np.random.default_rng(seed)
    lens = [1, 2, 1, 31, 32, 33, 700, 1, 5000, 64, 257, 3, 1023, 1024, 1025]
    n = sum(lens) + 11
    g, l = rb.f32(0.98), rb.f32(0.97)
    buf = B.ReplayBuffer(3, 2, n + 5, g, l, with_baseline)
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    offsets, last_vals = [0], []
    i = 0
    for k, L in enumerate(lens):
        for _ in range(L):
            buf.store(np.zeros(3), 0.0, np.ones(2), rew[i], val[i] if with_baseline else None, 0.0)
            i += 1
        lv = float(np.float32(rng.uniform(-1, 1)))
        buf.finish_path(lv)
        offsets.append(buf.path_start_idx)
        last_vals.append(lv)
        if k == 4:  # empty path
            lv = float(np.float32(rng.uniform(-1, 1)))
            buf.finish_path(lv)
            offsets.append(buf.path_start_idx)
            last_vals.append(lv)
    while i < n:  # unfinished tail keeps zeros (fresh buffer)
        buf.store(np.zeros(3), 0.0, np.ones(2), rew[i], val[i] if with_baseline else None, 0.0)
        i += 1
    adv_raw, ret_raw = buf.adv_buf[:n].copy(), buf.ret_buf[:n].copy()
    data = buf.get()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), rew=rew, val=val,
                        offsets=np.array(offsets, np.int64), last_vals=np.array(last_vals, np.float32),
                        gamma=g, lam=l, with_baseline=int(with_baseline), adv_raw=adv_raw,
                        ret_raw=ret_raw, adv_norm=data["adv"].numpy(), ret=data["ret"].numpy())
    print("wrote", name, n, len(offsets))


def actor_case(name, d, A, H, batch, seed):
    R, K, B = rb.load()
    torch.manual_seed(seed)
    m = K.PolicyWithBaseline(d, A, True, [H, H], torch.nn.ReLU)
    rng = np.random.default_rng(seed)
    obs = rng.standard_normal((batch, d)).astype(np.float32)
    mask = np.ones((batch, A), np.float32)
    for i in range(batch):
        k = int(rng.integers(0, min(7, A)))
        off = rng.choice(A, size=k, replace=False)
        mask[i, off] = 0.0
    to, tm = torch.from_numpy(obs), torch.from_numpy(mask)
    act, data = m.step(to, tm)
    probs, logits, _ = m.policy.forward(to, tm)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), d=d, A=A, H=H, obs=obs, mask=mask,
                        pi=flat(m.policy.pi_network), v=flat(m.baseline.v_network),
                        act=act.numpy(), logp_a=data["logp_a"].numpy(), vals=data["v"].numpy(),
                        probs=probs.detach().numpy(), logits=logits.detach().numpy())
    print("wrote", name)


if __name__ == "__main__":
    assert rb.available(), "needs /root/reference"
    scan_case("scan_with_baseline", True, 11)
    scan_case("scan_without_baseline", False, 12)
    # CartPole-like (C1): d=4, A=2, with baseline, H=64 and the reference default H=128
    run_case("epoch_vf_h128", 4, 2, 128, True, 2, [[40, 9, 120, 77, 1, 200, 33, 64],
                                                   [100, 8, 50, 61, 17, 90, 128, 5]], 80, 3,
             tail_epoch=1, double_done=2)
    run_case("epoch_vf_h64", 4, 2, 64, True, 1, [[60, 10, 111, 70, 2, 150, 43, 64]], 20, 4)
    # LunarLander-like, without baseline, masked actions
    run_case("epoch_novf_h128", 8, 4, 128, False, 2, [[90, 14, 130, 77, 3, 210],
                                                      [55, 8, 150, 31, 17, 99]], 0, 5, mask_p=0.2)
    actor_case("actor_d128_h256_a18", 128, 18, 256, 96, 6)
    actor_case("actor_d4_h128_a2", 4, 2, 128, 64, 7)
