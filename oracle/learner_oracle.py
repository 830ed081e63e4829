"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the RelayRL learner hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this module, and only as the checker or the timed CPU baseline; the product never does.

Every function restates one piece of the reference's REINFORCE learner (paths relative to
/root/reference/relayrl_framework/src/native/python/) in a batch/functional form, using the
same third-party arithmetic the reference calls (scipy.signal.lfilter, numpy f32 reductions,
torch CPU nn/autograd/optim.Adam).  Parity is PINNED: tests/test_oracle_vs_reference.py runs this
file against the unmodified reference modules imported in place (oracle/ref_bridge.py) and
tests/golden/*.npz hold outputs generated from the reference itself
(tests/golden/make_golden.py).  The reference ships no tests or golden vectors of its own
(SURVEY.md section 4), so those generated vectors are the pin.
"""
from __future__ import annotations

import numpy as np

try:  # scipy is what the reference calls (BaseReplayBuffer.py:27); the loop below is identical
    from scipy.signal import lfilter as _lfilter
except Exception:  # pragma: no cover
    _lfilter = None


# --------------------------------------------------------------------------------------
# a1: ingest + episode segmentation      (algorithms/REINFORCE/REINFORCE.py:70-95,
#                                          algorithms/REINFORCE/replay_buffer.py:34-46)
# --------------------------------------------------------------------------------------
def segment(done: np.ndarray, rew: np.ndarray, start_ptr: int = 0):
    """Walks a flat action stream the way receive_trajectory does.

    A non-done action is appended at `ptr`; a done action (terminal marker) stores nothing and
    closes the path: path_start_idx <- ptr, its reward is the path's `last_val`
    (REINFORCE.py:86 -> replay_buffer.py:48-79).

    Returns (keep_idx, offsets, last_vals): indices of the stored actions, the successive
    path_start_idx values as int64 (offsets[0] = start_ptr), and one last_val per closed path.
    """
    done = np.asarray(done).astype(bool)
    keep = np.flatnonzero(~done)
    # ptr after processing action i = start_ptr + number of non-done actions in [0, i]
    ptr_after = start_ptr + np.cumsum(~done)
    closes = np.flatnonzero(done)
    offsets = np.concatenate([[start_ptr], ptr_after[closes]]).astype(np.int64)
    last_vals = np.asarray(rew, dtype=np.float32)[closes]
    return keep, offsets, last_vals


def store_loop(obs, act, mask, rew, val, logp, done, gamma, lam, with_baseline):
    """The reference's ingest exactly as it runs: one Python iteration per action
    (REINFORCE.receive_trajectory, REINFORCE.py:72-87), ReplayBuffer.store per transition
    (replay_buffer.py:34-46: six element-wise assignments into f32 numpy buffers) and finish_path -- two lfilter
    calls -- at every terminal marker (replay_buffer.py:48-79).  Used as the timed CPU baseline of the ingest stage
    (BASELINE.md section 3); OracleLearner.ingest is the vectorised form of the same thing and far kinder to the CPU.
    Returns (ptr, adv_buf, ret_buf)."""
    n = len(done)
    d, A = obs.shape[1], mask.shape[1]
    obs_buf, mask_buf = np.zeros((n, d), np.float32), np.zeros((n, A), np.float32)
    act_buf, rew_buf, val_buf, logp_buf, adv_buf, ret_buf = (np.zeros(n, np.float32) for _ in range(6))
    ptr = path_start = 0
    for i in range(n):
        if done[i]:
            lv = float(rew[i])
            sl = slice(path_start, ptr)
            if with_baseline:
                rews, vals = np.append(rew_buf[sl], lv), np.append(val_buf[sl], lv)
                deltas = rews[:-1] + gamma * vals[1:] - vals[:-1]
                adv_buf[sl] = discount_cumsum(deltas, gamma * lam)
                ret_buf[sl] = discount_cumsum(rews, gamma)[:-1]
            else:
                adv_buf[sl] = discount_cumsum(rew_buf[sl], gamma * lam)
                ret_buf[sl] = discount_cumsum(rew_buf[sl], gamma)
            path_start = ptr
        else:
            obs_buf[ptr] = obs[i]
            act_buf[ptr] = act[i]
            mask_buf[ptr] = mask[i]
            rew_buf[ptr] = rew[i]
            val_buf[ptr] = val[i] if with_baseline else 0.0
            logp_buf[ptr] = logp[i]
            ptr += 1
    return ptr, adv_buf, ret_buf


# --------------------------------------------------------------------------------------
# a2: discounted reverse scan            (_common/_algorithms/BaseReplayBuffer.py:12-27,
#                                          algorithms/REINFORCE/replay_buffer.py:48-79)
# --------------------------------------------------------------------------------------
def discount_cumsum(x: np.ndarray, discount: float) -> np.ndarray:
    """y[t] = x[t] + discount * y[t+1], float64 whatever the input dtype."""
    x = np.asarray(x)
    if _lfilter is not None:
        return _lfilter([1], [1, float(-discount)], x[::-1], axis=0)[::-1]
    y = np.empty(len(x), dtype=np.float64)
    acc = 0.0
    for t in range(len(x) - 1, -1, -1):
        acc = float(x[t]) + float(discount) * acc
        y[t] = acc
    return y


def finish_paths(rew, val, offsets, last_vals, gamma, lam, with_baseline, adv_buf, ret_buf):
    """Applies ReplayBuffer.finish_path to every closed path [offsets[e], offsets[e+1]).

    With baseline: GAE-lambda on deltas and reward-to-go bootstrapped with last_val
    (replay_buffer.py:65-73).  Without: both are plain discounted sums of the rewards and
    last_val is ignored (replay_buffer.py:74-77).  Results are f64 rounded to f32
    on assignment into the f32 buffers.  adv_buf / ret_buf are modified in place; entries
    outside closed paths keep what they held.
    """
    rew = np.asarray(rew, dtype=np.float32)
    for e in range(len(offsets) - 1):
        s, t = int(offsets[e]), int(offsets[e + 1])
        if with_baseline:
            # get_rew() hands finish_path a Python float (o3_action.rs:131-135), so np.append
            # promotes both arrays to float64
            lv = float(last_vals[e])
            rews = np.append(rew[s:t], lv)
            vals = np.append(np.asarray(val, dtype=np.float32)[s:t], lv)
            deltas = rews[:-1] + gamma * vals[1:] - vals[:-1]
            adv_buf[s:t] = discount_cumsum(deltas, gamma * lam)
            ret_buf[s:t] = discount_cumsum(rews, gamma)[:-1]
        else:
            r = rew[s:t]
            adv_buf[s:t] = discount_cumsum(r, gamma * lam)
            ret_buf[s:t] = discount_cumsum(r, gamma)
    return adv_buf, ret_buf


# --------------------------------------------------------------------------------------
# a3: mean / std / normalise             (_common/_algorithms/BaseReplayBuffer.py:30-53,
#                                          algorithms/REINFORCE/replay_buffer.py:81-111)
# --------------------------------------------------------------------------------------
def adv_statistics(adv: np.ndarray):
    """Two-pass population mean/std, all float32, no epsilon."""
    x = np.array(adv, dtype=np.float32)
    n = len(x)
    mean = np.sum(x) / n
    std = np.sqrt(np.sum((x - mean) ** 2) / n)
    return mean, std


def normalise(adv: np.ndarray) -> np.ndarray:
    mean, std = adv_statistics(adv)
    return (np.array(adv, dtype=np.float32) - mean) / std


# --------------------------------------------------------------------------------------
# a4-a7: networks, losses, Adam          (algorithms/REINFORCE/kernel.py:12-46,78-84,
#                                          algorithms/REINFORCE/REINFORCE.py:97-125,141-160)
# torch is imported lazily so the scan/segment parts work without it.
# --------------------------------------------------------------------------------------
def param_shapes(obs_dim: int, hidden: int, out_dim: int):
    """torch.nn.Linear order W1,b1,W2,b2,W3,b3 with W as [out, in] (kernel.py:15-21)."""
    return [(hidden, obs_dim), (hidden,), (hidden, hidden), (hidden,), (out_dim, hidden), (out_dim,)]


def flat_size(obs_dim, hidden, out_dim):
    return sum(int(np.prod(s)) for s in param_shapes(obs_dim, hidden, out_dim))


def unflatten(flat, obs_dim, hidden, out_dim):
    import torch
    out, o = [], 0
    flat = torch.as_tensor(flat)
    for s in param_shapes(obs_dim, hidden, out_dim):
        n = int(np.prod(s))
        out.append(flat[o:o + n].reshape(s).clone())
        o += n
    return out


def flatten(params):
    import torch
    return torch.cat([p.detach().reshape(-1) for p in params])


def init_params(obs_dim, hidden, out_dim, seed):
    """torch.nn.Linear default init, as the reference gets from nn.Linear (kernel.py:15-21)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    ps = []
    for s in param_shapes(obs_dim, hidden, out_dim):
        fan_in = s[1] if len(s) == 2 else None
        if fan_in is not None:
            bound = 1.0 / np.sqrt(fan_in)
            last_fan = fan_in
        else:
            bound = 1.0 / np.sqrt(last_fan)
        ps.append((torch.rand(s, generator=g, dtype=torch.float32) * 2 - 1) * bound)
    return ps


def mlp3(params, x):
    """Linear-ReLU-Linear-ReLU-Linear (kernel.py:15-21; BaseKernel.py:25-39 with nn.ReLU)."""
    import torch
    W1, b1, W2, b2, W3, b3 = params
    h = torch.relu(torch.nn.functional.linear(x, W1, b1))
    h = torch.relu(torch.nn.functional.linear(h, W2, b2))
    return torch.nn.functional.linear(h, W3, b3)


def policy_forward(params, obs, mask, act=None):
    """DiscretePolicyNetwork.forward (kernel.py:23-46): logits get (mask-1)*1e8 added, probs are
    softmax(logits) and `logp` is the RAW masked logit gathered at act -- never a log-prob
    (SURVEY Q1)."""
    import torch
    logits = mlp3(params, obs) + (mask - 1) * 1e8
    probs = torch.softmax(logits, dim=-1)
    logp = None
    if act is not None:
        logp = logits.gather(-1, act.long().unsqueeze(-1)).squeeze(-1)
    return probs, logits, logp


def loss_pi(params, obs, mask, act, adv, logp_old):
    """REINFORCE.compute_loss_pi (REINFORCE.py:141-156).  Returns (loss, kl, entropy)."""
    import torch
    probs, logits, logp = policy_forward(params, obs, mask, act)
    loss = -(logp * adv).mean()
    kl = (logp_old - logp).mean().item()
    lg = torch.clamp(logits, min=torch.finfo(logits.dtype).min)
    ent = -(lg * probs).sum(-1).mean().item()
    return loss, kl, ent


def value_forward(params, obs):
    """BaselineValueNetwork.forward (kernel.py:78-84)."""
    import torch
    return torch.squeeze(mlp3(params, obs), -1)


def loss_v(params, obs, ret):
    """REINFORCE.compute_loss_vf (REINFORCE.py:158-160)."""
    return ((value_forward(params, obs) - ret) ** 2).mean()


def train_epoch(pi_params, v_params, data, pi_lr, vf_lr, train_vf_iters, pi_opt_state=None,
                v_opt_state=None, return_grads=False):
    """REINFORCE.train_model (REINFORCE.py:97-125) on `data` = dict of f32 torch tensors
    obs, act, mask, ret, adv (already normalised), logp.

    One policy Adam step, then train_vf_iters value Adam steps (torch.optim.Adam defaults,
    REINFORCE.py:47,50).  v_params may be None (without baseline).  Optimiser objects are
    created once and passed back in through *_opt_state to continue across epochs.

    Returns dict(stats..., pi_params, v_params, pi_opt, v_opt[, pi_grad, v_grad_first]).
    """
    import torch
    from torch.optim import Adam

    pi_params = [p.detach().clone().requires_grad_(True) for p in pi_params] \
        if pi_opt_state is None else pi_params
    pi_opt = pi_opt_state or Adam(pi_params, lr=pi_lr)
    with_v = v_params is not None
    if with_v:
        v_params = [p.detach().clone().requires_grad_(True) for p in v_params] \
            if v_opt_state is None else v_params
        v_opt = v_opt_state or Adam(v_params, lr=vf_lr)

    obs, mask, act = data["obs"], data["mask"], data["act"]
    adv, ret, logp_old = data["adv"], data["ret"], data["logp"]

    with torch.no_grad():
        pi_l_old = loss_pi(pi_params, obs, mask, act, adv, logp_old)[0].item()
        v_l_old = loss_v(v_params, obs, ret).item() if with_v else None

    out = {}
    pi_opt.zero_grad()
    l_pi, kl, ent = loss_pi(pi_params, obs, mask, act, adv, logp_old)
    l_pi.backward()
    if return_grads:
        out["pi_grad"] = torch.cat([p.grad.reshape(-1) for p in pi_params]).clone()
    pi_opt.step()

    l_v = None
    if with_v:
        for i in range(train_vf_iters):
            v_opt.zero_grad()
            l_v = loss_v(v_params, obs, ret)
            l_v.backward()
            if return_grads and i == 0:
                out["v_grad_first"] = torch.cat([p.grad.reshape(-1) for p in v_params]).clone()
            v_opt.step()

    out.update(LossPi=l_pi.item(), DeltaLossPi=l_pi.item() - pi_l_old, KL=kl, Entropy=ent,
               pi_params=pi_params, pi_opt=pi_opt)
    if with_v:
        out.update(LossV=l_v.item() if l_v is not None else v_l_old,
                   DeltaLossV=(l_v.item() - v_l_old) if l_v is not None else 0.0,
                   LossVOld=v_l_old, v_params=v_params, v_opt=v_opt)
    return out


# --------------------------------------------------------------------------------------
# a9: actor step                          (algorithms/REINFORCE/kernel.py:99-107,128-135)
# --------------------------------------------------------------------------------------
def actor_step(pi_params, v_params, obs, mask, act=None, generator=None):
    """PolicyWith[out]Baseline.step.  If `act` is given it is used instead of sampling (the
    multinomial draw cannot be bit-matched across RNGs; parity is on probs/logits/logp_a/v)."""
    import torch
    with torch.no_grad():
        probs, logits, _ = policy_forward(pi_params, obs, mask)
        if act is None:
            act = torch.multinomial(probs, 1, generator=generator)
        else:
            act = act.long().reshape(-1, 1)
        logp_a = logits.gather(-1, act).squeeze(-1)
        out = {"logp_a": logp_a, "probs": probs, "logits": logits}
        if v_params is not None:
            out["v"] = value_forward(v_params, obs)
    return act, out


# --------------------------------------------------------------------------------------
# whole learner, batch form (used as the timed CPU baseline and by the end-to-end tests)
# --------------------------------------------------------------------------------------
class OracleLearner:
    """Functional twin of REINFORCE + ReplayBuffer (REINFORCE.py:17-125, replay_buffer.py)."""

    def __init__(self, obs_dim, act_dim, capacity, hidden=128, with_baseline=True, gamma=0.98,
                 lam=0.97, pi_lr=3e-4, vf_lr=1e-3, train_vf_iters=80, seed=0):
        f = lambda x: float(np.float32(x))
        self.d, self.A, self.H, self.cap = obs_dim, act_dim, hidden, capacity
        self.with_baseline = with_baseline
        self.gamma, self.lam, self.pi_lr, self.vf_lr = f(gamma), f(lam), f(pi_lr), f(vf_lr)
        self.iters = train_vf_iters
        self.obs = np.zeros((capacity, obs_dim), np.float32)
        self.act = np.zeros(capacity, np.float32)
        self.mask = np.zeros((capacity, act_dim), np.float32)
        self.rew = np.zeros(capacity, np.float32)
        self.val = np.zeros(capacity, np.float32)
        self.logp = np.zeros(capacity, np.float32)
        self.adv = np.zeros(capacity, np.float32)
        self.ret = np.zeros(capacity, np.float32)
        self.ptr = 0
        self.offsets = [0]
        self.last_vals = []
        self.pi = [p.requires_grad_(True) for p in init_params(obs_dim, hidden, act_dim, seed)]
        self.v = [p.requires_grad_(True) for p in init_params(obs_dim, hidden, 1, seed + 1)] \
            if with_baseline else None
        self._pi_opt = None
        self._v_opt = None

    def ingest(self, obs, act, mask, rew, val, logp, done):
        keep, offs, lvs = segment(done, rew, self.ptr)
        n = len(keep)
        assert self.ptr + n < self.cap + 1
        s = slice(self.ptr, self.ptr + n)
        self.obs[s] = np.asarray(obs)[keep]
        self.act[s] = np.asarray(act).reshape(len(done), -1)[keep, 0]
        self.mask[s] = np.asarray(mask)[keep]
        self.rew[s] = np.asarray(rew)[keep]
        if val is not None:
            self.val[s] = np.asarray(val)[keep]
        self.logp[s] = np.asarray(logp)[keep]
        self.ptr += n
        first = len(self.offsets) - 1
        self.offsets.extend(int(o) for o in offs[1:])
        self.last_vals.extend(float(v) for v in lvs)
        # finish_path runs at each done marker (REINFORCE.py:86)
        o = np.asarray(self.offsets[first:], dtype=np.int64)
        o[0] = self.offsets[first]
        finish_paths(self.rew, self.val, o, np.asarray(self.last_vals[first:], np.float32),
                     self.gamma, self.lam, self.with_baseline, self.adv, self.ret)

    def get(self):
        import torch
        n = self.ptr
        self.ptr = 0
        offsets = np.asarray(self.offsets, np.int64)
        self.offsets, self.last_vals = [0], []
        adv = normalise(self.adv[:n])
        data = dict(obs=self.obs[:n], act=self.act[:n], mask=self.mask[:n], ret=self.ret[:n],
                    adv=adv, logp=self.logp[:n])
        return {k: torch.as_tensor(v, dtype=torch.float32) for k, v in data.items()}, offsets

    def train(self):
        data, offsets = self.get()
        out = train_epoch(self.pi, self.v, data, self.pi_lr, self.vf_lr, self.iters,
                          self._pi_opt, self._v_opt)
        self._pi_opt, self._v_opt = out["pi_opt"], out.get("v_opt")
        self.pi, self.v = out["pi_params"], out.get("v_params")
        out["offsets"] = offsets
        return out
