"""rrl_rollout_step / rrl_rollout_flag_last: many environments stepped per call, their
transitions appended to the learner's replay buffer on the device.  What lands in the buffer must be what the reference
would have built one RelayRLAction at a time (request_for_action / flag_last_action, agent_zmq.rs:458-571,605-610) and
fed through receive_trajectory: checked by replaying the same per-environment action streams through the host ingest
of a second learner and comparing offsets, returns, advantages and one full training epoch bit for bit."""
import ctypes as C

import numpy as np
import pytest

from oracle import learner_oracle as lo

pytestmark = pytest.mark.gpu


def test_rollout_fills_the_buffer_like_per_action_ingest():
    from relayrl_b200 import Actor, Learner, _lib
    L_ = _lib.lib()
    d, A, H, N, T = 4, 2, 128, 48, 37
    rng = np.random.default_rng(5)
    act = Actor(d, A, hidden=H, with_baseline=True, max_batch=64)
    pi0, v0 = lo.flatten(lo.init_params(d, H, A, 1)).numpy(), lo.flatten(lo.init_params(d, H, 1, 2)).numpy()
    act.set_params(0, pi0), act.set_params(1, v0)
    cap = N * T + 8
    Lr = Learner(d, A, cap, hidden=H, with_baseline=True, train_vf_iters=3)
    Lh = Learner(d, A, cap, hidden=H, with_baseline=True, train_vf_iters=3)
    for L in (Lr, Lh):
        L.set_params(0, pi0), L.set_params(1, v0)
    r = C.c_void_p()
    assert L_.rrl_rollout_create(act._h, Lr._h, N, T, C.byref(r)) == 0, L_.rrl_rollout_last_error(None)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    streams = [[] for _ in range(N)]          # per environment: the actions the reference agent would have accumulated
    order = []                                # episodes in the order they were flagged
    life = rng.integers(1, T + 1, N)          # steps until each environment's episode ends
    for t in range(T):
        obs = rng.standard_normal((N, d)).astype(np.float32)
        mask = np.ones((N, A), np.float32)
        mask[rng.random(N) < 0.2, 1] = 0.0
        rew = rng.uniform(-1, 1, N).astype(np.float32)
        a_out, lp, v = np.empty(N, np.int64), np.empty(N, np.float32), np.empty(N, np.float32)
        assert L_.rrl_rollout_step(r, p(obs), p(mask), p(rew), 7, t * N, p(a_out), p(lp), p(v)) == 0, L_.rrl_rollout_last_error(r)
        assert ((a_out >= 0) & (a_out < A)).all() and (mask[np.arange(N), a_out] == 1).all()
        for e in range(N):
            streams[e].append((obs[e], float(a_out[e]), mask[e], rew[e], v[e], lp[e]))
        ended = np.flatnonzero(life == t + 1)
        if len(ended):
            last = rng.integers(0, 2, len(ended)).astype(np.float32)
            assert L_.rrl_rollout_flag_last(r, p(ended.astype(np.int64)), p(last), len(ended)) == 0, L_.rrl_rollout_last_error(r)
            for e, lv in zip(ended, last):
                order.append((streams[e], float(lv)))
                streams[e] = []
                life[e] = t + 1 + rng.integers(1, T + 1)
    # the same episodes through the per-action host path
    for acts, lv in order:
        n = len(acts)
        done = np.r_[np.zeros(n, np.uint8), np.uint8(1)]
        col = lambda i, w=None: np.concatenate([np.stack([x[i] for x in acts]).astype(np.float32).reshape(n, -1) if w else
                                                np.array([x[i] for x in acts], np.float32), np.zeros((1, w) if w else 1, np.float32)])
        Lh.ingest(col(0, d), col(1), col(2, A), np.r_[np.array([x[3] for x in acts], np.float32), np.float32(lv)], col(4), col(5), done)
    nr, er = Lr.finish()
    nh, eh = Lh.finish()
    assert (nr, er) == (nh, eh) and er == len(order)
    assert np.array_equal(Lr.offsets(), Lh.offsets())
    assert np.array_equal(Lr.returns(nr), Lh.returns(nh)) and np.array_equal(Lr.advantages(nr, False), Lh.advantages(nh, False))
    sr_, sh_ = Lr.train_epoch(), Lh.train_epoch()
    assert sr_["loss_pi"] == sh_["loss_pi"] and sr_["loss_v"] == sh_["loss_v"] and sr_["kl"] == sh_["kl"]
    assert np.array_equal(Lr.get_params(0), Lh.get_params(0)) and np.array_equal(Lr.get_params(1), Lh.get_params(1))
    # an environment that never ends runs into max_len
    for t in range(T):
        st = L_.rrl_rollout_step(r, p(obs), p(mask), p(rew), 7, 0, p(a_out), p(lp), p(v))
        if st != 0:
            break
    assert st == -3 and b"max_len" in L_.rrl_rollout_last_error(r)
    assert L_.rrl_rollout_destroy(r) == 0
    for x in (Lr, Lh, act):
        x.close()
