"""Pins the oracle against the UNMODIFIED reference imported in place (build container only;
skipped where /root/reference is absent, e.g. on the GPU box)."""
import numpy as np
import pytest
import torch

from oracle import learner_oracle as lo
from oracle import ref_bridge as rb

pytestmark = pytest.mark.skipif(not rb.available(), reason="reference tree not present")


def test_replay_buffer_matches_reference():
    R, K, B = rb.load()
    rng = np.random.default_rng(0)
    for wb in (True, False):
        g, l = rb.f32(0.98), rb.f32(0.97)
        buf = B.ReplayBuffer(4, 2, 5000, g, l, wb)
        lens = list(rng.integers(1, 300, 20))
        n = int(sum(lens))
        rew = rng.uniform(-1, 1, n).astype(np.float32)
        val = rng.standard_normal(n).astype(np.float32)
        lvs, offs, i = [], [0], 0
        for L in lens:
            for _ in range(L):
                buf.store(np.zeros(4), 0, np.ones(2), rew[i], val[i] if wb else None, 0.0)
                i += 1
            lv = float(np.float32(rng.uniform(-1, 1)))
            buf.finish_path(lv)
            lvs.append(lv)
            offs.append(buf.path_start_idx)
        adv, ret = np.zeros(n, np.float32), np.zeros(n, np.float32)
        lo.finish_paths(rew, val, np.array(offs), np.array(lvs, np.float32), g, l, wb, adv, ret)
        assert np.array_equal(adv, buf.adv_buf[:n]) and np.array_equal(ret, buf.ret_buf[:n])
        data = buf.get()
        assert np.array_equal(lo.normalise(adv), data["adv"].numpy())


def test_policy_and_losses_match_reference_modules():
    R, K, B = rb.load()
    torch.manual_seed(0)
    m = K.PolicyWithBaseline(4, 3, True, [64, 64], torch.nn.ReLU)
    pi = [p.detach().clone() for p in m.policy.pi_network.parameters()]
    v = [p.detach().clone() for p in m.baseline.v_network.parameters()]
    obs = torch.randn(50, 4)
    mask = torch.ones(50, 3)
    mask[::5, 1] = 0
    act = torch.randint(0, 3, (50,)).float()
    act[::5] = 0
    probs, logits, logp = m.policy.forward(obs, mask, act)
    p2, l2, lp2 = lo.policy_forward(pi, obs, mask, act)
    assert torch.equal(probs, p2) and torch.equal(logits, l2) and torch.equal(logp, lp2)
    assert torch.equal(m.baseline.forward(obs, mask), lo.value_forward(v, obs))
