"""GPU parity of K8 (batched actor step) against the golden vectors from the reference's
PolicyWithBaseline.step and against the oracle; sampling checked by a chi-square test (the
reference's torch.multinomial stream cannot be bit-matched)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import learner_oracle as lo

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["actor_d128_h256_a18", "actor_d4_h128_a2"])
def test_actor_golden(name):
    from relayrl_b200 import Actor
    g = load_golden(name)
    d, A, H = int(g["d"]), int(g["A"]), int(g["H"])
    a = Actor(d, A, hidden=H, with_baseline=True, max_batch=256)
    a.set_params(0, g["pi"])
    a.set_params(1, g["v"])
    act, data, probs, logits = a.step(g["obs"], g["mask"], seed=4, offset=0, want_probs=True)
    np.testing.assert_allclose(logits, g["logits"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(probs, g["probs"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(data["v"], g["vals"], rtol=1e-5, atol=1e-6)
    B = len(act)
    assert (g["mask"][np.arange(B), act] == 1).all()                  # never a masked action
    np.testing.assert_array_equal(data["logp_a"], logits[np.arange(B), act])   # logp_a = logits[act]
    # same (seed, offset) -> same actions; different offset -> a different stream
    act2, _ = a.step(g["obs"], g["mask"], seed=4, offset=0)
    assert np.array_equal(act, act2)
    # forced actions reproduce the reference's logp_a diagonal
    dev = torch.device("cuda:0")
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    o_act = torch.empty(B, dtype=torch.int64, device=dev)
    o_lp, o_v = torch.empty(B, device=dev), torch.empty(B, device=dev)
    a.step_device(t(g["obs"]), t(g["mask"]), o_act, o_lp, o_v, force_act=t(g["act"].reshape(-1)))
    np.testing.assert_allclose(o_lp.cpu().numpy(), np.diagonal(g["logp_a"]), rtol=1e-5, atol=1e-6)
    a.close()


def test_actor_sampling_chi_square():
    from relayrl_b200 import Actor
    d, A, H, B = 4, 6, 64, 32
    pi = lo.init_params(d, H, A, 9)
    a = Actor(d, A, hidden=H, with_baseline=False, max_batch=1 << 16)
    a.set_params(0, lo.flatten(pi).numpy())
    rng = np.random.default_rng(0)
    obs1 = (rng.standard_normal((1, d)) * 3).astype(np.float32)
    mask1 = np.array([[1, 1, 0, 1, 1, 1]], np.float32)
    N = 1 << 16
    obs, mask = np.repeat(obs1, N, 0), np.repeat(mask1, N, 0)
    act, data, probs, _ = a.step(obs, mask, seed=123, offset=5, want_probs=True)
    p = probs[0].astype(np.float64)
    cnt = np.bincount(act, minlength=A).astype(np.float64)
    assert cnt[2] == 0
    keep = p > 0
    chi2 = (((cnt - N * p) ** 2)[keep] / (N * p[keep])).sum()
    assert chi2 < 30.0, chi2          # 4 dof: P(chi2 > 30) ~ 5e-6
    _, ref = lo.actor_step(pi, None, torch.from_numpy(obs1), torch.from_numpy(mask1), act=torch.zeros(1))
    np.testing.assert_allclose(p, ref["probs"].numpy()[0], rtol=1e-5, atol=1e-7)
    a.close()


def test_actor_c4_shape_runs():
    """BASELINE config C4: 65536 envs x 2x256 MLP, 18 actions (d = 128)."""
    from relayrl_b200 import Actor
    d, A, H, B = 128, 18, 256, 65536
    pi, v = lo.init_params(d, H, A, 1), lo.init_params(d, H, 1, 2)
    a = Actor(d, A, hidden=H, with_baseline=True, max_batch=B)
    a.set_params(0, lo.flatten(pi).numpy())
    a.set_params(1, lo.flatten(v).numpy())
    rng = np.random.default_rng(4)
    obs = rng.standard_normal((B, d)).astype(np.float32)
    mask = (rng.random((B, A)) > 0.2).astype(np.float32)
    mask[:, 0] = 1
    act, data = a.step(obs, mask, seed=4)
    idx = rng.choice(B, 512, replace=False)
    _, ref = lo.actor_step(pi, v, torch.from_numpy(obs[idx]), torch.from_numpy(mask[idx]),
                           act=torch.from_numpy(act[idx]))
    np.testing.assert_allclose(data["logp_a"][idx], ref["logp_a"].numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(data["v"][idx], ref["v"].numpy(), rtol=1e-5, atol=1e-6)
    assert (mask[np.arange(B), act] == 1).all()
    a.close()


@pytest.mark.parametrize("d,A,H,wb,B", [(20, 5, 256, True, 1000), (128, 32, 256, False, 129), (8, 1, 128, True, 128),
                                        (3, 2, 128, True, 5000)])
def test_actor_tensor_core_path_ragged_shapes(d, A, H, wb, B):
    """The tcgen05 kernel (csrc/actor_tc.cu: batch >= 128, hidden 128 or 256, obs_dim <= 128) against the oracle on shapes
    that exercise its edges: a partial last tile, obs_dim that is not a multiple of 8 or 16 (zero-padded K), the widest
    head, a single action, no value net, more tiles than SMs would take in one wave is covered by the C4 test."""
    from relayrl_b200 import Actor
    pi, v = lo.init_params(d, H, A, 11), lo.init_params(d, H, 1, 12)
    a = Actor(d, A, hidden=H, with_baseline=wb, max_batch=B)
    a.set_params(0, lo.flatten(pi).numpy())
    if wb:
        a.set_params(1, lo.flatten(v).numpy())
    rng = np.random.default_rng(d * 1000 + A)
    obs = (rng.standard_normal((B, d)) * 2).astype(np.float32)
    mask = (rng.random((B, A)) > 0.3).astype(np.float32)
    mask[:, 0] = 1
    act, data, probs, logits = a.step(obs, mask, seed=9, offset=3, want_probs=True)
    _, ref = lo.actor_step(pi, v if wb else None, torch.from_numpy(obs), torch.from_numpy(mask), act=torch.from_numpy(act))
    np.testing.assert_allclose(logits, ref["logits"].numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(probs, ref["probs"].numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(data["logp_a"], ref["logp_a"].numpy(), rtol=1e-5, atol=1e-6)
    if wb:
        np.testing.assert_allclose(data["v"], ref["v"].numpy(), rtol=1e-5, atol=1e-6)
    assert (mask[np.arange(B), act] == 1).all()
    # the same draw as the CUDA-core kernel makes from the same probabilities: both use Philox(seed, offset + row)
    small = Actor(d, A, hidden=H, with_baseline=wb, max_batch=B)
    small.set_params(0, lo.flatten(pi).numpy())
    if wb:
        small.set_params(1, lo.flatten(v).numpy())
    acts = np.concatenate([small.step(obs[i:i + 100], mask[i:i + 100], seed=9, offset=3 + i)[0] for i in range(0, B, 100)])
    assert (acts == act).mean() > 0.999      # (a draw within float32 rounding of a CDF step may fall either way)
    a.close(), small.close()
