"""The oracle (oracle/learner_oracle.py + oracle/scan_ref.c) against the golden vectors that
tests/golden/make_golden.py produced from the unmodified reference.  CPU only."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import learner_oracle as lo
from oracle import scan_ref as sr


@pytest.mark.parametrize("name", ["scan_with_baseline", "scan_without_baseline"])
def test_scan_golden_bit_exact(name):
    g = load_golden(name)
    wb = bool(g["with_baseline"])
    n = len(g["rew"])
    for impl in ("py", "c"):
        adv, ret = np.zeros(n, np.float32), np.zeros(n, np.float32)
        if impl == "py":
            lo.finish_paths(g["rew"], g["val"], g["offsets"], g["last_vals"], float(g["gamma"]),
                            float(g["lam"]), wb, adv, ret)
        else:
            sr.finish_paths(g["rew"], g["val"], g["offsets"], g["last_vals"], float(g["gamma"]),
                            float(g["lam"]), wb, adv, ret)
        assert np.array_equal(adv, g["adv_raw"]), impl
        assert np.array_equal(ret, g["ret_raw"]), impl
    assert np.array_equal(lo.normalise(g["adv_raw"]), g["adv_norm"])
    out, _, _ = sr.normalise(g["adv_raw"])
    assert np.array_equal(out, g["adv_norm"])


@pytest.mark.parametrize("name", ["epoch_vf_h128", "epoch_vf_h64", "epoch_novf_h128"])
def test_segment_golden(name):
    g = load_golden(name)
    for ep in range(int(g["epochs"])):
        p = "e%d_" % ep
        keep, offs, lvs = lo.segment(g[p + "in_done"], g[p + "in_rew"])
        assert np.array_equal(offs, g[p + "offsets"])
        k2, o2, l2 = sr.segment(g[p + "in_done"], g[p + "in_rew"])
        assert np.array_equal(k2, keep) and np.array_equal(o2, offs) and np.array_equal(l2, lvs)
        assert np.array_equal(g[p + "in_obs"][keep].astype(np.float32), g[p + "get_obs"])
        assert np.array_equal(g[p + "in_act"][keep], g[p + "get_act"])


def _epoch_inputs(g, ep):
    p = "e%d_" % ep
    return {k: torch.from_numpy(g[p + "get_" + k]) for k in ("obs", "act", "mask", "ret", "adv", "logp")}


@pytest.mark.parametrize("name", ["epoch_vf_h128", "epoch_vf_h64", "epoch_novf_h128"])
def test_train_epoch_golden(name):
    """The functional restatement reproduces the reference's updated weights.  Not bit-exact:
    MKL's reduction order depends on the thread count, which moves weights by ~7e-8 absolute after
    81 Adam steps -- hence rtol 1e-5 + atol 5e-7 (the same floor applies to the GPU parity tests)."""
    g = load_golden(name)
    d, A, H, wb = int(g["d"]), int(g["A"]), int(g["H"]), bool(g["with_baseline"])
    pi = lo.unflatten(g["pi0"], d, H, A)
    v = lo.unflatten(g["v0"], d, H, 1) if wb else None
    pi_opt = v_opt = None
    for ep in range(int(g["epochs"])):
        p = "e%d_" % ep
        out = lo.train_epoch(pi, v, _epoch_inputs(g, ep), float(g["pi_lr"]), float(g["vf_lr"]),
                             int(g["iters"]), pi_opt, v_opt)
        pi, v, pi_opt, v_opt = out["pi_params"], out.get("v_params"), out["pi_opt"], out.get("v_opt")
        np.testing.assert_allclose(lo.flatten(pi).numpy(), g[p + "pi"], rtol=1e-5, atol=5e-7)
        np.testing.assert_allclose(out["LossPi"], float(g[p + "LossPi"]), rtol=1e-6)
        np.testing.assert_allclose(out["KL"], float(g[p + "KL"]), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(out["Entropy"], float(g[p + "Entropy"]), rtol=1e-6)
        assert out["DeltaLossPi"] == 0.0
        if wb:
            np.testing.assert_allclose(lo.flatten(v).numpy(), g[p + "v"], rtol=1e-5, atol=5e-7)
            np.testing.assert_allclose(out["LossV"], float(g[p + "LossV"]), rtol=1e-6)
            np.testing.assert_allclose(out["DeltaLossV"], float(g[p + "DeltaLossV"]), rtol=1e-4, atol=1e-7)


def test_oracle_learner_end_to_end_golden():
    """OracleLearner (ingest -> scan -> normalise -> train) on the raw action stream."""
    g = load_golden("epoch_vf_h128")
    L = lo.OracleLearner(4, 2, 4096, hidden=128, with_baseline=True, train_vf_iters=int(g["iters"]))
    L.pi = [p.requires_grad_(True) for p in lo.unflatten(g["pi0"], 4, 128, 2)]
    L.v = [p.requires_grad_(True) for p in lo.unflatten(g["v0"], 4, 128, 1)]
    for ep in range(int(g["epochs"])):
        p = "e%d_" % ep
        L.ingest(g[p + "in_obs"], g[p + "in_act"], g[p + "in_mask"], g[p + "in_rew"],
                 g[p + "in_val"], g[p + "in_logp"], g[p + "in_done"])
        n = L.ptr
        assert np.array_equal(L.adv[:n], g[p + "adv_raw"])   # incl. the stale tail in epoch 1 (Q7)
        assert np.array_equal(L.ret[:n], g[p + "ret_raw"])
        out = L.train()
        assert np.array_equal(out["offsets"], g[p + "offsets"])
        np.testing.assert_allclose(lo.flatten(L.pi).numpy(), g[p + "pi"], rtol=1e-5, atol=5e-7)
        np.testing.assert_allclose(lo.flatten(L.v).numpy(), g[p + "v"], rtol=1e-5, atol=5e-7)


@pytest.mark.parametrize("name", ["actor_d128_h256_a18", "actor_d4_h128_a2"])
def test_actor_golden(name):
    g = load_golden(name)
    d, A, H = int(g["d"]), int(g["A"]), int(g["H"])
    pi, v = lo.unflatten(g["pi"], d, H, A), lo.unflatten(g["v"], d, H, 1)
    act, out = lo.actor_step(pi, v, torch.from_numpy(g["obs"]), torch.from_numpy(g["mask"]),
                             act=torch.from_numpy(g["act"]))
    np.testing.assert_allclose(out["probs"].numpy(), g["probs"], rtol=1e-6, atol=1e-9)
    # Reference quirk: with a batch B > 1, step() returns logp_a as a [B, B] matrix
    # (kernel.py:33-37 unsqueezes an already [B,1] action and broadcasts): entry [i, j] is
    # logits[j, act_i].  The per-env value is its diagonal, which is what the oracle returns.
    ref_lp = g["logp_a"]
    assert ref_lp.shape == (len(g["act"]), len(g["act"]))
    np.testing.assert_allclose(out["logp_a"].numpy(), np.diagonal(ref_lp), rtol=1e-6)
    np.testing.assert_allclose(ref_lp, g["logits"][:, g["act"].reshape(-1)].T, rtol=0, atol=0)
    np.testing.assert_allclose(out["v"].numpy(), g["vals"], rtol=1e-5, atol=5e-7)
    # a masked action is never the sampled one
    assert (g["mask"][np.arange(len(g["act"])), g["act"].reshape(-1)] == 1).all()


def test_c_scan_matches_python_random():
    rng = np.random.default_rng(5)
    lens = np.concatenate([rng.integers(1, 60, 200), [3000, 1, 1, 4097]])
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n = int(offs[-1])
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, len(lens)).astype(np.float32)
    for wb in (True, False):
        a1, r1 = np.zeros(n, np.float32), np.zeros(n, np.float32)
        lo.finish_paths(rew, val, offs, lvs, 0.99, 0.95, wb, a1, r1)
        a2, r2 = sr.finish_paths(rew, val, offs, lvs, 0.99, 0.95, wb)
        assert np.array_equal(a1, a2) and np.array_equal(r1, r2)
