"""GPU parity of the fused learner (K4-K7) through the C ABI and through the drop-in algorithm
class, against golden vectors produced by the unmodified reference and against the oracle.

Tolerance: rel 1e-5 (north star) with the absolute floor the reference itself shows between
thread counts (tests/test_oracle.py: ~7e-8 on weights after 81 Adam steps; 5e-7 used)."""
import os
import sys

import numpy as np
import pytest
import torch

from conftest import PKG, load_golden
from oracle import learner_oracle as lo

pytestmark = pytest.mark.gpu

W_RTOL, W_ATOL = 1e-5, 5e-7


def assert_weights_close(actual, desired, lr, learner=None, which=None, excused=None):
    """rel 1e-5 + abs 5e-7 on >= 99.9 % of the parameters, and every parameter within 2 % of one Adam step (0.02 lr).
    The few that miss the strict bound are counted, printed, and must be the ill-conditioned ones: Adam's update
    lr * m_hat / (sqrt(v_hat) + eps) stops being scale-invariant where sqrt(v_hat) ~ eps = 1e-8, so there a 1e-7 relative
    wobble of the gradient (float32 summation order) moves the weight by a visible fraction of lr.  Measured on B200
    (tools/diag_outliers.py, golden epochs): 0-2 such parameters of 17 k, sqrt(v_hat) between 3e-10 and 5e-8 against a
    largest sqrt(v_hat) of 2e-2 .. 9e-2, error <= 0.009 lr; the CUDA-core kernel: one weight at sqrt(v_hat) = 9.7e-5,
    error 0.0008 lr after 80 steps.  With `learner` given, a NEW outlier must be explained by its conditioning: float32
    summation noise of 3e-7 of the largest gradient entry per step moves a weight by at most
    lr * 3e-7 * max sqrt(v_hat) / sqrt(v_hat_i) per step, 80 steps an epoch; `excused` carries outliers of earlier
    epochs (the error stays, v_hat grows)."""
    actual, desired = np.asarray(actual), np.asarray(desired)
    err = np.abs(actual - desired)
    out = set(np.flatnonzero(err > W_ATOL + W_RTOL * np.abs(desired)).tolist())
    print("weights: %d of %d outside rel 1e-5 + abs 5e-7; largest error %.3g lr" % (len(out), len(actual), err.max() / lr))
    if out and learner is not None:   # what the offenders look like (shown with pytest -s or on failure)
        m_, v_, step_ = learner.get_opt_state(which)
        for i in sorted(out, key=lambda k: -err[k])[:8]:
            print("   parameter %d: got %.9g want %.9g err %.3g lr, m %.3g v %.3g (step %d)" % (
                i, actual[i], desired[i], err[i] / lr, m_[i], v_[i], step_))
    assert len(out) <= max(2, len(actual) // 1000), "%d parameters outside rel 1e-5 + abs 5e-7" % len(out)
    if learner is None:
        np.testing.assert_allclose(actual, desired, rtol=W_RTOL, atol=0.02 * lr)
    else:
        # a weight whose gradient is far below the layer's largest (or below Adam's eps = 1e-8) is ill-conditioned: the
        # float32 summation noise of 3e-7 of the largest gradient entry per step moves it by
        # lr * 3e-7 * max sqrt(v_hat) / (sqrt(v_hat_i) + eps) per step, at most 80 steps an epoch
        m, v, step = learner.get_opt_state(which)
        sv = np.sqrt(v / (1.0 - 0.999 ** step))
        cond = min(step, 80) * 3e-7 * sv.max() / (sv + 1e-8)
        over = err > W_RTOL * np.abs(desired) + lr * np.maximum(0.02, cond)
        if excused:        # outliers of earlier epochs keep their offset while v_hat grows: they stay within a tenth of a step
            ex = np.fromiter(excused, dtype=np.int64)
            assert (err[ex] <= 0.1 * lr).all(), "an earlier outlier grew beyond 0.1 lr: %s" % (err[ex] / lr)
            over[ex] = False
        assert not over.any(), "parameter %d is off by %.3g lr with sqrt(v_hat) = %.3g (max %.3g): more than 0.02 lr and " \
            "more than its conditioning explains (%.3g lr)" % (int(np.argmax(over)), err[np.argmax(over)] / lr,
                                                              sv[np.argmax(over)], sv.max(), cond[np.argmax(over)])
        for i in sorted(out - (excused or set())):
            assert err[i] / lr <= max(cond[i], 1e-3), "parameter %d is a new outlier at %.3g lr, sqrt(v_hat) = %.3g (max %.3g), " \
                "conditioning explains %.3g lr" % (i, err[i] / lr, sv[i], sv.max(), cond[i])
    return out | (excused or set())


def _mk(g, iters=None):
    from relayrl_b200 import Learner
    d, A, H, wb = int(g["d"]), int(g["A"]), int(g["H"]), bool(g["with_baseline"])
    L = Learner(d, A, 8192, hidden=H, with_baseline=wb, gamma=float(g["gamma"]), lam=float(g["lam"]),
                pi_lr=float(g["pi_lr"]), vf_lr=float(g["vf_lr"]),
                train_vf_iters=int(g["iters"]) if iters is None else iters)
    L.set_params(0, g["pi0"])
    if wb:
        L.set_params(1, g["v0"])
    return L, d, A, H, wb


def _ingest_epoch(L, g, ep):
    p = "e%d_" % ep
    L.ingest(g[p + "in_obs"], g[p + "in_act"], g[p + "in_mask"], g[p + "in_rew"], g[p + "in_val"],
             g[p + "in_logp"], g[p + "in_done"])


@pytest.mark.parametrize("name", ["epoch_vf_h128", "epoch_vf_h64", "epoch_novf_h128"])
def test_epoch_golden(name):
    g = load_golden(name)
    L, d, A, H, wb = _mk(g)
    seen_pi, seen_v = set(), set()
    for ep in range(int(g["epochs"])):
        p = "e%d_" % ep
        _ingest_epoch(L, g, ep)
        n, _ = L.finish()
        assert np.array_equal(L.offsets(), g[p + "offsets"])
        np.testing.assert_allclose(L.returns(n), g[p + "ret_raw"], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(L.advantages(n, False), g[p + "adv_raw"], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(L.advantages(n, True), g[p + "get_adv"], rtol=1e-5, atol=1e-5)
        st = L.train_epoch()
        assert st["n_transitions"] == n
        np.testing.assert_allclose(st["loss_pi"], float(g[p + "LossPi"]), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(st["kl"], float(g[p + "KL"]), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(st["entropy"], float(g[p + "Entropy"]), rtol=1e-5, atol=1e-7)
        assert st["delta_loss_pi"] == 0.0
        seen_pi = assert_weights_close(L.get_params(0), g[p + "pi"], float(g["pi_lr"]), L, 0, seen_pi)
        if wb:
            np.testing.assert_allclose(st["loss_v"], float(g[p + "LossV"]), rtol=1e-5)
            np.testing.assert_allclose(st["delta_loss_v"], float(g[p + "DeltaLossV"]), rtol=1e-4, atol=1e-6)
            seen_v = assert_weights_close(L.get_params(1), g[p + "v"], float(g["vf_lr"]), L, 1, seen_v)
    L.close()


@pytest.mark.parametrize("name", ["epoch_vf_h128", "epoch_vf_h64", "epoch_novf_h128"])
def test_gradients_vs_oracle_autograd(name):
    g = load_golden(name)
    L, d, A, H, wb = _mk(g)
    _ingest_epoch(L, g, 0)
    L.finish()
    data = {k: torch.from_numpy(g["e0_get_" + k]) for k in ("obs", "act", "mask", "ret", "adv", "logp")}
    ref = lo.train_epoch(lo.unflatten(g["pi0"], d, H, A), lo.unflatten(g["v0"], d, H, 1) if wb else None, data,
                         float(g["pi_lr"]), float(g["vf_lr"]), 1, return_grads=True)
    gp, (lpi, kl, ent) = L.eval_grad(0)
    scale = np.abs(ref["pi_grad"].numpy()).max()
    np.testing.assert_allclose(gp, ref["pi_grad"].numpy(), rtol=1e-5, atol=2e-6 * scale)
    np.testing.assert_allclose(lpi, ref["LossPi"], rtol=1e-5, atol=1e-7)
    if wb:
        gv, (lv, _, _) = L.eval_grad(1)
        scale = np.abs(ref["v_grad_first"].numpy()).max()
        np.testing.assert_allclose(gv, ref["v_grad_first"].numpy(), rtol=1e-5, atol=2e-6 * scale)
        np.testing.assert_allclose(lv, ref["LossVOld"], rtol=1e-5)
    L.close()


def test_larger_batch_vs_oracle_multi_tile():
    """~70k transitions (more tiles than SMs, several tiles per CTA), d=8, A=4, masks, 5 value steps.

    At this size the reference's own float32 arithmetic is the limiting factor: run in float64, the
    same reference code moves 1.3 % of the value weights by more than rel 1e-5 + abs 5e-7 (up to
    1.7e-5) away from its float32 result.  So the updated weights are judged against the float64 run:
    the CUDA path must be at least as close to it as the float32 reference is (the tensor-core path
    is closer: 99.96 % strict, max 3.8e-6 measured), and within 0.1 lr of the float32 reference."""
    rng = np.random.default_rng(7)
    d, A, H, iters = 8, 4, 128, 5
    lens = rng.integers(5, 400, 350)
    done = np.concatenate([np.r_[np.zeros(k, np.uint8), np.uint8(1)] for k in lens])
    n = len(done)
    obs = rng.standard_normal((n, d)).astype(np.float32)
    act = rng.integers(0, A, n).astype(np.float32)
    mask = np.ones((n, A), np.float32)
    off = rng.integers(0, A, n)
    sel = (rng.random(n) < 0.2) & (off != act)
    mask[np.flatnonzero(sel), off[sel]] = 0
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    logp = rng.standard_normal(n).astype(np.float32)
    O = lo.OracleLearner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, seed=3)
    pi0, v0 = [p.detach().clone() for p in O.pi], [p.detach().clone() for p in O.v]
    from relayrl_b200 import Learner
    L = Learner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters)
    L.set_params(0, lo.flatten(pi0).numpy())
    L.set_params(1, lo.flatten(v0).numpy())
    O.ingest(obs, act, mask, rew, val, logp, done)
    L.ingest(obs, act, mask, rew, val, logp, done)
    data, _ = O.get()
    ref = lo.train_epoch(pi0, v0, data, O.pi_lr, O.vf_lr, iters)
    ref64 = lo.train_epoch([p.double() for p in pi0], [p.double() for p in v0], {k: t.double() for k, t in data.items()},
                           O.pi_lr, O.vf_lr, iters)
    st = L.train_epoch()
    np.testing.assert_allclose(st["loss_pi"], ref["LossPi"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(st["loss_v"], ref["LossV"], rtol=1e-5)
    np.testing.assert_allclose(st["loss_v_old"], ref["LossVOld"], rtol=1e-5)
    np.testing.assert_allclose(st["entropy"], ref["Entropy"], rtol=1e-5)

    def strict(a, b):
        e = np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))
        return (e <= W_ATOL + W_RTOL * np.abs(b)).mean(), e.max()

    for which, key, lr in ((0, "pi_params", 3e-4), (1, "v_params", 1e-3)):
        w32, w64, wg = lo.flatten(ref[key]).numpy(), lo.flatten(ref64[key]).numpy(), L.get_params(which)
        f_ref, m_ref = strict(w32, w64)      # how far the float32 reference is from the float64 truth
        f_gpu, m_gpu = strict(wg, w64)
        assert f_gpu >= min(f_ref, 0.999) - 1e-3, (key, f_gpu, f_ref)
        assert m_gpu <= max(1.5 * m_ref, 1e-6), (key, m_gpu, m_ref)
        np.testing.assert_allclose(wg, w32, rtol=W_RTOL, atol=0.1 * lr)
    m, v, step = L.get_opt_state(1)
    assert step == iters
    sd = ref64["v_opt"].state_dict()["state"]
    m_ref = torch.cat([sd[i]["exp_avg"].reshape(-1) for i in range(6)]).numpy()
    # exp_avg after 5 steps is a trajectory quantity (weights that differ by 1e-5 give different later
    # gradients): agreement at the 1e-3 level checks the optimiser-state tap, not the arithmetic
    np.testing.assert_allclose(m, m_ref, rtol=1e-3, atol=1e-3 * np.abs(m_ref).max())
    L.close()


def test_log_softmax_mode_gradient():
    """Opt-in textbook REINFORCE (logp = log_softmax); checked against torch autograd."""
    from relayrl_b200 import Learner
    rng = np.random.default_rng(11)
    d, A, H, n = 4, 3, 64, 777
    obs = rng.standard_normal((n, d)).astype(np.float32)
    act = rng.integers(0, A, n).astype(np.float32)
    mask = np.ones((n, A), np.float32)
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    done = np.zeros(n + 1, np.uint8)
    done[-1] = 1
    pad = lambda a: np.concatenate([a, np.zeros((1,) + a.shape[1:], a.dtype)])
    L = Learner(d, A, n + 8, hidden=H, with_baseline=False, logp_mode=1)
    pi = lo.init_params(d, H, A, 5)
    L.set_params(0, lo.flatten(pi).numpy())
    L.ingest(pad(obs), pad(act), np.concatenate([mask, np.ones((1, A), np.float32)]), pad(rew), None, pad(rew), done)
    L.finish()
    adv = torch.from_numpy(L.advantages(n, True).copy())
    g, (loss, _, _) = L.eval_grad(0)
    ps = [p.clone().requires_grad_(True) for p in pi]
    logits = lo.mlp3(ps, torch.from_numpy(obs)) + (torch.from_numpy(mask) - 1) * 1e8
    lp = torch.log_softmax(logits, -1).gather(-1, torch.from_numpy(act).long()[:, None]).squeeze(-1)
    ref_loss = -(lp * adv).mean()
    ref_loss.backward()
    ref_g = torch.cat([p.grad.reshape(-1) for p in ps]).numpy()
    np.testing.assert_allclose(loss, ref_loss.item(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(g, ref_g, rtol=1e-4, atol=1e-6 * np.abs(ref_g).max())
    L.close()


def test_plugin_class_end_to_end(tmp_path):
    """The drop-in algorithm class fed duck-typed trajectories like the reference worker does
    (python_algorithm_reply.py:99-104), incl. the TorchScript hand-off contract
    (agent_wrapper.rs:88-168)."""
    sys.path.insert(0, os.path.join(PKG, "algorithms"))
    import importlib
    mod = importlib.import_module("REINFORCE_B200.REINFORCE_B200")
    from oracle.ref_bridge import DuckAction, DuckTrajectory
    g = load_golden("epoch_vf_h128")
    cfg = tmp_path / "relayrl_config.json"
    model = tmp_path / "server_model.pt"
    cfg.write_text('{"algorithms": {"REINFORCE": {"with_vf_baseline": true, "traj_per_epoch": 1, '
                   '"train_vf_iters": 80, "gamma": 0.98, "lam": 0.97, "pi_lr": 0.0003, "vf_lr": 0.001}}, '
                   '"model_paths": {"server_model": "%s"}}' % model)
    algo = mod.REINFORCE_B200(env_dir=str(tmp_path), config_path=str(cfg), obs_dim=4, act_dim=2, buf_size=8192)
    algo.learner.set_params(0, g["pi0"])
    algo.learner.set_params(1, g["v0"])
    for ep in range(int(g["epochs"])):
        p = "e%d_" % ep
        acts = []
        for i in range(len(g[p + "in_done"])):
            if g[p + "in_done"][i]:
                acts.append(DuckAction(rew=float(g[p + "in_rew"][i]), done=True))
            else:
                acts.append(DuckAction(g[p + "in_obs"][i], np.array([g[p + "in_act"][i]], np.float32),
                                       g[p + "in_mask"][i], float(g[p + "in_rew"][i]),
                                       {"logp_a": np.array([g[p + "in_logp"][i]], np.float32),
                                        "v": np.float32(g[p + "in_val"][i])}, False))
        assert algo.receive_trajectory(DuckTrajectory(acts)) is True
        np.testing.assert_allclose(algo.last_stats["loss_pi"], float(g[p + "LossPi"]), rtol=1e-5, atol=1e-7)
        assert_weights_close(algo.learner.get_params(1), g[p + "v"], 1e-3)
        algo.save()  # every epoch: the ScriptModule is scripted once and must still carry the fresh weights
        sd = torch.jit.load(str(model)).state_dict()
        assert np.array_equal(sd["policy.pi_network.0.weight"].numpy().reshape(-1), algo.learner.get_params(0)[:4 * 128])
        assert np.array_equal(sd["baseline.v_network.4.bias"].numpy().reshape(-1), algo.learner.get_params(1)[-1:])
    algo.save()
    m = torch.jit.load(str(model))
    assert m.get_input_dim() == 4 and m.get_output_dim() == 2
    act, data = m.step(torch.zeros(1, 4), torch.zeros(1, 2))
    assert isinstance(act, torch.Tensor) and len(data) > 0 and "logp_a" in data and "v" in data
    assert tuple(data["logp_a"].shape) == (1, 1)               # [1, d] in: the reference's unsqueeze of a [1, 1] action
    act1, data1 = m.step(torch.zeros(4), torch.zeros(2))       # the 1-D observation agents send (o3_action.rs:256-265)
    assert tuple(act1.shape) == (1,) and tuple(data1["logp_a"].shape) == (1,)   # as kernel.py:33-37 returns it
    sd = m.state_dict()
    np.testing.assert_allclose(sd["policy.pi_network.2.weight"].numpy().reshape(-1),
                               algo.learner.get_params(0)[4 * 128 + 128:4 * 128 + 128 + 128 * 128], rtol=0, atol=0)
    assert os.path.exists(os.path.join(str(tmp_path), "logs"))


def test_wire_action_ingest_safetensors():
    """rrl_ingest_wire_action on blobs laid out as safetensors 0.5 writes them (action.rs:345,372-381)."""
    import json
    import struct

    def blob(arr, name="tensor"):
        arr = np.ascontiguousarray(arr)
        dt = {np.dtype("float64"): "F64", np.dtype("float32"): "F32"}[arr.dtype]
        hdr = json.dumps({name: {"dtype": dt, "shape": list(arr.shape), "data_offsets": [0, arr.nbytes]}},
                         separators=(",", ":")).encode()
        hdr += b" " * ((8 - len(hdr) % 8) % 8)
        return struct.pack("<Q", len(hdr)) + hdr + arr.tobytes()

    from relayrl_b200 import Learner
    g = load_golden("epoch_vf_h128")
    L, d, A, H, wb = _mk(g, iters=1)
    p = "e0_"
    n_in = len(g[p + "in_done"])
    for i in range(n_in):
        if g[p + "in_done"][i]:
            L.ingest_wire_action(None, None, None, float(g[p + "in_rew"][i]), None, None, True)
        else:
            L.ingest_wire_action(blob(g[p + "in_obs"][i]), blob(np.array([g[p + "in_act"][i]], np.float32)),
                                 blob(g[p + "in_mask"][i]), float(g[p + "in_rew"][i]),
                                 blob(np.array(g[p + "in_val"][i], np.float32)),
                                 blob(np.array([g[p + "in_logp"][i]], np.float32)), False)
    n, _ = L.finish()
    assert np.array_equal(L.offsets(), g[p + "offsets"])
    np.testing.assert_allclose(L.returns(n), g[p + "ret_raw"], rtol=1e-6, atol=1e-6)
    from relayrl_b200 import RelayRLB200Error
    with pytest.raises(RelayRLB200Error) as e:
        L.ingest_wire_action(b"\x05\x00", b"", b"", 0.0, b"", b"", False)
    assert e.value.status == -6
    L.close()


def test_grpc_trajectory_ingest_matches_golden_epoch():
    """rrl_ingest_grpc_trajectory: the golden epoch's inputs sent as serialized proto Trajectory messages (one per
    episode, terminal marker last, as agent_grpc.rs flushes them) give the reference's offsets and returns."""
    from test_wire_grpc import _blob, _messages, _tensor_json
    g = load_golden("epoch_vf_h128")
    L, d, A, H, wb = _mk(g, iters=1)
    p = "e0_"
    _, Trajectory = _messages()
    t, sent, n_in = Trajectory(), 0, len(g[p + "in_done"])
    for i in range(n_in):
        if g[p + "in_done"][i]:
            t.actions.add(reward=float(g[p + "in_rew"][i]), done=True)
            sent += L.ingest_grpc_trajectory(t.SerializeToString())
            t = Trajectory()
        else:
            a = t.actions.add(obs=_blob(g[p + "in_obs"][i]), action=_blob(np.array([g[p + "in_act"][i]], np.float32)),
                              mask=_blob(g[p + "in_mask"][i]), reward=float(g[p + "in_rew"][i]))
            a.data["logp_a"] = _tensor_json(np.array([g[p + "in_logp"][i]], np.float32), "Float")
            a.data["v"] = _tensor_json(np.array(g[p + "in_val"][i], np.float32), "Float")
    if len(t.actions):
        sent += L.ingest_grpc_trajectory(t.SerializeToString())
    assert sent == n_in
    n, _ = L.finish()
    assert np.array_equal(L.offsets(), g[p + "offsets"])
    np.testing.assert_allclose(L.returns(n), g[p + "ret_raw"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(L.advantages(n, normalised=False), g[p + "adv_raw"], rtol=1e-6, atol=1e-6)
    from relayrl_b200 import RelayRLB200Error
    with pytest.raises(RelayRLB200Error) as e:
        L.ingest_grpc_trajectory(b"\x0a\x7f\x01")
    assert e.value.status == -6
    L.close()


@pytest.mark.parametrize("n,d,A", [(64 * 148 * 17 + 5, 4, 2), (64 * 148 * 33 + 63, 8, 4), (130, 1, 1 + 2), (63, 3, 2)])
def test_tc_gradients_many_tiles_vs_float64_autograd(n, d, A):
    """The tensor-core kernel's pipeline at scale: more than 16 (and 32) tiles per CTA, so the f32 window spills
    into the float64 partials inside the loop, ragged last tiles, and the deferred tail of the last tile.
    Gradients and losses of one policy and one value pass against the oracle run in float64 (autograd)."""
    rng = np.random.default_rng(n)
    H = 128
    lens = []
    while sum(lens) < n:
        lens.append(int(rng.integers(1, 3000)))
    lens[-1] -= sum(lens) - n
    done = np.concatenate([np.r_[np.zeros(k, np.uint8), np.uint8(1)] for k in lens])
    m = len(done)
    obs = rng.standard_normal((m, d)).astype(np.float32)
    act = rng.integers(0, A, m).astype(np.float32)
    mask = np.ones((m, A), np.float32)
    if A > 1:
        off = rng.integers(0, A, m)
        sel = (rng.random(m) < 0.2) & (off != act)
        mask[np.flatnonzero(sel), off[sel]] = 0
    rew = rng.uniform(-1, 1, m).astype(np.float32)
    val = rng.standard_normal(m).astype(np.float32)
    logp = rng.standard_normal(m).astype(np.float32)
    O = lo.OracleLearner(d, A, m + 1, hidden=H, with_baseline=True, train_vf_iters=1, seed=5)
    pi0, v0 = [p.detach().clone() for p in O.pi], [p.detach().clone() for p in O.v]
    from relayrl_b200 import Learner
    L = Learner(d, A, m + 1, hidden=H, with_baseline=True, train_vf_iters=1, flags=1)   # the tensor-core kernel at any size
    L.set_params(0, lo.flatten(pi0).numpy())
    L.set_params(1, lo.flatten(v0).numpy())
    O.ingest(obs, act, mask, rew, val, logp, done)
    L.ingest(obs, act, mask, rew, val, logp, done)
    assert L.finish()[0] == n
    data, _ = O.get()
    ref = lo.train_epoch([p.double() for p in pi0], [p.double() for p in v0], {k: t.double() for k, t in data.items()},
                         O.pi_lr, O.vf_lr, 1, return_grads=True)
    ref32 = lo.train_epoch(pi0, v0, data, O.pi_lr, O.vf_lr, 1, return_grads=True)   # the reference's own arithmetic
    for which, gkey, lkey in ((0, "pi_grad", "LossPi"), (1, "v_grad_first", "LossVOld")):
        g, (loss, _, _) = L.eval_grad(which)
        want, w32 = ref[gkey].numpy(), ref32[gkey].numpy().astype(np.float64)
        # judged against float64: at least as close as twice the float32 reference's own distance, or 3e-6 of the
        # largest entry (fp32 accumulation in windows of 1024 rows, float64 above), whichever is larger
        e_gpu, e_ref = np.abs(g - want).max(), np.abs(w32 - want).max()
        assert e_gpu <= max(2.0 * e_ref, 3e-6 * np.abs(want).max()), (gkey, e_gpu, e_ref, np.abs(want).max())
        np.testing.assert_allclose(g, want, rtol=1e-5, atol=max(2.0 * e_ref, 3e-6 * np.abs(want).max()), err_msg=gkey)
        np.testing.assert_allclose(loss, float(ref[lkey]), rtol=1e-5, atol=1e-7, err_msg=lkey)
    L.close()


def test_pickle_ingest_matches_golden_epoch_and_is_all_or_nothing():
    """rrl_ingest_pickle: the golden epoch's inputs sent as ZMQ trajectory messages (one pickled Vec<RelayRLAction> per
    episode, terminal marker last: src/types/trajectory.rs:183-193) give the reference's offsets, returns and
    advantages; a message with a bad action in the middle leaves the buffer exactly as it was."""
    import pickle
    from test_wire_pickle import _graph
    import pipe_driver as pd
    from relayrl_b200 import RelayRLB200Error
    g = load_golden("epoch_vf_h128")
    L, d, A, H, wb = _mk(g, iters=1)
    p = "e0_"
    acts, sent, n_in = [], 0, len(g[p + "in_done"])
    for i in range(n_in):
        if g[p + "in_done"][i]:
            acts.append(pd.rf.RelayRLAction(None, None, None, float(g[p + "in_rew"][i]), None, True, False))
            if sent == 0:   # first message: also try a corrupted copy -- nothing of it may stay in the buffer
                bad = _graph(acts)
                bad[len(bad) // 2]["mask"]["data"] = bad[len(bad) // 2]["mask"]["data"][:-3]
                with pytest.raises(RelayRLB200Error) as e:
                    L.ingest_pickle(pickle.dumps(bad, protocol=3))
                assert e.value.status == -6
                assert np.array_equal(L.offsets(), [0])
            sent += L.ingest_pickle(pickle.dumps(_graph(acts, bytes_as="bytes" if sent % 2 else "list"), protocol=3))
            acts = []
        else:
            acts.append(pd.rf.RelayRLAction(np.asarray(g[p + "in_obs"][i], np.float64), np.array([g[p + "in_act"][i]], np.float32),
                                            np.asarray(g[p + "in_mask"][i], np.float64), float(g[p + "in_rew"][i]),
                                            {"logp_a": np.array([g[p + "in_logp"][i]], np.float32),
                                             "v": np.array(g[p + "in_val"][i], np.float32)}, False, False))
    if acts:
        sent += L.ingest_pickle(pickle.dumps(_graph(acts), protocol=3))
    assert sent == n_in
    n, _ = L.finish()
    assert np.array_equal(L.offsets(), g[p + "offsets"])
    np.testing.assert_allclose(L.returns(n), g[p + "ret_raw"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(L.advantages(n, normalised=False), g[p + "adv_raw"], rtol=1e-6, atol=1e-6)
    L.close()


@pytest.mark.parametrize("H,n,iters", [(128, 2044, 80), (64, 2044, 80), (128, 40, 5), (128, 64 * 148, 7), (64, 64 * 100 + 1, 3)])
def test_fused_value_iterations_bit_identical(H, n, iters):
    """Small batches on one rank run all value steps of an epoch in ONE cooperative launch (csrc/mlp.cu,
    mlp_value_iters_kernel): the same kernels' arithmetic in the same order, so weights, Adam state and losses must equal
    the one-launch-per-step path bit for bit -- over two epochs (the second starts from step > 0 and a stale tail)."""
    from relayrl_b200 import Learner, _lib
    d, A = 4, 2
    rng = np.random.default_rng(H + n)
    out = []
    eps = []
    for _ in range(2):
        lens = rng.integers(1, 300, 400)
        lens = lens[np.cumsum(lens) <= n]
        if lens.sum() < n:
            lens = np.append(lens, n - lens.sum())
        done = np.concatenate([np.r_[np.zeros(k, np.uint8), np.uint8(1)] for k in lens])
        m = len(done)
        eps.append((rng.standard_normal((m, d)), rng.integers(0, A, m).astype(np.float32), np.ones((m, A)),
                    rng.uniform(-1, 1, m).astype(np.float32), rng.standard_normal(m).astype(np.float32),
                    rng.standard_normal(m).astype(np.float32), done))
    for flags in (0, _lib.RRL_FLAG_NO_FUSED_ITERS):
        L = Learner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, flags=flags)
        L.set_params(0, lo.flatten(lo.init_params(d, H, A, 1)).numpy())
        L.set_params(1, lo.flatten(lo.init_params(d, H, 1, 2)).numpy())
        res = []
        for e in eps:
            L.ingest(*e)
            st = L.train_epoch()
            res.append((st["loss_v"], st["loss_v_old"], st["loss_pi"], st["kernel_launches"]))
        res.append((L.get_params(0), L.get_params(1)) + tuple(L.get_opt_state(1)))
        out.append(res)
        L.close()
    fused, plain = out
    for a, b in zip(fused[:2], plain[:2]):
        assert a[:3] == b[:3], (a, b)
        assert a[3] < b[3] and b[3] - a[3] == 3 * iters - 1          # one launch instead of (pass + reduce + Adam) x iters
    for a, b in zip(fused[2], plain[2]):
        assert np.array_equal(np.asarray(a), np.asarray(b))


def test_host_alloc_pinned_view_ingest():
    """rrl_host_alloc / rrl_host_free: pinned host memory for the caller's packed columns (the e2e leg of bench.py uses
    torch's pinned tensors; a Rust or C host would use these).  A view over such memory ingests exactly like numpy arrays,
    with f64 obs/mask on the wire (Q11)."""
    import ctypes as C
    from relayrl_b200 import Learner, _lib, host_alloc, host_free
    from relayrl_b200._lib import BatchView
    d, A, n = 4, 2, 3000
    rng = np.random.default_rng(7)
    lens = [700, 1, 1299, 1000]
    done = np.concatenate([np.r_[np.zeros(k, np.uint8), np.uint8(1)] for k in lens])
    m = len(done)
    cols = dict(obs=rng.standard_normal((m, d)), act=rng.integers(0, A, m).astype(np.float32), mask=np.ones((m, A)),
                rew=rng.uniform(-1, 1, m).astype(np.float32), val=rng.standard_normal(m).astype(np.float32),
                logp=rng.standard_normal(m).astype(np.float32), done=done)
    ptrs = {}
    for k, arr in cols.items():
        arr = np.ascontiguousarray(arr)
        p = host_alloc(arr.nbytes)
        C.memmove(p, arr.ctypes.data, arr.nbytes)
        ptrs[k] = p
    view = BatchView(ptrs["obs"].value, ptrs["act"].value, ptrs["mask"].value, ptrs["rew"].value, ptrs["val"].value,
                     ptrs["logp"].value, ptrs["done"].value, m, _lib.RRL_F64, _lib.RRL_F64)
    a, b = Learner(d, A, n + 1, hidden=64, with_baseline=True), Learner(d, A, n + 1, hidden=64, with_baseline=True)
    a.ingest_raw(view)
    b.ingest(cols["obs"], cols["act"], cols["mask"], cols["rew"], cols["val"], cols["logp"], cols["done"])
    (na, ea), (nb, eb) = a.finish(), b.finish()
    assert (na, ea) == (nb, eb) == (n, len(lens))
    assert np.array_equal(a.offsets(), b.offsets())
    assert np.array_equal(a.returns(n), b.returns(n)) and np.array_equal(a.advantages(n, False), b.advantages(n, False))
    for p in ptrs.values():
        host_free(p)
    a.close(), b.close()
