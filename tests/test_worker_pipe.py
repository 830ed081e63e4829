"""Boundary B1 under the real worker: the UNMODIFIED reference worker
script (src/native/python/python_algorithm_reply.py) is spawned over a pipe exactly as the Rust server spawns it
(python_algorithm_request.rs:72-91), with tests/standin/relayrl_framework standing in for the PyO3 extension that
cannot be built here, and driven with the JSON lines the server would write (:235-253).

No GPU in this container, so three things are pinned here:
  1. the stand-in module + JSON encoder are accepted by the unmodified worker and the unmodified reference algorithm
     (REINFORCE): replies, the save_model hand-off, error replies;
  2. tests/worker_standin.py (the restatement the GPU box uses, where /root/reference does not exist) answers the same
     input identically;
  3. asked to load REINFORCE_B200 on a machine without a CUDA device the worker dies loudly (no CPU fallback) and never
     prints the ready marker.
The GPU leg (REINFORCE_B200 under the worker, config C1) is tests/test_gpu_c1.py."""
import json
import os

import numpy as np
import pytest
import torch

import pipe_driver as pd

needs_ref = pytest.mark.skipif(not os.path.exists(pd.REF_WORKER), reason="reference tree not present")


def _drive(script, tmp_path, tag):
    env_dir = tmp_path / tag
    env_dir.mkdir()
    model = str(env_dir / "server_model.pt")
    cfg = str(env_dir / "relayrl_config.json")
    pd.write_config(cfg, model, traj_per_epoch=4, train_vf_iters=3)
    w = pd.Worker(script, "REINFORCE", os.path.join(pd.REF_PY, "algorithms"), str(env_dir), cfg, 4, 2, 4096)
    assert w.ready, w.stderr()[-2000:]
    rng = np.random.default_rng(5)
    replies = []
    for _ in range(9):
        replies.append(w.receive_trajectory(pd.c1_trajectory(rng, hi=40)))
    replies.append(w.save_model())
    replies.append(w.request("this is not json"))
    replies.append(w.request(json.dumps({"command": "nope"})))
    # a trajectory whose action lacks a key: traj_from_json raises -> {"status": "error"}
    replies.append(w.request(json.dumps({"command": "receive_trajectory", "trajectory": {"inner": {
        "trajectory_server": "x", "max_length": 10, "actions": [{"obs": None}]}}})))
    rc = w.close()
    return replies, model, w, rc


@needs_ref
def test_unmodified_worker_with_standin_module_and_reference_algorithm(tmp_path):
    replies, model, w, rc = _drive(pd.REF_WORKER, tmp_path, "ref")
    st = [r["status"] for r in replies]
    assert st[:9] == ["not_updated"] * 3 + ["success"] + ["not_updated"] * 3 + ["success", "not_updated"], replies
    assert st[9] == "success" and os.path.exists(model)
    assert replies[10] == {"status": "error", "message": "Invalid JSON"}
    assert replies[11] == {"status": "error", "message": "Unknown command: nope"}
    assert replies[12]["status"] == "error"
    assert rc == 0
    m = torch.jit.load(model)                                   # what agents validate (agent_wrapper.rs:88-168)
    assert m.get_input_dim() == 4 and m.get_output_dim() == 2
    act, data = m.step(torch.zeros(1, 4), torch.zeros(1, 2))
    assert "logp_a" in data and "v" in data
    # the epoch table went to stdout as non-JSON lines (echoed by the Rust reader), never as {"status": ...}
    assert any("LossPi" in ln for ln in w.other_lines)


@needs_ref
def test_worker_restatement_answers_like_the_unmodified_worker(tmp_path):
    a, _, wa, rca = _drive(pd.REF_WORKER, tmp_path, "a")
    b, model_b, wb, rcb = _drive(pd.STANDIN_WORKER, tmp_path, "b")
    assert [r["status"] for r in a] == [r["status"] for r in b]
    assert a[10:12] == b[10:12]
    assert rca == rcb == 0 and os.path.exists(model_b)
    assert wa.other_lines[-1] == wb.other_lines[-1] == "[PythonWorker] Shutting down worker..."


@needs_ref
def test_b200_plugin_under_the_unmodified_worker_has_no_cpu_fallback(tmp_path):
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: covered by tests/test_gpu_c1.py")
    cfg = str(tmp_path / "relayrl_config.json")
    pd.write_config(cfg, str(tmp_path / "server_model.pt"))
    w = pd.Worker(pd.REF_WORKER, "REINFORCE_B200", pd.B200_ALGO_DIR, str(tmp_path), cfg, 4, 2, 4096,
                  startup_timeout=120)
    assert not w.ready
    assert w.close() not in (0, None)
    err = w.stderr()
    assert "CUDA" in err or "cuda" in err, err[-2000:]


def test_standin_json_round_trip_and_quirks():
    """traj_from_json(to_json) restores every getter; dtypes follow SURVEY.md Q11; a done action that fills the
    trajectory to max_length clears it (src/types/trajectory.rs:196-202)."""
    rng = np.random.default_rng(1)
    acts = pd.c1_trajectory(rng, hi=12)
    line = json.loads(pd.rf.receive_trajectory_line(acts))
    assert line["command"] == "receive_trajectory"
    t = pd.rf.RelayRLTrajectory.traj_from_json(line["trajectory"])
    back = t.get_actions()
    assert len(back) == len(acts)
    for x, y in zip(acts, back):
        assert x.get_done() == y.get_done() and x.get_rew() == y.get_rew()
        if not x.get_done():
            assert y.get_obs().dtype == np.float64 and y.get_act().dtype == np.float32 and y.get_act().shape == (1,)
            assert np.array_equal(x.get_obs(), y.get_obs()) and np.array_equal(x.get_mask(), y.get_mask())
            assert y.get_data()["v"].shape == () and y.get_data()["logp_a"].shape == (1,)
            assert np.array_equal(x.get_data()["logp_a"], y.get_data()["logp_a"])
        else:
            with pytest.raises(RuntimeError):
                y.get_obs()
            assert y.get_data() == {}
    d = pd.rf.RelayRLTrajectory(len(acts), "x")
    d._actions = list(acts)
    assert pd.rf.RelayRLTrajectory.traj_from_json(d.to_json_dict()).get_actions() == []
    cl = pd.rf.ConfigLoader("REINFORCE", "/nonexistent.json")
    hp = cl.get_algorithm_params()["REINFORCE"]
    assert hp["gamma"] == float(np.float32(0.98)) and hp["traj_per_epoch"] == 8 and hp["with_vf_baseline"] is False
