"""TEST INFRASTRUCTURE -- plays the Rust server's side of the worker pipe
(src/network/server/python_subprocesses/python_algorithm_request.rs:72-91 spawn, :165-196 stdout reader,
:235-262 request/reply pairing) and builds config C1's synthetic trajectories."""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, ".."))
STANDIN = os.path.join(HERE, "standin")
REF_PY = os.path.join(os.environ.get("RELAYRL_REFERENCE", "/root/reference"), "relayrl_framework", "src", "native",
                      "python")
REF_WORKER = os.path.join(REF_PY, "python_algorithm_reply.py")
STANDIN_WORKER = os.path.join(HERE, "worker_standin.py")
B200_ALGO_DIR = os.path.join(ROOT, "relayrl-prototype_b200", "algorithms")

sys.path.insert(0, STANDIN)
import relayrl_framework as rf  # noqa: E402  (the stand-in)


def write_config(path, model_path, **reinforce):
    hp = dict(discrete=True, with_vf_baseline=True, seed=1, traj_per_epoch=8, gamma=0.98, lam=0.97, pi_lr=3e-4,
              vf_lr=1e-3, train_vf_iters=80)
    hp.update(reinforce)
    with open(path, "w") as f:
        json.dump({"algorithms": {"REINFORCE": hp}, "model_paths": {"server_model": model_path,
                                                                     "client_model": model_path + ".client"}}, f)


class Worker:
    """One child `python -u <worker> --algorithm_name ... --hyperparams <json>` (python_algorithm_request.rs:83-91;
    arguments as training_zmq.rs:203-219 builds them: hyper-parameters arrive as strings)."""

    def __init__(self, script, algorithm_name, algorithm_dir, env_dir, config_path, obs_dim, act_dim, buf_size,
                 extra_hp=None, startup_timeout=180):
        hp = {"env_dir": env_dir, "obs_dim": str(obs_dim), "act_dim": str(act_dim), "buf_size": str(buf_size)}
        hp.update(extra_hp or {})
        env = dict(os.environ)
        # the stand-in `relayrl_framework` first; the reference's own tree (for _common / utils.logger) when it exists
        env["PYTHONPATH"] = os.pathsep.join([STANDIN, REF_PY, env.get("PYTHONPATH", "")])
        self.stderr_file = tempfile.NamedTemporaryFile("w+", suffix=".worker.err", delete=False)
        self.p = subprocess.Popen([sys.executable, "-u", script, "--algorithm_name", algorithm_name, "--env_dir", env_dir,
                                   "--resolved_algorithm_dir", algorithm_dir, "--config_path", config_path,
                                   "--hyperparams", json.dumps(hp)], stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                                  stderr=self.stderr_file, env=env, bufsize=0)
        self.other_lines = []
        self.ready = False
        self._buf = b""
        import time
        t0 = time.time()
        while time.time() - t0 < startup_timeout:
            line = self._readline(1.0)
            if line is None:
                if self.p.poll() is not None:
                    break
                continue
            if "[algorithm_pyscript_true]" in line:    # python_algorithm_request.rs:186-192
                self.ready = True
                break
            self.other_lines.append(line)

    def _readline(self, timeout):
        """One stdout line (without the newline), None on timeout or EOF; raw fd + own buffer so that select() and
        buffered reads cannot disagree."""
        import select
        fd = self.p.stdout.fileno()
        while b"\n" not in self._buf:
            r, _, _ = select.select([fd], [], [], timeout)
            if not r:
                return None
            chunk = os.read(fd, 1 << 16)
            if not chunk:
                if self._buf:
                    line, self._buf = self._buf, b""
                    return line.decode(errors="replace")
                return None
            self._buf += chunk
        line, self._buf = self._buf.split(b"\n", 1)
        return line.decode(errors="replace")

    def stderr(self):
        self.stderr_file.flush()
        with open(self.stderr_file.name) as f:
            return f.read()

    def request(self, line):
        """Write one line, read stdout until a line parses as {"status": ...} (python_algorithm_request.rs:169-181);
        everything else is what the Rust reader would just echo."""
        self.p.stdin.write((line + "\n").encode())
        self.p.stdin.flush()
        while True:
            out = self._readline(600.0)
            if out is None:
                raise RuntimeError("worker closed its stdout or timed out; stderr:\n" + self.stderr()[-3000:])
            try:
                msg = json.loads(out)
                if isinstance(msg, dict) and "status" in msg:
                    return msg
            except json.JSONDecodeError:
                pass
            self.other_lines.append(out)

    def receive_trajectory(self, actions):
        return self.request(rf.receive_trajectory_line(actions))

    def save_model(self):
        return self.request(json.dumps({"command": "save_model"}))

    def close(self):
        try:
            self.p.stdin.close()
            self.p.wait(timeout=60)
        except Exception:
            self.p.kill()
        while True:                                  # what the worker printed while shutting down
            line = self._readline(0.2)
            if line is None:
                break
            self.other_lines.append(line)
        return self.p.returncode


def c1_trajectory(rng, obs_dim=4, act_dim=2, lo=8, hi=500, with_baseline=True, policy=None):
    """CartPole-shaped synthetic trajectory: length ~ U{lo..hi}, reward 1.0 per step, obs ~ N(0,1) as f64 on the wire,
    mask = ones (f64), act f32 [1], logp_a f32 [1], v f32 [], then the terminal marker whose reward
    (0 or 1) is the path's last_val (agent_zmq.rs:605-610).  policy(obs[n,d] f32) -> (act[n], logp[n], v[n])."""
    n = int(rng.integers(lo, hi + 1))
    obs = rng.standard_normal((n, obs_dim))
    if policy is not None:
        act, logp, v = policy(obs.astype(np.float32))
    else:
        act = rng.integers(0, act_dim, n)
        logp, v = rng.standard_normal(n), rng.standard_normal(n)
    acts = []
    for i in range(n):
        data = {"logp_a": np.array([logp[i]], np.float32)}
        if with_baseline:
            data["v"] = np.array(v[i], np.float32)
        acts.append(rf.RelayRLAction(obs[i].astype(np.float64), np.array([act[i]], np.float32),
                                     np.ones(act_dim, np.float64), 1.0, data, False, False))
    acts.append(rf.RelayRLAction(None, None, None, float(rng.integers(0, 2)), None, True, False))
    return acts


def actions_to_columns(acts, obs_dim, act_dim):
    """The same actions as SoA arrays (rrl_batch_view_t convention), for feeding the oracle."""
    n = len(acts)
    obs, mask = np.zeros((n, obs_dim)), np.ones((n, act_dim))
    act, rew, val, logp = (np.zeros(n, np.float32) for _ in range(4))
    done = np.zeros(n, np.uint8)
    for i, a in enumerate(acts):
        rew[i] = a.get_rew()
        if a.get_done():
            done[i] = 1
            continue
        obs[i], mask[i], act[i] = a.get_obs(), a.get_mask(), a.get_act()[0]
        d = a.get_data()
        logp[i] = d["logp_a"][0]
        if "v" in d:
            val[i] = float(d["v"])
    return obs, act, mask, rew, val, logp, done
