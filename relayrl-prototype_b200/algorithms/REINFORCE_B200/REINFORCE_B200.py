"""REINFORCE_B200 -- drop-in algorithm plugin for RelayRL's Python worker.

Loaded exactly like the reference's own algorithm: the unmodified worker
(relayrl_framework/src/native/python/python_algorithm_reply.py:41-46) does
`importlib.import_module("REINFORCE_B200.REINFORCE_B200")` from `algorithm_dir` and calls
`REINFORCE_B200(**hyperparams)` with `env_dir, config_path, obs_dim, act_dim, buf_size`
(src/network/server/training_server_wrapper.rs:265-274).  It implements the AlgorithmAbstract
contract (_common/_algorithms/BaseAlgorithm.py:4-39) with the same semantics as
algorithms/REINFORCE/REINFORCE.py, but every number is produced by librelayrl_b200.so on a B200:
ingest -> segmented scan -> statistics -> fused MLP forward/backward -> Adam.  torch is used
only to export the TorchScript policy agents load (REINFORCE.py:64-68).

stdout is the worker's reply channel (python_algorithm_request.rs:169-196), so this module logs
to stderr only.
"""
import json
import os
import sys
import time

import numpy as np

_PKG = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
if _PKG not in sys.path:
    sys.path.insert(0, _PKG)

from relayrl_b200 import Learner, f32  # noqa: E402

try:  # the abstract base lives in the reference tree; present when run under the real worker
    from _common._algorithms.BaseAlgorithm import AlgorithmAbstract  # type: ignore
except Exception:  # pragma: no cover - standalone use
    class AlgorithmAbstract:  # minimal stand-in with the same four abstract names
        def save(self, *a, **k): raise NotImplementedError
        def receive_trajectory(self, trajectory): raise NotImplementedError
        def train_model(self): raise NotImplementedError
        def log_epoch(self): raise NotImplementedError


# Defaults the reference embeds (src/sys_utils/config_loader.rs:68-78), f32-rounded like the Rust loader
_DEFAULTS = dict(discrete=True, with_vf_baseline=False, seed=1, traj_per_epoch=8, gamma=f32(0.98), lam=f32(0.97),
                 pi_lr=f32(3e-4), vf_lr=f32(1e-3), train_vf_iters=80)


def _load_hyperparams(config_path):
    """ConfigLoader(algorithm_name='REINFORCE', ...) when the PyO3 module is importable
    (REINFORCE.py:25-27; the loader only knows the name "REINFORCE", config_loader.rs:406-432),
    else the same JSON file read directly, else the embedded defaults."""
    try:
        from relayrl_framework import ConfigLoader  # type: ignore
        cl = ConfigLoader(algorithm_name="REINFORCE", config_path=config_path)
        return dict(cl.get_algorithm_params()["REINFORCE"]), cl.get_server_model_path()
    except Exception:
        pass
    hp, model_path = dict(_DEFAULTS), "server_model.pt"
    try:
        with open(config_path) as f:
            cfg = json.load(f)
        sect = (cfg.get("algorithms") or {}).get("REINFORCE") or {}
        for k in hp:
            if k in sect:
                hp[k] = f32(sect[k]) if isinstance(hp[k], float) else type(hp[k])(sect[k])
        mp = (cfg.get("model_paths") or {}).get("server_model")
        if mp:
            model_path = mp
    except Exception:
        pass
    return hp, model_path


class _TsvLogger:
    """progress.txt writer with the reference's columns (utils/logger.py:328-376) for runs where
    the reference's EpochLogger is not importable."""

    def __init__(self, out_dir):
        os.makedirs(out_dir, exist_ok=True)
        self._f = open(os.path.join(out_dir, "progress.txt"), "w")
        self._first = True
        self._row = {}
        self._acc = {}

    def store(self, **kw):
        for k, v in kw.items():
            self._acc.setdefault(k, []).append(v)

    def log_tabular(self, key, val=None, with_min_and_max=False, average_only=False):
        if val is not None:
            self._row[key] = val
            return
        v = np.concatenate([np.atleast_1d(np.asarray(x, np.float32)) for x in self._acc.get(key, [0.0])])
        self._row[key if average_only else "Average" + key] = float(v.mean())
        if not average_only:
            self._row["Std" + key] = float(v.std())
        if with_min_and_max:
            self._row["Max" + key], self._row["Min" + key] = float(v.max()), float(v.min())
        self._acc[key] = []

    def dump_tabular(self):
        if self._first:
            self._f.write("\t".join(self._row.keys()) + "\n")
            self._first = False
        self._f.write("\t".join(str(v) for v in self._row.values()) + "\n")
        self._f.flush()
        print("[REINFORCE_B200] " + " ".join("%s=%.5g" % (k, v) for k, v in self._row.items()), file=sys.stderr)
        self._row = {}


class REINFORCE_B200(AlgorithmAbstract):
    def __init__(self, env_dir, config_path, obs_dim, act_dim, buf_size, hidden=128, device=0, checkpoint_path=None,
                 **extra):
        super().__init__()
        hp, self.save_model_path = _load_hyperparams(config_path)
        self._with_vf_baseline = bool(hp["with_vf_baseline"])
        self._traj_per_epoch = int(hp["traj_per_epoch"])
        self._gamma, self._lam = float(hp["gamma"]), float(hp["lam"])
        self._pi_lr, self._vf_lr = float(hp["pi_lr"]), float(hp["vf_lr"])
        self._train_vf_iters = int(hp["train_vf_iters"])
        self._obs_dim, self._act_dim, self._hidden = int(obs_dim), int(act_dim), int(hidden)
        seed = int(hp["seed"]) + 10000 * os.getpid()  # REINFORCE.py:40-42

        self.learner = Learner(self._obs_dim, self._act_dim, int(buf_size), hidden=self._hidden,
                               with_baseline=self._with_vf_baseline, gamma=self._gamma, lam=self._lam,
                               pi_lr=self._pi_lr, vf_lr=self._vf_lr, train_vf_iters=self._train_vf_iters,
                               device=int(device))
        # initial weights: torch.nn.Linear's default init, like the reference's modules (kernel.py:15-21)
        import torch
        torch.manual_seed(seed)
        self._module = build_module(self._obs_dim, self._act_dim, self._hidden, self._with_vf_baseline)
        self.learner.set_params(0, flat_params(self._module.policy.pi_network))
        if self._with_vf_baseline:
            self.learner.set_params(1, flat_params(self._module.baseline.v_network))

        exp = "relayrl-reinforce-vf-info" if self._with_vf_baseline else "relayrl-reinforce-info"
        out_dir = os.path.join(env_dir, "logs", exp, "%s_s%d" % (exp, seed))
        try:
            from utils.logger import EpochLogger, setup_logger_kwargs  # type: ignore  (reference tree)
            self.logger = EpochLogger(**setup_logger_kwargs(exp, seed=seed, data_dir=os.path.join(env_dir, "./logs/")))
        except Exception:
            self.logger = _TsvLogger(out_dir)
        self.traj = 0
        self.epoch = 0
        self.last_stats = None
        # training-state checkpoint (the reference restarts from fresh weights,
        # training_zmq.rs:322-465): an extra hyper-parameter names the file; it is loaded when present and rewritten
        # after every epoch
        self._checkpoint_path = checkpoint_path
        if checkpoint_path and os.path.exists(checkpoint_path):
            self.epoch = int(self.learner.checkpoint_load(checkpoint_path))
            self.traj = self.epoch * self._traj_per_epoch
            print("[REINFORCE_B200] resumed from %s at epoch %d" % (checkpoint_path, self.epoch), file=sys.stderr)
        print("[REINFORCE_B200] initialised on cuda:%d" % int(device), file=sys.stderr)

    # -- REINFORCE.save (REINFORCE.py:64-68) ------------------------------------------------------
    def save(self) -> None:
        import torch
        load_flat(self._module.policy.pi_network, self.learner.get_params(0))
        if self._with_vf_baseline:
            load_flat(self._module.baseline.v_network, self.learner.get_params(1))
        # hand-off fast path: the module is scripted once; the ScriptModule shares its
        # parameters with self._module, so later saves only serialise the fresh weights (35 ms -> 2 ms per save)
        if getattr(self, "_scripted", None) is None:
            self._module.eval()
            self._scripted = torch.jit.script(self._module)
            self._module.train()
        torch.jit.save(self._scripted, self.save_model_path)

    # -- REINFORCE.receive_trajectory (REINFORCE.py:70-95) ------------------------------------------
    def receive_trajectory(self, trajectory) -> bool:
        self.traj += 1
        packed = trajectory.packed() if hasattr(trajectory, "packed") else pack_actions(
            trajectory.get_actions(), self._obs_dim, self._act_dim, self._with_vf_baseline)
        obs, act, mask, rew, val, logp, done = packed
        self.learner.ingest(obs, act, mask, rew, val, logp, done)
        # the reference never resets ep_ret / ep_len inside one trajectory
        if len(done):
            cr, cl = np.cumsum(rew, dtype=np.float64), np.arange(1, len(done) + 1)
            for i in np.flatnonzero(done):
                self.logger.store(EpRet=float(cr[i]), EpLen=int(cl[i]))
            if self._with_vf_baseline and (~done.astype(bool)).any():
                self.logger.store(VVals=val[~done.astype(bool)])
        if self.traj > 0 and self.traj % self._traj_per_epoch == 0:
            self.epoch += 1
            self.train_model()
            self.log_epoch()
            if self._checkpoint_path:
                self.learner.checkpoint_save(self._checkpoint_path, tag=self.epoch)
            return True
        return False

    # -- REINFORCE.train_model (REINFORCE.py:97-125) --------------------------------------------
    def train_model(self) -> None:
        st = self.learner.train_epoch()
        self.last_stats = st
        self.logger.store(LossPi=st["loss_pi"], DeltaLossPi=st["delta_loss_pi"], KL=st["kl"], Entropy=st["entropy"])
        if self._with_vf_baseline:
            self.logger.store(LossV=st["loss_v"], DeltaLossV=st["delta_loss_v"])

    # -- REINFORCE.log_epoch (REINFORCE.py:127-139) -----------------------------------------------
    def log_epoch(self) -> None:
        lg = self.logger
        lg.log_tabular("Epoch", self.epoch)
        lg.log_tabular("EpRet", with_min_and_max=True)
        lg.log_tabular("EpLen", average_only=True)
        lg.log_tabular("LossPi", average_only=True)
        lg.log_tabular("DeltaLossPi", average_only=True)
        if self._with_vf_baseline:
            lg.log_tabular("VVals", with_min_and_max=True)
            lg.log_tabular("LossV", average_only=True)
            lg.log_tabular("DeltaLossV", average_only=True)
        lg.log_tabular("KL", average_only=True)
        lg.log_tabular("Entropy", average_only=True)
        lg.dump_tabular()


def pack_actions(actions, obs_dim, act_dim, with_baseline):
    """Duck-typed action objects (get_obs/get_act/get_mask/get_rew/get_data/get_done,
    src/bindings/python/o3_action.rs:96-141) -> SoA arrays for one rrl_ingest_packed call."""
    n = len(actions)
    obs = np.zeros((n, obs_dim), np.float64)
    mask = np.zeros((n, act_dim), np.float64)
    act, rew = np.zeros(n, np.float32), np.zeros(n, np.float32)
    val, logp = np.zeros(n, np.float32), np.zeros(n, np.float32)
    done = np.zeros(n, np.uint8)
    for i, a in enumerate(actions):
        rew[i] = a.get_rew()
        if a.get_done():
            done[i] = 1
            continue
        data = a.get_data()
        obs[i] = np.asarray(a.get_obs()).reshape(-1)
        act[i] = np.asarray(a.get_act()).reshape(-1)[0]
        mask[i] = np.asarray(a.get_mask()).reshape(-1)
        logp[i] = np.asarray(data["logp_a"]).reshape(-1)[0]
        if with_baseline:
            val[i] = np.asarray(data["v"]).reshape(-1)[0]
    return obs, act, mask, rew, (val if with_baseline else None), logp, done


# ---- TorchScript hand-off: a module agents can validate and run (agent_wrapper.rs:88-168) ----------
def build_module(obs_dim, act_dim, hidden, with_baseline):
    import torch
    import torch.nn as nn
    from typing import Dict, Tuple

    def trunk(out):
        return nn.Sequential(nn.Linear(obs_dim, hidden), nn.ReLU(), nn.Linear(hidden, hidden), nn.ReLU(),
                             nn.Linear(hidden, out))

    class _Pi(nn.Module):  # state_dict keys match the reference: policy.pi_network.{0,2,4}.*
        def __init__(self):
            super().__init__()
            self.pi_network = trunk(act_dim)

        def forward(self, obs: torch.Tensor, mask: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
            logits = self.pi_network(obs) + (mask - 1) * 1e8
            return torch.softmax(logits, dim=-1), logits

    class _V(nn.Module):  # baseline.v_network.{0,2,4}.*
        def __init__(self):
            super().__init__()
            self.v_network = trunk(1)

        def forward(self, obs: torch.Tensor) -> torch.Tensor:
            return torch.squeeze(self.v_network(obs), -1)

    class Stepper(nn.Module):
        def __init__(self):
            super().__init__()
            self.policy = _Pi()
            self.baseline = _V() if with_baseline else nn.Identity()
            self.has_v = with_baseline
            self.input_dim, self.output_dim = obs_dim, act_dim

        def forward(self, obs: torch.Tensor, mask: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
            return self.policy(obs, mask)

        @torch.jit.export
        def step(self, obs: torch.Tensor, mask: torch.Tensor) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
            with torch.no_grad():
                probs, logits = self.policy(obs, mask)
                act = torch.multinomial(probs, 1)
                # raw masked logit of the sampled action, with the reference's shapes
                # (kernel.py:33-37: the 1-D obs agents send gives [1], a [1, d] batch gives [1, 1])
                a = act.long().unsqueeze(-1)
                a, lp = torch.broadcast_tensors(a, logits)
                logp_a = lp.gather(-1, a[..., :1]).squeeze(-1)
                data = {"logp_a": logp_a}
                if self.has_v:
                    data["v"] = self.baseline(obs)
            return act, data

        @torch.jit.export
        def get_input_dim(self) -> int:
            return self.input_dim

        @torch.jit.export
        def get_output_dim(self) -> int:
            return self.output_dim

    return Stepper()


def flat_params(seq):
    import torch
    return torch.cat([p.detach().reshape(-1) for p in seq.parameters()]).cpu().numpy()


def load_flat(seq, flat):
    import torch
    o = 0
    with torch.no_grad():
        for p in seq.parameters():
            n = p.numel()
            p.copy_(torch.from_numpy(np.asarray(flat[o:o + n])).reshape(p.shape))
            o += n
