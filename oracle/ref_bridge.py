"""TEST INFRASTRUCTURE ONLY -- loads the *unmodified* reference learner in place.

Works only where /root/reference exists (the build container); nothing on the GPU box may
import this.  It is used by tests/golden/make_golden.py (to produce the committed golden
vectors) and by tests/test_oracle_vs_reference.py (to pin oracle/learner_oracle.py against
the real reference code).

Recipe: the reference Python learner
(relayrl_framework/src/native/python/algorithms/REINFORCE/REINFORCE.py) imports two names from
the compiled PyO3 module `relayrl_framework` -- `ConfigLoader` and `RelayRLTrajectory`.  No Rust
toolchain exists here, so a tiny stand-in module is placed in sys.modules that returns the
hyper-parameters the real loader would return: REINFORCE params are `f32` in Rust
(src/sys_utils/config_loader.rs:148-159) and reach Python as f32-rounded doubles
(src/bindings/python/o3_config_loader.rs:76-84).
"""
import os
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("RELAYRL_REFERENCE", "/root/reference")
_PY = os.path.join(REF_ROOT, "relayrl_framework", "src", "native", "python")


def available() -> bool:
    return os.path.isdir(_PY)


def f32(x: float) -> float:
    """A Rust f32 widened to a Python float."""
    return float(np.float32(x))


DEFAULT_HP = dict(discrete=True, with_vf_baseline=True, seed=1, traj_per_epoch=8,
                  gamma=f32(0.98), lam=f32(0.97), pi_lr=f32(3e-4), vf_lr=f32(1e-3),
                  train_vf_iters=80)


_STATE = {"hp": dict(DEFAULT_HP), "model_path": "/tmp/rrl_ref_server_model.pt"}


def install_stub(hp: dict, model_path: str):
    # the reference binds `ConfigLoader` at import time, so the stand-in is created once and
    # reads its answers from _STATE
    _STATE["hp"], _STATE["model_path"] = dict(hp), model_path
    if "relayrl_framework" in sys.modules and getattr(sys.modules["relayrl_framework"], "_rrl_stub", False):
        return sys.modules["relayrl_framework"]
    mod = types.ModuleType("relayrl_framework")
    mod._rrl_stub = True

    class ConfigLoader:  # stands in for src/bindings/python/o3_config_loader.rs
        def __init__(self, algorithm_name=None, config_path=None):
            pass

        def get_algorithm_params(self):
            return {"REINFORCE": dict(_STATE["hp"])}

        def get_server_model_path(self):
            return _STATE["model_path"]

    class RelayRLTrajectory:  # annotation only (REINFORCE.py:13,70)
        pass

    mod.ConfigLoader = ConfigLoader
    mod.RelayRLTrajectory = RelayRLTrajectory
    sys.modules["relayrl_framework"] = mod
    return mod


def load(hp=None, model_path="/tmp/rrl_ref_server_model.pt"):
    """Returns the reference modules: (REINFORCE module, kernel module, replay_buffer module)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % _PY)
    hp = dict(DEFAULT_HP, **(hp or {}))
    install_stub(hp, model_path)
    for p in (os.path.join(_PY, "algorithms"), _PY):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib
    R = importlib.import_module("REINFORCE.REINFORCE")
    K = importlib.import_module("REINFORCE.kernel")
    B = importlib.import_module("REINFORCE.replay_buffer")
    return R, K, B


class DuckAction:
    """Same getters as the PyO3 action (src/bindings/python/o3_action.rs:96-141)."""

    def __init__(self, obs=None, act=None, mask=None, rew=0.0, data=None, done=False):
        self._o, self._a, self._m, self._r, self._d, self._done = obs, act, mask, rew, data, done

    def get_obs(self):
        return self._o

    def get_act(self):
        return self._a

    def get_mask(self):
        return self._m

    def get_rew(self):
        return self._r

    def get_data(self):
        return self._d

    def get_done(self):
        return self._done


class DuckTrajectory:
    """src/bindings/python/o3_trajectory.rs:49-59"""

    def __init__(self, actions):
        self._actions = list(actions)

    def get_actions(self):
        return self._actions
