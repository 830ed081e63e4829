"""TEST INFRASTRUCTURE -- stand-in for the compiled PyO3 module `relayrl_framework`.

The reference's Python worker (src/native/python/python_algorithm_reply.py:13,99) imports
`RelayRLTrajectory` from the Rust extension module and the algorithm classes import `ConfigLoader`
(REINFORCE.py:13,25-27).  No Rust toolchain exists in this image, so the extension cannot be built; this
pure-Python module restates exactly the parts the worker and an algorithm plug-in touch, so that the UNMODIFIED
worker script can be spawned over a pipe in tests:

  RelayRLTrajectory.traj_from_json    src/bindings/python/o3_trajectory.rs:115-166
  RelayRLAction getters                src/bindings/python/o3_action.rs:96-141 (+ :177-235 action_from_json,
                                       :301-371 tensor -> numpy, :392-469 JSON data dict, :544-598 data -> dict)
  TensorData blob                      src/types/action.rs:196-201,238-273,345-418 (safetensors, one tensor per blob)
  ConfigLoader                         src/bindings/python/o3_config_loader.rs:60-90, src/sys_utils/config_loader.rs:
                                       66-113 (embedded default), 148-159 (f32 fields), 406-432 (fallback values)

It also has the encoder side the Rust server would run (serde_json of RelayRLTrajectory,
src/network/server/python_subprocesses/python_algorithm_request.rs:235-253) so tests can produce the JSON lines.
Nothing in the product imports this.
"""
import json
import struct

import numpy as np

_DTYPES = {"Byte": ("U8", np.uint8), "Short": ("I16", np.int16), "Int": ("I32", np.int32), "Long": ("I64", np.int64),
           "Float": ("F32", np.float32), "Double": ("F64", np.float64), "Bool": ("BOOL", np.uint8)}
_ST2NP = {"U8": np.uint8, "I16": np.int16, "I32": np.int32, "I64": np.int64, "F32": np.float32, "F64": np.float64,
          "BOOL": np.uint8}


def f32(x):
    """A Rust f32 widened to a Python float (o3_config_loader.rs:76-84)."""
    return float(np.float32(x))


# ---- TensorData: {"shape": [...], "dtype": "Float", "data": [safetensors blob bytes]} -----------------------
def blob_encode(arr):
    """safetensors 0.5 serialisation of one tensor named "tensor" (action.rs:345: serialize([("tensor", view)]))."""
    arr = np.asarray(arr)
    if not arr.flags.c_contiguous:
        arr = arr.copy()                      # (np.ascontiguousarray would turn a 0-d scalar into shape [1])
    st = {np.dtype(v[1]): v[0] for k, v in _DTYPES.items() if k != "Bool"}[arr.dtype]
    hdr = json.dumps({"tensor": {"dtype": st, "shape": list(arr.shape), "data_offsets": [0, arr.nbytes]}},
                     separators=(",", ":")).encode()
    hdr += b" " * ((8 - len(hdr) % 8) % 8)
    return struct.pack("<Q", len(hdr)) + hdr + arr.tobytes()


def blob_decode(data):
    """bytes of one safetensors blob -> numpy array (action.rs:372-418)."""
    data = bytes(data)
    (hlen,) = struct.unpack_from("<Q", data, 0)
    hdr = json.loads(data[8:8 + hlen].decode())
    (name, meta), = [(k, v) for k, v in hdr.items() if k != "__metadata__"]
    b0, b1 = meta["data_offsets"]
    a = np.frombuffer(data, dtype=_ST2NP[meta["dtype"]], count=(b1 - b0) // np.dtype(_ST2NP[meta["dtype"]]).itemsize,
                      offset=8 + hlen + b0)
    return a.reshape(meta["shape"]).copy()


def tensor_data_json(arr):
    """serde_json of TensorData (action.rs:196-201): `data` is the blob as a JSON array of byte values."""
    arr = np.asarray(arr)
    name = {np.dtype(np.float32): "Float", np.dtype(np.float64): "Double", np.dtype(np.int64): "Long",
            np.dtype(np.int32): "Int", np.dtype(np.uint8): "Byte", np.dtype(np.int16): "Short"}[arr.dtype]
    return {"shape": list(arr.shape), "dtype": name, "data": list(blob_encode(arr))}


def _tensor_from_json(d):
    if not isinstance(d, dict):
        return None
    if d["dtype"] not in _DTYPES:
        raise ValueError("unsupported dtype %r" % (d["dtype"],))
    return blob_decode(bytes(bytearray(d["data"])))


# ---- RelayRLAction ---------------------------------------------------------------------------------------------
class RelayRLAction:
    def __init__(self, obs=None, act=None, mask=None, rew=0.0, data=None, done=False, reward_updated=False):
        self._obs, self._act, self._mask = obs, act, mask
        self._rew, self._data, self._done, self._ru = float(np.float32(rew)), data, bool(done), bool(reward_updated)

    # the Rust getters panic on a terminal marker (`expect("Could not get obs")`, o3_action.rs:98)
    def get_obs(self):
        if self._obs is None:
            raise RuntimeError("Could not get obs")
        return self._obs

    def get_act(self):
        if self._act is None:
            raise RuntimeError("Could not get act")
        return self._act

    def get_mask(self):
        if self._mask is None:
            raise RuntimeError("Could not get mask")
        return self._mask

    def get_rew(self):
        return self._rew

    def get_data(self):
        return dict(self._data) if self._data is not None else {}

    def get_done(self):
        return self._done

    @staticmethod
    def action_from_json(d):
        """o3_action.rs:177-235: every key must be present; a non-dict tensor field (null) is None."""
        for k in ("obs", "act", "mask", "rew", "data", "done", "reward_updated"):
            if k not in d:
                raise KeyError("Failed to get item: %s" % k)
        data = None
        if isinstance(d["data"], dict):
            data = {}
            for key, sub in d["data"].items():
                if not isinstance(sub, dict):
                    data = None
                    break
                for tag, val in sub.items():      # serde's externally tagged enum: {"Tensor": {...}} / {"Float": 1.0}
                    if tag in ("torch.Tensor", "Tensor", "tensor", "numpy.ndarray", "ndarray"):
                        data[key] = _tensor_from_json(val)
                    elif tag in ("Float", "float"):
                        data[key] = float(np.float32(val))
                    elif tag in ("Byte", "byte", "Int", "int", "Long", "long"):
                        data[key] = int(val)
                    elif tag in ("Double", "double"):
                        data[key] = float(val)
                    elif tag in ("String", "str"):
                        data[key] = str(val)
                    elif tag == "bool":
                        data[key] = bool(val)
                    else:
                        data = None
                        break
                if data is None:
                    break
        return RelayRLAction(_tensor_from_json(d["obs"]), _tensor_from_json(d["act"]), _tensor_from_json(d["mask"]),
                             d["rew"], data, d["done"], d["reward_updated"])

    def to_json_dict(self):
        def t(a):
            return tensor_data_json(a) if a is not None else None
        data = None
        if self._data is not None:
            data = {}
            for k, v in self._data.items():
                if isinstance(v, np.ndarray):
                    data[k] = {"Tensor": tensor_data_json(v)}
                elif isinstance(v, float):
                    data[k] = {"Float": v}
                else:
                    data[k] = {"Long": int(v)}
        return {"obs": t(self._obs), "act": t(self._act), "mask": t(self._mask), "rew": self._rew, "data": data,
                "done": self._done, "reward_updated": self._ru}


# ---- RelayRLTrajectory -------------------------------------------------------------------------------------------
class RelayRLTrajectory:
    def __init__(self, max_length=None, trajectory_server=None):
        self.max_length = 1000 if max_length is None else int(max_length)
        self.trajectory_server = trajectory_server
        self._actions = []

    def get_actions(self):
        return list(self._actions)

    def add_action(self, action):
        # src/types/trajectory.rs:172-203 (the ZMQ send is transport, out of scope): a done action that brings the
        # trajectory to max_length clears it
        self._actions.append(action)
        if action.get_done() and len(self._actions) >= self.max_length:
            self._actions.clear()

    @staticmethod
    def traj_from_json(trajectory_dict):
        inner = trajectory_dict["inner"]
        if not isinstance(inner, dict):
            raise TypeError("Expected 'inner' to be a dictionary")
        max_length = inner.get("max_length")
        if "trajectory_server" not in inner:
            raise KeyError("Missing 'trajectory_server' field")
        ts = inner["trajectory_server"] if isinstance(inner["trajectory_server"], str) else None
        t = RelayRLTrajectory(1000 if max_length is None else max_length, ts)
        acts = inner.get("actions")
        if acts is not None:
            if not isinstance(acts, list):
                raise TypeError("'actions' must be a list")
            for a in acts:
                t.add_action(RelayRLAction.action_from_json(a))
        return t

    def to_json_dict(self):
        """{"inner": {...}}: serde_json of PyRelayRLTrajectory as the Rust server writes it on the worker's stdin."""
        return {"inner": {"trajectory_server": self.trajectory_server, "max_length": self.max_length,
                          "actions": [a.to_json_dict() for a in self._actions]}}


def receive_trajectory_line(actions, max_length=100000, trajectory_server="tcp://127.0.0.1:7776"):
    """One JSON line of the Rust -> Python pipe (python_algorithm_request.rs:235-253)."""
    t = RelayRLTrajectory(max_length, trajectory_server)
    t._actions = list(actions)
    return json.dumps({"command": "receive_trajectory", "trajectory": t.to_json_dict()}, separators=(",", ":"))


# ---- ConfigLoader ------------------------------------------------------------------------------------------------
_EMBEDDED = dict(discrete=True, with_vf_baseline=False, seed=1, traj_per_epoch=8, gamma=0.98, lam=0.97, pi_lr=3e-4,
                 vf_lr=1e-3, train_vf_iters=80)                       # config_loader.rs:66-78 (default_config.json)
_FALLBACK = dict(discrete=True, with_vf_baseline=True, seed=0, traj_per_epoch=12, gamma=0.99, lam=0.97, pi_lr=3e-4,
                 vf_lr=1e-3, train_vf_iters=80)                       # config_loader.rs:412-422 (section missing)


class ConfigLoader:
    def __init__(self, algorithm_name=None, config_path=None):
        cfg = None
        try:
            with open(config_path) as f:
                cfg = json.load(f)
        except Exception:
            cfg = {"algorithms": {"REINFORCE": dict(_EMBEDDED)}, "model_paths": {"server_model": "server_model.pt",
                                                                                 "client_model": "client_model.pt"}}
        self._algo = algorithm_name
        sect = (cfg.get("algorithms") or {}).get("REINFORCE")
        hp = dict(_FALLBACK) if sect is None else {k: sect.get(k, _EMBEDDED[k]) for k in _EMBEDDED}
        for k in ("gamma", "lam", "pi_lr", "vf_lr"):
            hp[k] = f32(hp[k])                                         # f32 fields widened to Python floats
        self._hp = hp
        self._server_model = (cfg.get("model_paths") or {}).get("server_model", "server_model.pt")

    def get_algorithm_params(self):
        return {"REINFORCE": dict(self._hp)} if self._algo == "REINFORCE" else None

    def get_server_model_path(self):
        return self._server_model
