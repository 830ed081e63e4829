"""Host-side mirror of the reference learner objects (REINFORCE + ReplayBuffer) over the C ABI.

`Learner` owns one `rrl_learner_t` (one GPU).  Method names follow the reference:
store-like ingest (`ingest`), `finish` (= every pending ReplayBuffer.finish_path), `get`-like taps
(`offsets`, `returns`, `advantages`), `train_epoch` (= REINFORCE.train_model).  Reference paths are
relative to relayrl_framework/src/native/python/algorithms/REINFORCE/.
"""
import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import BatchView, Config, EpochStats, check


def f32(x):
    """Hyper-parameters are f32 in the reference's Rust config (src/sys_utils/config_loader.rs:148-159)."""
    return float(np.float32(x))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _dev_ptr(t):
    """Device pointer of a torch CUDA tensor (or an int)."""
    if t is None:
        return None
    if isinstance(t, int):
        return C.c_void_p(t)
    assert t.is_cuda and t.is_contiguous(), "device inputs must be contiguous CUDA tensors"
    return C.c_void_p(t.data_ptr())


class Learner:
    def __init__(self, obs_dim, act_dim, capacity, hidden=128, with_baseline=True, gamma=0.98, lam=0.97,
                 pi_lr=3e-4, vf_lr=1e-3, train_vf_iters=80, device=0, logp_mode=_lib.RRL_LOGP_REFERENCE,
                 rank=0, nranks=1, nccl_id=None, flags=0):
        self._L = _lib.lib()
        cfg = Config()
        self._L.rrl_config_defaults(C.byref(cfg))
        cfg.obs_dim, cfg.act_dim, cfg.hidden = obs_dim, act_dim, hidden
        cfg.with_baseline, cfg.train_vf_iters, cfg.logp_mode = int(with_baseline), train_vf_iters, logp_mode
        cfg.device, cfg.rank, cfg.nranks, cfg.capacity, cfg.flags = device, rank, nranks, capacity, flags
        cfg.gamma, cfg.lam, cfg.pi_lr, cfg.vf_lr = f32(gamma), f32(lam), f32(pi_lr), f32(vf_lr)
        if nranks > 1:
            assert nccl_id is not None and len(nccl_id) == _lib.NCCL_ID_BYTES
            C.memmove(cfg.nccl_id, bytes(nccl_id), _lib.NCCL_ID_BYTES)
        self.cfg = cfg
        self._h = C.c_void_p()
        check(self._L.rrl_learner_create(C.byref(cfg), C.byref(self._h)))
        self.obs_dim, self.act_dim, self.hidden, self.with_baseline = obs_dim, act_dim, hidden, bool(with_baseline)

    # -- life cycle ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.rrl_learner_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, status):
        check(status, self._h)

    @property
    def stream(self):
        """cudaStream_t (as an int) that every call on this learner is ordered on."""
        st = C.c_void_p()
        self._ck(self._L.rrl_learner_stream(self._h, C.byref(st)))
        return st.value or 0

    # -- parameters (torch.nn.Linear order W1,b1,W2,b2,W3,b3; kernel.py:15-21) -----------------
    def param_count(self, which):
        n = C.c_int64()
        self._ck(self._L.rrl_param_count(self._h, which, C.byref(n)))
        return n.value

    def set_params(self, which, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        self._ck(self._L.rrl_set_params(self._h, which, _ptr(flat), flat.size))

    def get_params(self, which):
        out = np.empty(self.param_count(which), np.float32)
        self._ck(self._L.rrl_get_params(self._h, which, _ptr(out), out.size))
        return out

    def get_opt_state(self, which):
        n = self.param_count(which)
        m, v, step = np.empty(n, np.float32), np.empty(n, np.float32), C.c_int64()
        self._ck(self._L.rrl_get_opt_state(self._h, which, _ptr(m), _ptr(v), n, C.byref(step)))
        return m, v, step.value

    def set_opt_state(self, which, m, v, step):
        m, v = np.ascontiguousarray(m, np.float32), np.ascontiguousarray(v, np.float32)
        self._ck(self._L.rrl_set_opt_state(self._h, which, _ptr(m), _ptr(v), m.size, step))

    def checkpoint_save(self, path, tag=0):
        """Parameters + Adam state of both networks + `tag` (e.g. the epoch counter) -> one file, atomically."""
        self._ck(self._L.rrl_checkpoint_save(self._h, os.fsencode(path), int(tag)))

    def checkpoint_load(self, path):
        """Restores what checkpoint_save wrote; returns the tag."""
        tag = C.c_uint64()
        self._ck(self._L.rrl_checkpoint_load(self._h, os.fsencode(path), C.byref(tag)))
        return tag.value

    # -- ingest (REINFORCE.receive_trajectory's loop + ReplayBuffer.store, REINFORCE.py:70-87) ---
    def ingest(self, obs, act, mask, rew, val, logp, done):
        """Host SoA arrays; rows with done != 0 are terminal markers.  obs/mask may be f32 or f64."""
        done = np.ascontiguousarray(done, np.uint8)
        n = done.shape[0]

        def col(a, width):
            a = np.asarray(a)
            if a.dtype != np.float64:
                a = a.astype(np.float32, copy=False)
            a = np.ascontiguousarray(a).reshape(n, width) if n else np.zeros((0, width), a.dtype)
            return a

        obs, mask = col(obs, self.obs_dim), col(mask, self.act_dim)
        act = np.ascontiguousarray(np.asarray(act, np.float32).reshape(n))
        rew = np.ascontiguousarray(rew, np.float32)
        logp = np.ascontiguousarray(np.asarray(logp, np.float32).reshape(n))
        val = np.ascontiguousarray(np.asarray(val, np.float32).reshape(n)) if val is not None else None
        if self.with_baseline and val is None:
            raise ValueError("with_baseline learner needs data['v'] (REINFORCE.py:81)")
        b = BatchView(_ptr(obs), _ptr(act), _ptr(mask), _ptr(rew), _ptr(val), _ptr(logp), _ptr(done), n,
                      _lib.RRL_F64 if obs.dtype == np.float64 else _lib.RRL_F32,
                      _lib.RRL_F64 if mask.dtype == np.float64 else _lib.RRL_F32)
        self._ck(self._L.rrl_ingest_packed(self._h, C.byref(b)))

    def ingest_raw(self, view: BatchView):
        """Pre-built view (e.g. over pinned memory from host_alloc)."""
        self._ck(self._L.rrl_ingest_packed(self._h, C.byref(view)))

    def ingest_device(self, obs, act, mask, rew, val, logp, count, offsets, last_vals):
        """Device-resident f32 columns (torch CUDA tensors) + host episode offsets / last_vals."""
        offsets = np.ascontiguousarray(offsets, np.int64)
        last_vals = np.ascontiguousarray(last_vals, np.float32)
        self._ck(self._L.rrl_ingest_device(self._h, _dev_ptr(obs), _dev_ptr(act), _dev_ptr(mask), _dev_ptr(rew),
                                           _dev_ptr(val), _dev_ptr(logp), count, _ptr(offsets), _ptr(last_vals),
                                           len(offsets) - 1))

    def ingest_wire_action(self, obs_blob, act_blob, mask_blob, rew, v_blob, logp_blob, done):
        """One action whose tensors are safetensors blobs (src/types/action.rs:196-201)."""
        def b(x):
            return (x, len(x)) if x is not None else (None, 0)
        (o, ol), (a, al), (m, ml), (v, vl), (lp, ll) = b(obs_blob), b(act_blob), b(mask_blob), b(v_blob), b(logp_blob)
        self._ck(self._L.rrl_ingest_wire_action(self._h, o, ol, a, al, m, ml, float(rew), v, vl, lp, ll, int(bool(done))))

    def ingest_grpc_trajectory(self, msg):
        """One serialized proto Trajectory as SendActions receives it (proto/relayrl_grpc.proto:5-18,
        src/sys_utils/grpc_utils.rs:83-158); returns the number of actions it held."""
        n = C.c_int64()
        self._ck(self._L.rrl_ingest_grpc_trajectory(self._h, msg, len(msg), C.byref(n)))
        return n.value

    def ingest_pickle(self, msg):
        """One ZMQ trajectory message: serde-pickle of Vec<RelayRLAction> (src/types/trajectory.rs:50-55,
        src/network/server/training_zmq.rs:994-1001); returns the number of actions it held."""
        n = C.c_int64()
        self._ck(self._L.rrl_ingest_pickle(self._h, msg, len(msg), C.byref(n)))
        return n.value

    # -- finish_path for every closed path + statistics (replay_buffer.py:48-79) ------------------
    def finish(self):
        n, e = C.c_int64(), C.c_int64()
        self._ck(self._L.rrl_finish(self._h, C.byref(n), C.byref(e)))
        return n.value, e.value

    def offsets(self):
        cap = 1 << 16
        while True:
            out, nw = np.empty(cap, np.int64), C.c_int64()
            st = self._L.rrl_get_offsets(self._h, _ptr(out), cap, C.byref(nw))
            if st == _lib.RRL_OK:
                return out[:nw.value].copy()
            cap *= 16
            if cap > 1 << 36:
                self._ck(st)

    def _tap(self, fn, n, *extra):
        out = np.empty(max(n, 1), np.float32)
        self._ck(fn(self._h, _ptr(out), out.size, *extra))
        return out[:n]

    def returns(self, n):
        return self._tap(self._L.rrl_get_returns, n)

    def advantages(self, n, normalised=True):
        return self._tap(self._L.rrl_get_advantages, n, int(normalised))

    def adv_stats(self):
        m, s = C.c_float(), C.c_float()
        self._ck(self._L.rrl_get_adv_stats(self._h, C.byref(m), C.byref(s)))
        return np.float32(m.value), np.float32(s.value)

    # -- REINFORCE.train_model (REINFORCE.py:97-125) -------------------------------------------
    def train_epoch(self):
        st = EpochStats()
        self._ck(self._L.rrl_train_epoch(self._h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in EpochStats._fields_ if k != "reserved"}

    def eval_grad(self, which):
        g = np.empty(self.param_count(which), np.float32)
        out3 = (C.c_double * 3)()
        self._ck(self._L.rrl_eval_grad(self._h, which, _ptr(g), g.size, out3))
        return g, tuple(out3)

    def reset_buffer(self):
        self._ck(self._L.rrl_reset_buffer(self._h))


class Actor:
    """PolicyWith[out]Baseline.step for a batch of environments (kernel.py:99-107,128-135)."""

    def __init__(self, obs_dim, act_dim, hidden=128, with_baseline=True, device=0, max_batch=65536):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        check(self._L.rrl_actor_create(obs_dim, act_dim, hidden, int(with_baseline), device, max_batch,
                                       C.byref(self._h)))
        self.obs_dim, self.act_dim, self.hidden, self.with_baseline = obs_dim, act_dim, hidden, bool(with_baseline)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.rrl_actor_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        st = C.c_void_p()
        check(self._L.rrl_actor_stream(self._h, C.byref(st)), self._h, actor=True)
        return st.value or 0

    def set_params(self, which, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        check(self._L.rrl_actor_set_params(self._h, which, _ptr(flat), flat.size), self._h, actor=True)

    def step(self, obs, mask, seed=0, offset=0, want_probs=False):
        """Returns (act int64 [B], {'logp_a': f32 [B], 'v': f32 [B]}[, probs, logits])."""
        obs = np.ascontiguousarray(obs, np.float32).reshape(-1, self.obs_dim)
        mask = np.ascontiguousarray(mask, np.float32).reshape(-1, self.act_dim)
        B = obs.shape[0]
        act, logp = np.empty(B, np.int64), np.empty(B, np.float32)
        v = np.empty(B, np.float32) if self.with_baseline else None
        probs = np.empty((B, self.act_dim), np.float32) if want_probs else None
        logits = np.empty((B, self.act_dim), np.float32) if want_probs else None
        check(self._L.rrl_actor_step(self._h, _ptr(obs), _ptr(mask), B, seed, offset, _ptr(act), _ptr(logp), _ptr(v),
                                     _ptr(probs), _ptr(logits)), self._h, actor=True)
        data = {"logp_a": logp}
        if v is not None:
            data["v"] = v
        return (act, data, probs, logits) if want_probs else (act, data)

    def step_device(self, obs, mask, act, logp_a, v=None, probs=None, logits=None, seed=0, offset=0, force_act=None):
        """All arguments are contiguous torch CUDA tensors (f32; act/force_act int64)."""
        B = obs.shape[0]
        check(self._L.rrl_actor_step_device(self._h, _dev_ptr(obs), _dev_ptr(mask), B, seed, offset, _dev_ptr(force_act),
                                            _dev_ptr(act), _dev_ptr(logp_a), _dev_ptr(v), _dev_ptr(probs),
                                            _dev_ptr(logits)), self._h, actor=True)


def nccl_unique_id():
    buf = (C.c_uint8 * _lib.NCCL_ID_BYTES)()
    check(_lib.lib().rrl_nccl_unique_id(buf))
    return bytes(buf)


def host_alloc(nbytes):
    p = C.c_void_p()
    check(_lib.lib().rrl_host_alloc(C.byref(p), nbytes))
    return p


def host_free(p):
    check(_lib.lib().rrl_host_free(p))


def scan_device(rew, val, offsets, last_vals, n_total, n_ep, gamma, lam, with_baseline, adv, ret, stats, workspace,
                stream=None):
    """Stateless K2 on torch CUDA tensors (see rrl_scan_device)."""
    check(_lib.lib().rrl_scan_device(_dev_ptr(rew), _dev_ptr(val), _dev_ptr(offsets), _dev_ptr(last_vals), n_total,
                                     n_ep, float(gamma), float(lam), int(with_baseline), _dev_ptr(adv), _dev_ptr(ret),
                                     _dev_ptr(stats), _dev_ptr(workspace), workspace.numel() * workspace.element_size(),
                                     C.c_void_p(stream) if stream else None))


SCAN_NOALLOC_MAX = 16 * 1024 * 1024


def scan_device_noalloc(rew, val, offsets, last_vals, n_total, n_ep, gamma, lam, with_baseline, adv, ret, stats,
                        stream=None):
    """rrl_scan_device_noalloc: the scan for hosts that may not allocate device memory (a locked
    memory pool, an exhausted quota, a plug-in inside an application that bans allocation after
    start-up).  Its scratch is a fixed array inside the library, hence n_total <= SCAN_NOALLOC_MAX;
    calls on one device must not overlap; make the first call per device outside stream capture."""
    check(_lib.lib().rrl_scan_device_noalloc(_dev_ptr(rew), _dev_ptr(val), _dev_ptr(offsets), _dev_ptr(last_vals),
                                             n_total, n_ep, float(gamma), float(lam), int(with_baseline),
                                             _dev_ptr(adv), _dev_ptr(ret), _dev_ptr(stats),
                                             C.c_void_p(stream) if stream else None))


def scan_workspace_bytes(n):
    b = C.c_size_t()
    check(_lib.lib().rrl_scan_workspace_bytes(n, C.byref(b)))
    return b.value


def normalise_device(adv, n, stats, out, stream=None):
    check(_lib.lib().rrl_normalise_device(_dev_ptr(adv), n, _dev_ptr(stats), _dev_ptr(out),
                                          C.c_void_p(stream) if stream else None))


def decode_pickle_trajectory(msg, obs_dim, act_dim, with_baseline, cap=None):
    """Host-only decode of a serde-pickle Vec<RelayRLAction> (the ZMQ transport's message) into SoA f32 columns."""
    return _decode(_lib.lib().rrl_decode_pickle_trajectory, msg, obs_dim, act_dim, with_baseline, cap)


def decode_grpc_trajectory(msg, obs_dim, act_dim, with_baseline, cap=None):
    """Host-only decode of a serialized proto Trajectory into SoA f32 columns (rrl_batch_view_t convention:
    a terminal marker is a row with done = 1 that carries only its reward).  Needs no GPU."""
    return _decode(_lib.lib().rrl_decode_grpc_trajectory, msg, obs_dim, act_dim, with_baseline, cap)


def _decode(fn, msg, obs_dim, act_dim, with_baseline, cap):
    L = _lib.lib()
    cap = cap if cap is not None else len(msg) // 2 + 1
    cols = {"obs": np.empty((cap, obs_dim), np.float32), "act": np.empty(cap, np.float32),
            "mask": np.empty((cap, act_dim), np.float32), "rew": np.empty(cap, np.float32),
            "val": np.empty(cap, np.float32), "logp": np.empty(cap, np.float32), "done": np.empty(cap, np.uint8)}
    n, err = C.c_int64(), C.create_string_buffer(256)
    st = fn(msg, len(msg), obs_dim, act_dim, int(bool(with_baseline)), _ptr(cols["obs"]),
                                      _ptr(cols["act"]), _ptr(cols["mask"]), _ptr(cols["rew"]), _ptr(cols["val"]),
                                      _ptr(cols["logp"]), _ptr(cols["done"]), cap, C.byref(n), err, len(err))
    if st != _lib.RRL_OK:
        raise _lib.RelayRLB200Error(st, err.value.decode(errors="replace"))
    return {k: v[:n.value] for k, v in cols.items()}
