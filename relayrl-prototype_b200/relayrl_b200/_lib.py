"""ctypes binding of librelayrl_b200.so (declared in include/relayrl_b200.h).

There is no fallback: if the shared library is missing, or no CUDA device is visible, the first
call raises.  Build the library with `make -C relayrl-prototype_b200/csrc` (or
`__graft_entry__.build()`).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("RELAYRL_B200_SO") or os.path.join(_HERE, "librelayrl_b200.so")

RRL_OK = 0
RRL_F32, RRL_F64 = 0, 1
RRL_LOGP_REFERENCE, RRL_LOGP_LOG_SOFTMAX = 0, 1
RRL_FLAG_TC_ANY_SIZE = 1
RRL_FLAG_NO_FUSED_ITERS = 2
NCCL_ID_BYTES = 128


class Config(C.Structure):
    _fields_ = [("obs_dim", C.c_int32), ("act_dim", C.c_int32), ("hidden", C.c_int32),
                ("with_baseline", C.c_int32), ("train_vf_iters", C.c_int32), ("logp_mode", C.c_int32),
                ("device", C.c_int32), ("rank", C.c_int32), ("nranks", C.c_int32), ("flags", C.c_int32),
                ("capacity", C.c_int64), ("gamma", C.c_double), ("lam", C.c_double),
                ("pi_lr", C.c_double), ("vf_lr", C.c_double), ("adam_beta1", C.c_double),
                ("adam_beta2", C.c_double), ("adam_eps", C.c_double),
                ("nccl_id", C.c_uint8 * NCCL_ID_BYTES)]


class BatchView(C.Structure):
    _fields_ = [("obs", C.c_void_p), ("act", C.c_void_p), ("mask", C.c_void_p), ("rew", C.c_void_p),
                ("val", C.c_void_p), ("logp", C.c_void_p), ("done", C.c_void_p), ("count", C.c_int64),
                ("obs_dtype", C.c_int32), ("mask_dtype", C.c_int32)]


class EpochStats(C.Structure):
    _fields_ = [("loss_pi", C.c_double), ("delta_loss_pi", C.c_double), ("kl", C.c_double),
                ("entropy", C.c_double), ("loss_v", C.c_double), ("delta_loss_v", C.c_double),
                ("loss_v_old", C.c_double), ("adv_mean", C.c_double), ("adv_std", C.c_double),
                ("n_transitions", C.c_int64), ("n_episodes", C.c_int64), ("ms_scan", C.c_float),
                ("ms_policy", C.c_float), ("ms_value", C.c_float), ("ms_total", C.c_float),
                ("kernel_launches", C.c_int32), ("reserved", C.c_int32)]


# every symbol include/relayrl_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    "rrl_abi_version": (C.c_int32, []),
    "rrl_status_string": (C.c_char_p, [C.c_int32]),
    "rrl_last_error": (C.c_char_p, [_P]),
    "rrl_device_count": (C.c_int32, [C.POINTER(C.c_int32)]),
    "rrl_host_alloc": (C.c_int32, [C.POINTER(_P), C.c_size_t]),
    "rrl_host_free": (C.c_int32, [_P]),
    "rrl_nccl_unique_id": (C.c_int32, [C.POINTER(C.c_uint8)]),
    "rrl_config_defaults": (None, [C.POINTER(Config)]),
    "rrl_learner_create": (C.c_int32, [C.POINTER(Config), C.POINTER(_P)]),
    "rrl_learner_destroy": (C.c_int32, [_P]),
    "rrl_learner_stream": (C.c_int32, [_P, C.POINTER(_P)]),
    "rrl_param_count": (C.c_int32, [_P, C.c_int32, C.POINTER(C.c_int64)]),
    "rrl_set_params": (C.c_int32, [_P, C.c_int32, _P, C.c_int64]),
    "rrl_get_params": (C.c_int32, [_P, C.c_int32, _P, C.c_int64]),
    "rrl_get_opt_state": (C.c_int32, [_P, C.c_int32, _P, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "rrl_set_opt_state": (C.c_int32, [_P, C.c_int32, _P, _P, C.c_int64, C.c_int64]),
    "rrl_checkpoint_save": (C.c_int32, [_P, C.c_char_p, C.c_uint64]),
    "rrl_checkpoint_load": (C.c_int32, [_P, C.c_char_p, C.POINTER(C.c_uint64)]),
    "rrl_ingest_packed": (C.c_int32, [_P, C.POINTER(BatchView)]),
    "rrl_ingest_device": (C.c_int32, [_P, _P, _P, _P, _P, _P, _P, C.c_int64, _P, _P, C.c_int64]),
    "rrl_ingest_wire_action": (C.c_int32, [_P, _P, C.c_size_t, _P, C.c_size_t, _P, C.c_size_t, C.c_float,
                                           _P, C.c_size_t, _P, C.c_size_t, C.c_uint8]),
    "rrl_ingest_grpc_trajectory": (C.c_int32, [_P, _P, C.c_size_t, C.POINTER(C.c_int64)]),
    "rrl_decode_grpc_trajectory": (C.c_int32, [_P, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P, _P,
                                               C.c_int64, C.POINTER(C.c_int64), C.c_char_p, C.c_size_t]),
    "rrl_ingest_pickle": (C.c_int32, [_P, _P, C.c_size_t, C.POINTER(C.c_int64)]),
    "rrl_decode_pickle_trajectory": (C.c_int32, [_P, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P, _P,
                                                 C.c_int64, C.POINTER(C.c_int64), C.c_char_p, C.c_size_t]),
    "rrl_finish": (C.c_int32, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "rrl_get_offsets": (C.c_int32, [_P, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "rrl_get_returns": (C.c_int32, [_P, _P, C.c_int64]),
    "rrl_get_advantages": (C.c_int32, [_P, _P, C.c_int64, C.c_int32]),
    "rrl_get_adv_stats": (C.c_int32, [_P, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "rrl_train_epoch": (C.c_int32, [_P, C.POINTER(EpochStats)]),
    "rrl_eval_grad": (C.c_int32, [_P, C.c_int32, _P, C.c_int64, C.POINTER(C.c_double)]),
    "rrl_reset_buffer": (C.c_int32, [_P]),
    "rrl_actor_create": (C.c_int32, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64,
                                     C.POINTER(_P)]),
    "rrl_actor_destroy": (C.c_int32, [_P]),
    "rrl_actor_stream": (C.c_int32, [_P, C.POINTER(_P)]),
    "rrl_actor_set_params": (C.c_int32, [_P, C.c_int32, _P, C.c_int64]),
    "rrl_actor_step": (C.c_int32, [_P, _P, _P, C.c_int64, C.c_uint64, C.c_uint64, _P, _P, _P, _P, _P]),
    "rrl_actor_step_device": (C.c_int32, [_P, _P, _P, C.c_int64, C.c_uint64, C.c_uint64, _P, _P, _P, _P,
                                          _P, _P]),
    "rrl_actor_last_error": (C.c_char_p, [_P]),
    "rrl_rollout_create": (C.c_int32, [_P, _P, C.c_int64, C.c_int64, C.POINTER(_P)]),
    "rrl_rollout_destroy": (C.c_int32, [_P]),
    "rrl_rollout_step": (C.c_int32, [_P, _P, _P, _P, C.c_uint64, C.c_uint64, _P, _P, _P]),
    "rrl_rollout_flag_last": (C.c_int32, [_P, _P, _P, C.c_int64]),
    "rrl_rollout_last_error": (C.c_char_p, [_P]),
    "rrl_scan_workspace_bytes": (C.c_int32, [C.c_int64, C.POINTER(C.c_size_t)]),
    "rrl_scan_device": (C.c_int32, [_P, _P, _P, _P, C.c_int64, C.c_int64, C.c_double, C.c_double, C.c_int32,
                                    _P, _P, _P, _P, C.c_size_t, _P]),
    "rrl_scan_device_noalloc": (C.c_int32, [_P, _P, _P, _P, C.c_int64, C.c_int64, C.c_double, C.c_double,
                                            C.c_int32, _P, _P, _P, _P]),
    "rrl_normalise_device": (C.c_int32, [_P, C.c_int64, _P, _P, _P]),
}

DEV_SO_PATH = os.path.join(_HERE, "librelayrl_b200_dev.so")
_lib = None


def dev_lib():
    """The developer library (tcgen05 self-test, rate probes, RRL_DISABLE_TC switch): tests/test_gpu_tc.py and tools/
    only.  Loaded on its own handle; the product library is what everything else uses."""
    return C.CDLL(DEV_SO_PATH)


class RelayRLB200Error(RuntimeError):
    def __init__(self, status, message):
        super().__init__("librelayrl_b200: %s (status %d)" % (message, status))
        self.status = status


def lib():
    """Loads the CUDA library; raises if it was not built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RelayRLB200Error(-2, "%s not found: build it with `make -C relayrl-prototype_b200/csrc`; "
                                       "there is no CPU fallback" % SO_PATH)
        L = C.CDLL(SO_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(status, handle=None, actor=False):
    if status != RRL_OK:
        L = lib()
        msg = (L.rrl_actor_last_error(handle) if actor else L.rrl_last_error(handle)) or b""
        raise RelayRLB200Error(status, "%s: %s" % (L.rrl_status_string(status).decode(), msg.decode()))
