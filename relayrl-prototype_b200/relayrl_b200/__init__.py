"""relayrl_b200: B200-native (sm_100a) learner hot path of RelayRL behind a C ABI.

The compute lives in librelayrl_b200.so (hand-written CUDA, csrc/); this package is the thin
host-side mirror of the reference's Python objects.  No CPU fallback exists.
"""
from ._lib import RelayRLB200Error, SO_PATH, lib  # noqa: F401
from .learner import (Actor, Learner, decode_grpc_trajectory, decode_pickle_trajectory, f32, host_alloc, host_free, nccl_unique_id,  # noqa: F401
                      normalise_device, scan_device, scan_device_noalloc, scan_workspace_bytes)
