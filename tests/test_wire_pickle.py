"""rrl_decode_pickle_trajectory: the ZMQ transport's message -- serde_pickle of Vec<RelayRLAction>
(src/types/trajectory.rs:50-55, decoded by the server at src/network/server/training_zmq.rs:994-1001) -- decoded
natively into the packed columns.  Host only: no GPU needed.

serde-pickle 1.2.0 is a third-party crate that is not in the reference tree and cannot be run here, so the encoder used
as the independent check is Python's own pickler on the object graph serde's data model prescribes for these types
(src/types/action.rs:92-101,196-218,428-437), under every representation the crate could choose:
  Vec<u8> as a list of ints or as bytes; DType as a str or a 1-tuple; RelayRLData::Tensor / ::Float as a 1-entry dict
  or a 2-tuple; pickle protocols 2, 3 and 4 (memoised keys, MARK batches, FRAME).
"""
import pickle
import pickletools
import sys

import numpy as np
import pytest

from conftest import ROOT  # noqa: F401
import pipe_driver as pd

rf = pd.rf


def _tensor(arr, bytes_as, dtype_as):
    arr = np.asarray(arr)
    name = {np.dtype(np.float32): "Float", np.dtype(np.float64): "Double", np.dtype(np.int64): "Long"}[arr.dtype]
    blob = rf.blob_encode(arr)
    return {"shape": list(arr.shape), "dtype": name if dtype_as == "str" else (name,),
            "data": list(blob) if bytes_as == "list" else blob}


def _variant(tag, payload, variant_as):
    return {tag: payload} if variant_as == "dict" else (tag, payload)


def _graph(acts, bytes_as="list", dtype_as="str", variant_as="dict", scalar_v=False):
    out = []
    for a in acts:
        if a.get_done():
            out.append({"obs": None, "act": None, "mask": None, "rew": a.get_rew(), "data": None, "done": True,
                        "reward_updated": False})
            continue
        d = a.get_data()
        data = {"logp_a": _variant("Tensor", _tensor(d["logp_a"], bytes_as, dtype_as), variant_as)}
        if "v" in d:
            data["v"] = _variant("Float", float(d["v"]), variant_as) if scalar_v else \
                _variant("Tensor", _tensor(d["v"], bytes_as, dtype_as), variant_as)
        out.append({"obs": _tensor(a.get_obs(), bytes_as, dtype_as), "act": _tensor(a.get_act(), bytes_as, dtype_as),
                    "mask": _tensor(a.get_mask(), bytes_as, dtype_as), "rew": a.get_rew(), "data": data, "done": False,
                    "reward_updated": False})
    return out


def _trajectory(seed, n_paths=3):
    rng = np.random.default_rng(seed)
    acts = []
    for _ in range(n_paths):
        acts += pd.c1_trajectory(rng, hi=30)
    return acts


@pytest.mark.parametrize("protocol", [2, 3, 4])
@pytest.mark.parametrize("bytes_as,dtype_as,variant_as,scalar_v", [("list", "str", "dict", False),
                                                                   ("bytes", "str", "dict", False),
                                                                   ("list", "tuple", "tuple", False),
                                                                   ("bytes", "tuple", "tuple", True),
                                                                   ("list", "str", "dict", True)])
def test_decode_matches_the_actions(protocol, bytes_as, dtype_as, variant_as, scalar_v):
    from relayrl_b200 import decode_pickle_trajectory
    if protocol == 2 and bytes_as == "bytes":
        pytest.skip("protocol 2 has no bytes opcode: Python falls back to a _codecs.encode GLOBAL, which is not data")
    acts = _trajectory(protocol * 10 + len(bytes_as))
    msg = pickle.dumps(_graph(acts, bytes_as, dtype_as, variant_as, scalar_v), protocol=protocol)
    got = decode_pickle_trajectory(msg, 4, 2, True, cap=len(acts))
    obs, act, mask, rew, val, logp, done = pd.actions_to_columns(acts, 4, 2)
    assert np.array_equal(got["done"], done)
    assert np.array_equal(got["rew"], rew)
    keep = done == 0
    assert np.array_equal(got["obs"][keep], obs[keep].astype(np.float32))      # f64 on the wire -> f32 rows (Q11)
    assert np.array_equal(got["mask"][keep], mask[keep].astype(np.float32))
    assert np.array_equal(got["act"][keep], act[keep])
    assert np.array_equal(got["logp"][keep], logp[keep])
    assert np.array_equal(got["val"][keep], val[keep])
    # rows of terminal markers carry nothing but their reward
    assert not got["obs"][~keep].any() and not got["val"][~keep].any()


def test_without_baseline_ignores_a_missing_v_and_tuple_top_level():
    from relayrl_b200 import decode_pickle_trajectory
    rng = np.random.default_rng(3)
    acts = pd.c1_trajectory(rng, hi=10, with_baseline=False)
    msg = pickle.dumps(tuple(_graph(acts)), protocol=3)
    got = decode_pickle_trajectory(msg, 4, 2, False, cap=64)
    assert len(got["done"]) == len(acts) and got["done"][-1] == 1
    with pytest.raises(Exception) as e:                        # the same message for a learner that needs data['v']
        decode_pickle_trajectory(msg, 4, 2, True, cap=64)
    assert e.value.status == -6 and "data['v'] missing" in str(e.value)


def test_one_at_a_time_opcodes_and_long_memo():
    """A pickler that never batches (APPEND / SETITEM, like protocol-1-style writers) and uses LONG_BINPUT."""
    from relayrl_b200 import decode_pickle_trajectory
    acts = _trajectory(77, n_paths=1)
    msg = pickle.dumps(_graph(acts), protocol=2)
    ops = list(pickletools.genops(msg))
    assert any(o.name == "BINPUT" for o, _, _ in ops)
    slow = pickletools.optimize(msg)                            # drops unused memo puts: another legal stream
    a = decode_pickle_trajectory(msg, 4, 2, True, cap=len(acts))
    b = decode_pickle_trajectory(slow, 4, 2, True, cap=len(acts))
    for k in a:
        assert np.array_equal(a[k], b[k])


@pytest.mark.parametrize("mutate", ["truncate", "global", "not_a_list", "bad_blob", "odd_action", "capacity", "garbage"])
def test_malformed_streams_are_refused(mutate):
    from relayrl_b200 import RelayRLB200Error, decode_pickle_trajectory
    acts = _trajectory(5, n_paths=1)
    g = _graph(acts)
    status = -6
    if mutate == "truncate":
        msg = pickle.dumps(g, protocol=3)[:-40]
    elif mutate == "global":                                    # what a hostile pickle would use to run code
        msg = pickle.dumps([np.float32(1.0)], protocol=3)
    elif mutate == "not_a_list":
        msg = pickle.dumps({"actions": g}, protocol=3)
    elif mutate == "bad_blob":
        g[0]["obs"]["data"] = g[0]["obs"]["data"][:20]
        msg = pickle.dumps(g, protocol=3)
    elif mutate == "odd_action":
        g[1] = 7
        msg = pickle.dumps(g, protocol=3)
    elif mutate == "capacity":
        msg, status = pickle.dumps(g, protocol=3), -3
    else:
        msg = bytes(np.random.default_rng(0).integers(0, 256, 500, dtype=np.uint8))
    with pytest.raises(RelayRLB200Error) as e:
        decode_pickle_trajectory(msg, 4, 2, True, cap=2 if mutate == "capacity" else 4096)
    assert e.value.status == status, str(e.value)
    if mutate == "odd_action":
        assert "action 1" in str(e.value)


def test_every_prefix_and_random_corruption_never_crashes():
    """Robustness: all prefixes of a valid message and 300 single-byte corruptions either decode or return a status."""
    from relayrl_b200 import RelayRLB200Error, decode_pickle_trajectory
    acts = _trajectory(9, n_paths=1)[:6] + [rf.RelayRLAction(None, None, None, 1.0, None, True, False)]
    msg = pickle.dumps(_graph(acts, bytes_as="bytes"), protocol=3)
    for cut in range(0, len(msg), 7):
        try:
            decode_pickle_trajectory(msg[:cut], 4, 2, True, cap=64)
        except RelayRLB200Error as e:
            assert e.status in (-6, -3)
    rng = np.random.default_rng(1)
    for _ in range(300):
        b = bytearray(msg)
        b[int(rng.integers(0, len(b)))] = int(rng.integers(0, 256))
        try:
            decode_pickle_trajectory(bytes(b), 4, 2, True, cap=64)
        except RelayRLB200Error as e:
            assert e.status in (-6, -3)


@pytest.mark.parametrize("hdr", [b'{"tensor":{"dtype":"', b'{"tensor":{"dtype":"F64","shape":[4],"data_offsets":[0',
                                 b'{"tensor":{"dtype":"F64","shape":[4],"data_offsets":[0,32',
                                 b'{"tensor":{"dtype":"F64","shape":[5],"data_offsets":[0,32]}}',
                                 b'{"tensor":{"dtype":"F64","shape":[4,"data_offsets":[0,32]}}',
                                 b'{"tensor":{"dtype":"F6","shape":[4],"data_offsets":[0,32]}}'])
def test_safetensors_header_is_parsed_within_its_bounds(hdr):
    """headers that end inside a key or a number, a shape that disagrees with data_offsets -- every one is
    refused with RRL_EFORMAT, with the blob LAST in its buffer so that an over-read would leave the allocation."""
    import struct
    from relayrl_b200 import RelayRLB200Error, decode_pickle_trajectory
    acts = _trajectory(2, n_paths=1)[:1] + [rf.RelayRLAction(None, None, None, 1.0, None, True, False)]
    g = _graph(acts, bytes_as="bytes")
    g[0]["obs"]["data"] = struct.pack("<Q", len(hdr)) + hdr + (b"\x00" * 32 if hdr.endswith(b"}}") else b"")
    with pytest.raises(RelayRLB200Error) as e:
        decode_pickle_trajectory(pickle.dumps(g, protocol=3), 4, 2, True, cap=8)
    assert e.value.status == -6 and "safetensors" in str(e.value)
