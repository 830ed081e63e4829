"""gRPC wire-format ingest: the native decoder of a serialized proto `Trajectory`
against messages produced by an independent encoder (the Python protobuf runtime, message types restated from
relayrl_framework/proto/relayrl_grpc.proto:5-18), safetensors blobs written by the safetensors package (the
reference's codec, src/types/action.rs:345,372-381) and the serde_json shape of RelayRLData
(src/types/action.rs:196-218; src/sys_utils/grpc_utils.rs:45-55).  Host only: no GPU needed."""
import json

import numpy as np
import pytest

from relayrl_b200 import RelayRLB200Error, decode_grpc_trajectory


def _messages():
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    F = descriptor_pb2.FieldDescriptorProto
    fd = descriptor_pb2.FileDescriptorProto(name="relayrl_grpc_test.proto", package="relayrl_grpc_test", syntax="proto3")
    act = fd.message_type.add(name="RelayRLAction")
    for name, num, typ in (("obs", 1, F.TYPE_BYTES), ("action", 2, F.TYPE_BYTES), ("mask", 3, F.TYPE_BYTES),
                           ("reward", 4, F.TYPE_FLOAT), ("done", 6, F.TYPE_BOOL), ("reward_update_flag", 7, F.TYPE_BOOL),
                           ("future_field", 15, F.TYPE_STRING)):          # an unknown field the decoder must skip
        act.field.add(name=name, number=num, type=typ, label=F.LABEL_OPTIONAL)
    entry = act.nested_type.add(name="DataEntry")
    entry.options.map_entry = True
    entry.field.add(name="key", number=1, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL)
    entry.field.add(name="value", number=2, type=F.TYPE_BYTES, label=F.LABEL_OPTIONAL)
    act.field.add(name="data", number=5, type=F.TYPE_MESSAGE, label=F.LABEL_REPEATED,
                  type_name=".relayrl_grpc_test.RelayRLAction.DataEntry")
    traj = fd.message_type.add(name="Trajectory")
    traj.field.add(name="actions", number=1, type=F.TYPE_MESSAGE, label=F.LABEL_REPEATED,
                   type_name=".relayrl_grpc_test.RelayRLAction")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = message_factory.GetMessageClass
    return get(pool.FindMessageTypeByName("relayrl_grpc_test.RelayRLAction")), get(
        pool.FindMessageTypeByName("relayrl_grpc_test.Trajectory"))


def _blob(arr):
    from safetensors.numpy import save
    return save({"tensor": np.ascontiguousarray(arr)})


def _tensor_json(arr, dtype_name):
    return json.dumps({"Tensor": {"shape": list(arr.shape), "dtype": dtype_name, "data": list(_blob(arr))}},
                      separators=(",", ":")).encode()


def _trajectory(rng, n, d, A, with_v=True, scalar_data=False, marker_rew=0.5):
    Action, Trajectory = _messages()
    t = Trajectory()
    want = {k: [] for k in ("obs", "act", "mask", "rew", "val", "logp", "done")}
    for i in range(n):
        obs = rng.standard_normal(d)                                   # f64 on the wire
        mask = (rng.random(A) > 0.2).astype(np.float64)
        act = np.array([rng.integers(0, A)], np.float32)
        logp, v = np.array([rng.standard_normal()], np.float32), np.array(rng.standard_normal(), np.float32)
        rew = 0.0 if i == 1 else float(np.float32(rng.standard_normal()))    # 0.0 is left out by proto3
        a = t.actions.add(obs=_blob(obs), action=_blob(act), mask=_blob(mask), reward=rew, future_field="x" * (i % 3))
        if scalar_data:
            a.data["logp_a"] = json.dumps({"Float": float(logp[0])}).encode()
            a.data["v"] = json.dumps({"Double": float(v)}, indent=1).encode()
        else:
            a.data["logp_a"] = _tensor_json(logp, "Float")
            if with_v:
                a.data["v"] = _tensor_json(v, "Float")
        a.data["extra"] = json.dumps({"String": "ignored"}).encode()
        for k, x in (("obs", obs.astype(np.float32)), ("act", act[0]), ("mask", mask.astype(np.float32)), ("rew", np.float32(rew)),
                     ("val", v if (with_v or scalar_data) else np.float32(0)), ("logp", logp[0]), ("done", 0)):
            want[k].append(x)
    t.actions.add(reward=marker_rew, done=True)                          # terminal marker: no tensors, no data
    for k, x in (("obs", np.zeros(d, np.float32)), ("act", 0.0), ("mask", np.zeros(A, np.float32)), ("rew", np.float32(marker_rew)),
                 ("val", 0.0), ("logp", 0.0), ("done", 1)):
        want[k].append(x)
    return t.SerializeToString(), {k: np.asarray(v) for k, v in want.items()}


@pytest.mark.parametrize("n,d,A", [(1, 4, 2), (37, 4, 2), (5, 128, 18)])
def test_decode_matches_encoder(n, d, A):
    msg, want = _trajectory(np.random.default_rng(n), n, d, A)
    got = decode_grpc_trajectory(msg, d, A, True)
    assert len(got["rew"]) == n + 1
    for k in want:
        assert np.array_equal(got[k], want[k].astype(got[k].dtype)), k


def test_scalar_variants_and_missing_value_head():
    msg, want = _trajectory(np.random.default_rng(3), 6, 4, 2, scalar_data=True)
    got = decode_grpc_trajectory(msg, 4, 2, True)
    for k in want:
        assert np.array_equal(got[k], want[k].astype(got[k].dtype)), k
    # without a baseline data['v'] may be absent
    msg, want = _trajectory(np.random.default_rng(4), 6, 4, 2, with_v=False)
    got = decode_grpc_trajectory(msg, 4, 2, False)
    assert np.array_equal(got["logp"], want["logp"].astype(np.float32)) and not got["val"].any()
    with pytest.raises(RelayRLB200Error, match="data\\['v'\\] missing"):
        decode_grpc_trajectory(msg, 4, 2, True)


def test_empty_and_marker_only():
    assert len(decode_grpc_trajectory(b"", 4, 2, True)["rew"]) == 0
    _, Trajectory = _messages()
    t = Trajectory()
    t.actions.add(done=True)                    # reward 0.0 and nothing else: a 2-byte action
    got = decode_grpc_trajectory(t.SerializeToString(), 4, 2, True)
    assert got["done"].tolist() == [1] and got["rew"].tolist() == [0.0]


def test_malformed_inputs_fail_loudly():
    msg, _ = _trajectory(np.random.default_rng(5), 3, 4, 2)
    for cut in (1, 7, len(msg) // 2, len(msg) - 1):
        with pytest.raises(RelayRLB200Error) as e:
            decode_grpc_trajectory(msg[:cut], 4, 2, True)
        assert e.value.status == -6
    with pytest.raises(RelayRLB200Error, match="obs has 4 values \\(want 8\\)"):
        decode_grpc_trajectory(msg, 8, 2, True)
    with pytest.raises(RelayRLB200Error) as e:
        decode_grpc_trajectory(msg, 4, 2, True, cap=2)
    assert e.value.status == -3


def test_large_message_decodes_on_several_threads():
    """Above 2048 actions per thread the decoder splits the action list over host threads; the result and the
    error reporting (lowest-numbered bad action) must not depend on that."""
    msg1, want1 = _trajectory(np.random.default_rng(9), 1500, 4, 2)
    reps = 6                                     # 9006 actions: four threads
    msg = msg1 * reps                            # concatenated serialisations of a repeated field = one message with all
    got = decode_grpc_trajectory(msg, 4, 2, True)
    assert len(got["rew"]) == reps * 1501
    for k in want1:
        assert np.array_equal(got[k], np.concatenate([want1[k].astype(got[k].dtype)] * reps)), k
    # two bad actions (obs of the wrong width) in different threads' ranges: the first one is reported
    Action, Trajectory = _messages()
    bad = Trajectory()
    a = bad.actions.add(obs=_blob(np.zeros(7)), action=_blob(np.zeros(1, np.float32)), mask=_blob(np.ones(2)), reward=1.0)
    a.data["logp_a"] = _tensor_json(np.zeros(1, np.float32), "Float")
    a.data["v"] = _tensor_json(np.zeros((), np.float32), "Float")
    bad = bad.SerializeToString()
    msg = msg1 * 2 + bad + msg1 * 3 + bad + msg1
    with pytest.raises(RelayRLB200Error, match="action 3002: obs has 7 values"):
        decode_grpc_trajectory(msg, 4, 2, True)
