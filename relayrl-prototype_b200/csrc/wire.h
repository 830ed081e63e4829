// Shared between the wire-format decoders (wire.cu: safetensors blobs + the proto3 Trajectory; wire_pickle.cu: the
// serde-pickle Vec<RelayRLAction> of the ZMQ transport).  Host code only.
#pragma once
#include <algorithm>
#include <memory>
#include <new>
#include <stdexcept>

#include "learner.cuh"

namespace rrl {

struct BlobView {
    int dtype;  // 0 F32, 1 F64, 2 I64, 3 I32, 4 U8/BOOL, 5 I16
    const uint8_t *data;
    size_t nbytes;
    size_t count;
};
// One safetensors blob holding one tensor (TensorData.data, src/types/action.rs:196-201,345).  Every read is bounded by
// `len`; the header's shape must agree with its data_offsets.
int32_t parse_blob(const uint8_t *blob, size_t len, BlobView &out, ErrBuf &err);
float blob_at(const BlobView &b, size_t i);

// caller-provided (or exactly sized) SoA rows in the rrl_batch_view_t convention
struct HostColumns {
    float *obs, *act, *mask, *rew, *val, *logp;
    uint8_t *done;
    int64_t cap;
};

// No exception may cross the C ABI (the header promises status codes): every entry point that touches the heap runs
// its body through this.
template <typename F>
int32_t abi_guard(ErrBuf &err, const char *where, F &&body)
{
    try {
        return body();
    } catch (const std::bad_alloc &) {
        err.set("%s: out of host memory", where);
        return RRL_ENOMEM;
    } catch (const std::exception &e) {
        err.set("%s: %s", where, e.what());
        return RRL_EINVAL;
    } catch (...) {
        err.set("%s: unexpected failure", where);
        return RRL_EINVAL;
    }
}

constexpr int64_t INGEST_CHUNK_ROWS = 1 << 16;

// exactly-sized, uninitialised host rows for one decoded message (the decoders write every row they report)
struct OwnedColumns {
    std::unique_ptr<float[]> obs, act, mask, rew, val, logp;
    std::unique_ptr<uint8_t[]> done;
    HostColumns view;
    OwnedColumns(int64_t rows, int d, int A)
    {
        const size_t n = (size_t)std::max<int64_t>(rows, 1);
        obs.reset(new float[n * d]), mask.reset(new float[n * A]);
        act.reset(new float[n]), rew.reset(new float[n]), val.reset(new float[n]), logp.reset(new float[n]);
        done.reset(new uint8_t[n]);
        view = HostColumns{obs.get(), act.get(), mask.get(), rew.get(), val.get(), logp.get(), done.get(), rows};
    }
};

}  // namespace rrl
