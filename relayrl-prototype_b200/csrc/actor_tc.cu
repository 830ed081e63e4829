// placeholder while the kernel is being written
#include "actor_internal.cuh"
namespace rrl {
bool actor_tc_supported(int, int, int, int64_t) { return false; }
size_t actor_tc_pack_bytes(int, int) { return 0; }
int32_t actor_tc_pack_launch(const float *, int, int, int, unsigned char *, cudaStream_t, ErrBuf &) { return RRL_OK; }
int32_t actor_tc_forward_launch(const ActorArgs &, int, cudaStream_t, ErrBuf &err) { err.set("not built"); return RRL_EINVAL; }
size_t actor_pack_bytes(int d, int H) { return actor_tc_pack_bytes(d, H); }
int32_t actor_pack_launch(const float *params, int d, int H, int out, unsigned char *pack, cudaStream_t stream, ErrBuf &err) { return actor_tc_pack_launch(params, d, H, out, pack, stream, err); }
}
