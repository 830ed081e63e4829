#include "nccl_dl.h"

#include <dlfcn.h>

#include <mutex>

namespace rrl {

NcclApi *nccl_api(const char **why)
{
    static NcclApi api;
    static const char *fail = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        // RTLD_NOLOAD first: inside a torch process this picks the copy torch already mapped.
        void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) {
            fail = "libnccl.so.2 not found (dlopen)";
            return;
        }
        api.lib = lib;
#define RRL_SYM(field, name)                                        \
    *reinterpret_cast<void **>(&api.field) = dlsym(lib, name);     \
    if (!api.field) {                                               \
        fail = "libnccl.so.2 lacks " name;                          \
        return;                                                     \
    }
        RRL_SYM(GetUniqueId, "ncclGetUniqueId")
        RRL_SYM(CommInitRank, "ncclCommInitRank")
        RRL_SYM(CommDestroy, "ncclCommDestroy")
        RRL_SYM(AllReduce, "ncclAllReduce")
        RRL_SYM(AllGather, "ncclAllGather")
        RRL_SYM(GetErrorString, "ncclGetErrorString")
#undef RRL_SYM
    });
    if (fail) {
        if (why) *why = fail;
        return nullptr;
    }
    return &api;
}

}  // namespace rrl
