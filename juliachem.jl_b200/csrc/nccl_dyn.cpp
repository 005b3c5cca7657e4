#include "nccl_dyn.h"

#include <dlfcn.h>
#include <mutex>

namespace jcdf {

static NcclApi g_api;
static std::once_flag g_once;

static void load() {
    // RTLD_NOLOAD first: reuse an already-mapped NCCL (torch's) so that one process never holds two copies.
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
        g_api.why = "libnccl.so.2 could not be loaded (dlopen)";
        return;
    }
#define BIND(field, sym)                                                   \
    g_api.field = reinterpret_cast<decltype(g_api.field)>(dlsym(lib, sym)); \
    if (!g_api.field) {                                                    \
        g_api.why = "missing NCCL symbol " sym;                            \
        return;                                                            \
    }
    BIND(GetUniqueId, "ncclGetUniqueId")
    BIND(CommInitRank, "ncclCommInitRank")
    BIND(CommInitAll, "ncclCommInitAll")
    BIND(CommDestroy, "ncclCommDestroy")
    BIND(AllReduce, "ncclAllReduce")
    BIND(Broadcast, "ncclBroadcast")
    BIND(GroupStart, "ncclGroupStart")
    BIND(GroupEnd, "ncclGroupEnd")
    BIND(GetErrorString, "ncclGetErrorString")
    BIND(GetVersion, "ncclGetVersion")
#undef BIND
    g_api.ok = true;
}

const NcclApi& nccl() {
    std::call_once(g_once, load);
    return g_api;
}

}  // namespace jcdf
