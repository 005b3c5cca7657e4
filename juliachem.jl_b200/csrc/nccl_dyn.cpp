#include "nccl_dyn.h"

#include <dlfcn.h>
#include <cstdlib>
#include <mutex>

namespace jcdf {

static NcclApi g_api;
static std::once_flag g_once;

static void load() {
    // RTLD_NOLOAD first: reuse an already-mapped NCCL (torch's) so that one process never holds two copies.
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    // Nothing mapped yet: the loader would now take the first libnccl.so.2 on its search path, and whatever is mapped first
    // serves every later request for that soname in the process.  A host that may load another NCCL user afterwards (PyTorch
    // ships its own, newer NCCL next to its libraries and fails to import over an older one) names the copy to share in
    // JCDF_NCCL_LIBRARY — juliachem.jl_b200/_lib.py points it at the pip package's library when there is one.
    if (!lib) {
        const char* named = getenv("JCDF_NCCL_LIBRARY");
        if (named && *named) lib = dlopen(named, RTLD_NOW | RTLD_GLOBAL);
    }
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
