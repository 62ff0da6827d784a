// sail_comm.cpp -- run-time binding of NCCL (see sail_comm.h).
#include "sail_comm.h"

#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

namespace {
SailNccl g_nccl;
std::once_flag g_once;

void loadNccl() {
    SailNccl& n = g_nccl;
    memset(&n, 0, sizeof(n));
    // The soname first: when the host process has already mapped a libnccl.so.2 (PyTorch bundles one), dlopen returns that
    // copy instead of loading a second NCCL into the process.
    const char* names[] = {getenv("SAIL_B200_NCCL"), "libnccl.so.2", "/usr/lib/x86_64-linux-gnu/libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    char tried[200];
    tried[0] = 0;
    for (const char* name : names) {
        if (!name || !name[0]) continue;
        h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
        snprintf(tried + strlen(tried), sizeof(tried) - strlen(tried), "%s ", name);
    }
    if (!h) { snprintf(n.why, sizeof(n.why), "NCCL not found (tried %s)", tried); return; }
#define SYM(field, name)                                                            \
    *(void**)(&n.field) = dlsym(h, name);                                           \
    if (!n.field) { snprintf(n.why, sizeof(n.why), "NCCL: missing symbol %s", name); return; }
    SYM(getUniqueId, "ncclGetUniqueId")
    SYM(commInitRank, "ncclCommInitRank")
    SYM(commInitAll, "ncclCommInitAll")
    SYM(commDestroy, "ncclCommDestroy")
    SYM(reduce, "ncclReduce")
    SYM(allReduce, "ncclAllReduce")
    SYM(allGather, "ncclAllGather")
    SYM(send, "ncclSend")
    SYM(recv, "ncclRecv")
    SYM(groupStart, "ncclGroupStart")
    SYM(groupEnd, "ncclGroupEnd")
    SYM(getErrorString, "ncclGetErrorString")
#undef SYM
    if (int (*ver)(int*) = (int (*)(int*))dlsym(h, "ncclGetVersion")) ver(&n.version);
    n.ok = true;
}
}  // namespace

const SailNccl* sail_nccl(void) {
    std::call_once(g_once, loadNccl);
    return &g_nccl;
}
