// sail_comm.h -- NCCL, opened at run time (dlopen("libnccl.so.2"), like NVRTC in sail_jit.cpp): libsail_b200.so has no
// link-time dependency on it, and a single-GPU host never loads it.  Used by the sharded contexts of sail_api.cu for the
// one exchange the path has: combining the members' accumulation buffers on the root when an image is asked for.
#ifndef SAIL_COMM_H
#define SAIL_COMM_H
#include <stddef.h>

typedef struct ncclComm* sailNcclComm_t;
struct SailNcclId { char internal[128]; };  // ncclUniqueId (passed by value)

enum { SAIL_NCCL_INT8 = 0, SAIL_NCCL_UINT8 = 1, SAIL_NCCL_FLOAT32 = 7 };  // ncclDataType_t
enum { SAIL_NCCL_SUM = 0 };                                                // ncclRedOp_t

struct SailNccl {
    bool ok;
    char why[256];
    int version;  // e.g. 22809
    int (*getUniqueId)(SailNcclId*);
    int (*commInitRank)(sailNcclComm_t*, int, SailNcclId, int);
    int (*commInitAll)(sailNcclComm_t*, int, const int*);
    int (*commDestroy)(sailNcclComm_t);
    int (*reduce)(const void*, void*, size_t, int, int, int, sailNcclComm_t, void*);
    int (*allReduce)(const void*, void*, size_t, int, int, sailNcclComm_t, void*);
    int (*allGather)(const void*, void*, size_t, int, sailNcclComm_t, void*);
    int (*send)(const void*, size_t, int, int, sailNcclComm_t, void*);
    int (*recv)(void*, size_t, int, int, sailNcclComm_t, void*);
    int (*groupStart)(void);
    int (*groupEnd)(void);
    const char* (*getErrorString)(int);
};
// Loads the library once per process (SAIL_B200_NCCL=<path> overrides the search).  Never null; check ->ok.
const SailNccl* sail_nccl(void);
#endif
