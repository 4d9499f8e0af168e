// z-slab halo exchange over NCCL point-to-point (NVLink 5 / NVSwitch), one rank per GPU.
//
// Replaces the reference's per-stage path  copy_br_to_gc -> cudaMemcpy2DAsync D2H -> MPI_Isend /
// MPI_Irecv -> cudaMemcpy2DAsync H2D -> copy_gc_to_br  (MultiGPU/Burgers3d_Baseline/main.c:208-302,
// Kernels.cu:40-105).  A z-plane is contiguous in memory, so the three boundary planes are sent
// device-to-device straight from the field and received straight into the ghost planes: no pack /
// unpack kernels and no host staging.  Plane indices are the reference's (Kernels.cu:50-52,84-86):
//   send right: planes [nzs-6, nzs-3)  -> right neighbour's ghost planes [0, 3)        (tag 1)
//   send left : planes [3, 6)          -> left  neighbour's ghost planes [nzs-3, nzs)  (tag 5)
// The reference ships the RHS array Lu and re-applies RK on ghosts; the fused kernels have no Lu,
// so the stage OUTPUT u is shipped instead -- the same bits.
//
// NCCL is resolved with dlopen at first use so that the library loads (and single-GPU runs work)
// on hosts without NCCL, and so that inside a PyTorch process the already-loaded NCCL is reused.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <mutex>
#include <stdio.h>
#include <string.h>

#include "../../include/advdiff.h"

namespace adv {

struct NcclApi {
    typedef struct ncclComm* comm_t;
    struct unique_id { char internal[128]; };
    int (*GetUniqueId)(unique_id*);
    int (*CommInitRank)(comm_t*, int, unique_id, int);
    int (*CommDestroy)(comm_t);
    int (*Send)(const void*, size_t, int, int, comm_t, cudaStream_t);
    int (*Recv)(void*, size_t, int, int, comm_t, cudaStream_t);
    int (*AllReduce)(const void*, void*, size_t, int, int, comm_t, cudaStream_t);
    int (*GroupStart)();
    int (*GroupEnd)();
    const char* (*GetErrorString)(int);
    void* handle;

    static NcclApi* get(char* err, size_t errlen)
    {
        static NcclApi api;
        static int state = 0;   // 0 untried, 1 ok, -1 failed
        static std::mutex mu;   // ranks may be threads of one process (drivers/mgpu*.c)
        std::lock_guard<std::mutex> lock(mu);
        if (state == 1) return &api;
        if (state == -1) { snprintf(err, errlen, "NCCL is not loadable (libnccl.so.2)"); return nullptr; }
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { state = -1; snprintf(err, errlen, "dlopen(libnccl.so.2): %s", dlerror()); return nullptr; }
        api.handle = h;
#define NCCL_SYM(field, name) \
        *(void**)(&api.field) = dlsym(h, name); \
        if (!api.field) { state = -1; snprintf(err, errlen, "dlsym(%s) failed", name); return nullptr; }
        NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        NCCL_SYM(CommInitRank, "ncclCommInitRank")
        NCCL_SYM(CommDestroy, "ncclCommDestroy")
        NCCL_SYM(Send, "ncclSend")
        NCCL_SYM(Recv, "ncclRecv")
        NCCL_SYM(AllReduce, "ncclAllReduce")
        NCCL_SYM(GroupStart, "ncclGroupStart")
        NCCL_SYM(GroupEnd, "ncclGroupEnd")
        NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef NCCL_SYM
        state = 1;
        return &api;
    }
};

struct HaloComm {
    NcclApi* api;
    NcclApi::comm_t comm;
    int nranks, rank;

    static int unique_id(unsigned char id[128], char* err, size_t errlen)
    {
        NcclApi* a = NcclApi::get(err, errlen);
        if (!a) return ADV_ERR_COMM;
        NcclApi::unique_id u;
        int rc = a->GetUniqueId(&u);
        if (rc) { snprintf(err, errlen, "ncclGetUniqueId: %s", a->GetErrorString(rc)); return ADV_ERR_COMM; }
        memcpy(id, u.internal, 128);
        return ADV_OK;
    }

    int init(int nranks_, int rank_, const unsigned char id[128], char* err, size_t errlen)
    {
        api = NcclApi::get(err, errlen);
        if (!api) return ADV_ERR_COMM;
        nranks = nranks_; rank = rank_;
        NcclApi::unique_id u;
        memcpy(u.internal, id, 128);
        int rc = api->CommInitRank(&comm, nranks, u, rank);
        if (rc) { comm = nullptr; snprintf(err, errlen, "ncclCommInitRank: %s", api->GetErrorString(rc)); return ADV_ERR_COMM; }
        return ADV_OK;
    }

    void destroy()
    {
        if (api && comm) api->CommDestroy(comm);
        comm = nullptr;
    }

    // max over ranks of one non-negative double given by its IEEE bits (ordered like the values): the dynamic
    // CFL step's max|u| (LFWENO5FDM3d.m:71 takes it over the whole field)
    int allreduce_max_u64(const unsigned long long* src, unsigned long long* dst, cudaStream_t st, char* err, size_t errlen)
    {
        if (!comm) { snprintf(err, errlen, "all-reduce without a communicator"); return ADV_ERR_COMM; }
        const int kUint64 = 5, kMax = 2;    // ncclUint64, ncclMax
        int rc = api->AllReduce(src, dst, 1, kUint64, kMax, comm, st);
        if (rc) { snprintf(err, errlen, "ncclAllReduce: %s", api->GetErrorString(rc)); return ADV_ERR_COMM; }
        return ADV_OK;
    }

    // field: this rank's stored slab (nzs planes of plane_elems elements); exchanges 3 planes per side
    int exchange(void* field, size_t plane_elems, int nzs, size_t esz, cudaStream_t st, char* err, size_t errlen)
    {
        if (!comm) { snprintf(err, errlen, "halo exchange without a communicator"); return ADV_ERR_COMM; }
        const int R = ADV_RADIUS;
        const size_t bytes = (size_t)R * plane_elems * esz;
        char* base = (char*)field;
        const size_t plane_b = plane_elems * esz;
        const int kChar = 0;    // ncclChar: ship raw bytes (REAL is float or double)
        int rc = api->GroupStart();
        if (!rc && rank < nranks - 1) {
            rc = api->Send(base + (size_t)(nzs - 2 * R) * plane_b, bytes, kChar, rank + 1, comm, st);
            if (!rc) rc = api->Recv(base + (size_t)(nzs - R) * plane_b, bytes, kChar, rank + 1, comm, st);
        }
        if (!rc && rank > 0) {
            rc = api->Send(base + (size_t)R * plane_b, bytes, kChar, rank - 1, comm, st);
            if (!rc) rc = api->Recv(base, bytes, kChar, rank - 1, comm, st);
        }
        int rc2 = api->GroupEnd();
        if (rc || rc2) { snprintf(err, errlen, "NCCL halo exchange: %s", api->GetErrorString(rc ? rc : rc2)); return ADV_ERR_COMM; }
        return ADV_OK;
    }
};

}  // namespace adv
