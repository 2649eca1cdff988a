// NCCL comparison path (IGG_MODE_NCCL_3PHASE): the reference's structure -- one phase per dimension, pack both sides,
// exchange, unpack (ref: src/update_halo.jl:51-88) -- with the MPI messages replaced by a grouped ncclSend/ncclRecv on
// device buffers.  It exists to be measured against the fused peer-store kernel, not as the product path.
// NCCL is dlopen'ed (libnccl.so.2) so that libigg_b200 has no link-time dependency on it.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <memory>

#include "igg_internal.h"

namespace igg {

namespace {

typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId_ { char internal[128]; };
constexpr int kNcclUint8 = 1;   // ncclDataType_t: ncclInt8 = 0, ncclUint8 = 1 (nccl.h)

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId_*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (api.handle) {
            auto sym = [&](const char* s) { return dlsym(api.handle, s); };
            api.GetUniqueId = (int (*)(ncclUniqueId_*))sym("ncclGetUniqueId");
            api.CommInitRank = (int (*)(ncclComm_t*, int, ncclUniqueId_, int))sym("ncclCommInitRank");
            api.CommDestroy = (int (*)(ncclComm_t))sym("ncclCommDestroy");
            api.GroupStart = (int (*)())sym("ncclGroupStart");
            api.GroupEnd = (int (*)())sym("ncclGroupEnd");
            api.Send = (int (*)(const void*, size_t, int, int, ncclComm_t, cudaStream_t))sym("ncclSend");
            api.Recv = (int (*)(void*, size_t, int, int, ncclComm_t, cudaStream_t))sym("ncclRecv");
            api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
            if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.GroupStart || !api.GroupEnd || !api.Send ||
                !api.Recv) {
                dlclose(api.handle);
                api.handle = nullptr;
            }
        }
    }
    return api.handle ? &api : nullptr;
}

int nccl_fail(NcclApi* a, int rc, const char* what) {
    return fail(IGG_ERR_NCCL, "%s failed: %s", what, (a && a->GetErrorString) ? a->GetErrorString(rc) : "NCCL error");
}

struct PhaseEntry {
    std::vector<FieldKey> key;
    int d = 0;
    TileRec *d_pack = nullptr, *d_unpack = nullptr;
    int n_pack = 0, n_unpack = 0;
    int64_t bytes[2] = {0, 0};
};

}  // namespace

struct NcclPath {
    NcclApi* api = nullptr;
    ncclComm_t comm = nullptr;
    char* sendbuf[2] = {nullptr, nullptr};
    char* recvbuf[2] = {nullptr, nullptr};
    int64_t cap = 0;
    std::vector<std::unique_ptr<PhaseEntry>> cache;
};

size_t nccl_unique_id_size() { return sizeof(ncclUniqueId_); }

int nccl_get_unique_id(void* out) {
    NcclApi* a = nccl_api();
    if (!a) return fail(IGG_ERR_NCCL, "libnccl.so.2 could not be loaded (%s).", dlerror() ? dlerror() : "missing symbols");
    ncclUniqueId_ id;
    const int rc = a->GetUniqueId(&id);
    if (rc) return nccl_fail(a, rc, "ncclGetUniqueId");
    std::memcpy(out, &id, sizeof(id));
    return IGG_OK;
}

int nccl_create(NcclPath** out, int nprocs, int me, const void* id_bytes) {
    NcclApi* a = nccl_api();
    if (!a) return fail(IGG_ERR_NCCL, "libnccl.so.2 could not be loaded.");
    std::unique_ptr<NcclPath> np(new NcclPath());
    np->api = a;
    ncclUniqueId_ id;
    std::memcpy(&id, id_bytes, sizeof(id));
    const int rc = a->CommInitRank(&np->comm, nprocs, id, me);
    if (rc) return nccl_fail(a, rc, "ncclCommInitRank");
    *out = np.release();
    return IGG_OK;
}

void nccl_free(NcclPath* np) {
    if (!np) return;
    cudaDeviceSynchronize();
    for (auto& e : np->cache) { if (e->d_pack) cudaFree(e->d_pack); if (e->d_unpack) cudaFree(e->d_unpack); }
    for (int n = 0; n < 2; ++n) { if (np->sendbuf[n]) cudaFree(np->sendbuf[n]); if (np->recvbuf[n]) cudaFree(np->recvbuf[n]); }
    if (np->comm && np->api) np->api->CommDestroy(np->comm);
    delete np;
}

static PhaseEntry* phase_entry(NcclPath* np, const igg_grid_desc& g, int nfields, const igg_field* fields, int d, int* rc_out) {
    for (auto& e : np->cache) {
        if (e->d != d || (int)e->key.size() != nfields) continue;
        bool same = true;
        for (int i = 0; i < nfields && same; ++i) {
            const FieldKey& k = e->key[i];
            const igg_field& f = fields[i];
            same = k.elsize == f.elsize && k.ndims == f.ndims && k.align == (int32_t)((uintptr_t)f.ptr % 16);
            for (int q = 0; q < 3 && same; ++q) same = k.size[q] == f.size[q] && k.hw[q] == f.halowidths[q];
        }
        if (same) { *rc_out = IGG_OK; return e.get(); }
    }
    std::unique_ptr<PhaseEntry> e(new PhaseEntry());
    e->d = d;
    for (int i = 0; i < nfields; ++i) {
        FieldKey k;
        std::memset(&k, 0, sizeof(k));
        for (int q = 0; q < 3; ++q) { k.size[q] = fields[i].size[q]; k.hw[q] = fields[i].halowidths[q]; }
        k.elsize = fields[i].elsize; k.ndims = fields[i].ndims; k.align = (int32_t)((uintptr_t)fields[i].ptr % 16);
        e->key.push_back(k);
    }
    std::vector<TileRec> pack, unpack;
    int rc = build_dim_phase(g, nfields, fields, d, pack, unpack, e->bytes);
    if (rc) { *rc_out = rc; return nullptr; }
    e->n_pack = (int)pack.size();
    e->n_unpack = (int)unpack.size();
    cudaError_t ce = cudaSuccess;
    if (e->n_pack) {
        ce = cudaMalloc(&e->d_pack, pack.size() * sizeof(TileRec));
        if (ce == cudaSuccess) ce = cudaMemcpy(e->d_pack, pack.data(), pack.size() * sizeof(TileRec), cudaMemcpyHostToDevice);
    }
    if (ce == cudaSuccess && e->n_unpack) {
        ce = cudaMalloc(&e->d_unpack, unpack.size() * sizeof(TileRec));
        if (ce == cudaSuccess) ce = cudaMemcpy(e->d_unpack, unpack.data(), unpack.size() * sizeof(TileRec), cudaMemcpyHostToDevice);
    }
    if (ce != cudaSuccess) { *rc_out = fail(IGG_ERR_CUDA, "NCCL path: tile upload failed: %s", cudaGetErrorString(ce)); return nullptr; }
    np->cache.push_back(std::move(e));
    *rc_out = IGG_OK;
    return np->cache.back().get();
}

// One phase (dimension d, 0-based) whose neighbours are other ranks.  `proto` carries the status word / time-out.
int nccl_phase(NcclPath* np, const igg_grid_desc& g, int nfields, const igg_field* fields, int d, int max_blocks,
               const KParams& proto, void* stream_v, int64_t* launches) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    int rc;
    PhaseEntry* e = phase_entry(np, g, nfields, fields, d, &rc);
    if (!e) return rc;
    if (e->n_pack == 0) return IGG_OK;          // no field has a halo in this dimension
    const int64_t need = std::max(e->bytes[0], e->bytes[1]);
    if (need > np->cap) {
        cudaDeviceSynchronize();
        for (int n = 0; n < 2; ++n) {
            if (np->sendbuf[n]) cudaFree(np->sendbuf[n]);
            if (np->recvbuf[n]) cudaFree(np->recvbuf[n]);
            if (cudaMalloc(&np->sendbuf[n], (size_t)need) != cudaSuccess || cudaMalloc(&np->recvbuf[n], (size_t)need) != cudaSuccess)
                return fail(IGG_ERR_CUDA, "NCCL path: buffer allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        }
        np->cap = need;
    }
    KParams p = proto;
    for (int i = 0; i < nfields; ++i) p.field[i] = fields[i].ptr;
    p.n_push_tiles = 0;
    for (int n = 0; n < 2; ++n) { p.copy_dst[n] = np->sendbuf[n]; p.copy_src[n] = np->recvbuf[n]; }
    // pack (replaces iwrite_sendbufs!, ref: src/CUDAExt/update_halo.jl:161-174)
    p.tiles = e->d_pack;
    p.n_tiles = e->n_pack;
    rc = launch_halo_kernel(p, std::min(e->n_pack, max_blocks * 4), stream);
    if (rc) return rc;
    // exchange (replaces irecv_halo!/isend_halo, ref: src/update_halo.jl:343-367): receives n = 2,1 then sends n = 1,2
    // (ref: src/update_halo.jl:58,61) so that the pairing is right when both neighbours are the same rank.
    NcclApi* a = np->api;
    int nr = a->GroupStart();
    if (nr) return nccl_fail(a, nr, "ncclGroupStart");
    for (int n = 1; n >= 0; --n)
        if (g.neighbors[d][n] >= 0 && e->bytes[n] > 0) {
            nr = a->Recv(np->recvbuf[n], (size_t)e->bytes[n], kNcclUint8, (int)g.neighbors[d][n], np->comm, stream);
            if (nr) return nccl_fail(a, nr, "ncclRecv");
        }
    for (int n = 0; n < 2; ++n)
        if (g.neighbors[d][n] >= 0 && e->bytes[n] > 0) {
            nr = a->Send(np->sendbuf[n], (size_t)e->bytes[n], kNcclUint8, (int)g.neighbors[d][n], np->comm, stream);
            if (nr) return nccl_fail(a, nr, "ncclSend");
        }
    nr = a->GroupEnd();
    if (nr) return nccl_fail(a, nr, "ncclGroupEnd");
    // unpack (replaces iread_recvbufs!, ref: src/CUDAExt/update_halo.jl:190-203)
    p.tiles = e->d_unpack;
    p.n_tiles = e->n_unpack;
    rc = launch_halo_kernel(p, std::min(e->n_unpack, max_blocks * 4), stream);
    if (rc) return rc;
    if (launches) *launches += 2;
    return IGG_OK;
}

}  // namespace igg
