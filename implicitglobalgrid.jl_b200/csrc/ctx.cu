// Context, rank bootstrap (CUDA IPC), update_halo and gather entry points of libigg_b200.
//
// Ownership mirrors the reference's module-level buffer/stream singletons (ref: src/update_halo.jl:103-107,
// src/CUDAExt/update_halo.jl:19-25,148-151) but lives in an explicit context: the context owns the control segment
// (arrival flags + counters), the double-buffered receive staging, cached plans and peer mappings; user arrays are
// only borrowed for the duration of a call.
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdlib>
#include <memory>
#include <mutex>

#include "ctrl_shm.h"
#include "igg_internal.h"

#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) return igg::fail(IGG_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

namespace igg {

constexpr uint64_t kBlobMagic = 0x31424747495f4742ull;   // "BG_IGGB1"

struct BootBlob {
    uint64_t magic;
    uint64_t proc_nonce;          // equal for ranks living in the same process (raw pointers are valid then)
    int32_t rank, device;
    cudaIpcMemHandle_t ctrl_handle;
    uint64_t ctrl_ptr;
    char shm_name[64];
    char pad[96];
};
static_assert(sizeof(BootBlob) == 256, "BootBlob size");
static_assert(sizeof(DeviceState) <= kFlagsOff, "DeviceState must fit before the flags");

struct SegPayload {               // travels through the mailbox when a segment is (re)allocated
    uint64_t proc_nonce;
    int32_t device, status;
    cudaIpcMemHandle_t handle;
    uint64_t ptr;
    int64_t bytes;                // staging: bytes of ONE half; gather: buffer bytes
    char msg[kMailPayload - 96];
};
static_assert(sizeof(SegPayload) == kMailPayload, "SegPayload size");

struct PeerSeg {
    char* ptr = nullptr;
    bool ipc = false;
    int64_t bytes = 0;
    uint64_t raw = 0;             // the owner's own address of the segment (identity check)
};

struct Peer {
    uint64_t proc_nonce = 0;
    int device = -1;
    PeerSeg ctrl, staging, gather;
};

static uint64_t process_nonce() {
    static uint64_t nonce = [] {
        uint64_t n = 0;
        FILE* f = fopen("/dev/urandom", "rb");
        if (f) { if (fread(&n, sizeof(n), 1, f) != 1) n = 0; fclose(f); }
        if (!n) n = ((uint64_t)getpid() << 32) ^ 0x9e3779b97f4a7c15ull;
        return n;
    }();
    return nonce;
}

}  // namespace igg

using namespace igg;

struct igg_ctx {
    igg_grid_desc g;
    int device = 0;
    int mode = IGG_MODE_IPC_LL128;              // default protocol (IGG_MODE=fence selects IGG_MODE_IPC_1PHASE)
    char* ctrl = nullptr;                       // device: flags at 0, counters at kCountersOff
    char* staging = nullptr;                    // device: two halves
    int64_t staging_half = 0;
    int* h_status = nullptr;                    // host-mapped status word
    int* d_status = nullptr;
    ControlPlane cp;
    std::string shm_name;
    std::vector<Peer> peers;
    bool bootstrapped = false;
    std::vector<std::unique_ptr<Plan>> plans;
    int max_blocks = 0;
    int64_t launches = 0;
    cudaStream_t last_stream = nullptr;
    cudaEvent_t last_event = nullptr;
    bool have_last = false;
    // Staging / flag REGIONS, one per neighbour offset, identical on every rank (they only depend on the sequence of
    // plans seen, and update_halo is collective): bytes inside one staging half and arrival-flag words.  Grow-only.
    int64_t  region_cap[IGG_NOFFSETS] = {0}, region_off[IGG_NOFFSETS] = {0};
    uint32_t flag_cap[IGG_NOFFSETS] = {0}, flag_off[IGG_NOFFSETS] = {0};
    cudaStream_t side_stream = nullptr;         // host-resident fields: the wrap-lines kernel runs beside the exchange
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    unsigned long long timeout_ns = 20ull * 1000000000ull;
    struct PtrEntry { const void* user = nullptr; void* resolved = nullptr; bool host = false; };
    PtrEntry ptr_cache[64];
    // gather
    char* gather_buf = nullptr;   int64_t gather_bytes = 0;       // root: device image of A_global (peer mapped)
    char* gather_src = nullptr;   int64_t gather_src_bytes = 0;   // staging for host-resident A
    TileRec* d_gather_tiles = nullptr; int gather_tiles_cap = 0;
    // NCCL comparison path (nccl_path.cu)
    NcclPath* nccl = nullptr;
    // optional kernel timeline (igg_trace_enable)
    unsigned long long* d_trace = nullptr;
    int last_grid = 0;
};

// ---- helpers -----------------------------------------------------------------------------------------------------

static int open_segment(igg_ctx* ctx, const Peer& peer, const cudaIpcMemHandle_t& h, uint64_t raw_ptr, PeerSeg& out) {
    if (peer.proc_nonce == process_nonce()) {         // same process: the raw pointer is valid, just enable P2P
        if (peer.device != ctx->device) {
            cudaError_t e = cudaDeviceEnablePeerAccess(peer.device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return fail(IGG_ERR_IPC, "cudaDeviceEnablePeerAccess(%d) failed: %s", peer.device, cudaGetErrorString(e));
        }
        out.ptr = reinterpret_cast<char*>(raw_ptr);
        out.ipc = false;
        return IGG_OK;
    }
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return fail(IGG_ERR_IPC, "cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
    out.ptr = static_cast<char*>(p);
    out.ipc = true;
    return IGG_OK;
}

static void close_segment(PeerSeg& s) {
    if (s.ptr && s.ipc) cudaIpcCloseMemHandle(s.ptr);
    s = PeerSeg();
}

// Collective (re)allocation of a peer-visible segment: everyone frees the old one, allocates `bytes`, and learns
// everyone else's new mapping through the mailbox.  `which`: 0 = staging, 1 = gather buffer (root only allocates).
static int exchange_segment(igg_ctx* ctx, int which, int64_t my_bytes, int status, const char* msg,
                            std::vector<SegPayload>& all_out) {
    SegPayload mine;
    std::memset(&mine, 0, sizeof(mine));
    mine.proc_nonce = process_nonce();
    mine.device = ctx->device;
    mine.status = status;
    if (msg) snprintf(mine.msg, sizeof(mine.msg), "%s", msg);
    char* ptr = (which == 0) ? ctx->staging : ctx->gather_buf;
    if (ptr && my_bytes > 0) {
        CU(cudaIpcGetMemHandle(&mine.handle, ptr));
        mine.ptr = reinterpret_cast<uint64_t>(ptr);
    }
    mine.bytes = my_bytes;
    std::vector<char> all;
    int rc = ctx->cp.allgather(&mine, all);
    if (rc) return rc;
    all_out.resize(ctx->g.nprocs);
    std::memcpy(all_out.data(), all.data(), all.size());
    return IGG_OK;
}

static int grow_staging(igg_ctx* ctx, int64_t need_half) {
    // Nobody may still write into / read from the old staging: my kernels are done, then everyone's are.
    CU(cudaDeviceSynchronize());
    int rc = ctx->cp.barrier();
    if (rc) return rc;
    for (auto& p : ctx->peers) close_segment(p.staging);
    rc = ctx->cp.barrier();                     // every mapping of my old staging is closed before I free it
    if (rc) return rc;
    if (ctx->staging) { CU(cudaFree(ctx->staging)); ctx->staging = nullptr; }
    int64_t half = std::max<int64_t>(need_half + need_half / 4, 1 << 20);
    half = (half + 0xfffff) & ~int64_t(0xfffff);
    CU(cudaMalloc(&ctx->staging, (size_t)(2 * half)));
    // Touch the new memory and drain the device before any peer learns its handle: nothing of the allocation's
    // set-up may still be pending when a neighbour's kernel starts storing into it.
    CU(cudaMemset(ctx->staging, 0, (size_t)(2 * half)));
    CU(cudaDeviceSynchronize());
    ctx->staging_half = half;
    std::vector<SegPayload> all;
    rc = exchange_segment(ctx, 0, half, 0, nullptr, all);
    if (rc) return rc;
    for (int r = 0; r < ctx->g.nprocs; ++r) {
        if (r == ctx->g.me) continue;
        rc = open_segment(ctx, ctx->peers[r], all[r].handle, all[r].ptr, ctx->peers[r].staging);
        if (rc) return rc;
        ctx->peers[r].staging.bytes = all[r].bytes;
    }
    return IGG_OK;
}

// Makes sure every neighbour offset's region (staging bytes, arrival flags) is large enough for `plan`.  Regions keep
// their place from call to call, so data a neighbour pushes for one offset can never land on memory another
// offset's neighbour writes or this rank still reads -- also when calls with different field sets or `dims=`
// alternate and the ranks are not in lock-step.  Growing is collective and rare (first use of a bigger plan): every
// rank reaches the same decision at the same call because the needs depend on the call's arguments only.
static int ensure_layout(igg_ctx* ctx, const Plan& plan) {
    bool grow = false;
    for (int o = 0; o < IGG_NOFFSETS; ++o)
        grow = grow || plan.region_need[o] > ctx->region_cap[o] || plan.flag_need[o] > ctx->flag_cap[o];
    if (!grow) return IGG_OK;
    int64_t cap[IGG_NOFFSETS], total = 0;
    uint64_t fcap[IGG_NOFFSETS], ftotal = 0;
    for (int o = 0; o < IGG_NOFFSETS; ++o) {
        cap[o] = ctx->region_cap[o];
        if (plan.region_need[o] > cap[o]) cap[o] = (plan.region_need[o] + plan.region_need[o] / 4 + 127) / 128 * 128;
        fcap[o] = ctx->flag_cap[o];
        if (plan.flag_need[o] > fcap[o]) fcap[o] = plan.flag_need[o];
        total += cap[o];
        ftotal += fcap[o];
    }
    if (ftotal > (uint64_t)kMaxFlags)
        return fail(IGG_ERR_ARG, "update_halo: the calls made so far need %llu arrival flags, the control segment holds %d; "
                    "use the LL128 protocol (default) or split the call.", (unsigned long long)ftotal, kMaxFlags);
    int rc;
    if (total > ctx->staging_half) {
        rc = grow_staging(ctx, total);              // quiesces all ranks, re-allocates, re-maps
    } else {
        CU(cudaDeviceSynchronize());                // nobody exchanges while the regions move
        rc = ctx->cp.barrier();
    }
    if (rc) return rc;
    total = 0; ftotal = 0;
    for (int o = 0; o < IGG_NOFFSETS; ++o) {
        ctx->region_cap[o] = cap[o];  ctx->region_off[o] = total;          total += cap[o];
        ctx->flag_cap[o] = (uint32_t)fcap[o]; ctx->flag_off[o] = (uint32_t)ftotal; ftotal += fcap[o];
    }
    return IGG_OK;
}

// Resolves a user pointer to an address the kernels can use: device memory as is, mapped page-locked host memory
// through its device alias.  Pageable host memory is rejected (no CPU fallback, no hidden copies in update_halo).
static int resolve_pointer(igg_ctx* ctx, const void* p, void** out, bool* is_host) {
    igg_ctx::PtrEntry& ce = ctx->ptr_cache[((uintptr_t)p >> 8) & 63];
    if (ce.user == p) { *out = ce.resolved; if (is_host) *is_host = ce.host; return IGG_OK; }
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(IGG_ERR_CUDA, "cudaPointerGetAttributes failed: %s", cudaGetErrorString(e)); }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) {
        *out = const_cast<void*>(p);
        if (is_host) *is_host = false;
        ce.user = p; ce.resolved = *out; ce.host = false;
        return IGG_OK;
    }
    if (a.type == cudaMemoryTypeHost && a.devicePointer) {
        *out = a.devicePointer;
        if (is_host) *is_host = true;
        ce.user = p; ce.resolved = *out; ce.host = true;
        return IGG_OK;
    }
    return fail(IGG_ERR_UNSUPPORTED,
                "pointer %p is pageable host memory: pass device memory or page-locked (pinned/registered) host memory.", p);
}

static Plan* find_or_build_plan(igg_ctx* ctx, int nfields, const igg_field* fields, int dims_mask, const uint8_t* host_lines,
                                int* rc_out) {
    const bool ll128 = ctx->mode == IGG_MODE_IPC_LL128;
    for (auto& up : ctx->plans) {
        Plan* p = up.get();
        if (p->dims_mask != dims_mask || (int)p->key.size() != nfields || p->ll128 != ll128) continue;
        bool same = true;
        for (int i = 0; i < nfields && same; ++i) {
            const FieldKey& k = p->key[i];
            const igg_field& f = fields[i];
            same = k.elsize == f.elsize && k.ndims == f.ndims && k.align == (int32_t)((uintptr_t)f.ptr % 16) &&
                   k.lines == ((host_lines && host_lines[i]) ? 1 : 0);
            for (int d = 0; d < 3 && same; ++d) same = k.size[d] == f.size[d] && k.hw[d] == f.halowidths[d];
        }
        if (same) { *rc_out = IGG_OK; return p; }
    }
    std::unique_ptr<Plan> plan(new Plan());
    int rc = build_plan(ctx->g, nfields, fields, dims_mask, *plan, host_lines, ll128);
    if (rc) { *rc_out = rc; return nullptr; }
    if (!plan->tiles.empty()) {
        cudaError_t e = cudaMalloc(&plan->d_tiles, plan->tiles.size() * sizeof(TileRec));
        if (e == cudaSuccess)
            e = cudaMemcpy(plan->d_tiles, plan->tiles.data(), plan->tiles.size() * sizeof(TileRec), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { *rc_out = fail(IGG_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e)); return nullptr; }
        plan->grid = (int)std::min<int64_t>((int64_t)plan->tiles.size(), ctx->max_blocks);
    }
    if (!plan->wrap_tiles.empty()) {
        cudaError_t e = cudaMalloc(&plan->d_wrap_tiles, plan->wrap_tiles.size() * sizeof(TileRec));
        if (e == cudaSuccess)
            e = cudaMemcpy(plan->d_wrap_tiles, plan->wrap_tiles.data(), plan->wrap_tiles.size() * sizeof(TileRec), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { *rc_out = fail(IGG_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e)); return nullptr; }
    }
    if (ctx->plans.size() >= 64) {   // bound the cache (fields of ever-changing shapes)
        if (ctx->plans.front()->d_tiles || ctx->plans.front()->d_wrap_tiles) {
            cudaDeviceSynchronize();
            cudaFree(ctx->plans.front()->d_tiles);
            cudaFree(ctx->plans.front()->d_wrap_tiles);
        }
        ctx->plans.erase(ctx->plans.begin());
    }
    ctx->plans.push_back(std::move(plan));
    *rc_out = IGG_OK;
    return ctx->plans.back().get();
}

static void order_after_previous(igg_ctx* ctx, cudaStream_t stream) {
    // Calls on one context must execute in issue order (the staging double-buffer relies on it).  Same stream: stream
    // order does it for free.  Different stream: make the new stream wait for the tail of the previous one.
    if (ctx->have_last && ctx->last_stream != stream) {
        cudaEventRecord(ctx->last_event, ctx->last_stream);
        cudaStreamWaitEvent(stream, ctx->last_event, 0);
    }
}

static void remember_launch(igg_ctx* ctx, cudaStream_t stream) {
    ctx->last_stream = stream;
    ctx->have_last = true;
    ctx->launches += 1;
}

static int check_status(igg_ctx* ctx) {
    if (*ctx->h_status != 0) {
        const int s = *ctx->h_status;
        *ctx->h_status = 0;
        return fail(s, "update_halo: a neighbour rank did not deliver its halo within %.1f s (device-side time-out).",
                    ctx->timeout_ns * 1e-9);
    }
    return IGG_OK;
}

// ---- C ABI -------------------------------------------------------------------------------------------------------
extern "C" {

int igg_create(const igg_grid_desc* desc, igg_ctx** out) {
    if (!desc || !out) return fail(IGG_ERR_ARG, "igg_create: NULL argument.");
    int rc = check_grid(*desc);
    if (rc) return rc;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(IGG_ERR_CUDA, "No usable CUDA device (%s); libigg_b200 has no CPU fallback.", cudaGetErrorString(e));
    }
    // ref: src/select_device.jl:30
    if (desc->device < 0 || desc->device >= ndev)
        return fail(IGG_ERR_ARG, "More processes have been launched per node than there are GPUs available.");
    std::unique_ptr<igg_ctx> ctx(new igg_ctx());
    ctx->g = *desc;
    ctx->device = desc->device;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMalloc(&ctx->ctrl, kCtrlBytes));
    CU(cudaMemset(ctx->ctrl, 0, kCtrlBytes));
    CU(cudaHostAlloc(&ctx->h_status, sizeof(int), cudaHostAllocMapped));
    *ctx->h_status = 0;
    CU(cudaHostGetDevicePointer(&ctx->d_status, ctx->h_status, 0));
    CU(cudaEventCreateWithFlags(&ctx->last_event, cudaEventDisableTiming));
    ctx->max_blocks = halo_kernel_max_coresident_blocks(ctx->device);
    if (ctx->max_blocks <= 0) return fail(IGG_ERR_CUDA, "occupancy query for the halo kernel failed (was the library built for sm_100a?): %s", cudaGetErrorString(cudaGetLastError()));
    if (const char* t = getenv("IGG_DEVICE_TIMEOUT_S")) ctx->timeout_ns = (unsigned long long)(atof(t) * 1e9);
    // IGG_MODE=fence / ll128: the IPC protocol a new context starts in (default ll128; the same on every rank).
    if (const char* m = getenv("IGG_MODE")) {
        if (!strcmp(m, "ll128")) ctx->mode = IGG_MODE_IPC_LL128;
        else if (!strcmp(m, "fence") || !strcmp(m, "ipc")) ctx->mode = IGG_MODE_IPC_1PHASE;
    }
    ctx->peers.resize(desc->nprocs);
    if (desc->nprocs > 1) {
        rc = ctx->cp.create(desc->me, desc->nprocs, ctx->shm_name);
        if (rc) return rc;
    } else {
        ctx->bootstrapped = true;
    }
    CU(cudaDeviceSynchronize());
    *out = ctx.release();
    return IGG_OK;
}

int igg_destroy(igg_ctx* ctx) {
    if (!ctx) return IGG_OK;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    if (ctx->g.nprocs > 1 && ctx->cp.attached()) ctx->cp.barrier();   // nobody is still writing into my segments
    nccl_free(ctx->nccl);
    ctx->nccl = nullptr;
    for (auto& p : ctx->peers) { close_segment(p.ctrl); close_segment(p.staging); close_segment(p.gather); }
    for (auto& pl : ctx->plans) { cudaFree(pl->d_tiles); cudaFree(pl->d_wrap_tiles); }
    if (ctx->staging) cudaFree(ctx->staging);
    if (ctx->gather_buf) cudaFree(ctx->gather_buf);
    if (ctx->gather_src) cudaFree(ctx->gather_src);
    if (ctx->d_gather_tiles) cudaFree(ctx->d_gather_tiles);
    if (ctx->d_trace) cudaFree(ctx->d_trace);
    if (ctx->ctrl) cudaFree(ctx->ctrl);
    if (ctx->h_status) cudaFreeHost(ctx->h_status);
    if (ctx->last_event) cudaEventDestroy(ctx->last_event);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->side_stream) cudaStreamDestroy(ctx->side_stream);
    ctx->cp.close();
    delete ctx;
    return IGG_OK;
}

size_t igg_bootstrap_size(void) { return sizeof(BootBlob); }

int igg_bootstrap_export(igg_ctx* ctx, void* blob) {
    if (!ctx || !blob) return fail(IGG_ERR_ARG, "igg_bootstrap_export: NULL argument.");
    CU(cudaSetDevice(ctx->device));
    BootBlob b;
    std::memset(&b, 0, sizeof(b));
    b.magic = kBlobMagic;
    b.proc_nonce = process_nonce();
    b.rank = ctx->g.me;
    b.device = ctx->device;
    CU(cudaIpcGetMemHandle(&b.ctrl_handle, ctx->ctrl));
    b.ctrl_ptr = reinterpret_cast<uint64_t>(ctx->ctrl);
    snprintf(b.shm_name, sizeof(b.shm_name), "%s", ctx->shm_name.c_str());
    std::memcpy(blob, &b, sizeof(b));
    return IGG_OK;
}

int igg_bootstrap_import(igg_ctx* ctx, const void* all_blobs, int nprocs) {
    if (!ctx || !all_blobs) return fail(IGG_ERR_ARG, "igg_bootstrap_import: NULL argument.");
    if (nprocs != ctx->g.nprocs) return fail(IGG_ERR_ARG, "igg_bootstrap_import: got %d blobs for %d ranks.", nprocs, ctx->g.nprocs);
    if (ctx->bootstrapped) return IGG_OK;
    CU(cudaSetDevice(ctx->device));
    const BootBlob* blobs = static_cast<const BootBlob*>(all_blobs);
    std::vector<std::string> names(nprocs);
    for (int r = 0; r < nprocs; ++r) {
        BootBlob b;
        std::memcpy(&b, blobs + r, sizeof(b));
        if (b.magic != kBlobMagic || b.rank != r) return fail(IGG_ERR_ARG, "igg_bootstrap_import: blob %d is corrupt or out of rank order.", r);
        names[r] = b.shm_name;
        if (r == ctx->g.me) continue;
        ctx->peers[r].proc_nonce = b.proc_nonce;
        ctx->peers[r].device = b.device;
        int rc = open_segment(ctx, ctx->peers[r], b.ctrl_handle, b.ctrl_ptr, ctx->peers[r].ctrl);
        if (rc) return rc;
    }
    int rc = ctx->cp.attach(names);
    if (rc) return rc;
    ctx->bootstrapped = true;
    return IGG_OK;
}

int igg_set_mode(igg_ctx* ctx, int mode) {
    if (!ctx) return fail(IGG_ERR_ARG, "igg_set_mode: NULL context.");
    if (mode != IGG_MODE_IPC_1PHASE && mode != IGG_MODE_NCCL_3PHASE && mode != IGG_MODE_IPC_LL128)
        return fail(IGG_ERR_ARG, "igg_set_mode: unknown mode %d.", mode);
    if (mode == IGG_MODE_NCCL_3PHASE && ctx->g.nprocs > 1 && !ctx->nccl)
        return fail(IGG_ERR_STATE, "igg_set_mode: call igg_nccl_init before selecting IGG_MODE_NCCL_3PHASE.");
    const bool was_ll = ctx->mode == IGG_MODE_IPC_LL128, is_ll = mode == IGG_MODE_IPC_LL128;
    if (was_ll != is_ll && ctx->g.nprocs > 1 && ctx->bootstrapped) {
        // Entering or leaving the line format is collective: the other protocol's payload may sit where LL128 expects
        // a line's epoch word, so the staging memory of every rank is cleared while nobody exchanges.
        CU(cudaSetDevice(ctx->device));
        CU(cudaDeviceSynchronize());
        int rc = ctx->cp.barrier();
        if (rc) return rc;
        if (ctx->staging) CU(cudaMemset(ctx->staging, 0, (size_t)(2 * ctx->staging_half)));
        CU(cudaDeviceSynchronize());
        rc = ctx->cp.barrier();
        if (rc) return rc;
    }
    ctx->mode = mode;
    return IGG_OK;
}

// The fused single-phase exchange for the dimensions in `dims_mask` (plan lookup, staging growth, one launch).
static int launch_exchange(igg_ctx* ctx, int nfields, const igg_field* fields, int dims_mask, cudaStream_t stream,
                           const uint8_t* host_lines = nullptr) {
    int rc;
    Plan* plan = find_or_build_plan(ctx, nfields, fields, dims_mask, host_lines, &rc);
    if (!plan) return rc;
    bool join_side = false;
    if (!plan->wrap_tiles.empty()) {
        // Whole-line x self-wrap of host-resident fields: its own kernel, independent of the other tiles of the call.
        // It runs before the exchange kernel on the caller's stream.  IGG_HOST_FORK=1 runs it beside the exchange
        // kernel on a side stream instead (fork / join with events); measured SLOWER on cfg2 (e2e 674 vs 570 us: the
        // two kernels fight over the same PCIe read tags), so it is off by default.
        static const bool fork = std::getenv("IGG_HOST_FORK") && std::atoi(std::getenv("IGG_HOST_FORK")) == 1;
        void* ptrs[IGG_MAX_FIELDS];
        for (int i = 0; i < nfields; ++i) ptrs[i] = fields[i].ptr;
        order_after_previous(ctx, stream);
        cudaStream_t ws = stream;
        if (fork && !plan->tiles.empty()) {
            if (!ctx->side_stream) {
                CU(cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking));
                CU(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
                CU(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
            }
            CU(cudaEventRecord(ctx->ev_fork, stream));
            CU(cudaStreamWaitEvent(ctx->side_stream, ctx->ev_fork, 0));
            ws = ctx->side_stream;
            join_side = true;
        }
        rc = launch_wrap_lines_kernel(plan->d_wrap_tiles, (int)plan->wrap_tiles.size(), ptrs, nfields, ws);
        if (rc) return rc;
        if (join_side) CU(cudaEventRecord(ctx->ev_join, ws));
        remember_launch(ctx, stream);
    }
    if (plan->tiles.empty()) return IGG_OK;                       // e.g. one rank, nothing periodic
    if (plan->n_flags > (uint32_t)kMaxFlags)
        return fail(IGG_ERR_ARG, "update_halo: this call needs %u arrival flags, the control segment holds %d; split the call.",
                    plan->n_flags, kMaxFlags);
    if (plan->active_mask) {
        rc = ensure_layout(ctx, *plan);
        if (rc) return rc;
    }

    KParams p;
    std::memset(&p, 0, sizeof(p));
    p.tiles = plan->d_tiles;
    p.n_tiles = (int32_t)plan->tiles.size();
    p.n_push_tiles = plan->n_push_tiles;
    for (int i = 0; i < nfields; ++i) p.field[i] = fields[i].ptr;
    p.local_flags = reinterpret_cast<const unsigned long long*>(ctx->ctrl + kFlagsOff);
    p.state = reinterpret_cast<DeviceState*>(ctx->ctrl);       // epoch + staging parity live on the device
    p.active_mask = plan->active_mask;
    p.unpack_half = ctx->staging_half;
    p.status = ctx->d_status;
    p.timeout_ns = ctx->timeout_ns;
    for (int oi = 0; oi < IGG_NOFFSETS; ++oi) {
        if (!((plan->active_mask >> oi) & 1u)) continue;
        const Peer& peer = ctx->peers[plan->peer[oi]];
        // The tiles carry offsets / flag numbers of the plan's own packed layout; shift every offset's base to that
        // offset's region (same table on every rank).  What I push at offset oi lands in the receiver's region of
        // the opposite offset.
        const int ro = opposite_index(oi);
        p.push_base[oi] = peer.staging.ptr + (ctx->region_off[ro] - plan->region_pack[ro]);
        p.push_half[oi] = peer.staging.bytes;
        p.unpack_base[oi] = ctx->staging + (ctx->region_off[oi] - plan->region_pack[oi]);
        p.remote_flags[oi] = reinterpret_cast<unsigned long long*>(peer.ctrl.ptr + kFlagsOff) +
                             ((long long)ctx->flag_off[ro] - (long long)plan->flag_pack[ro]);
        p.local_flag_delta[oi] = (int32_t)((long long)ctx->flag_off[oi] - (long long)plan->flag_pack[oi]);
    }
    order_after_previous(ctx, stream);
    if (ctx->d_trace) {
        p.trace = ctx->d_trace;
        ctx->last_grid = plan->grid;
        CU(cudaMemsetAsync(ctx->d_trace, 0, (size_t)ctx->max_blocks * 8 * sizeof(unsigned long long), stream));
    }
    rc = launch_halo_kernel(p, plan->grid, stream, plan->ll128);
    if (join_side) CU(cudaStreamWaitEvent(stream, ctx->ev_join, 0));   // the call is complete when both kernels are
    if (rc) return rc;
    remember_launch(ctx, stream);
    return IGG_OK;
}

int igg_update_halo(igg_ctx* ctx, int nfields, const igg_field* fields_in, const int* dims_order, int ndims_order,
                    void* stream_v, int flags) {
    if (!ctx || !fields_in) return fail(IGG_ERR_ARG, "igg_update_halo: NULL argument.");
    if (!ctx->bootstrapped) return fail(IGG_ERR_STATE, "igg_update_halo: call igg_bootstrap_import first (nprocs > 1).");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    int dims_mask;
    int rc = dims_mask_from_order(dims_order, ndims_order, &dims_mask);
    if (rc) return rc;
    rc = check_fields(ctx->g, nfields, fields_in);
    if (rc) return rc;
    CU(cudaSetDevice(ctx->device));

    igg_field fields[IGG_MAX_FIELDS];
    uint8_t host_lines[IGG_MAX_FIELDS];
    static const bool host_lines_enabled = !(std::getenv("IGG_HOST_LINES") && std::atoi(std::getenv("IGG_HOST_LINES")) == 0);
    for (int i = 0; i < nfields; ++i) {
        fields[i] = fields_in[i];
        bool is_host = false;
        rc = resolve_pointer(ctx, fields_in[i].ptr, &fields[i].ptr, &is_host);
        if (rc) return rc;
        // Fields in page-locked host memory: whole-line x self-wrap (kOpWrapLines) unless IGG_HOST_LINES=0.
        host_lines[i] = (is_host && host_lines_enabled && (uintptr_t)fields[i].ptr % 64 == 0) ? 1 : 0;
    }

    if (ctx->mode == IGG_MODE_NCCL_3PHASE && ctx->g.nprocs > 1) {
        // Comparison path: the reference's dependent phases, in the caller's order (default z, x, y).
        static const int kDefaultOrder[3] = {3, 1, 2};
        const int* order = (dims_order && ndims_order > 0) ? dims_order : kDefaultOrder;
        const int norder = (dims_order && ndims_order > 0) ? ndims_order : 3;
        KParams proto;
        std::memset(&proto, 0, sizeof(proto));
        proto.local_flags = reinterpret_cast<const unsigned long long*>(ctx->ctrl + kFlagsOff);
        proto.status = ctx->d_status;
        proto.timeout_ns = ctx->timeout_ns;
        order_after_previous(ctx, stream);
        for (int q = 0; q < norder; ++q) {
            const int d = order[q] - 1;
            const int64_t l = ctx->g.neighbors[d][0], r = ctx->g.neighbors[d][1];
            if (l < 0 && r < 0) continue;
            if (l == ctx->g.me && r == ctx->g.me) rc = launch_exchange(ctx, nfields, fields, 1 << d, stream, host_lines);   // local wrap
            else rc = nccl_phase(ctx->nccl, ctx->g, nfields, fields, d, ctx->max_blocks, proto, stream, &ctx->launches);
            if (rc) return rc;
        }
        ctx->last_stream = stream;
        ctx->have_last = true;
    } else {
        rc = launch_exchange(ctx, nfields, fields, dims_mask, stream, host_lines);
        if (rc) return rc;
    }
    if (!(flags & IGG_ASYNC)) {
        CU(cudaStreamSynchronize(stream));
        return check_status(ctx);
    }
    return IGG_OK;
}

// gather! -- ref: src/gather.jl:25-54.  Every rank stores its block straight into root's device image of A_global
// through a CUDA-IPC peer mapping (3-D strided stores, contiguous x-rows); root then hands the image to the caller's
// array (device-to-device or device-to-host).
int igg_gather(igg_ctx* ctx, const void* A, const int64_t sizeA_in[3], int ndimsA, void* A_global,
               const int64_t sizeG_in[3], int ndimsG, int elsize, int root, void* stream_v) {
    if (!ctx || !A || !sizeA_in) return fail(IGG_ERR_ARG, "igg_gather: NULL argument.");
    if (!ctx->bootstrapped) return fail(IGG_ERR_STATE, "igg_gather: call igg_bootstrap_import first (nprocs > 1).");
    if (root < 0 || root >= ctx->g.nprocs) return fail(IGG_ERR_ARG, "igg_gather: invalid root %d.", root);
    if (elsize < 1 || ndimsA < 1 || ndimsA > 3) return fail(IGG_ERR_ARG, "igg_gather: invalid elsize / ndims.");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    CU(cudaSetDevice(ctx->device));
    const igg_grid_desc& g = ctx->g;
    const bool is_root = (g.me == root);
    int64_t sizeA[3] = {1, 1, 1};
    for (int d = 0; d < ndimsA; ++d) sizeA[d] = sizeA_in[d];
    const int64_t bytesA = sizeA[0] * sizeA[1] * sizeA[2] * elsize;

    // Root validates (ref: src/gather.jl:27-33) and tells everyone the verdict, so that no rank hangs.
    int status = 0;
    char msg[kMailPayload - 96] = {0};
    int64_t sizeG[3] = {1, 1, 1};
    if (is_root) {
        if (!A_global) { status = IGG_ERR_ARG; snprintf(msg, sizeof(msg), "The input argument `A_global` can't be `nothing` on the root."); }
        else if (!sizeG_in || ndimsG < 1) { status = IGG_ERR_ARG; snprintf(msg, sizeof(msg), "igg_gather: sizeG missing on root."); }
        else if (ndimsA > ndimsG) { status = IGG_ERR_ARG; snprintf(msg, sizeof(msg), "The number of dimension of `A` must be less than or equal to the number of dimensions of `A_global`."); }
        else if (ndimsG > 3) { status = IGG_ERR_ARG; snprintf(msg, sizeof(msg), "The number of dimensions of `A_global` must be less than or equal to the number of dimensions of the Cartesian grid of MPI processes."); }
        else {
            for (int d = 0; d < ndimsG; ++d) sizeG[d] = sizeG_in[d];
            for (int d = 0; d < 3; ++d) {
                const int64_t want = (d < ndimsG ? g.dims[d] : 1) * sizeA[d];
                if (d < ndimsG ? sizeG[d] != want : false) status = IGG_ERR_ARG;
            }
            // ranks beyond ndimsG dimensions would overwrite each other, the reference requires dims[1:N] only; blocks of
            // ranks with a non-zero coordinate in a dimension >= ndimsG land on the same place (as with MPI displacements).
            if (status) snprintf(msg, sizeof(msg), "The size of the global array `size(A_global)` must be equal to the product of `size(A)` and `dims`.");
            for (int d = ndimsG; d < 3 && !status; ++d)
                if (g.dims[d] > 1) { status = IGG_ERR_ARG; snprintf(msg, sizeof(msg), "gather!: the process grid has more than one rank in a dimension beyond ndims(A_global)."); }
        }
    }

    const int64_t bytesG = sizeG[0] * sizeG[1] * sizeG[2] * elsize;
    if (g.nprocs == 1) {
        if (status) return fail(status, "%s", msg);
        CU(cudaMemcpyAsync(A_global, A, (size_t)bytesA, cudaMemcpyDefault, stream));
        CU(cudaStreamSynchronize(stream));
        return IGG_OK;
    }

    // Round 1 of the hand-shake: root's verdict and whether its device image of A_global must grow.
    const bool regrow_mine = is_root && !status && bytesG > ctx->gather_bytes;
    std::vector<SegPayload> all;
    int rc;
    {
        const int64_t keep = ctx->gather_bytes;
        rc = exchange_segment(ctx, 1, regrow_mine ? -1 : keep, status, msg, all);
        if (rc) return rc;
    }
    if (all[root].status) return fail(all[root].status, "%s", all[root].msg[0] ? all[root].msg : "gather!: root reported an error.");
    if (all[root].bytes < 0) {
        // Root replaces its buffer: peers drop the old mapping first, then root reallocates and publishes the new one.
        if (!is_root) close_segment(ctx->peers[root].gather);
        rc = ctx->cp.barrier();
        if (rc) return rc;
        if (is_root) {
            CU(cudaDeviceSynchronize());
            if (ctx->gather_buf) { CU(cudaFree(ctx->gather_buf)); ctx->gather_buf = nullptr; }
            CU(cudaMalloc(&ctx->gather_buf, (size_t)bytesG));
            ctx->gather_bytes = bytesG;
        }
        rc = exchange_segment(ctx, 1, ctx->gather_bytes, 0, nullptr, all);
        if (rc) return rc;
    }
    const SegPayload& rp = all[root];
    char* dst_base = nullptr;
    if (is_root) {
        dst_base = ctx->gather_buf;
    } else {
        Peer& pr = ctx->peers[root];
        if (pr.gather.ptr && pr.gather.raw != rp.ptr) close_segment(pr.gather);
        if (!pr.gather.ptr) {
            rc = open_segment(ctx, pr, rp.handle, rp.ptr, pr.gather);
            if (rc) return rc;
            pr.gather.bytes = rp.bytes;
            pr.gather.raw = rp.ptr;
        }
        dst_base = pr.gather.ptr;
    }

    // Source block: device memory as is, host memory through a device staging copy.
    const char* src = nullptr;
    {
        cudaPointerAttributes a;
        cudaError_t e = cudaPointerGetAttributes(&a, A);
        if (e != cudaSuccess) { cudaGetLastError(); return fail(IGG_ERR_CUDA, "cudaPointerGetAttributes failed: %s", cudaGetErrorString(e)); }
        if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) {
            src = static_cast<const char*>(A);
        } else {
            if (bytesA > ctx->gather_src_bytes) {
                CU(cudaDeviceSynchronize());
                if (ctx->gather_src) { CU(cudaFree(ctx->gather_src)); ctx->gather_src = nullptr; }
                CU(cudaMalloc(&ctx->gather_src, (size_t)bytesA));
                ctx->gather_src_bytes = bytesA;
            }
            CU(cudaMemcpyAsync(ctx->gather_src, A, (size_t)bytesA, cudaMemcpyHostToDevice, stream));
            src = ctx->gather_src;
        }
    }

    // Global sizes are known to every rank from dims and sizeA (ref: src/gather.jl:33).
    int64_t gsz[2] = {sizeA[0] * g.dims[0], sizeA[1] * g.dims[1]};
    int64_t origin[3] = {g.coords[0] * sizeA[0], g.coords[1] * sizeA[1], g.coords[2] * sizeA[2]};
    std::vector<TileRec> tiles;
    build_block_copy_tiles(elsize, sizeA, gsz, origin, (int64_t)((uintptr_t)src % 16), 0, tiles);
    if ((int)tiles.size() > ctx->gather_tiles_cap) {
        CU(cudaDeviceSynchronize());
        if (ctx->d_gather_tiles) CU(cudaFree(ctx->d_gather_tiles));
        CU(cudaMalloc(&ctx->d_gather_tiles, tiles.size() * sizeof(TileRec)));
        ctx->gather_tiles_cap = (int)tiles.size();
    }
    CU(cudaMemcpyAsync(ctx->d_gather_tiles, tiles.data(), tiles.size() * sizeof(TileRec), cudaMemcpyHostToDevice, stream));
    KParams p;
    std::memset(&p, 0, sizeof(p));
    p.tiles = ctx->d_gather_tiles;
    p.n_tiles = (int32_t)tiles.size();
    p.n_push_tiles = 0;
    p.field[0] = const_cast<char*>(src);
    p.push_base[0] = dst_base;
    p.local_flags = reinterpret_cast<const unsigned long long*>(ctx->ctrl + kFlagsOff);
    p.status = ctx->d_status;
    p.timeout_ns = ctx->timeout_ns;
    order_after_previous(ctx, stream);
    rc = launch_halo_kernel(p, (int)std::min<int64_t>((int64_t)tiles.size(), (int64_t)ctx->max_blocks * 4), stream);
    if (rc) return rc;
    remember_launch(ctx, stream);
    CU(cudaStreamSynchronize(stream));       // my block has landed in root's memory
    rc = ctx->cp.barrier();                  // ... and so has everyone else's
    if (rc) return rc;
    if (is_root) {
        CU(cudaMemcpyAsync(A_global, ctx->gather_buf, (size_t)bytesG, cudaMemcpyDefault, stream));
        CU(cudaStreamSynchronize(stream));
    }
    return ctx->cp.barrier();                // root's image may be overwritten by the next gather only after this
}

int igg_debug_state(igg_ctx* ctx, uint64_t out[32]) {
    if (!ctx || !out) return fail(IGG_ERR_ARG, "igg_debug_state: NULL argument.");
    CU(cudaSetDevice(ctx->device));
    CU(cudaDeviceSynchronize());
    DeviceState st;
    CU(cudaMemcpy(&st, ctx->ctrl, sizeof(st), cudaMemcpyDeviceToHost));
    out[0] = st.call_epoch;
    for (int o = 0; o < IGG_NOFFSETS; ++o) out[1 + o] = st.exchanges[o];
    out[28] = st.blocks_done;
    out[29] = (uint64_t)ctx->staging_half;
    out[30] = (uint64_t)ctx->launches;
    out[31] = 0;
    return IGG_OK;
}

int igg_trace_enable(igg_ctx* ctx, int on) {
    if (!ctx) return fail(IGG_ERR_ARG, "igg_trace_enable: NULL context.");
    CU(cudaSetDevice(ctx->device));
    CU(cudaDeviceSynchronize());
    if (on && !ctx->d_trace) CU(cudaMalloc(&ctx->d_trace, (size_t)ctx->max_blocks * 8 * sizeof(unsigned long long)));
    if (!on && ctx->d_trace) { CU(cudaFree(ctx->d_trace)); ctx->d_trace = nullptr; }
    return IGG_OK;
}

int igg_trace_read(igg_ctx* ctx, uint64_t* out, int cap_blocks, int* nblocks_out) {
    if (!ctx || !out || !nblocks_out) return fail(IGG_ERR_ARG, "igg_trace_read: NULL argument.");
    if (!ctx->d_trace) return fail(IGG_ERR_STATE, "igg_trace_read: tracing is not enabled.");
    CU(cudaSetDevice(ctx->device));
    CU(cudaDeviceSynchronize());
    const int n = ctx->last_grid < cap_blocks ? ctx->last_grid : cap_blocks;
    CU(cudaMemcpy(out, ctx->d_trace, (size_t)n * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    *nblocks_out = n;
    return IGG_OK;
}

size_t igg_nccl_unique_id_size(void) { return nccl_unique_id_size(); }

int igg_nccl_get_unique_id(void* id_out) {
    if (!id_out) return fail(IGG_ERR_ARG, "igg_nccl_get_unique_id: NULL argument.");
    return nccl_get_unique_id(id_out);
}

int igg_nccl_init(igg_ctx* ctx, const void* id) {
    if (!ctx || !id) return fail(IGG_ERR_ARG, "igg_nccl_init: NULL argument.");
    if (ctx->nccl) return IGG_OK;
    CU(cudaSetDevice(ctx->device));
    return nccl_create(&ctx->nccl, ctx->g.nprocs, ctx->g.me, id);
}

int64_t igg_launch_count(const igg_ctx* ctx) { return ctx ? ctx->launches : 0; }
const char* igg_kernel_name(void) {
    // the kernel a new context launches: LL128 unless IGG_MODE selects the fence protocol (as in igg_create)
    const char* m = getenv("IGG_MODE");
    const bool fence = m && (!strcmp(m, "fence") || !strcmp(m, "ipc"));
    return fence ? halo_kernel_name() : "igg_halo_exchange_ll128_kernel";
}

}  // extern "C"
