// Context, rank bootstrap (CUDA IPC), update_halo and gather entry points of libigg_b200.
//
// Ownership mirrors the reference's module-level buffer/stream singletons (ref: src/update_halo.jl:103-107,
// src/CUDAExt/update_halo.jl:19-25,148-151) but lives in an explicit context: the context owns the control segment
// (arrival flags + counters), the double-buffered receive staging, cached plans and peer mappings; user arrays are
// only borrowed for the duration of a call.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <time.h>
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
    cudaIpcMemHandle_t handle = {};   // ... and its IPC handle (the same address may be handed out again for another allocation)
};

struct Peer {
    uint64_t proc_nonce = 0;
    int device = -1;
    PeerSeg ctrl, staging, gather;
};

// Base address and size of the device allocation `p` lies in (driver API, looked up at run time: the library links
// the CUDA runtime only).  false if it cannot be told.
static bool device_address_range(const void* p, void** base, size_t* bytes) {
    typedef int (*fn_t)(unsigned long long*, size_t*, unsigned long long);
    static fn_t fn = [] {
        void* h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        return h ? reinterpret_cast<fn_t>(dlsym(h, "cuMemGetAddressRange_v2")) : nullptr;
    }();
    if (!fn) return false;
    unsigned long long b = 0;
    size_t n = 0;
    if (fn(&b, &n, (unsigned long long)(uintptr_t)p) != 0 || !b) return false;
    *base = reinterpret_cast<void*>((uintptr_t)b);
    *bytes = n;
    return true;
}

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

struct igg_prepared;

struct igg_ctx {
    igg_grid_desc g;
    int device = 0;
    int mode = IGG_MODE_IPC_AUTO;               // default: protocol chosen per call by size (IGG_MODE=fence|ll128 forces one)
    char* ctrl = nullptr;                       // device: exchange state at 0, arrival flags at kFlagsOff
    char* staging = nullptr;                    // device: two halves
    int64_t staging_half = 0;
    int* h_status = nullptr;                    // host-mapped status word
    int* d_status = nullptr;
    unsigned long long* h_done = nullptr;       // host-mapped completion word of host-synchronous calls
    unsigned long long* d_done = nullptr;
    unsigned long long ticket = 0;              // value the next host-synchronous launch stores into *h_done
    bool spin_wait = true;                      // IGG_BLOCKING=sync: wait with cudaStreamSynchronize instead
    bool ll128_ok = true;                       // false: link check or self-test said LL128 lines may tear -> fence protocol only
    unsigned long long ll128_test[3] = {0, 0, 0};   // self-test: torn line images, time-outs, line images checked
    ControlPlane cp;
    std::string shm_name;
    std::vector<Peer> peers;
    bool bootstrapped = false;
    std::vector<std::unique_ptr<Plan>> plans;
    std::vector<igg_prepared*> prepared;        // live prepared calls (they hold Plan pointers of this context)
    int max_blocks = 0;
    int64_t launches = 0;
    cudaStream_t last_stream = nullptr;
    cudaEvent_t last_event = nullptr;
    bool have_last = false;
    // Staging / flag REGIONS, one per neighbour offset and protocol ([0] fence, [1] LL128), identical on every rank
    // (they only depend on the sequence of plans seen, and update_halo is collective): bytes inside one staging half
    // and arrival-flag words.  Grow-only.  A half is [LL128 regions | fence regions]: LL128 lines are only ever
    // polled where LL128 lines are written, so payload of the other protocol cannot be mistaken for an epoch word.
    int64_t  region_cap[2][IGG_NOFFSETS] = {{0}}, region_off[2][IGG_NOFFSETS] = {{0}};
    uint32_t flag_cap[IGG_NOFFSETS] = {0}, flag_off[IGG_NOFFSETS] = {0};
    char* hostx_cache = nullptr; int64_t hostx_cache_cap = 0;     // row-end lines of host-resident fields (exchanged x-faces)
    std::vector<char*> hostx_old_caches;        // outgrown caches (captured launches may still use them)
    cudaStream_t side_stream = nullptr;         // host-resident fields: the wrap-lines kernel runs beside the exchange
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    unsigned long long timeout_ns = 20ull * 1000000000ull;
    // Pointer classification cache.  Only entries whose device alias equals the user pointer are kept, so a stale
    // entry (memory freed and the address handed out again as another kind of memory) can at worst pick the other of
    // two equivalent kernels for an x self-wrap, never a wrong address.
    struct PtrEntry { const void* user = nullptr; bool host = false; };
    PtrEntry ptr_cache[64];
    // Symmetric field allocations (igg_alloc_field): the k-th one is "the same array" on every rank, mapped into every
    // rank's address space, so that a neighbour can store into its copy of a field directly (direct-to-halo).
    struct FieldReg { char* base = nullptr; int64_t bytes = 0; std::vector<PeerSeg> peer; };   // peer[r]: rank r's copy, mapped here
    std::vector<FieldReg> regs;                 // index = slot (the same on every rank: allocation is collective)
    char** d_reg_table = nullptr;               // device: [kMaxFieldRegs][nprocs] base addresses
    bool direct_enabled = true;                 // IGG_DIRECT=0: never store into a neighbour's field
    // gather
    char* gather_buf = nullptr;   int64_t gather_bytes = 0;       // root: device image of A_global (peer mapped)
    char* gather_src = nullptr;   int64_t gather_src_bytes = 0;   // staging for host-resident A
    TileRec* d_gather_tiles = nullptr; int gather_tiles_cap = 0;
    int64_t gather_tile_key[8] = {0}; int gather_ntiles = 0;     // what the table on the device was built for
    // NCCL comparison path (nccl_path.cu)
    NcclPath* nccl = nullptr;
    // optional kernel timeline (igg_trace_enable)
    unsigned long long* d_trace = nullptr;
    int last_grid = 0;
};

// A validated update_halo call kept for reuse (igg_prepare): resolved field pointers and the plan.
struct igg_prepared {
    igg_ctx* ctx = nullptr;
    int nfields = 0;
    igg_field fields[IGG_MAX_FIELDS];           // pointers resolved to device addresses
    igg_field user_fields[IGG_MAX_FIELDS];      // as passed in (for the NCCL comparison path)
    uint8_t host_lines[IGG_MAX_FIELDS];
    uint8_t direct[IGG_MAX_FIELDS];
    int dims_mask = 7;
    int dims_order[3] = {3, 1, 2};
    int ndims_order = 3;
    int mode = -1;                              // ctx->mode the plan was chosen under
    Plan* plan = nullptr;
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

// The per-plan launch constants for the context's current staging layout.
static void fill_plan_dev(const igg_ctx* ctx, const Plan& plan, PlanDev& d) {
    std::memset(&d, 0, sizeof(d));
    const int pr = plan.ll128 ? 1 : 0;
    d.active_mask = plan.active_mask;
    d.direct_mask = plan.direct_mask;
    d.unpack_half = ctx->staging_half;
    for (int oi = 0; oi < IGG_NOFFSETS; ++oi) {
        if (!((plan.active_mask >> oi) & 1u)) continue;
        const Peer& peer = ctx->peers[plan.peer[oi]];
        d.peer_rank[oi] = plan.peer[oi];
        d.remote_ready[oi] = reinterpret_cast<unsigned long long*>(peer.ctrl.ptr + kReadyOff) + opposite_index(oi);
        // The tiles carry offsets / flag numbers of the plan's own packed layout; shift every offset's base to that
        // offset's region (same table on every rank).  What I push at offset oi lands in the receiver's region of
        // the opposite offset.
        const int ro = opposite_index(oi);
        d.push_base[oi] = peer.staging.ptr + (ctx->region_off[pr][ro] - plan.region_pack[ro]);
        d.push_half[oi] = peer.staging.bytes;
        d.unpack_base[oi] = ctx->staging + (ctx->region_off[pr][oi] - plan.region_pack[oi]);
        d.remote_flags[oi] = reinterpret_cast<unsigned long long*>(peer.ctrl.ptr + kFlagsOff) +
                             ((long long)ctx->flag_off[ro] - (long long)plan.flag_pack[ro]);
        d.local_flag_delta[oi] = (int32_t)((long long)ctx->flag_off[oi] - (long long)plan.flag_pack[oi]);
    }
}

// Host -> device copy of a small table that is complete when the call returns (a pageable-memory cudaMemcpy may
// return while its DMA is still in flight, and the caller's stream need not be ordered against the legacy stream).
static cudaError_t upload_now(void* dst, const void* src, size_t bytes) {
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, cudaStreamLegacy);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);
    return e;
}

static int upload_plan_dev(igg_ctx* ctx, Plan& plan) {
    if (!plan.d_plan) return IGG_OK;
    PlanDev d;
    fill_plan_dev(ctx, plan, d);
    const cudaError_t e = upload_now(plan.d_plan, &d, sizeof(d));
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "plan constants upload failed: %s", cudaGetErrorString(e));
    return IGG_OK;
}

// Makes sure every neighbour offset's region (staging bytes, arrival flags) is large enough for `plan`.  Regions keep
// their place from call to call, so data a neighbour pushes for one offset can never land on memory another
// offset's neighbour writes or this rank still reads -- also when calls with different field sets or `dims=`
// alternate and the ranks are not in lock-step.  Growing is collective and rare (first use of a bigger plan): every
// rank reaches the same decision at the same call because the needs depend on the call's arguments only.  It blocks
// the host (device synchronisation + barrier across ranks), also for IGG_ASYNC calls.
static int ensure_layout(igg_ctx* ctx, const Plan& plan) {
    const int pr = plan.ll128 ? 1 : 0;
    bool grow = false;
    for (int o = 0; o < IGG_NOFFSETS; ++o)
        grow = grow || plan.region_need[o] > ctx->region_cap[pr][o] || plan.flag_need[o] > ctx->flag_cap[o];
    if (!grow) return IGG_OK;
    int64_t cap[2][IGG_NOFFSETS], total = 0;
    uint64_t fcap[IGG_NOFFSETS], ftotal = 0;
    for (int q = 0; q < 2; ++q)
        for (int o = 0; o < IGG_NOFFSETS; ++o) {
            cap[q][o] = ctx->region_cap[q][o];
            if (q == pr && plan.region_need[o] > cap[q][o])
                cap[q][o] = (plan.region_need[o] + plan.region_need[o] / 4 + 127) / 128 * 128;
            total += cap[q][o];
        }
    for (int o = 0; o < IGG_NOFFSETS; ++o) {
        fcap[o] = ctx->flag_cap[o];
        if (plan.flag_need[o] > fcap[o]) fcap[o] = plan.flag_need[o];
        ftotal += fcap[o];
    }
    if (ftotal > (uint64_t)kMaxFlags)
        return fail(IGG_ERR_ARG, "update_halo: the calls made so far need %llu arrival flags, the control segment holds %d; "
                    "use the LL128 protocol or split the call.", (unsigned long long)ftotal, kMaxFlags);
    int rc;
    if (total > ctx->staging_half) {
        rc = grow_staging(ctx, total);              // quiesces all ranks, re-allocates (zeroed), re-maps
        if (rc) return rc;
    } else {
        CU(cudaDeviceSynchronize());                // nobody exchanges while the regions move
        rc = ctx->cp.barrier();
        if (rc) return rc;
        // An LL128 region may now lie where another protocol's payload was: no stale word may look like an epoch.
        if (ctx->staging) CU(cudaMemset(ctx->staging, 0, (size_t)(2 * ctx->staging_half)));
        CU(cudaDeviceSynchronize());
        rc = ctx->cp.barrier();
        if (rc) return rc;
    }
    total = 0; ftotal = 0;
    for (int q = 1; q >= 0; --q)                    // [LL128 regions | fence regions]
        for (int o = 0; o < IGG_NOFFSETS; ++o) { ctx->region_cap[q][o] = cap[q][o]; ctx->region_off[q][o] = total; total += cap[q][o]; }
    for (int o = 0; o < IGG_NOFFSETS; ++o) { ctx->flag_cap[o] = (uint32_t)fcap[o]; ctx->flag_off[o] = (uint32_t)ftotal; ftotal += fcap[o]; }
    // Every cached plan's launch constants follow the new layout (launches captured into CUDA graphs stay valid:
    // they hold the pointer to the constants, not the constants).
    for (auto& up : ctx->plans) {
        rc = upload_plan_dev(ctx, *up);
        if (rc) return rc;
    }
    return IGG_OK;
}

// Resolves a user pointer to an address the kernels can use: device memory as is, mapped page-locked host memory
// through its device alias.  Pageable host memory is rejected (no CPU fallback, no hidden copies in update_halo).
static int resolve_pointer(igg_ctx* ctx, const void* p, void** out, bool* is_host) {
    igg_ctx::PtrEntry& ce = ctx->ptr_cache[((uintptr_t)p >> 8) & 63];
    if (ce.user == p) { *out = const_cast<void*>(p); if (is_host) *is_host = ce.host; return IGG_OK; }
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(IGG_ERR_CUDA, "cudaPointerGetAttributes failed: %s", cudaGetErrorString(e)); }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) {
        *out = const_cast<void*>(p);
        if (is_host) *is_host = false;
        ce.user = p; ce.host = false;
        return IGG_OK;
    }
    if (a.type == cudaMemoryTypeHost && a.devicePointer) {
        *out = a.devicePointer;
        if (is_host) *is_host = true;
        if (a.devicePointer == p) { ce.user = p; ce.host = true; }
        return IGG_OK;
    }
    return fail(IGG_ERR_UNSUPPORTED,
                "pointer %p is pageable host memory: pass device memory or page-locked (pinned/registered) host memory.", p);
}

static int proto_of_mode(int mode) {
    return mode == IGG_MODE_IPC_LL128 ? kProtoLL128 : mode == IGG_MODE_IPC_AUTO ? kProtoAuto : kProtoFence;
}

// Which symmetric allocation (igg_alloc_field) a device address lies in: slot index or -1; *off = offset inside it.
static int find_reg(const igg_ctx* ctx, const void* p, long long* off) {
    const char* c = static_cast<const char*>(p);
    for (size_t k = 0; k < ctx->regs.size(); ++k) {
        const igg_ctx::FieldReg& r = ctx->regs[k];
        if (r.base && c >= r.base && c < r.base + r.bytes) { if (off) *off = c - r.base; return (int)k; }
    }
    return -1;
}

static Plan* find_or_build_plan(igg_ctx* ctx, int nfields, const igg_field* fields, int dims_mask, const uint8_t* host_lines,
                                int* rc_out, const uint8_t* direct = nullptr) {
    const int proto = proto_of_mode(ctx->mode);
    for (auto& up : ctx->plans) {
        Plan* p = up.get();
        if (p->dims_mask != dims_mask || (int)p->key.size() != nfields || p->proto_req != proto) continue;
        bool same = true;
        for (int i = 0; i < nfields && same; ++i) {
            const FieldKey& k = p->key[i];
            const igg_field& f = fields[i];
            same = k.elsize == f.elsize && k.ndims == f.ndims && k.align == (int32_t)((uintptr_t)f.ptr % 16) &&
                   k.lines == (((host_lines && host_lines[i]) ? 1 : 0) | ((direct && direct[i]) ? 2 : 0));
            for (int d = 0; d < 3 && same; ++d) same = k.size[d] == f.size[d] && k.hw[d] == f.halowidths[d];
        }
        if (same) { *rc_out = IGG_OK; return p; }
    }
    std::unique_ptr<Plan> plan(new Plan());
    int rc = build_plan(ctx->g, nfields, fields, dims_mask, *plan, host_lines, proto, direct);
    if (rc) { *rc_out = rc; return nullptr; }
    plan->proto_req = proto;
    {
        // One allocation: [PlanDev | tile table].
        const size_t head = (sizeof(PlanDev) + 255) & ~size_t(255);
        char* mem = nullptr;
        cudaError_t e = cudaMalloc(&mem, head + std::max<size_t>(plan->tiles.size(), 1) * sizeof(TileRec));
        if (e == cudaSuccess && !plan->tiles.empty())
            e = upload_now(mem + head, plan->tiles.data(), plan->tiles.size() * sizeof(TileRec));
        if (e != cudaSuccess) { if (mem) cudaFree(mem); *rc_out = fail(IGG_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e)); return nullptr; }
        plan->d_plan = reinterpret_cast<PlanDev*>(mem);
        plan->d_tiles = reinterpret_cast<TileRec*>(mem + head);
        plan->grid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)plan->tiles.size(), ctx->max_blocks));
    }
    if (!plan->wrap_tiles.empty()) {
        cudaError_t e = cudaMalloc(&plan->d_wrap_tiles, plan->wrap_tiles.size() * sizeof(TileRec));
        if (e == cudaSuccess) e = upload_now(plan->d_wrap_tiles, plan->wrap_tiles.data(), plan->wrap_tiles.size() * sizeof(TileRec));
        if (e != cudaSuccess) { cudaFree(plan->d_plan); *rc_out = fail(IGG_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e)); return nullptr; }
    }
    if (!plan->hostx_tiles.empty()) {
        cudaError_t e = cudaMalloc(&plan->d_hostx_tiles, plan->hostx_tiles.size() * sizeof(TileRec));
        if (e == cudaSuccess) e = upload_now(plan->d_hostx_tiles, plan->hostx_tiles.data(), plan->hostx_tiles.size() * sizeof(TileRec));
        if (e != cudaSuccess) { cudaFree(plan->d_plan); cudaFree(plan->d_wrap_tiles); *rc_out = fail(IGG_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e)); return nullptr; }
    }
    // Bound the cache (fields of ever-changing shapes): drop the oldest plan that nothing can still refer to.  Plans
    // behind a prepared call, or launched while their stream was being captured into a CUDA graph, are never dropped.
    if (ctx->plans.size() >= 256) {
        for (size_t k = 0; k < ctx->plans.size(); ++k) {
            if (ctx->plans[k]->pinned) continue;
            cudaDeviceSynchronize();
            cudaFree(ctx->plans[k]->d_plan);
            cudaFree(ctx->plans[k]->d_wrap_tiles);
            cudaFree(ctx->plans[k]->d_hostx_tiles);
            ctx->plans.erase(ctx->plans.begin() + (long)k);
            break;
        }
    }
    // Launch constants for the layout as it is now (ensure_layout uploads them again whenever the layout moves).
    rc = upload_plan_dev(ctx, *plan);
    if (rc) { cudaFree(plan->d_plan); cudaFree(plan->d_wrap_tiles); *rc_out = rc; return nullptr; }
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

// A device-side time-out of an earlier launch (also of a stream-ordered one nobody waited for) is reported by the
// next call of any kind.
static int check_status(igg_ctx* ctx) {
    if (*reinterpret_cast<volatile int*>(ctx->h_status) != 0) {
        const int s = *ctx->h_status;
        *ctx->h_status = 0;
        return fail(s, "update_halo: a neighbour rank did not deliver its halo within %.1f s (device-side time-out).",
                    ctx->timeout_ns * 1e-9);
    }
    return IGG_OK;
}

static int make_current(igg_ctx* ctx) {
    int cur = -1;
    if (cudaGetDevice(&cur) == cudaSuccess && cur == ctx->device) return IGG_OK;
    CU(cudaSetDevice(ctx->device));
    return IGG_OK;
}

// Host-synchronous completion.  `ticket` != 0: the launch stores it into the mapped word *h_done when its last block
// has finished, and the host polls that word (a stream synchronisation costs several microseconds more); the stream
// is queried now and then so that a failed launch or a fault cannot hang the caller.
static int wait_done(igg_ctx* ctx, cudaStream_t stream, unsigned long long ticket) {
    if (ticket) {
        for (unsigned spins = 1;; ++spins) {
            if (__atomic_load_n(ctx->h_done, __ATOMIC_ACQUIRE) == ticket) return check_status(ctx);
            if ((spins & 0xfffu) == 0) {
                const cudaError_t q = cudaStreamQuery(stream);
                if (q == cudaSuccess) return check_status(ctx);
                if (q != cudaErrorNotReady) return fail(IGG_ERR_CUDA, "update_halo: %s", cudaGetErrorString(q));
            }
        }
    }
    CU(cudaStreamSynchronize(stream));
    return check_status(ctx);
}

// Is LL128 safe between these ranks?  (1) Every peer must sit behind NVLink: a PCIe peer path may split a 128-byte
// warp store (native peer atomics are the attribute that tells the two apart).  (2) The assumption itself is put to
// the test between ring neighbours (ll128_check.cu).  The verdicts are combined over all ranks, so that every rank
// ends up with the same protocol; a failure downgrades the context to the fence protocol and says so once.
// IGG_LL128_SELFTEST=0 skips the test, IGG_LL128_SELFTEST_FAIL=1 makes it fail on purpose (to exercise the fallback).
static int check_ll128(igg_ctx* ctx) {
    const int n = ctx->g.nprocs, me = ctx->g.me;
    bool ok = true, same_gpu = false;
    char why[160] = "";
    for (int r = 0; r < n; ++r) {
        if (r == me) continue;
        const Peer& p = ctx->peers[r];
        if (p.device == ctx->device && p.proc_nonce == process_nonce()) { same_gpu = true; continue; }
        int native = 0;
        if (p.device != ctx->device &&
            cudaDeviceGetP2PAttribute(&native, cudaDevP2PAttrNativeAtomicSupported, ctx->device, p.device) == cudaSuccess) {
            if (!native) { ok = false; snprintf(why, sizeof(why), "the peer path to rank %d (device %d) is not NVLink", r, p.device); }
        } else {
            cudaGetLastError();
        }
    }
    const char* st = getenv("IGG_LL128_SELFTEST");
    const bool run_test = ok && n > 1 && !same_gpu && !(st && st[0] == '0');
    const bool corrupt = getenv("IGG_LL128_SELFTEST_FAIL") && getenv("IGG_LL128_SELFTEST_FAIL")[0] == '1';
    int rc = ctx->cp.barrier();                 // every rank has mapped every control segment
    if (rc) return rc;
    if (run_test) {
        const int lines = 4096, rounds = 32;    // 512 KiB of the (still unused) arrival-flag area as scratch
        const int right = (me + 1) % n;
        rc = ll128_selftest(ctx->peers[right].ctrl.ptr + kFlagsOff, ctx->ctrl + kFlagsOff, lines, rounds, corrupt,
                            5ull * 1000000000ull, ctx->ll128_test);
        if (rc) return rc;
        if (ctx->ll128_test[0] || ctx->ll128_test[1]) {
            ok = false;
            snprintf(why, sizeof(why), "self-test: %llu torn line images, %llu time-outs in %llu line images",
                     ctx->ll128_test[0], ctx->ll128_test[1], ctx->ll128_test[2]);
        }
        rc = ctx->cp.barrier();                 // nobody writes into my scratch any more
        if (rc) return rc;
        CU(cudaMemset(ctx->ctrl + kFlagsOff, 0, (size_t)lines * 128));
        CU(cudaDeviceSynchronize());
    }
    SegPayload mine;
    std::memset(&mine, 0, sizeof(mine));
    mine.status = ok ? 1 : 0;
    snprintf(mine.msg, sizeof(mine.msg), "%s", why);
    std::vector<char> all;
    rc = ctx->cp.allgather(&mine, all);
    if (rc) return rc;
    for (int r = 0; r < n; ++r) {
        SegPayload q;
        std::memcpy(&q, all.data() + (size_t)r * sizeof(SegPayload), sizeof(q));
        if (!q.status) {
            if (ctx->ll128_ok && me == 0)
                fprintf(stderr, "libigg_b200: LL128 is not safe on this system (rank %d: %s); using the fence protocol.\n", r, q.msg);
            ctx->ll128_ok = false;
        }
    }
    if (!ctx->ll128_ok && (ctx->mode == IGG_MODE_IPC_AUTO || ctx->mode == IGG_MODE_IPC_LL128)) ctx->mode = IGG_MODE_IPC_1PHASE;
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
    // Whatever fails below, igg_destroy releases what has been set up so far (it copes with a half-built context).
    struct Guard { igg_ctx* c; ~Guard() { if (c) { const std::string keep = last_error_cstr(); igg_destroy(c); set_error("%s", keep.c_str()); } } } guard{new igg_ctx()};
    igg_ctx* ctx = guard.c;
    ctx->g = *desc;
    ctx->device = desc->device;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMalloc(&ctx->ctrl, kCtrlBytes));
    CU(cudaMemset(ctx->ctrl, 0, kCtrlBytes));
    CU(cudaHostAlloc(&ctx->h_status, 64, cudaHostAllocMapped));
    *ctx->h_status = 0;
    CU(cudaHostGetDevicePointer(&ctx->d_status, ctx->h_status, 0));
    ctx->h_done = reinterpret_cast<unsigned long long*>(ctx->h_status) + 4;      // same page-locked block, own 32 bytes
    ctx->d_done = reinterpret_cast<unsigned long long*>(ctx->d_status) + 4;
    *ctx->h_done = 0;
    CU(cudaEventCreateWithFlags(&ctx->last_event, cudaEventDisableTiming));
    ctx->max_blocks = halo_kernel_max_coresident_blocks(ctx->device);
    if (ctx->max_blocks <= 0) return fail(IGG_ERR_CUDA, "occupancy query for the halo kernel failed (was the library built for sm_100a?): %s", cudaGetErrorString(cudaGetLastError()));
    if (const char* t = getenv("IGG_DEVICE_TIMEOUT_S")) ctx->timeout_ns = (unsigned long long)(atof(t) * 1e9);
    if (const char* b = getenv("IGG_BLOCKING")) ctx->spin_wait = strcmp(b, "sync") != 0;
    if (const char* dd = getenv("IGG_DIRECT")) ctx->direct_enabled = dd[0] != '0';
    // IGG_MODE=fence / ll128 / auto: the IPC protocol a new context starts in (default auto; the same on every rank).
    if (const char* m = getenv("IGG_MODE")) {
        if (!strcmp(m, "ll128")) ctx->mode = IGG_MODE_IPC_LL128;
        else if (!strcmp(m, "fence") || !strcmp(m, "ipc")) ctx->mode = IGG_MODE_IPC_1PHASE;
        else if (!strcmp(m, "auto")) ctx->mode = IGG_MODE_IPC_AUTO;
    }
    ctx->peers.resize(desc->nprocs);
    if (desc->nprocs > 1) {
        rc = ctx->cp.create(desc->me, desc->nprocs, ctx->shm_name);
        if (rc) return rc;
    } else {
        ctx->bootstrapped = true;
    }
    CU(cudaDeviceSynchronize());
    *out = ctx;
    guard.c = nullptr;
    return IGG_OK;
}

int igg_destroy(igg_ctx* ctx) {
    if (!ctx) return IGG_OK;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    if (ctx->g.nprocs > 1 && ctx->cp.attached()) ctx->cp.barrier();   // nobody is still writing into my segments
    for (igg_prepared* h : ctx->prepared) { h->ctx = nullptr; h->plan = nullptr; }   // handles outlive the context harmlessly
    nccl_free(ctx->nccl);
    ctx->nccl = nullptr;
    for (auto& p : ctx->peers) { close_segment(p.ctrl); close_segment(p.staging); close_segment(p.gather); }
    for (auto& pl : ctx->plans) { cudaFree(pl->d_plan); cudaFree(pl->d_wrap_tiles); cudaFree(pl->d_hostx_tiles); }
    if (ctx->hostx_cache) cudaFree(ctx->hostx_cache);
    for (char* c : ctx->hostx_old_caches) cudaFree(c);
    for (auto& reg : ctx->regs) { for (auto& ps : reg.peer) close_segment(ps); }
    if (ctx->g.nprocs > 1 && ctx->cp.attached() && !ctx->regs.empty()) ctx->cp.barrier();   // mappings closed before the frees
    for (auto& reg : ctx->regs) if (reg.base) cudaFree(reg.base);
    if (ctx->d_reg_table) cudaFree(ctx->d_reg_table);
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
    rc = check_ll128(ctx);
    if (rc) return rc;
    ctx->bootstrapped = true;
    return IGG_OK;
}

int igg_set_mode(igg_ctx* ctx, int mode) {
    if (!ctx) return fail(IGG_ERR_ARG, "igg_set_mode: NULL context.");
    if (mode != IGG_MODE_IPC_1PHASE && mode != IGG_MODE_NCCL_3PHASE && mode != IGG_MODE_IPC_LL128 && mode != IGG_MODE_IPC_AUTO)
        return fail(IGG_ERR_ARG, "igg_set_mode: unknown mode %d.", mode);
    if (mode == IGG_MODE_NCCL_3PHASE && ctx->g.nprocs > 1 && !ctx->nccl)
        return fail(IGG_ERR_STATE, "igg_set_mode: call igg_nccl_init before selecting IGG_MODE_NCCL_3PHASE.");
    if (!ctx->ll128_ok && mode == IGG_MODE_IPC_LL128)
        return fail(IGG_ERR_UNSUPPORTED, "igg_set_mode: LL128 was found unsafe on this system at bootstrap (link check / self-test).");
    if (!ctx->ll128_ok && mode == IGG_MODE_IPC_AUTO) mode = IGG_MODE_IPC_1PHASE;
    // The two IPC protocols stage in regions of their own (ensure_layout), so switching needs no clean-up.
    ctx->mode = mode;
    return IGG_OK;
}

static int upload_reg_table(igg_ctx* ctx) {
    const int n = ctx->g.nprocs;
    std::vector<char*> host((size_t)kMaxFieldRegs * n, nullptr);
    for (size_t k = 0; k < ctx->regs.size(); ++k)
        for (int r = 0; r < n; ++r)
            host[k * n + r] = r == ctx->g.me ? ctx->regs[k].base : (ctx->regs[k].base ? ctx->regs[k].peer[r].ptr : nullptr);
    if (!ctx->d_reg_table) CU(cudaMalloc(&ctx->d_reg_table, host.size() * sizeof(char*)));
    const cudaError_t e = upload_now(ctx->d_reg_table, host.data(), host.size() * sizeof(char*));
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "field table upload failed: %s", cudaGetErrorString(e));
    return IGG_OK;
}

// Symmetric allocation: collective, every rank asks for the same `bytes` in the same order.  cudaMalloc + one CUDA-IPC
// handle per rank, exchanged through the mailbox; every rank maps every other rank's copy.
int igg_alloc_field(igg_ctx* ctx, size_t bytes, void** ptr_out) {
    if (!ctx || !ptr_out || bytes == 0) return fail(IGG_ERR_ARG, "igg_alloc_field: invalid argument.");
    if (!ctx->bootstrapped) return fail(IGG_ERR_STATE, "igg_alloc_field: call igg_bootstrap_import first (nprocs > 1).");
    int rc = make_current(ctx);
    if (rc) return rc;
    const int n = ctx->g.nprocs, me = ctx->g.me;
    size_t slot = 0;
    while (slot < ctx->regs.size() && ctx->regs[slot].base) ++slot;       // first free slot: the same on every rank
    if (slot >= (size_t)kMaxFieldRegs) return fail(IGG_ERR_ARG, "igg_alloc_field: at most %d symmetric allocations per context.", kMaxFieldRegs);
    if (slot == ctx->regs.size()) ctx->regs.emplace_back();
    igg_ctx::FieldReg& reg = ctx->regs[slot];
    char* mem = nullptr;
    cudaError_t e = cudaMalloc(&mem, bytes);
    int status = 0;
    if (e == cudaSuccess) e = cudaMemset(mem, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaGetLastError(); status = IGG_ERR_CUDA; }
    reg.peer.assign(n, PeerSeg());
    if (n > 1) {
        SegPayload mine;
        std::memset(&mine, 0, sizeof(mine));
        mine.proc_nonce = process_nonce();
        mine.device = ctx->device;
        mine.status = status;
        mine.bytes = (int64_t)bytes;
        mine.ptr = reinterpret_cast<uint64_t>(mem);
        if (!status && cudaIpcGetMemHandle(&mine.handle, mem) != cudaSuccess) { cudaGetLastError(); mine.status = status = IGG_ERR_IPC; }
        std::vector<char> all;
        rc = ctx->cp.allgather(&mine, all);
        if (rc) { if (mem) cudaFree(mem); return rc; }
        bool bad = false, differ = false;
        std::vector<SegPayload> pay(n);
        std::memcpy(pay.data(), all.data(), all.size());
        for (int r = 0; r < n; ++r) { bad = bad || pay[r].status != 0; differ = differ || pay[r].bytes != (int64_t)bytes; }
        for (int r = 0; r < n && !bad && !differ; ++r) {
            if (r == me) continue;
            rc = open_segment(ctx, ctx->peers[r], pay[r].handle, pay[r].ptr, reg.peer[r]);
            if (rc) bad = true;
            reg.peer[r].bytes = pay[r].bytes;
        }
        // everyone learns whether everyone could map everything (a half-registered allocation must not be used)
        SegPayload verdict;
        std::memset(&verdict, 0, sizeof(verdict));
        verdict.status = (bad || differ) ? 1 : 0;
        rc = ctx->cp.allgather(&verdict, all);
        if (rc) return rc;
        std::memcpy(pay.data(), all.data(), all.size());
        for (int r = 0; r < n; ++r) bad = bad || pay[r].status != 0;
        if (bad || differ) {
            for (auto& ps : reg.peer) close_segment(ps);
            ctx->cp.barrier();
            if (mem) cudaFree(mem);
            return fail(differ ? IGG_ERR_ARG : IGG_ERR_IPC, differ ? "igg_alloc_field: the ranks asked for different sizes (the call is collective and symmetric)."
                                                                   : "igg_alloc_field: allocation or CUDA-IPC mapping failed on some rank.");
        }
    } else if (status) {
        return fail(IGG_ERR_CUDA, "igg_alloc_field: cudaMalloc of %zu bytes failed.", bytes);
    }
    reg.base = mem;
    reg.bytes = (int64_t)bytes;
    rc = upload_reg_table(ctx);
    if (rc) return rc;
    *ptr_out = mem;
    return IGG_OK;
}

int igg_free_field(igg_ctx* ctx, void* ptr) {
    if (!ctx) return fail(IGG_ERR_ARG, "igg_free_field: NULL context.");
    if (!ptr) return IGG_OK;
    int rc = make_current(ctx);
    if (rc) return rc;
    const int slot = find_reg(ctx, ptr, nullptr);
    if (slot < 0 || ctx->regs[slot].base != ptr) return fail(IGG_ERR_ARG, "igg_free_field: %p was not returned by igg_alloc_field.", ptr);
    igg_ctx::FieldReg& reg = ctx->regs[slot];
    CU(cudaDeviceSynchronize());                    // my kernels are done with everyone's copies ...
    if (ctx->g.nprocs > 1) { rc = ctx->cp.barrier(); if (rc) return rc; }      // ... and everyone's with mine
    for (auto& ps : reg.peer) close_segment(ps);
    if (ctx->g.nprocs > 1) { rc = ctx->cp.barrier(); if (rc) return rc; }      // every mapping of my copy is closed
    CU(cudaFree(reg.base));
    reg = igg_ctx::FieldReg();
    return upload_reg_table(ctx);
}

}  // extern "C"

// Launches `plan` for `fields` (pointers already resolved).  `sync`: the call is host-synchronous.
static int launch_plan(igg_ctx* ctx, Plan* plan, int nfields, const igg_field* fields, cudaStream_t stream, bool sync,
                       bool* waited) {
    int rc;
    *waited = false;
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
    // Nothing to copy and no rank exchanges in this call (e.g. one rank, nothing periodic): no launch.  A rank that
    // has nothing to copy while others exchange still launches: the exchange epoch advances on every rank alike.
    if (plan->tiles.empty() && !plan->global_active) return IGG_OK;
    if (plan->n_flags > (uint32_t)kMaxFlags)
        return fail(IGG_ERR_ARG, "update_halo: this call needs %u arrival flags, the control segment holds %d; split the call.",
                    plan->n_flags, kMaxFlags);
    if (plan->active_mask) {
        rc = ensure_layout(ctx, *plan);
        if (rc) return rc;
    }

    KParams p;
    p.tiles = plan->d_tiles;
    p.n_tiles = (int32_t)plan->tiles.size();
    p.n_push_tiles = plan->n_push_tiles;
    p.plan = plan->active_mask ? plan->d_plan : nullptr;
    // Epoch + staging parity live on the device.  A call in which no rank exchanges with another one (one rank, or
    // only self-wrapped dimensions updated: the same verdict on every rank) needs neither and skips the prologue.
    DeviceState* state = reinterpret_cast<DeviceState*>(ctx->ctrl);
    p.state = plan->global_active ? state : nullptr;
    p.local_flags = reinterpret_cast<const unsigned long long*>(ctx->ctrl + kFlagsOff);
    p.status = ctx->d_status;
    p.timeout_ns = ctx->timeout_ns;
    p.trace = nullptr;
    p.notify = nullptr;
    p.notify_value = 0;
    p.fin_counter = &state->blocks_finished;
    p.bulk = plan->has_bulk ? 1 : 0;
    p.reg_table = ctx->d_reg_table;
    p.nprocs = ctx->g.nprocs;
    p.local_ready = reinterpret_cast<const unsigned long long*>(ctx->ctrl + kReadyOff);
    for (int i = 0; i < IGG_MAX_FIELDS; ++i) { p.field_reg[i] = -1; p.field_reg_off[i] = 0; }
    if (plan->direct_mask)
        for (int i = 0; i < nfields; ++i) {
            long long off = 0;
            const int slot = find_reg(ctx, fields[i].ptr, &off);
            if (slot < 0 && (plan->key[i].lines & 2))
                return fail(IGG_ERR_STATE, "update_halo: field %d was in a symmetric allocation when the call was planned and is not any more.", i + 1);
            p.field_reg[i] = (int16_t)slot;
            p.field_reg_off[i] = off;
        }
    static const bool wx_params = !(std::getenv("IGG_WX_PARAMS") && std::atoi(std::getenv("IGG_WX_PARAMS")) == 0);
    // (Measured and dropped: letting the WRAPX blocks sleep 400-1200 ns so that the contiguous tiles' loads get out
    // before the strided flood -- 8.62 / 8.76 / 9.01 us against 8.64 us per cfg2 update, cold L2.)
    p.wx_first = plan->wx_first;
    p.wx_count = 0;
    p.n_wx_ops = 0;
    if (wx_params && plan->wx_count > 0 && plan->wx_ops.size() <= (size_t)kMaxWxOps) {
        p.wx_count = plan->wx_count;
        p.n_wx_ops = (int32_t)plan->wx_ops.size();
        for (int k = 0; k < p.n_wx_ops; ++k) p.wx[k] = plan->wx_ops[k];
    }
    p.copy_dst[0] = p.copy_dst[1] = nullptr;
    p.copy_src[0] = p.copy_src[1] = nullptr;
    for (int i = 0; i < nfields; ++i) p.field[i] = fields[i].ptr;
    for (int i = nfields; i < IGG_MAX_FIELDS; ++i) p.field[i] = nullptr;
    order_after_previous(ctx, stream);
    const bool hostx = !plan->hostx_tiles.empty();
    void* host_ptrs[IGG_MAX_FIELDS];
    if (hostx) {
        // Exchanged x-faces of host-resident fields: fetch the row-end lines into the device cache; the exchange
        // kernel sees every such field's lines as one more field (index nfields + k).
        if (plan->hostx_cache_bytes > ctx->hostx_cache_cap) {
            // A bigger cache is a new allocation; the old one stays until igg_destroy: launches recorded into a CUDA
            // graph keep using the cache they were captured with (all three kernels of a call carry the same pointer).
            if (ctx->hostx_cache) ctx->hostx_old_caches.push_back(ctx->hostx_cache);
            ctx->hostx_cache = nullptr;
            ctx->hostx_cache_cap = 0;
            CU(cudaMalloc(&ctx->hostx_cache, (size_t)plan->hostx_cache_bytes));
            ctx->hostx_cache_cap = plan->hostx_cache_bytes;
        }
        for (int i = 0; i < nfields; ++i) host_ptrs[i] = fields[i].ptr;
        rc = launch_host_lines_kernel(plan->d_hostx_tiles, (int)plan->hostx_tiles.size(), host_ptrs, nfields, ctx->hostx_cache, false, stream);
        if (rc) return rc;
        ctx->launches += 1;
        for (size_t k = 0; k < plan->hostx.size(); ++k) p.field[nfields + (int)k] = ctx->hostx_cache + plan->hostx[k].cache_off;
    }
    if (ctx->d_trace) {
        p.trace = ctx->d_trace;
        ctx->last_grid = plan->grid;
        CU(cudaMemsetAsync(ctx->d_trace, 0, (size_t)ctx->max_blocks * 8 * sizeof(unsigned long long), stream));
    }
    unsigned long long ticket = 0;
    if (sync && ctx->spin_wait && !join_side && !hostx) {
        ticket = ++ctx->ticket;
        p.notify = ctx->d_done;
        p.notify_value = ticket;
    }
    rc = launch_halo_kernel(p, plan->grid, stream, plan->ll128);
    if (join_side) CU(cudaStreamWaitEvent(stream, ctx->ev_join, 0));   // the call is complete when both kernels are
    if (rc) return rc;
    remember_launch(ctx, stream);
    if (hostx) {                                   // the cached lines, halos patched, go back to the host whole
        rc = launch_host_lines_kernel(plan->d_hostx_tiles, (int)plan->hostx_tiles.size(), host_ptrs, nfields, ctx->hostx_cache, true, stream);
        if (rc) return rc;
        ctx->launches += 1;
    }
    if (ticket) {
        *waited = true;
        return wait_done(ctx, stream, ticket);
    }
    return IGG_OK;
}

// The fused single-phase exchange for the dimensions in `dims_mask` (plan lookup, staging growth, one launch).
static int launch_exchange(igg_ctx* ctx, int nfields, const igg_field* fields, int dims_mask, cudaStream_t stream,
                           const uint8_t* host_lines, bool sync, bool* waited, const uint8_t* direct = nullptr) {
    int rc;
    Plan* plan = find_or_build_plan(ctx, nfields, fields, dims_mask, host_lines, &rc, direct);
    if (!plan) return rc;
    if (!plan->pinned) {
        // A launch recorded into a CUDA graph keeps pointing at the plan's device tables: never evict such a plan.
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone) plan->pinned = true;
        else cudaGetLastError();
    }
    return launch_plan(ctx, plan, nfields, fields, stream, sync, waited);
}

static int resolve_fields(igg_ctx* ctx, int nfields, const igg_field* fields_in, igg_field* fields, uint8_t* host_lines,
                          uint8_t* direct) {
    static const bool host_lines_enabled = !(std::getenv("IGG_HOST_LINES") && std::atoi(std::getenv("IGG_HOST_LINES")) == 0);
    for (int i = 0; i < nfields; ++i) {
        fields[i] = fields_in[i];
        bool is_host = false;
        int rc = resolve_pointer(ctx, fields_in[i].ptr, &fields[i].ptr, &is_host);
        if (rc) return rc;
        // Fields in page-locked host memory: whole-line x self-wrap (kOpWrapLines) unless IGG_HOST_LINES=0.
        host_lines[i] = (is_host && host_lines_enabled && (uintptr_t)fields[i].ptr % 64 == 0) ? 1 : 0;
        // Fields inside a symmetric allocation may be written by the neighbours directly (direct-to-halo).  The whole
        // array must lie inside the allocation.  (Symmetric: the same verdict on every rank.)
        long long off = 0;
        const int slot = (ctx->direct_enabled && ctx->g.nprocs > 1 && !is_host) ? find_reg(ctx, fields[i].ptr, &off) : -1;
        const long long bytes = (long long)fields[i].size[0] * fields[i].size[1] * fields[i].size[2] * fields[i].elsize;
        direct[i] = (slot >= 0 && off + bytes <= ctx->regs[slot].bytes) ? 1 : 0;
    }
    return IGG_OK;
}

// Comparison path: the reference's dependent phases, in the caller's order (default z, x, y).
static int nccl_update(igg_ctx* ctx, int nfields, const igg_field* fields, const uint8_t* host_lines, const int* dims_order,
                       int ndims_order, cudaStream_t stream) {
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
        int rc;
        bool waited;
        if (l == ctx->g.me && r == ctx->g.me) rc = launch_exchange(ctx, nfields, fields, 1 << d, stream, host_lines, false, &waited);   // local wrap
        else rc = nccl_phase(ctx->nccl, ctx->g, nfields, fields, d, ctx->max_blocks, proto, stream, &ctx->launches);
        if (rc) return rc;
    }
    ctx->last_stream = stream;
    ctx->have_last = true;
    return IGG_OK;
}

extern "C" {

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
    rc = check_status(ctx);
    if (rc) return rc;
    rc = make_current(ctx);
    if (rc) return rc;

    igg_field fields[IGG_MAX_FIELDS];
    uint8_t host_lines[IGG_MAX_FIELDS], direct[IGG_MAX_FIELDS];
    rc = resolve_fields(ctx, nfields, fields_in, fields, host_lines, direct);
    if (rc) return rc;

    const bool sync = !(flags & IGG_ASYNC);
    bool waited = false;
    if (ctx->mode == IGG_MODE_NCCL_3PHASE && ctx->g.nprocs > 1) rc = nccl_update(ctx, nfields, fields, host_lines, dims_order, ndims_order, stream);
    else rc = launch_exchange(ctx, nfields, fields, dims_mask, stream, host_lines, sync, &waited, direct);
    if (rc) return rc;
    if (sync && !waited) return wait_done(ctx, stream, 0);
    return IGG_OK;
}

int igg_prepare(igg_ctx* ctx, int nfields, const igg_field* fields_in, const int* dims_order, int ndims_order,
                igg_prepared** out) {
    if (!ctx || !fields_in || !out) return fail(IGG_ERR_ARG, "igg_prepare: NULL argument.");
    if (!ctx->bootstrapped) return fail(IGG_ERR_STATE, "igg_prepare: call igg_bootstrap_import first (nprocs > 1).");
    if (ndims_order > 3) return fail(IGG_ERR_ARG, "igg_prepare: at most 3 entries in dims.");
    std::unique_ptr<igg_prepared> h(new igg_prepared());
    int rc = dims_mask_from_order(dims_order, ndims_order, &h->dims_mask);
    if (rc) return rc;
    rc = check_fields(ctx->g, nfields, fields_in);
    if (rc) return rc;
    rc = make_current(ctx);
    if (rc) return rc;
    h->ctx = ctx;
    h->nfields = nfields;
    if (dims_order && ndims_order > 0) { h->ndims_order = ndims_order; for (int i = 0; i < ndims_order; ++i) h->dims_order[i] = dims_order[i]; }
    for (int i = 0; i < nfields; ++i) h->user_fields[i] = fields_in[i];
    rc = resolve_fields(ctx, nfields, fields_in, h->fields, h->host_lines, h->direct);
    if (rc) return rc;
    // Building the plan here allocates its device tables now, not in the time loop.  (Staging for exchanges between
    // ranks still grows at the first call that needs it: that step is collective.)
    h->plan = find_or_build_plan(ctx, nfields, h->fields, h->dims_mask, h->host_lines, &rc, h->direct);
    if (!h->plan) return rc;
    h->plan->pinned = true;
    h->mode = ctx->mode;
    ctx->prepared.push_back(h.get());
    *out = h.release();
    return IGG_OK;
}

int igg_update_halo_prepared(igg_prepared* h, void* stream_v, int flags) {
    if (!h) return fail(IGG_ERR_ARG, "igg_update_halo_prepared: NULL handle.");
    igg_ctx* ctx = h->ctx;
    if (!ctx) return fail(IGG_ERR_STATE, "igg_update_halo_prepared: the context of this prepared call has been destroyed.");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    int rc = check_status(ctx);
    if (rc) return rc;
    rc = make_current(ctx);
    if (rc) return rc;
    const bool sync = !(flags & IGG_ASYNC);
    bool waited = false;
    if (ctx->mode == IGG_MODE_NCCL_3PHASE && ctx->g.nprocs > 1) {
        rc = nccl_update(ctx, h->nfields, h->fields, h->host_lines, h->dims_order, h->ndims_order, stream);
    } else {
        if (h->mode != ctx->mode) {            // the exchange mode changed since: the plan of the new mode
            h->plan = find_or_build_plan(ctx, h->nfields, h->fields, h->dims_mask, h->host_lines, &rc, h->direct);
            if (!h->plan) return rc;
            h->plan->pinned = true;
            h->mode = ctx->mode;
        }
        rc = launch_plan(ctx, h->plan, h->nfields, h->fields, stream, sync, &waited);
    }
    if (rc) return rc;
    if (sync && !waited) return wait_done(ctx, stream, 0);
    return IGG_OK;
}

// Host cost of issuing `n` stream-ordered prepared calls from C (no foreign-call overhead in between): ns per call.
int igg_debug_issue_loop(igg_prepared* h, void* stream, int n, double* ns_per_call_out) {
    if (!h || !ns_per_call_out || n < 1) return fail(IGG_ERR_ARG, "igg_debug_issue_loop: invalid argument.");
    timespec a, b;
    clock_gettime(CLOCK_MONOTONIC, &a);
    for (int i = 0; i < n; ++i) {
        const int rc = igg_update_halo_prepared(h, stream, IGG_ASYNC);
        if (rc) return rc;
    }
    clock_gettime(CLOCK_MONOTONIC, &b);
    *ns_per_call_out = ((b.tv_sec - a.tv_sec) * 1e9 + (b.tv_nsec - a.tv_nsec)) / n;
    return IGG_OK;
}

// The timed loop of a compiled-language caller (what a Julia time loop over `ccall`s would be): `steps` prepared calls,
// call i using calls[i % ncalls].  IGG_ASYNC: every call sits between its own pair of CUDA events on `stream`
// (us_out[i] = event time of call i, *loop_us_out = first event to last event / steps).  IGG_SYNC: host clock around
// every host-synchronous call.
int igg_debug_timed_calls(igg_prepared** calls, int ncalls, void* stream_v, int steps, int flags, float* us_out,
                          double* loop_us_out) {
    if (!calls || ncalls < 1 || steps < 1 || !us_out) return fail(IGG_ERR_ARG, "igg_debug_timed_calls: invalid argument.");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!(flags & IGG_ASYNC)) {
        for (int i = 0; i < steps; ++i) {
            timespec a, b;
            clock_gettime(CLOCK_MONOTONIC, &a);
            const int rc = igg_update_halo_prepared(calls[i % ncalls], stream_v, IGG_SYNC);
            clock_gettime(CLOCK_MONOTONIC, &b);
            if (rc) return rc;
            us_out[i] = (float)(((b.tv_sec - a.tv_sec) * 1e9 + (b.tv_nsec - a.tv_nsec)) * 1e-3);
        }
        if (loop_us_out) *loop_us_out = 0;
        return IGG_OK;
    }
    std::vector<cudaEvent_t> ev(2 * (size_t)steps, nullptr);
    int rc = IGG_OK;
    for (auto& e : ev)
        if (cudaEventCreate(&e) != cudaSuccess) rc = fail(IGG_ERR_CUDA, "igg_debug_timed_calls: cudaEventCreate failed.");
    const bool per_call = !(flags & 2);        // flags | 2: only the first and the last event are recorded
    for (int i = 0; i < steps && !rc; ++i) {
        if (per_call || i == 0) cudaEventRecord(ev[2 * i], stream);
        rc = igg_update_halo_prepared(calls[i % ncalls], stream_v, IGG_ASYNC);
        if (per_call || i == steps - 1) cudaEventRecord(ev[2 * i + 1], stream);
    }
    if (!rc && cudaStreamSynchronize(stream) != cudaSuccess) rc = fail(IGG_ERR_CUDA, "igg_debug_timed_calls: %s", cudaGetErrorString(cudaGetLastError()));
    if (!rc) {
        for (int i = 0; i < steps && per_call; ++i) {
            float ms = 0;
            cudaEventElapsedTime(&ms, ev[2 * i], ev[2 * i + 1]);
            us_out[i] = ms * 1e3f;
        }
        float ms = 0;
        cudaEventElapsedTime(&ms, ev[0], ev[2 * (size_t)steps - 1]);
        if (loop_us_out) *loop_us_out = ms * 1e3 / steps;
    }
    for (auto& e : ev) if (e) cudaEventDestroy(e);
    return rc;
}

int igg_debug_peer_copy(igg_ctx* ctx, void* sym_ptr, int peer, size_t bytes, int reps, double* gbs_out) {
    if (!ctx || !sym_ptr || !gbs_out || reps < 1) return fail(IGG_ERR_ARG, "igg_debug_peer_copy: invalid argument.");
    long long off = 0;
    const int slot = find_reg(ctx, sym_ptr, &off);
    if (slot < 0 || peer < 0 || peer >= ctx->g.nprocs || peer == ctx->g.me || off + (long long)bytes > ctx->regs[slot].bytes)
        return fail(IGG_ERR_ARG, "igg_debug_peer_copy: not a symmetric allocation of this context, or a bad peer / size.");
    int rc = make_current(ctx);
    if (rc) return rc;
    char* dst = ctx->regs[slot].peer[peer].ptr + off;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r <= reps; ++r) {               // the first one warms up
        CU(cudaEventRecord(e0, 0));
        CU(cudaMemcpyAsync(dst, sym_ptr, bytes, cudaMemcpyDeviceToDevice, 0));
        CU(cudaEventRecord(e1, 0));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        if (r > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *gbs_out = (double)bytes / (best * 1e-3) / 1e9;
    return IGG_OK;
}

int igg_release_prepared(igg_prepared* h) {
    if (!h) return IGG_OK;
    if (h->ctx) {
        auto& v = h->ctx->prepared;
        for (size_t k = 0; k < v.size(); ++k)
            if (v[k] == h) { v.erase(v.begin() + (long)k); break; }
    }
    delete h;      // its plan stays cached (and pinned: a captured graph may refer to it) until the context goes
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

    // Round 1 of the hand-shake: root's verdict and where the blocks go.  If root's A_global is device memory whose
    // allocation can be exported over CUDA IPC (cudaMalloc'ed: torch's caching allocator, igg_alloc_field, CUDA.jl with
    // the non-pool allocator), every rank stores its block straight into A_global (as the reference's Gatherv places
    // the blocks, ref: src/gather.jl:36-49).  Otherwise (host memory, stream-ordered pool memory) the blocks go into
    // a device image owned by the context and root copies it out afterwards.
    SegPayload mine;
    std::memset(&mine, 0, sizeof(mine));
    mine.proc_nonce = process_nonce();
    mine.device = ctx->device;
    mine.status = status;
    snprintf(mine.msg, sizeof(mine.msg), "%s", msg);
    bool direct = false;
    long long direct_off = 0;
    if (is_root && !status) {
        static const bool gather_direct_on = !(getenv("IGG_GATHER_DIRECT") && getenv("IGG_GATHER_DIRECT")[0] == '0');
        cudaPointerAttributes a;
        void* base = nullptr;
        size_t alloc_bytes = 0;
        if (gather_direct_on && cudaPointerGetAttributes(&a, A_global) == cudaSuccess && a.type == cudaMemoryTypeDevice &&
            a.device == ctx->device && device_address_range(A_global, &base, &alloc_bytes) &&
            cudaIpcGetMemHandle(&mine.handle, base) == cudaSuccess) {
            direct = true;
            direct_off = static_cast<char*>(A_global) - static_cast<char*>(base);
            mine.ptr = reinterpret_cast<uint64_t>(base);
            mine.bytes = (int64_t)alloc_bytes;
            snprintf(mine.msg, sizeof(mine.msg), "D%lld", direct_off);
        } else {
            cudaGetLastError();
        }
    }
    const bool regrow_mine = is_root && !status && !direct && bytesG > ctx->gather_bytes;
    if (is_root && !direct) {
        if (ctx->gather_buf && !regrow_mine) {
            CU(cudaIpcGetMemHandle(&mine.handle, ctx->gather_buf));
            mine.ptr = reinterpret_cast<uint64_t>(ctx->gather_buf);
        }
        mine.bytes = regrow_mine ? -1 : ctx->gather_bytes;
    }
    std::vector<SegPayload> all(g.nprocs);
    int rc;
    {
        std::vector<char> raw;
        rc = ctx->cp.allgather(&mine, raw);
        if (rc) return rc;
        std::memcpy(all.data(), raw.data(), raw.size());
    }
    if (all[root].status) return fail(all[root].status, "%s", all[root].msg[0] ? all[root].msg : "gather!: root reported an error.");
    direct = all[root].msg[0] == 'D';
    if (direct) direct_off = atoll(all[root].msg + 1);
    if (!direct && all[root].bytes < 0) {
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
        dst_base = direct ? static_cast<char*>(A_global) : ctx->gather_buf;
    } else {
        Peer& pr = ctx->peers[root];
        if (pr.gather.ptr && (pr.gather.raw != rp.ptr || std::memcmp(&pr.gather.handle, &rp.handle, sizeof(rp.handle)) != 0))
            close_segment(pr.gather);                                                 // another allocation than last time
        if (!pr.gather.ptr) {
            rc = open_segment(ctx, pr, rp.handle, rp.ptr, pr.gather);
            if (rc) return rc;
            pr.gather.bytes = rp.bytes;
            pr.gather.raw = rp.ptr;
            pr.gather.handle = rp.handle;
        }
        dst_base = pr.gather.ptr + (direct ? direct_off : 0);
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
    // The tile table only depends on the block shape, the element size, my place in the global array and the source
    // alignment: kept on the device from call to call.
    const int64_t tkey[8] = {sizeA[0], sizeA[1], sizeA[2], elsize, origin[0] + (origin[1] << 20) + (origin[2] << 40), gsz[0], gsz[1],
                             (int64_t)((uintptr_t)src % 16)};
    if (std::memcmp(tkey, ctx->gather_tile_key, sizeof(tkey)) != 0 || ctx->gather_ntiles == 0) {
        std::vector<TileRec> tiles;
        build_block_copy_tiles(elsize, sizeA, gsz, origin, (int64_t)((uintptr_t)src % 16), 0, tiles);
        CU(cudaDeviceSynchronize());             // an earlier gather's kernel may still read the old table
        if ((int)tiles.size() > ctx->gather_tiles_cap) {
            if (ctx->d_gather_tiles) CU(cudaFree(ctx->d_gather_tiles));
            CU(cudaMalloc(&ctx->d_gather_tiles, tiles.size() * sizeof(TileRec)));
            ctx->gather_tiles_cap = (int)tiles.size();
        }
        const cudaError_t e = upload_now(ctx->d_gather_tiles, tiles.data(), tiles.size() * sizeof(TileRec));
        if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "gather!: tile table upload failed: %s", cudaGetErrorString(e));
        std::memcpy(ctx->gather_tile_key, tkey, sizeof(tkey));
        ctx->gather_ntiles = (int)tiles.size();
    }
    const int ntiles = ctx->gather_ntiles;
    KParams p;
    std::memset(&p, 0, sizeof(p));
    p.tiles = ctx->d_gather_tiles;
    p.n_tiles = (int32_t)ntiles;
    p.n_push_tiles = 0;
    p.field[0] = const_cast<char*>(src);
    p.copy_dst[0] = dst_base;
    p.local_flags = reinterpret_cast<const unsigned long long*>(ctx->ctrl + kFlagsOff);
    p.status = ctx->d_status;
    p.timeout_ns = ctx->timeout_ns;
    order_after_previous(ctx, stream);
    rc = launch_halo_kernel(p, (int)std::min<int64_t>((int64_t)ntiles, (int64_t)ctx->max_blocks * 4), stream);
    if (rc) return rc;
    remember_launch(ctx, stream);
    CU(cudaStreamSynchronize(stream));       // my block has landed in root's memory
    rc = ctx->cp.barrier();                  // ... and so has everyone else's
    if (rc || direct) return rc;             // straight into A_global: done (the next gather starts with a barrier)
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
    out[31] = (uint64_t)ctx->mode | ((uint64_t)(ctx->ll128_ok ? 1 : 0) << 8) | (ctx->ll128_test[2] << 16);
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
    // The kernel a new context launches for calls below the bulk threshold (all single-rank calls): the LL128
    // instantiation unless IGG_MODE selects the fence protocol (as in igg_create).  Same body, see halo_kernel.cu.
    const char* m = getenv("IGG_MODE");
    const bool fence = m && (!strcmp(m, "fence") || !strcmp(m, "ipc"));
    return fence ? halo_kernel_name() : "igg_halo_exchange_ll128_kernel";
}

}  // extern "C"
