// igg_update_halo_noalloc: a halo update for callers that must not allocate device memory.
//
// Who needs this: code that runs where cudaMalloc is off limits -- a host that has handed the whole GPU to its own
// pool allocator and forbids foreign allocations, a library called from inside a region where allocation would
// serialise or deadlock the caller's streams, or a time loop that has promised its users "no allocation after
// start-up".  The regular igg_update_halo allocates on first use (its context, the device copy of the tile table
// of every new plan, staging when it grows).  This entry point allocates nothing, on the device or in page-locked
// host memory, at any call: its only scratch -- the tile table the kernel walks and a few control words -- lives in
// a fixed array that is part of the library image.
//
// Limits that follow from that:
//  * one rank only (igg_grid_desc.nprocs == 1): every exchange is a periodic wrap inside the caller's own arrays, so
//    no peer staging is needed;
//  * the plan may have at most IGG_NOALLOC_MAX_TILES tiles (8 KiB of halo data each, about 25 MiB per call);
//    larger calls are refused with IGG_ERR_ARG and a message that says so;
//  * the array is reserved on every GPU that loads the library, for as long as the process lives;
//  * all calls share the one array and it keeps the previous call's contents.  The control words are cleared at
//    the start of every call and the table is rewritten, in stream order.  Two calls whose work overlaps on the GPU
//    (different streams, or different host threads) would overwrite each other's table: do not do that.
//  * the first call on a device looks up the array's device address (once per device); make that first call
//    before starting any stream capture;
//  * a call on a stream that is being captured is refused (IGG_ERR_UNSUPPORTED): the table is copied from host
//    memory that only lives for the duration of the call;
//  * the call makes desc->device the current device and leaves it so; a call whose plan differs from the previous
//    call's on that device copies the new table from ordinary host memory, which makes even an IGG_ASYNC call wait
//    for earlier work on the stream -- repeated calls with the same shapes skip the copy and are truly stream-ordered.
#include <cuda_runtime.h>

#include <mutex>

#include "igg_internal.h"

namespace {

constexpr int kMaxTiles = IGG_NOALLOC_MAX_TILES;
constexpr size_t kControlBytes = 4096;   // arrival flags (unused with one rank) and the status word
constexpr size_t kScratchBytes = kControlBytes + (size_t)kMaxTiles * sizeof(igg::TileRec);
constexpr int kMaxDevices = 64;

// The scratch itself.  In host code this name is only a stand-in; the device address comes from
// cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

std::mutex g_mutex;
unsigned char* g_scratch_ptr[kMaxDevices] = {nullptr};   // per device, filled at set-up
int g_max_blocks[kMaxDevices] = {0};
// What the table in the array was built from, per device: a call with the same key finds its table already there.
struct TableKey {
    igg_grid_desc desc;
    int nfields, mask;
    igg::FieldKey field[IGG_MAX_FIELDS];
    bool valid;
};
TableKey g_last[kMaxDevices];

// Once per device, with that device current.
int setup_device(int device, unsigned char** ptr_out, int* max_blocks_out) {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[device]) {
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return igg::fail(IGG_ERR_CUDA, "igg_update_halo_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                             cudaGetErrorString(e));
        const int mb = igg::halo_kernel_max_coresident_blocks(device);
        if (mb <= 0) return igg::fail(IGG_ERR_CUDA, "igg_update_halo_noalloc: occupancy query failed on device %d.", device);
        g_scratch_ptr[device] = p;
        g_max_blocks[device] = mb;
    }
    *ptr_out = g_scratch_ptr[device];
    *max_blocks_out = g_max_blocks[device];
    return IGG_OK;
}

}  // namespace

extern "C" int igg_update_halo_noalloc(const igg_grid_desc* desc, int nfields, const igg_field* fields,
                                       const int* dims_order, int ndims_order, void* stream_v, int flags) {
    using namespace igg;
    if (!desc || !fields) return fail(IGG_ERR_ARG, "igg_update_halo_noalloc: NULL argument.");
    if (desc->nprocs != 1)
        return fail(IGG_ERR_UNSUPPORTED, "igg_update_halo_noalloc: only single-rank grids (nprocs == 1) are supported; "
                                         "exchanges between ranks need peer staging, use igg_update_halo.");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(IGG_ERR_CUDA, "No usable CUDA device (%s); libigg_b200 has no CPU fallback.", cudaGetErrorString(e));
    }
    if (desc->device < 0 || desc->device >= ndev || desc->device >= kMaxDevices)
        return fail(IGG_ERR_ARG, "igg_update_halo_noalloc: invalid device %d.", desc->device);
    int mask;
    int rc = dims_mask_from_order(dims_order, ndims_order, &mask);
    if (rc) return rc;
    {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(static_cast<cudaStream_t>(stream_v), &cs) != cudaSuccess) cudaGetLastError();
        else if (cs != cudaStreamCaptureStatusNone)
            return fail(IGG_ERR_UNSUPPORTED, "igg_update_halo_noalloc: the stream is being captured; this entry point cannot "
                                             "be recorded into a CUDA graph (use igg_update_halo / igg_update_halo_prepared).");
    }
    e = cudaSetDevice(desc->device);
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "cudaSetDevice(%d) failed: %s", desc->device, cudaGetErrorString(e));

    // Device memory only: resolve mapped host pointers the same way igg_update_halo does (no allocation involved).
    igg_field f[IGG_MAX_FIELDS];
    if (nfields < 1 || nfields > IGG_MAX_FIELDS) return fail(IGG_ERR_ARG, "igg_update_halo_noalloc: nfields must be in 1..%d.", IGG_MAX_FIELDS);
    for (int i = 0; i < nfields; ++i) {
        f[i] = fields[i];
        if (!f[i].ptr) return fail(IGG_ERR_ARG, "The field at position %d has a NULL pointer.", i + 1);
        cudaPointerAttributes a;
        e = cudaPointerGetAttributes(&a, f[i].ptr);
        if (e != cudaSuccess) { cudaGetLastError(); return fail(IGG_ERR_CUDA, "cudaPointerGetAttributes failed: %s", cudaGetErrorString(e)); }
        if (a.type == cudaMemoryTypeHost && a.devicePointer) f[i].ptr = a.devicePointer;
        else if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged)
            return fail(IGG_ERR_UNSUPPORTED, "pointer %p is pageable host memory: pass device memory or page-locked "
                                             "(pinned/registered) host memory.", fields[i].ptr);
    }

    Plan plan;                                  // host-side only
    plan.small_tiles = true;                    // the limit below counts tiles of 8 KiB
    rc = build_plan(*desc, nfields, f, mask, plan);
    if (rc) return rc;
    if (plan.tiles.empty()) return IGG_OK;
    if ((int64_t)plan.tiles.size() > kMaxTiles)
        return fail(IGG_ERR_ARG, "igg_update_halo_noalloc: this call needs %lld tiles of halo data but the built-in "
                                 "scratch holds at most %d (about %d MiB per call); use igg_update_halo for larger calls.",
                    (long long)plan.tiles.size(), kMaxTiles, (int)((int64_t)kMaxTiles * kTileBytes >> 20));

    unsigned char* scratch = nullptr;
    int max_blocks = 0;
    rc = setup_device(desc->device, &scratch, &max_blocks);
    if (rc) return rc;

    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    // The array keeps whatever the last call left.  If that is this very table (same grid, shapes, halowidths, element
    // size, alignment and dims as the previous call on this device), it is used as it is; else the control words are
    // cleared and the new table is written, in stream order (earlier calls' kernels are done with the old one by then).
    TableKey key;
    std::memset(&key, 0, sizeof(key));
    key.desc = *desc;
    key.nfields = nfields;
    key.mask = mask;
    for (int i = 0; i < nfields; ++i) key.field[i] = plan.key[i];
    key.valid = true;
    bool same;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        same = g_last[desc->device].valid && std::memcmp(&g_last[desc->device], &key, sizeof(key)) == 0;
        g_last[desc->device].valid = false;     // until the new table is known to be on its way
    }
    if (!same) {
        e = cudaMemsetAsync(scratch, 0, kControlBytes, stream);
        if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "igg_update_halo_noalloc: clearing the scratch failed: %s", cudaGetErrorString(e));
        e = cudaMemcpyAsync(scratch + kControlBytes, plan.tiles.data(), plan.tiles.size() * sizeof(TileRec),
                            cudaMemcpyHostToDevice, stream);
        if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "igg_update_halo_noalloc: writing the tile table failed: %s", cudaGetErrorString(e));
    }
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        g_last[desc->device] = key;
    }

    KParams p;
    std::memset(&p, 0, sizeof(p));
    p.tiles = reinterpret_cast<const TileRec*>(scratch + kControlBytes);
    p.n_tiles = (int32_t)plan.tiles.size();
    p.n_push_tiles = 0;                                   // one rank: LOCAL tiles only
    for (int i = 0; i < nfields; ++i) p.field[i] = f[i].ptr;
    // One rank: the plan has only LOCAL / WRAPX tiles, nothing ever waits, so the flag array and the status word
    // below are never touched by the kernel; they point into the scratch so that they are valid device addresses.
    p.local_flags = reinterpret_cast<const unsigned long long*>(scratch);
    p.status = reinterpret_cast<int*>(scratch + 3968);
    p.timeout_ns = 1000000000ull;
    rc = launch_halo_kernel(p, (int)std::min<int64_t>((int64_t)plan.tiles.size(), max_blocks), stream);
    if (rc) return rc;
    if (!(flags & IGG_ASYNC)) {
        e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "igg_update_halo_noalloc: %s", cudaGetErrorString(e));
    }
    return IGG_OK;
}

extern "C" int64_t igg_noalloc_scratch_bytes(void) { return (int64_t)kScratchBytes; }
