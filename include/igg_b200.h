/* igg_b200.h -- C ABI of libigg_b200.so: the B200-native halo-update path of ImplicitGlobalGrid.jl v0.17.1.
 *
 * This library replaces, below the reference's `_update_halo!` / `gather!`, everything that moves data:
 * buffer allocation, pack/unpack kernels, stream handling and the MPI messages.  The host language (Julia via
 * `ccall`, Python via `ctypes`, C/C++) keeps the reference's API surface (init_global_grid, update_halo!,
 * gather!, finalize_global_grid, x_g/nx_g ...) and the process topology (MPI.jl Cart_create / Cart_shift or an
 * equivalent), and hands this library a plain description of the grid and raw device pointers.
 *
 * All functions return 0 (IGG_OK) on success and a negative error code otherwise; `igg_last_error()` returns a
 * thread-local message whose text mirrors the reference's `error("...")` strings where one exists.
 * The library does not link MPI or NCCL (NCCL is dlopen'ed only by the comparison path).  There is no CPU
 * fallback: every data-moving entry point fails with IGG_ERR_CUDA if no CUDA device is usable.
 *
 * Citations `ref:` are relative to the reference repository root.
 */
#ifndef IGG_B200_H
#define IGG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGG_VERSION 200          /* 0.2.0: prepared calls, symmetric fields, IGG_MODE_IPC_AUTO */
#define IGG_PROC_NULL (-1)       /* stand-in for MPI.PROC_NULL in igg_grid_desc.neighbors */
#define IGG_MAX_FIELDS 32        /* fields per update_halo call */
#define IGG_NOFFSETS 27          /* neighbour offsets (-1,0,1)^3, index = (ox+1) + 3*(oy+1) + 9*(oz+1) */

enum igg_status {
    IGG_OK = 0,
    IGG_ERR_ARG = -1,          /* invalid argument (NULL pointer, bad size ...) */
    IGG_ERR_CUDA = -2,         /* a CUDA runtime call failed (message has the CUDA error string) */
    IGG_ERR_STATE = -3,        /* call order violated (e.g. multi-rank update before bootstrap import) */
    IGG_ERR_FIELDS = -4,       /* check_fields failure; message = reference text (ref: src/update_halo.jl:418-480) */
    IGG_ERR_IPC = -5,          /* CUDA-IPC / shared-memory control plane failure */
    IGG_ERR_NCCL = -6,         /* NCCL comparison path failure */
    IGG_ERR_TIMEOUT = -7,      /* a neighbour did not arrive (device flag or host mailbox) within the time limit */
    IGG_ERR_UNSUPPORTED = -8,  /* valid in the reference but not supported here (e.g. disp != 1) */
    IGG_ERR_GRID = -9          /* incoherent grid description (ref: src/update_halo.jl:76, src/shared.jl:156-207) */
};

enum igg_flags {
    IGG_SYNC = 0,   /* default: block the host until the halos are written (the reference's semantics,
                       ref: src/update_halo.jl:82-87) */
    IGG_ASYNC = 1   /* stream-ordered: return after the launch; completion is ordered on `stream` */
};

enum igg_mode {
    IGG_MODE_IPC_1PHASE = 0,  /* one fused kernel, all 26 neighbour regions in one phase, peer stores over NVLink with
                                 16-byte accesses, one release fence + arrival flag per tile of 8-128 KiB ("fence"
                                 protocol, the bandwidth regime; env IGG_MODE=fence makes new contexts start in it) */
    IGG_MODE_NCCL_3PHASE = 1, /* comparison path: pack kernel -> grouped ncclSend/ncclRecv -> unpack kernel, one
                                 phase per dimension in the reference's order (ref: src/update_halo.jl:51-88) */
    IGG_MODE_IPC_LL128 = 2,   /* Like IGG_MODE_IPC_1PHASE, but the staged data carries its own arrival flags: every
                                 128-byte line of staging is 120 bytes of payload + the call's 8-byte epoch, stored by
                                 one warp instruction, so the sender needs neither a release fence nor a separate flag
                                 and the receiver unpacks each line as it lands (the latency regime; IGG_MODE=ll128). */
    IGG_MODE_IPC_AUTO = 3     /* DEFAULT.  Per call: LL128 when the call stages less than IGG_BULK_THRESHOLD bytes per
                                 rank (default 64 MiB: measured, LL128 is the faster one up to at least 16 MiB per rank), the fence protocol above.  The choice is made from the field
                                 sizes, so every rank makes the same one.  Each protocol has staging of its own. */
};
/* igg_set_mode is collective in the sense that every rank must select the same mode before the next update. */

/* Mirror of the reference's GlobalGrid (ref: src/shared.jl:58-81), reduced to what the halo path reads.
 * `neighbors[d][s]` is the rank of the neighbour in dimension d (0..2) on side s (0 = left/low, 1 = right/high)
 * or IGG_PROC_NULL; this is the memory order of the Julia 2x3 matrix `gg.neighbors`, so it can be passed as is.
 * Ranks must follow the MPI Cartesian convention rank = (cx*dims[1] + cy)*dims[2] + cz (row-major,
 * ref: src/gather.jl:40-46). */
typedef struct igg_grid_desc {
    int64_t nxyz[3];        /* local base grid size                      (gg.nxyz)       */
    int64_t dims[3];        /* process grid                              (gg.dims)       */
    int64_t coords[3];      /* my Cartesian coordinates                  (gg.coords)     */
    int64_t neighbors[3][2];/* see above                                 (gg.neighbors)  */
    int64_t periods[3];     /* 0/1                                       (gg.periods)    */
    int64_t overlaps[3];    /*                                           (gg.overlaps)   */
    int64_t halowidths[3];  /* default halowidths                        (gg.halowidths) */
    int32_t me;             /* my rank                                   (gg.me)         */
    int32_t nprocs;         /*                                           (gg.nprocs)     */
    int32_t device;         /* CUDA device ordinal to use (ref: src/select_device.jl:28-35: node-local rank) */
    int32_t reserved;       /* must be 0 */
} igg_grid_desc;

/* One field of an update_halo call = the reference's GGField (A, halowidths) (ref: src/shared.jl:50,252-256).
 * `ptr` addresses a dense column-major array of `size[0] x size[1] x size[2]` elements of `elsize` bytes
 * (size[d] = 1 for d >= ndims).  It may be device memory of the context's device, or page-locked host memory
 * that is mapped into the device address space (then the kernels read/write it over PCIe). */
typedef struct igg_field {
    void*   ptr;
    int64_t size[3];
    int64_t halowidths[3];
    int32_t elsize;         /* bytes per element, any value >= 1 (isbits types) */
    int32_t ndims;          /* 1..3 */
} igg_field;

typedef struct igg_ctx igg_ctx;

/* ---- library ---------------------------------------------------------------------------------------------- */
int         igg_version(void);
const char* igg_last_error(void);

/* ---- context life cycle ------------------------------------------------------------------------------------
 * igg_create   replaces: allocate_bufs / allocate_custreams / init_cubufs_arrays (ref: src/update_halo.jl:126-148,
 *              src/CUDAExt/update_halo.jl:60-76,155-159) -- but allocates lazily and grow-only inside the context.
 * igg_destroy  replaces: free_update_halo_buffers / free_update_halo_cubuffers (ref: src/update_halo.jl:109-114,
 *              src/CUDAExt/update_halo.jl:27-33), called from finalize_global_grid (ref: src/finalize_global_grid.jl:17).
 */
int igg_create(const igg_grid_desc* desc, igg_ctx** out);
int igg_destroy(igg_ctx* ctx);

/* ---- rank bootstrap (only needed when nprocs > 1) -----------------------------------------------------------
 * Replaces the MPI message transport of the halo path (ref: src/update_halo.jl:343-367 MPI.Isend / MPI.Irecv!)
 * by CUDA-IPC peer mappings that are set up once.  The library is transport agnostic: every rank exports an
 * opaque blob, the CALLER all-gathers the blobs in rank order (MPI.Allgather in Julia, torch.distributed in the
 * Python harness) and hands the concatenation back.  igg_bootstrap_import is collective over all ranks. */
size_t igg_bootstrap_size(void);
int    igg_bootstrap_export(igg_ctx* ctx, void* blob);
int    igg_bootstrap_import(igg_ctx* ctx, const void* all_blobs, int nprocs);

/* ---- update_halo! -------------------------------------------------------------------------------------------
 * Replaces `_update_halo!(fields...; dims)` (ref: src/update_halo.jl:44-89) including iwrite_sendbufs! /
 * iread_recvbufs! / wait_iwrite / wait_iread (ref: src/CUDAExt/update_halo.jl:143-203), the kernels write_d2x! /
 * read_x2d! (:210-227), write_d2h_async! / read_h2d_async! (:230-251), sendrecv_halo_local / gpumemcopy!
 * (ref: src/update_halo.jl:369-386, src/CUDAExt/update_halo.jl:257-259) and check_fields (:418-480).
 * `dims_order` lists the dimensions (1..3) to update, the reference's `dims=` keyword (default (3,1,2)); the result
 * does not depend on the order.  Collective over all ranks, same field list (sizes) on every rank.
 * `stream` is a cudaStream_t (NULL = default stream); the call is ordered after prior work on it. */
int igg_update_halo(igg_ctx* ctx, int nfields, const igg_field* fields,
                    const int* dims_order, int ndims_order, void* stream, int flags);

/* Prepared calls: validate the arguments, resolve the pointers and build the exchange plan ONCE, then repeat the same
 * update_halo! with one cheap call per time step.  The reference has no such object; what it corresponds to is the
 * steady state of `update_halo!` after its first call has allocated the buffers (ref: src/update_halo.jl:103-107,
 * 126-148 allocate_bufs; src/CUDAExt/update_halo.jl:60-76), minus the per-call check_fields (:418-480), which a
 * prepared call skips.  The fields (addresses, sizes, halowidths) must stay as they were when igg_prepare was called.
 * igg_update_halo_prepared has the semantics of igg_update_halo (collective, `flags`, stream order) and may be
 * captured into a CUDA graph with IGG_ASYNC.  Handles are released with igg_release_prepared (or simply dropped at
 * igg_destroy: using a handle after its context is gone returns IGG_ERR_STATE). */
typedef struct igg_prepared igg_prepared;
int igg_prepare(igg_ctx* ctx, int nfields, const igg_field* fields, const int* dims_order, int ndims_order,
                igg_prepared** out);
int igg_update_halo_prepared(igg_prepared* call, void* stream, int flags);
int igg_release_prepared(igg_prepared* call);

/* ---- direct-to-halo: fields the neighbours may store into -------------------------------------------------------
 * igg_alloc_field / igg_free_field replace, for the arrays themselves, what the reference leaves to CUDA.jl
 * (`CUDA.zeros`, e.g. ref: examples/diffusion3D_multigpu_CuArrays_novis.jl:27-33) -- and make the reference's receive
 * buffers and unpack step (ref: src/CUDAExt/update_halo.jl:190-203 iread_recvbufs!, :220-227 read_x2d!) unnecessary
 * for them: an allocation made here is SYMMETRIC (collective: every rank calls with the same size in the same order)
 * and mapped into every rank's address space through CUDA IPC, so that update_halo! of a field inside it stores every
 * send box that spans x (the y/z faces and the y-z edges: rows of at least 1 KiB) straight into the neighbour's halo
 * over NVLink -- for calls that stage at least 32 MiB per rank (IGG_DIRECT_THRESHOLD, bytes, the same on every rank;
 * below that LL128 staging is the faster way, measured on 2 and 8 B200s) --; the strided x-faces and fields in foreign
 * memory keep going through staging.  A field must lie at
 * the same offset of the same allocation on every rank (e.g. the k-th allocation holds array k everywhere).
 * Zero-initialised.  IGG_DIRECT=0 in the environment keeps everything staged (for comparison). */
int igg_alloc_field(igg_ctx* ctx, size_t bytes, void** ptr_out);
int igg_free_field(igg_ctx* ctx, void* ptr);

/* igg_update_halo_noalloc -- no counterpart in the reference.
 * The same halo update for callers that are not allowed to allocate device memory (a host application whose own pool
 * owns the GPU, a library entered where an allocation would stall the caller's streams, a time loop that guarantees
 * "no allocation after start-up").  It needs no context and never calls cudaMalloc / cudaFree / cudaHostAlloc: the
 * kernel's scratch (its tile table and control words) is a fixed array inside the library image, reserved on every
 * GPU that loads the library until the process exits (igg_noalloc_scratch_bytes() bytes, about 254 KiB).
 * Restrictions: desc->nprocs must be 1 (periodic self-exchange only; ranks talking to each other need peer staging);
 * a call may need at most IGG_NOALLOC_MAX_TILES tiles of 8 KiB halo data (about 25 MiB) -- larger calls return
 * IGG_ERR_ARG with a message naming the limit.  All calls share the one array, which keeps the last call's contents:
 * every call clears its control words and rewrites the table in stream order, so calls whose GPU work overlaps
 * (different streams or host threads) race -- serialise them.  The first call on a device resolves the array's
 * device address once; issue that first call before any stream capture begins.  A call on a stream that is being
 * captured is refused with IGG_ERR_UNSUPPORTED (the table is copied from host memory that only lives during the
 * call).  The call makes desc->device current and leaves it current.  With IGG_ASYNC a call whose plan differs from the
 * previous one on that device still waits for earlier work on the stream (the new table comes from ordinary host
 * memory); repeating the same shapes reuses the table and is stream-ordered.  `flags` as for igg_update_halo. */
#define IGG_NOALLOC_MAX_TILES 3200
int     igg_update_halo_noalloc(const igg_grid_desc* desc, int nfields, const igg_field* fields,
                                const int* dims_order, int ndims_order, void* stream, int flags);
int64_t igg_noalloc_scratch_bytes(void);

/* Select the exchange implementation (enum igg_mode).  IGG_MODE_NCCL_3PHASE needs igg_nccl_init first. */
int igg_set_mode(igg_ctx* ctx, int mode);
/* NCCL comparison path: rank 0 obtains an id (igg_nccl_unique_id_size bytes), the caller broadcasts it, every rank
 * calls igg_nccl_init (collective). */
size_t igg_nccl_unique_id_size(void);
int    igg_nccl_get_unique_id(void* id_out);
int    igg_nccl_init(igg_ctx* ctx, const void* id);

/* ---- gather! ------------------------------------------------------------------------------------------------
 * Replaces `gather!(A, A_global; root)` (ref: src/gather.jl:18-54, MPI.Gatherv! with a subarray datatype).
 * Every rank passes its block `A` (sizeA, device or host memory); on `root`, `A_global` (sizeG, device or host
 * memory) receives the block of the rank with coordinates (cx,cy,cz) at element offset (cx*sizeA[0], cy*sizeA[1],
 * cz*sizeA[2]).  `A_global` may be NULL on non-root ranks.  Collective and host-synchronous like the reference. */
int igg_gather(igg_ctx* ctx, const void* A, const int64_t sizeA[3], int ndimsA,
               void* A_global, const int64_t sizeG[3], int ndimsG,
               int elsize, int root, void* stream);

/* ---- the caller of the hot path: diffusion3D time step ------------------------------------------------------------
 * One step of the reference's example (ref: examples/diffusion3D_multigpu_CuArrays_novis.jl:41-49, the five
 * broadcasts qx, qy, qz, dTedt, T update) fused into one kernel: T2 = T + dt/Cp * div(lam grad T) on the inner cells,
 * boundary cells copied, same operation order, IEEE double without FMA (bit-identical to the example's arithmetic).
 * T, Cp, T2: dense column-major n[0] x n[1] x n[2] doubles on the current device.  part = 0 all cells, 1 = the two
 * outermost layers (what update_halo! sends; do it first, then exchange while part 2 runs), 2 = the rest; adding 4 to
 * `part` selects the bandwidth-bound arithmetic (divisions replaced by reciprocal multiplications, FMA allowed --
 * agrees with the exact variant to ~1e-15 relative, not bit for bit).
 * The caller then calls igg_update_halo on T2 and swaps T and T2. */
int igg_diffusion3d_step(const double* T, const double* Cp, double* T2, const int64_t n[3], double lam, double dt,
                         double dx, double dy, double dz, int part, void* stream);

/* ---- host-only helpers (no GPU needed) ----------------------------------------------------------------------
 * igg_sendranges / igg_recvranges replace sendranges / recvranges (ref: src/update_halo.jl:281-302): n = 1|2,
 * dim = 1..3, out = {x_lo,x_hi,y_lo,y_hi,z_lo,z_hi}, 1-based inclusive.  IGG_ERR_FIELDS if the field has no
 * halo in `dim` ("Incoherent arguments: ol(A,dim)<2*halowidths[dim]."). */
int igg_sendranges(const igg_grid_desc* desc, const igg_field* f, int n, int dim, int64_t out[6]);
int igg_recvranges(const igg_grid_desc* desc, const igg_field* f, int n, int dim, int64_t out[6]);
/* check_fields (ref: src/update_halo.jl:418-480): halowidth < 1, no halo at all, duplicates, differing types. */
int igg_check_fields(const igg_grid_desc* desc, int nfields, const igg_field* fields);

/* Plan inspection: the exact copy operations igg_update_halo would launch for this rank.  Used by the CPU tests
 * to execute the plan of every simulated rank with numpy and compare against the oracle. */
enum igg_op_kind { IGG_OP_PUSH = 0, IGG_OP_LOCAL = 1, IGG_OP_UNPACK = 2 };
typedef struct igg_plan_op {
    int32_t kind;           /* enum igg_op_kind */
    int32_t field;          /* index into the field list */
    int32_t offset[3];      /* neighbour offset o in {-1,0,1}^3 this op talks to */
    int32_t peer;           /* rank at coords+o (== me for IGG_OP_LOCAL) */
    int64_t src_lo[3], src_hi[3];  /* PUSH/LOCAL: source box in my field (1-based inclusive) */
    int64_t dst_lo[3], dst_hi[3];  /* LOCAL/UNPACK: destination box in my field; PUSH: box in the peer's field */
    int64_t staging_off;    /* PUSH: byte offset in the peer's staging half; UNPACK: in mine; LOCAL: -1 */
    int64_t nbytes;         /* bytes moved by this op */
    /* merged low-level form the kernel executes: `nz` planes of `ny` rows of `row_bytes` contiguous bytes */
    int64_t src_off, dst_off, src_sy, src_sz, dst_sy, dst_sz;
    int64_t row_bytes, ny, nz;
    int32_t vec_bytes;      /* access width chosen from the alignment of all of the above */
    int32_t ntiles;
    int32_t flag_base;      /* PUSH: first arrival flag set at the receiver (tile k sets flag_base + k); UNPACK: first
                               flag waited for; LOCAL: -1 */
    int32_t reserved;
} igg_plan_op;
int igg_plan_describe(const igg_grid_desc* desc, int nfields, const igg_field* fields,
                      const int* dims_order, int ndims_order,
                      igg_plan_op* ops_out, int ops_cap, int* nops_out, int64_t* staging_bytes_out);
/* The tile table the exchange kernel would walk in `mode` (IGG_MODE_IPC_1PHASE or IGG_MODE_IPC_LL128) for this call
 * on this rank: `*ntiles_out` records of 80 bytes each (layout: struct TileRec in csrc/igg_internal.h; PUSH tiles
 * first, then the LOCAL tiles and the x self-wraps (WRAPX), then UNPACK).  `mode` may also be
 * IGG_MODE_IPC_AUTO; adding 0x100 to it treats every field as living in a symmetric allocation (igg_alloc_field):
 * boxes that span x then go straight into the neighbour's field (PUSH_DIRECT tiles) and the receiver only waits.
 * Adding 0x200 treats every field as page-locked host memory on a 64-byte boundary: the table is then followed by the
 * whole-line tiles the host-field kernels run (kind 5: x self-wraps; kind 11: fetch / write-back of the row-end lines
 * of x-faces that go to another rank, whose cache the exchange tiles address as fields nfields, nfields + 1, ...).  `regions_out` (may be NULL) receives 4 x 27 numbers, indexed by neighbour offset
 * (ox+1) + 3(oy+1) + 9(oz+1): where that offset's staging slots start in the table's packed layout, how many bytes
 * of a staging half the offset needs, and the same two for its arrival flags.  (At run time every offset owns a
 * region whose position does not depend on the call; the launch shifts each offset's base accordingly.)
 * Host only, for tests that execute the table without a GPU.  No reference counterpart. */
int igg_plan_tiles(const igg_grid_desc* desc, int nfields, const igg_field* fields,
                   const int* dims_order, int ndims_order, int mode,
                   void* tiles_out, int tiles_cap, int* ntiles_out, int64_t* staging_bytes_out,
                   int64_t* regions_out);

/* Control-plane self test (host only).  After the bootstrap the library talks between ranks through one POSIX
 * shared-memory mailbox per rank (segment re-allocation hand-shakes, the barrier of gather!) -- the places where the
 * reference would call MPI (ref: src/tools.jl:896-897 MPI.Barrier, src/gather.jl:49 MPI.Gatherv!).  These two calls
 * run that mailbox across real processes without a GPU: create on every rank, exchange the 64-byte names, run. */
int igg_ctrl_selftest_create(int me, int nprocs, char name_out[64], void** handle);
int igg_ctrl_selftest_run(void* handle, int me, int nprocs, const char* names, int rounds);

/* ---- introspection used by tests / bench ---------------------------------------------------------------------*/
/* Kernel timeline for profiling: when enabled, thread 0 of every block of the next update_halo launch records
 * 8 words {globaltimer ns at start, SM clock at: start, first tile record loaded, first copy issued, pushes
 * signalled, first arrival flag seen, end, unused}; igg_trace_read synchronises and copies the last launch's
 * records (8 words per block, `*nblocks_out` blocks). */
/* Debug: the device-resident exchange state after synchronising the device: out[0] = exchange launches completed,
 * out[1..27] = exchanges completed per neighbour offset, out[28] = blocks-done counter (0 between launches),
 * out[29] = bytes of one staging half, out[30] = launches issued by the host, out[31] = current igg_mode (bits 0-7),
 * "LL128 passed the bootstrap link check and self-test" (bit 8), line images the self-test checked (bits 16..). */
int igg_debug_state(igg_ctx* ctx, uint64_t out[32]);
int igg_trace_enable(igg_ctx* ctx, int on);
int igg_trace_read(igg_ctx* ctx, uint64_t* out, int cap_blocks, int* nblocks_out);
/* Host cost of issuing `n` stream-ordered prepared calls back to back from C: nanoseconds per call (for profiles). */
int igg_debug_issue_loop(igg_prepared* call, void* stream, int n, double* ns_per_call_out);
/* The timed loop of bench.py, issued from C as a compiled-language host (Julia `ccall` loop) would issue it: `steps`
 * prepared calls, call i = calls[i % ncalls].  flags = IGG_ASYNC: one pair of CUDA events on `stream` around every
 * call, us_out[i] = event time of call i in microseconds, *loop_us_out = (first event to last event) / steps.
 * flags = IGG_ASYNC | 2: the same loop with only the first and the last event (us_out is left alone).
 * flags = IGG_SYNC: host clock around every host-synchronous call. */
int igg_debug_timed_calls(igg_prepared** calls, int ncalls, void* stream, int steps, int flags, float* us_out,
                          double* loop_us_out);
/* Peer-copy bandwidth over the mappings the exchange itself uses: copies `bytes` from this rank's copy of the symmetric
 * allocation `sym_ptr` (igg_alloc_field) into rank `peer`'s copy with cudaMemcpyAsync, `reps` times, and returns the
 * best GB/s (CUDA events).  The practical NVLink roof of this box for the roofline.  Not collective by itself, but
 * every rank copying to its right neighbour at once is the exchange's traffic pattern. */
int igg_debug_peer_copy(igg_ctx* ctx, void* sym_ptr, int peer, size_t bytes, int reps, double* gbs_out);
/* Number of kernels this context has launched so far (the bench's `gpu_launches`). */
int64_t igg_launch_count(const igg_ctx* ctx);
/* Name of the kernel symbol the update path launches (for profiles). */
const char* igg_kernel_name(void);

#ifdef __cplusplus
}
#endif
#endif /* IGG_B200_H */
