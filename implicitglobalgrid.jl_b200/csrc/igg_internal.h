// Internal declarations shared by the host (plan, context) and device (kernel) parts of libigg_b200.
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "igg_b200.h"

namespace igg {

// ---- error reporting ---------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);          // thread-local message behind igg_last_error()
int  fail(int code, const char* fmt, ...);     // set_error + return code

// ---- device tables -----------------------------------------------------------------------------------------
// One tile = the unit of work of one thread block iteration: `nrows` rows starting at flattened row `row0`
// (row -> (y,z) = (row % ny, row / ny)), each row `ncols` vectors of `1 << vec_log2` bytes.  The column origin is
// already folded into src_off / dst_off.  80 bytes, read with five 16-byte loads.
struct alignas(16) TileRec {
    int64_t  src_off, dst_off;                  // byte offsets from the op's source / destination base pointer
    int64_t  src_sy, src_sz, dst_sy, dst_sz;    // byte strides between rows (y) and planes (z)
    uint32_t row0, nrows, ncols, ny;
    uint16_t field;                             // index into KParams::field
    uint8_t  vec_log2;                          // 0..4
    uint8_t  kind;                              // igg_op_kind
    uint8_t  nbr;                               // neighbour offset index 0..26 (PUSH: target, UNPACK: source)
    uint8_t  bulk;                              // 1: large contiguous rows -- may go through the bulk-copy (TMA) pipeline
    uint8_t  pad[2];
    uint32_t flag_idx;                          // PUSH: arrival flag to set at the receiver; UNPACK: flag to wait for
    uint32_t pad2;
};
static_assert(sizeof(TileRec) == 80, "TileRec layout");

constexpr int kThreads      = 256;      // threads per block of the halo kernel (128 measured slower: 11.6 vs 9.7 us at
                                        // 1 GPU, 22.5 vs 20.7 us at 8 GPUs on cfg2)
constexpr int kBlocksPerSM  = 1024 / kThreads;  // launch bound: 1024 resident threads per SM at <= 64 registers
constexpr int kTileBytes    = 8192;     // bytes moved per LOCAL / COPY tile (bulk PUSH / UNPACK tiles: Plan::tile_bytes)
constexpr int kMaxBulkTileBytes = 131072;
// Bulk-copy pipeline of the fence kernel (halo_kernel.cu): tiles of at least kBulkMinTileBytes whose rows are at least
// kBulkMinRowBytes long and 16-byte aligned are moved global -> shared -> global by cp.async.bulk through a ring of
// kBulkStages shared-memory stages of kBulkStageBytes, driven by one thread of the block.
constexpr int kBulkStageBytes = 16384, kBulkStages = 3, kBulkMinTileBytes = 16384, kBulkMinRowBytes = 1024;
constexpr size_t kFlagsOff   = 4096;    // control segment: arrival flags start here, one 8-byte word per tile
constexpr int kMaxFlags      = 1 << 17; // 131 072 tiles per call (about 1 GiB of halo data)
constexpr size_t kCtrlBytes  = kFlagsOff + (size_t)kMaxFlags * 8;

// Per-context exchange state, first bytes of the control segment (private to the owning GPU).
struct DeviceState {
    unsigned long long call_epoch;             // exchange launches completed on this context
    unsigned int exchanges[IGG_NOFFSETS];      // exchanges completed with the neighbour at offset o (staging parity)
    unsigned int blocks_done;                  // blocks of the running launch that have read the state
    unsigned int blocks_finished;              // blocks of the running launch that have finished (blocking calls only)
};

// Per-plan launch constants that live in DEVICE memory (uploaded when the plan is built and again whenever the
// context's staging layout changes): where each neighbour offset's staging and arrival flags are.  Keeping them out
// of the kernel parameters means that nothing per-offset is recomputed or copied per call (the parameters are about 1.2 KB,
// most of it the field pointers and the x self-wrap descriptors, filled from the call's arguments alone) and
// keeps launches that were captured into a CUDA graph valid when the layout moves (the graph only holds the pointer).
struct PlanDev {
    char*       push_base[IGG_NOFFSETS];       // staging (half 0) of the peer at offset o, shifted to o's region
    long long   push_half[IGG_NOFFSETS];       // bytes of one staging half on the peer at offset o
    const char* unpack_base[IGG_NOFFSETS];     // my staging (half 0), shifted to offset o's region
    unsigned long long* remote_flags[IGG_NOFFSETS];  // arrival flags in the control segment of the peer at offset o
    int32_t     local_flag_delta[IGG_NOFFSETS];      // UNPACK tile of offset o waits for local_flags[flag_idx + delta[o]]
    long long   unpack_half;                   // bytes of one staging half of mine
    uint32_t    active_mask;                   // bit o: launches of this plan exchange with the neighbour at offset o
    uint32_t    direct_mask;                   // bit o: some box goes straight into the field of the neighbour at offset o
    // Direct-to-halo: where my "launch k has started, my halos may be overwritten" word lives on the neighbour at
    // offset o (its ready[26 - o]), and which rank that neighbour is (row of the context's field table).
    unsigned long long* remote_ready[IGG_NOFFSETS];
    int32_t     peer_rank[IGG_NOFFSETS];
    int32_t     pad;
};
constexpr size_t kReadyOff = 2048;             // control segment: ready[27] words, written by the neighbours
constexpr int kMaxFieldRegs = 64;              // symmetric field allocations per context (igg_alloc_field)

// One periodic x self-wrap of a field (both x-faces), as the kernel needs it to derive a kOpWrapX tile WITHOUT a
// tile-table load: tile j of the op covers rows [j * rows_per_tile, ...) (see kOpWrapX below for the geometry).
struct WrapXOp {
    int64_t  src_off, dst_off, sy, sz;         // low send box, high halo box (byte offsets), row / plane strides
    uint32_t delta;                            // (ol - hw) * elsize
    uint32_t ny, rows, ncols, rows_per_tile;
    uint32_t tile0;                            // first tile of this op among the launch's WRAPX tiles
    uint16_t field;
    uint8_t  vec_log2, pad;
};
constexpr int kMaxWxOps = 8;

// Kernel parameters (passed by value).
struct KParams {
    const TileRec* tiles;
    int32_t  n_tiles;
    int32_t  n_push_tiles;                     // tiles [0, n_push_tiles) push to a neighbour
    const PlanDev* plan;                       // NULL for launches that exchange nothing (gather!, NCCL pack/unpack)
    // Exchange state in my control segment (NULL for launches that exchange nothing).  It lives on the device so that
    // the launch parameters do not change from call to call and an update_halo call can be captured into a CUDA
    // graph and replayed: the kernel derives this call's epoch (flag value) and each neighbour's staging half from
    // it and the last block to read it advances it.
    DeviceState* state;
    const unsigned long long* local_flags;     // my arrival-flag array (written by my neighbours)
    int*     status;                           // host-mapped status word (set on time-out)
    unsigned long long timeout_ns;
    unsigned long long* trace;                 // optional per-block timeline (8 words per block), NULL = off
    // Host-synchronous calls: the last block to finish stores `notify_value` into this host-mapped word, which the
    // host polls instead of paying a stream synchronisation (NULL = off).
    unsigned long long* notify;
    unsigned long long notify_value;
    unsigned int* fin_counter;                 // blocks finished so far (device word, reset by the last block)
    // Tiles [wx_first, wx_first + wx_count) are kOpWrapX tiles the kernel derives from `wx` below instead of loading
    // them from the table: the strided x-face loads -- the longest latency chain of a call -- start a memory round
    // trip earlier, and a tile-record fetch cannot queue behind the flood of strided loads (wx_count = 0: off).
    int32_t  bulk;                             // the launch has kBulkStages * kBulkStageBytes of dynamic shared memory
    // Direct-to-halo: field i of the call lives at byte field_reg_off[i] of symmetric allocation field_reg[i] (-1:
    // not in one); reg_table[field_reg[i] * nprocs + r] is where that allocation is mapped for rank r's copy.
    char* const* reg_table;
    int32_t  nprocs;
    const unsigned long long* local_ready;     // my ready[27] words
    int16_t  field_reg[IGG_MAX_FIELDS];
    long long field_reg_off[IGG_MAX_FIELDS];
    int32_t  wx_first, wx_count, n_wx_ops;
    WrapXOp  wx[kMaxWxOps];
    char*       copy_dst[2];                   // kOpCopy tiles: destination buffer t.nbr (gather!, NCCL pack)
    const char* copy_src[2];                   // kOpCopyIn tiles: source buffer t.nbr (NCCL unpack)
    void*    field[IGG_MAX_FIELDS];            // base pointers of the fields of this call
};

constexpr int kProtoFenceDefault = 0;
// ---- host-side plan ----------------------------------------------------------------------------------------
struct FieldKey {
    int64_t size[3];
    int64_t hw[3];
    int32_t elsize;
    int32_t ndims;
    int32_t align;       // ptr % 16
    int32_t lines;       // 1: page-locked host field on a 64-byte boundary (x self-wrap may go through whole lines)
    bool operator==(const FieldKey& o) const { return std::memcmp(this, &o, sizeof(*this)) == 0; }
};

// A page-locked host field whose x-faces are exchanged through a device cache of its row-end lines (plan.cpp).
struct HostXField {
    int      field = 0;                    // index into the call's field list; the cache is field nfields + k of the launch
    int64_t  rows = 0, ny = 0;             // rows of the x-face box (ny per plane)
    int64_t  cache_off = 0;                // byte offset of this field's lines in the context's cache (128 B per row)
    int64_t  row_org = 0, pitch = 0, plane = 0;   // first row of the box in the host field, byte strides
    uint32_t sides = 0;                    // bit 0: low end exchanged, bit 1: high end
};

struct Plan {
    std::vector<FieldKey>    key;
    int                      dims_mask = 0;          // bit d set: dimension d (0..2) is updated
    std::vector<igg_plan_op> ops;                    // PUSH ops, then LOCAL, then UNPACK
    std::vector<TileRec>     tiles;
    int      proto_req = kProtoFenceDefault;         // what the caller asked for (fence, LL128 or auto)
    bool     ll128 = false;                          // remote tiles are kOpPushLL / kOpUnpackLL (else PUSH / UNPACK)
    int32_t  tile_bytes = kTileBytes;                // bytes per PUSH / UNPACK tile of the fence protocol (per plan)
    int64_t  remote_bytes = 0;                       // payload bytes of all staging slots (the same on every rank)
    bool     global_active = false;                  // some rank exchanges with another rank in this call (rank-independent)
    bool     has_bulk = false;                       // some tile is flagged for the bulk-copy pipeline
    bool     small_tiles = false;                    // input: keep every tile at kTileBytes (igg_update_halo_noalloc's limit counts those)
    int64_t  local_bytes = 0;                        // bytes of all self-wrap boxes of this rank
    int32_t  n_push_tiles = 0;
    uint32_t n_flags = 0;                           // arrival flags this plan uses (must be <= kMaxFlags)
    uint32_t active_mask = 0;                        // bit o set: this plan exchanges with the neighbour at offset o
    uint32_t direct_mask = 0;                        // bit o set: some box goes straight into that neighbour's field
    int32_t  peer[IGG_NOFFSETS];                     // rank at offset o (or -1)
    int64_t  staging_bytes = 0;                      // bytes of one staging half this plan needs
    // Per neighbour offset o: where this plan's slots / flags of offset o start when the offsets are packed one after
    // the other (the numbers the tiles carry), and how much offset o needs.  At run time every offset has a REGION of
    // its own whose position does not depend on the plan (igg_ctx::region_off / flag_off), and the launch shifts each
    // offset's base pointer by (region_off[o] - region_pack[o]): memory an offset's neighbour writes is then never
    // shared with another offset's neighbour, whatever sequence of plans (field sets, dims=) the caller runs.
    int64_t  region_pack[IGG_NOFFSETS] = {0}, region_need[IGG_NOFFSETS] = {0};
    uint32_t flag_pack[IGG_NOFFSETS] = {0}, flag_need[IGG_NOFFSETS] = {0};
    int64_t  bytes_sent = 0;                         // algorithmic halo bytes this rank sends (local + remote)
    // device copy (owned by the context): [PlanDev | tiles] in one allocation
    PlanDev* d_plan = nullptr;
    TileRec* d_tiles = nullptr;
    int      grid = 0;
    bool     pinned = false;                         // referenced by a prepared call or a stream capture: never evicted
    // kOpWrapX tiles: where they are in `tiles` and the per-op descriptors they were cut from
    int32_t  wx_first = 0, wx_count = 0;
    std::vector<WrapXOp>     wx_ops;
    // exchanged x-faces of host-resident fields: cached lines (kOpHostLinesX tiles: fetched before, written back after)
    std::vector<HostXField>  hostx;
    std::vector<TileRec>     hostx_tiles;
    TileRec* d_hostx_tiles = nullptr;
    int64_t  hostx_cache_bytes = 0;
    // kOpWrapLines tiles of host-resident fields (run by igg_wrap_lines_kernel, one block each)
    std::vector<TileRec>     wrap_tiles;
    TileRec* d_wrap_tiles = nullptr;
};

inline int offset_index(const int o[3]) { return (o[0] + 1) + 3 * (o[1] + 1) + 9 * (o[2] + 1); }
inline void index_offset(int idx, int o[3]) { o[0] = idx % 3 - 1; o[1] = (idx / 3) % 3 - 1; o[2] = idx / 9 - 1; }
inline int opposite_index(int idx) { return 26 - idx; }

// Validates the grid description (coherent neighbours, disp == 1).
int check_grid(const igg_grid_desc& g);
// ol(dim,A) = overlaps[dim] + (size(A,dim) - nxyz[dim])      (ref: src/shared.jl:220)
inline int64_t field_ol(const igg_grid_desc& g, const igg_field& f, int d) {
    return g.overlaps[d] + (f.size[d] - g.nxyz[d]);
}
int check_fields(const igg_grid_desc& g, int nfields, const igg_field* fields);
int ranges(const igg_grid_desc& g, const igg_field& f, int n, int dim, bool send, int64_t out[6]);
// Builds ops + tiles for this rank.  Pure host code.  `host_lines[i]` != 0 (optional): field i lives in page-locked
// host memory on a 64-byte boundary, so its periodic x self-wrap may be emitted as kOpWrapLines tiles.
// `proto`: how data that goes to another rank is staged -- kProtoFence: PUSH / UNPACK tiles, one release fence and
// arrival flag per tile (tile size grows with the call: the bandwidth regime); kProtoLL128: kOpPushLL / kOpUnpackLL
// tiles over line-formatted staging (the latency regime); kProtoAuto: LL128 below auto_bulk_threshold() payload bytes
// per call, the fence protocol above (decided from numbers that are the same on every rank).
constexpr int kProtoFence = 0, kProtoLL128 = 1, kProtoAuto = 2;
int64_t auto_bulk_threshold();
// `direct[i]` != 0 (optional): field i lives in a symmetric allocation on every rank (direct-to-halo allowed).
int build_plan(const igg_grid_desc& g, int nfields, const igg_field* fields, int dims_mask, Plan& plan,
               const uint8_t* host_lines = nullptr, int proto = kProtoFence, const uint8_t* direct = nullptr);
void build_block_copy_tiles(int64_t el, const int64_t ext[3], const int64_t dst_size[2], const int64_t dst_origin[3],
                            int64_t src_align, int64_t dst_align, std::vector<TileRec>& tiles);
constexpr int kOpCopy = 3;   // internal tile kind: field[t.field] -> push_base[t.nbr], no signalling (gather!, NCCL pack)
constexpr int kOpCopyIn = 4; // internal tile kind: unpack_base[t.nbr] -> field[t.field], no waiting (NCCL unpack)
// Internal tile kind for fields in page-locked HOST memory: the periodic x self-wrap of a row done through whole
// 64-byte lines -- read the first and the last line of the row, copy low send -> high halo and high send -> low halo
// in registers/shared memory, write both lines back whole.  Partial writes cost the host a read-modify-write per
// cell (measured over PCIe: 65 536 strided 8-byte stores 328 us, the same rows as 64-byte lines 129 us).
// TileRec use: src_off/src_sy/src_sz/row0/nrows/ny address the rows (x = 0); dst_off = row pitch in bytes,
// dst_sy = halo width in bytes, dst_sz = overlap in bytes (<= 64); ncols = 8 lanes per row.
constexpr int kOpWrapLines = 5;
// Exchanged x-faces of page-locked HOST fields: one tile = up to kWrapRowsPerTile rows whose first and last 64-byte
// line are copied into / out of a device cache (128 bytes per row) by igg_host_lines_kernel.  TileRec use: src_off /
// src_sy / src_sz / row0 / nrows / ny address the rows in the host field, dst_off = row pitch, dst_sy = byte offset of
// the field's lines in the cache, dst_sz = sides (bit 0 low end, bit 1 high end).
constexpr int kOpHostLinesX = 11;
constexpr int kWrapRowsPerTile = 128;
// IGG_MODE_IPC_LL128: PUSH / UNPACK through line-formatted staging.  A staging slot of n payload bytes is
// ceil(n / 120) lines of 128 bytes: bytes [0,120) payload (the box in column-major order, as a stream of W-byte
// words, W = widest power of two <= 8 that divides element size and alignments), bytes [120,128) the epoch of the
// call that wrote the line.  Eight lanes own a line (16 bytes each; lane 7: 8 payload bytes + the epoch); a warp
// stores / loads 4 lines per instruction, and each line travels as one 128-byte transaction (what NCCL's LL128
// protocol relies on over NVLink).
// TileRec use: the FIELD side of either kind is in src_off / src_sy / src_sz / ny and ncols = bytes of a merged row;
// row0 = first line of the tile inside the slot, nrows = lines in the tile; dst_off = byte offset of the slot in a
// staging half, dst_sy = payload bytes of the slot; vec_log2 = log2(W); nbr as for PUSH / UNPACK.
constexpr int kOpPushLL = 6;
constexpr int kOpUnpackLL = 7;
// Periodic x self-wrap of device-resident fields, one thread per (row, word): the thread loads the row's low send
// word and its high send word together and stores them into the high halo and the low halo.  Replaces the two LOCAL
// ops (-x and +x) of a field, which walked the same rows twice; a send cell and the halo cell beside it share a
// 32-byte sector, so the row's two sectors are fetched once (profiles/xface_floor.cu: 6.8 us vs 9.2 us per 256^3 f64
// x-wrap, cold L2).
// TileRec use: src_off = byte offset of the low send box, dst_off = of the high halo box (row 0 of the tile's box),
// src_sy / src_sz = row / plane strides, dst_sy = (ol - hw) * elsize (high send = high halo - dst_sy, low halo =
// low send - dst_sy), ncols = words per side per row, row0 / nrows / ny as for copy tiles.
constexpr int kOpWrapX = 8;
// Direct-to-halo (fields inside symmetric allocations, igg_alloc_field): a send box whose rows span x goes straight
// into the receive box of the neighbour's copy of the field -- no staging, no unpack.  kOpPushDirect: like PUSH, but
// dst_off / dst_sy / dst_sz address the NEIGHBOUR's field (same shape as mine); the block first waits until the
// neighbour's launch of the same epoch has started (ready word: its earlier kernels are done with the halo), and
// releases the tile like the fence protocol does (barrier, fence.acq_rel.sys, arrival flag).  kOpWait: the
// receiver's side of such a tile: wait for the arrival flag, copy nothing.
constexpr int kOpPushDirect = 9;
constexpr int kOpWait = 10;
constexpr int64_t kDirectMinRowBytes = 1024;   // direct only for boxes that span x of fields at least this wide
constexpr int kWrapXRowsPerTile = 256;
constexpr int kLinePayload = 120;
constexpr int kLinesPerTile = 64;
int build_dim_phase(const igg_grid_desc& g, int nfields, const igg_field* fields, int d, std::vector<TileRec>& pack,
                    std::vector<TileRec>& unpack, int64_t bytes[2]);
// dims=(...) keyword -> bit mask of updated dimensions (NULL / empty = all three).
int dims_mask_from_order(const int* dims_order, int n, int* mask_out);
const char* last_error_cstr();

// ---- NCCL comparison path (nccl_path.cu) ------------------------------------------------------------------
struct NcclPath;
size_t nccl_unique_id_size();
int  nccl_get_unique_id(void* out);
int  nccl_create(NcclPath** out, int nprocs, int me, const void* id_bytes);
void nccl_free(NcclPath* np);
int  nccl_phase(NcclPath* np, const igg_grid_desc& g, int nfields, const igg_field* fields, int d, int max_blocks,
                const KParams& proto, void* stream, int64_t* launches);

// ---- LL128 self-test (ll128_check.cu) ----------------------------------------------------------------------
int  ll128_selftest(char* to_peer, const char* mine, int lines, int rounds, bool corrupt, unsigned long long timeout_ns,
                    unsigned long long out[3]);

// ---- kernel launcher (halo_kernel.cu) ----------------------------------------------------------------------
int  halo_kernel_max_coresident_blocks(int device);     // occupancy * SM count
int  launch_halo_kernel(const KParams& p, int grid, void* stream, bool ll128 = false);
const char* halo_kernel_name();
// kOpWrapLines tiles (host-resident fields), one block per tile, stream-ordered.
int  launch_wrap_lines_kernel(const TileRec* d_tiles, int n_tiles, void* const* fields, int nfields, void* stream);
// kOpHostLinesX tiles: host lines -> cache (write_back = false) or cache -> host lines (true), one block per tile.
int  launch_host_lines_kernel(const TileRec* d_tiles, int n_tiles, void* const* fields, int nfields, char* cache,
                              bool write_back, void* stream);

}  // namespace igg
