// The fused halo-exchange kernel of libigg_b200 (sm_100a).
//
// One launch per update_halo call does, for all fields and all 26 neighbour offsets:
//   PUSH   pack a send box of my field and store it straight into the neighbour GPU's staging slot through its
//          CUDA-IPC peer mapping (NVLink).  Two ways, chosen per plan (igg_internal.h, kProto*):
//            LL128 (latency regime)   staging lines of 120 payload bytes + the call's 8-byte epoch, no fence, no flag;
//            fence (bandwidth regime) 16-byte loads / stores, tiles of up to 128 KiB, each released on its own:
//                                     barrier, fence.acq_rel.sys, then the epoch goes into the tile's arrival flag.
//   LOCAL  periodic wrap inside my own array (the rank is its own neighbour): send box -> opposite receive box;
//   WRAPX  the same for the two x-faces of a field, one thread per row doing both directions (strided cells).
//   UNPACK wait for the tile's lines (LL128) or arrival flag (ld.acquire.sys), then copy it into my receive box.
// It replaces the reference's write_d2x! / read_x2d! kernels, the cuMemcpy3DAsync staging copies, the D2D
// self-copy and the 12 host-blocking stream synchronisations per field
// (ref: src/CUDAExt/update_halo.jl:153-259, src/update_halo.jl:51-88).
//
//   PUSH_DIRECT / WAIT  fields in symmetric allocations: a send box that spans x goes straight into the neighbour's
//          halo (after its "ready" word), the receiver only waits for the arrival flags.
// Work is cut into tiles (TileRec, built on the host in plan.cpp); block b handles tiles b, b+grid, ...  Tiles are
// ordered PUSH, LOCAL, WRAPX, UNPACK, so every block finishes (and flags) its pushes before it may spin on a flag.
// The grid never exceeds the number of co-resident blocks, so spinning blocks cannot starve pushing ones.
//
// Pure data movement: bytes are copied as opaque 1/2/4/8/16-byte words, so results are bit-exact for any element
// type.  No tensor cores.  Each copy-tile row is a contiguous run in both source and destination; strided x-faces
// degenerate to one element per row, and what they cost is DRAM's 128-byte line per cell in both directions
// (profiles/xface_floor.cu), not instructions.
#include <cuda_runtime.h>

#include <cstdlib>

#include "igg_internal.h"

namespace igg {

namespace {

template <typename V> struct Word;

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Loads.  Field data was written by earlier kernels of this stream: plain loads (measured on cfg2, cold L2: .cs
// 12.5 us per update, plain / .nc / .cg 9.6 us).  Staging data is written by a peer GPU during this kernel: read it
// at L2 (ld.cg), the point of coherence.
template <typename V> __device__ __forceinline__ V load_stream(const V* p) { return *p; }
template <typename V> __device__ __forceinline__ V load_l2(const V* p) { return __ldcg(p); }
template <> __device__ __forceinline__ unsigned char load_l2(const unsigned char* p) {
    return *reinterpret_cast<const volatile unsigned char*>(p);
}
template <> __device__ __forceinline__ unsigned short load_l2(const unsigned short* p) {
    return *reinterpret_cast<const volatile unsigned short*>(p);
}

// Loads kept in flight per thread: two 16-byte or four 8-byte ones (the strided x-face cells of 8-byte fields are
// latency-bound: x-only update 8.08 vs 8.25 us with four).  More, or four for the narrower words as well, spills
// registers at 4 blocks per SM.
template <typename V> struct Unroll { static constexpr int value = sizeof(V) == 8 ? 4 : 2; };

__device__ __forceinline__ int64_t i64(unsigned lo, unsigned hi) { return (int64_t)(((unsigned long long)hi << 32) | lo); }

// Reads a TileRec with five 16-byte read-only loads and rebuilds it field by field, so that it stays in registers.
__device__ __forceinline__ TileRec load_tile(const TileRec* g) {
    const uint4* q = reinterpret_cast<const uint4*>(g);
    const uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3), e = __ldg(q + 4);
    TileRec t;
    t.src_off = i64(a.x, a.y); t.dst_off = i64(a.z, a.w);
    t.src_sy = i64(b.x, b.y);  t.src_sz = i64(b.z, b.w);
    t.dst_sy = i64(c.x, c.y);  t.dst_sz = i64(c.z, c.w);
    t.row0 = d.x; t.nrows = d.y; t.ncols = d.z; t.ny = d.w;
    t.field = (uint16_t)(e.x & 0xffffu);
    t.vec_log2 = (uint8_t)((e.x >> 16) & 0xffu);
    t.kind = (uint8_t)(e.x >> 24);
    t.nbr = (uint8_t)(e.y & 0xffu);
    t.bulk = (uint8_t)((e.y >> 8) & 0xffu);
    t.flag_idx = e.z;
    return t;
}

// Tile `ti` of the launch: derived from the kernel parameters if it is one of the WRAPX tiles described there, else
// loaded from the table.  (The choice is uniform over the block.)
__device__ __forceinline__ TileRec get_tile(const KParams& p, int ti) {
    const int j = ti - p.wx_first;
    if (j >= 0 && j < p.wx_count) {
        int k = 0;
        while (k + 1 < p.n_wx_ops && (uint32_t)j >= p.wx[k + 1].tile0) ++k;
        const WrapXOp& w = p.wx[k];
        TileRec t;
        t.src_off = w.src_off; t.dst_off = w.dst_off; t.src_sy = w.sy; t.src_sz = w.sz; t.dst_sy = w.delta; t.dst_sz = 0;
        t.row0 = ((uint32_t)j - w.tile0) * w.rows_per_tile;
        t.nrows = min(w.rows_per_tile, w.rows - t.row0);
        t.ncols = w.ncols; t.ny = w.ny; t.field = w.field; t.vec_log2 = w.vec_log2; t.kind = (uint8_t)kOpWrapX; t.nbr = 0;
        t.bulk = 0; t.flag_idx = 0;
        return t;
    }
    return load_tile(p.tiles + ti);
}

// Byte offsets of item i of a tile: i -> (row r = i / ncols, column c = i % ncols) -> (y, z) of the flattened row.
template <typename V>
__device__ __forceinline__ void item_offsets(const TileRec& t, uint32_t i, int64_t& soff, int64_t& doff) {
    uint32_t r = 0, c = i;
    if (t.nrows != 1) { r = i / t.ncols; c = i - r * t.ncols; }
    const uint32_t row = t.row0 + r;
    uint32_t z = 0, y = row;
    if (row >= t.ny) { z = row / t.ny; y = row - z * t.ny; }
    const int64_t col = (int64_t)c * (int64_t)sizeof(V);
    soff = (int64_t)y * t.src_sy + (int64_t)z * t.src_sz + col;
    doff = (int64_t)y * t.dst_sy + (int64_t)z * t.dst_sz + col;
}

// Copies one tile; all loads of an unrolled group are issued before the first store (offsets are recomputed for
// the stores instead of being kept live, which keeps the kernel within 64 registers = 4 blocks per SM).
template <typename V, bool FROM_STAGING>
__device__ __forceinline__ void copy_tile(const TileRec& t, const char* __restrict__ src, char* __restrict__ dst) {
    constexpr int kUnroll = Unroll<V>::value;
    const uint32_t items = t.ncols * t.nrows;
    // (An up-front prefetch.global.L2 of strided x-face rows was measured and did not help: those rows are bound by
    // DRAM random-row throughput, 131 072 lines of 128 B per update at cfg2, not by the SM's miss queue.)
    for (uint32_t base = threadIdx.x; base < items; base += kThreads * kUnroll) {
        V v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const uint32_t i = base + u * kThreads;
            if (i < items) {
                int64_t soff, doff;
                item_offsets<V>(t, i, soff, doff);
                const V* sp = reinterpret_cast<const V*>(src + soff);
                v[u] = FROM_STAGING ? load_l2<V>(sp) : load_stream<V>(sp);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const uint32_t i = base + u * kThreads;
            if (i < items) {
                int64_t soff, doff;
                item_offsets<V>(t, i, soff, doff);
                *reinterpret_cast<V*>(dst + doff) = v[u];
            }
        }
    }
}

template <bool FROM_STAGING>
__device__ __forceinline__ void copy_tile_dispatch(const TileRec& t, const char* src, char* dst) {
    switch (t.vec_log2) {
        case 4: copy_tile<uint4, FROM_STAGING>(t, src, dst); break;
        case 3: copy_tile<uint2, FROM_STAGING>(t, src, dst); break;
        case 2: copy_tile<unsigned int, FROM_STAGING>(t, src, dst); break;
        case 1: copy_tile<unsigned short, FROM_STAGING>(t, src, dst); break;
        default: copy_tile<unsigned char, FROM_STAGING>(t, src, dst); break;
    }
}

// kOpWrapX: periodic x self-wrap, one thread per (row, word), both directions at once.  The four accesses of a row
// touch two 32-byte sectors (send cell and halo cell are neighbours at either end of the row).
// Strided single words: ask L2 for 64 bytes, not for the whole 128-byte line (the default L1 fill requests all four
// sectors: 524 k sector reads for 131 k cells; with the hint 262 k, profiles/xface_floor.cu: 3.4 -> 2.6 us L2-hot).
__device__ __forceinline__ unsigned long long ld_strided(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.global.L2::64B.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
template <typename V> __device__ __forceinline__ V ld_strided(const V* p) { return *p; }
template <> __device__ __forceinline__ uint2 ld_strided(const uint2* p) {
    const unsigned long long v = ld_strided(reinterpret_cast<const unsigned long long*>(p));
    return make_uint2((unsigned)v, (unsigned)(v >> 32));
}
template <> __device__ __forceinline__ unsigned int ld_strided(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.global.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// One item per thread and pass (more rows per thread measured no faster: profiles/xface_floor.cu, r1 / r2 / r4).
template <typename V>
__device__ __forceinline__ void wrapx_tile(const TileRec& t, char* __restrict__ base) {
    const uint32_t items = t.ncols * t.nrows;
    for (uint32_t i = threadIdx.x; i < items; i += kThreads) {
        uint32_t r = i, c = 0;
        if (t.ncols != 1) { r = i / t.ncols; c = i - r * t.ncols; }
        const uint32_t row = t.row0 + r;
        uint32_t z = 0, y = row;
        if (row >= t.ny) { z = row / t.ny; y = row - z * t.ny; }
        char* lo_send = base + t.src_off + ((int64_t)y * t.src_sy + (int64_t)z * t.src_sz + (int64_t)c * (int64_t)sizeof(V));
        char* hi_halo = lo_send + (t.dst_off - t.src_off);
        const V a = ld_strided(reinterpret_cast<const V*>(lo_send));
        const V b = ld_strided(reinterpret_cast<const V*>(hi_halo - t.dst_sy));
        *reinterpret_cast<V*>(hi_halo) = a;
        *reinterpret_cast<V*>(lo_send - t.dst_sy) = b;
    }
}
__device__ __forceinline__ void wrapx_tile_dispatch(const TileRec& t, char* base) {
    switch (t.vec_log2) {
        case 4: wrapx_tile<uint4>(t, base); break;
        case 3: wrapx_tile<uint2>(t, base); break;
        case 2: wrapx_tile<unsigned int>(t, base); break;
        case 1: wrapx_tile<unsigned short>(t, base); break;
        default: wrapx_tile<unsigned char>(t, base); break;
    }
}

// kOpWrapLines (fields in page-locked host memory): periodic x self-wrap of `nrows` rows through whole 64-byte lines.
// Eight neighbouring lanes own one row: lanes 0-3 its first line, lanes 4-7 its last line, so that every line is one
// 64-byte request on the way in and on the way out (no partial write reaches the host's memory controller).
// The halo cells are patched in shared memory: low halo <- high send, high halo <- low send.
// A kernel of its own (one tile per block, launched on the same stream next to the exchange kernel): as a branch of
// igg_halo_exchange_kernel it cost that kernel registers (spills at 4 blocks per SM).  It needs no ordering against
// the other tiles of the call: they neither write these rows nor read the halo cells written here.
struct WrapFields { void* field[IGG_MAX_FIELDS]; };

__global__ void __launch_bounds__(kThreads) igg_wrap_lines_kernel(const TileRec* __restrict__ tiles, const __grid_constant__ WrapFields f) {
    __shared__ uint4 s_lines[kThreads];               // one 16-byte piece of a host line per thread
    const TileRec t = load_tile(tiles + blockIdx.x);
    char* rows = static_cast<char*>(f.field[t.field]) + t.src_off;
    const int64_t sy = t.src_sy, sz = t.src_sz, pitch = t.dst_off;
    const int hwb = (int)t.dst_sy, olb = (int)t.dst_sz;
    const uint32_t row0 = t.row0, ny = t.ny;
    const uint32_t items = t.nrows * 8u;
    for (uint32_t i0 = 0; i0 < items; i0 += kThreads) {
        const uint32_t i = i0 + threadIdx.x;
        const bool on = i < items;                      // whole groups of 8 lanes are on or off together
        uint4* addr = nullptr;
        if (on) {
            const uint32_t row = row0 + (i >> 3), j = i & 7u;
            uint32_t z = 0, y = row;
            if (row >= ny) { z = row / ny; y = row - z * ny; }
            addr = reinterpret_cast<uint4*>(rows + (int64_t)y * sy + (int64_t)z * sz + (j < 4 ? 0 : pitch - 128)) + j;
            s_lines[threadIdx.x] = *addr;
        }
        __syncwarp();
        if (on) {
            unsigned char* L = reinterpret_cast<unsigned char*>(s_lines + (threadIdx.x & ~7u));   // [0,64) first line, [64,128) last
            for (int b = (int)(i & 7u); b < hwb; b += 8) {
                L[b] = L[128 - olb + b];                // low halo  <- high send
                L[128 - hwb + b] = L[olb - hwb + b];    // high halo <- low send
            }
        }
        __syncwarp();
        if (on) *addr = s_lines[threadIdx.x];
        __syncwarp();
    }
}

// kOpHostLinesX (fields in page-locked host memory whose x-faces go to another rank): copies the first and the last
// 64-byte line of `nrows` rows between the host field and a device cache (128 bytes per row: first line, last line).
// Eight lanes own a row (lanes 0-3 the first line, 4-7 the last), so every line is one 64-byte PCIe request in either
// direction and no partial write reaches the host's memory controller.  WRITE_BACK = false: host -> cache, in front of
// the exchange kernel; true: cache -> host, behind it.  Only the ends that are exchanged (t.dst_sz) are touched.
template <bool WRITE_BACK>
__global__ void __launch_bounds__(kThreads) igg_host_lines_kernel(const TileRec* __restrict__ tiles, const __grid_constant__ WrapFields f,
                                                                  char* __restrict__ cache) {
    const TileRec t = load_tile(tiles + blockIdx.x);
    char* rows = static_cast<char*>(f.field[t.field]) + t.src_off;
    const int64_t sy = t.src_sy, sz = t.src_sz, pitch = t.dst_off;
    const uint32_t sides = (uint32_t)t.dst_sz, ny = t.ny;
    uint4* lines = reinterpret_cast<uint4*>(cache + t.dst_sy);
    for (uint32_t i = threadIdx.x; i < t.nrows * 8u; i += kThreads) {
        const uint32_t row = t.row0 + (i >> 3), j = i & 7u;
        if (!((sides >> (j >> 2)) & 1u)) continue;
        uint32_t z = 0, y = row;
        if (row >= ny) { z = row / ny; y = row - z * ny; }
        uint4* host = reinterpret_cast<uint4*>(rows + (int64_t)y * sy + (int64_t)z * sz + (j < 4 ? 0 : pitch - 128)) + j;
        uint4* dev = lines + (size_t)row * 8 + j;
        if (WRITE_BACK) *host = *dev;
        else *dev = *host;
    }
}

// ---- bulk-copy pipeline (fence kernel): global -> shared -> global with cp.async.bulk (TMA unit) ------------------
// For tiles of long contiguous rows (the y/z faces of the bandwidth regime: 2 KiB - 512 KiB runs) one thread of the
// block moves the tile through a ring of shared-memory stages: bulk loads complete on an mbarrier per stage, bulk
// stores are committed as groups.  No register traffic, 48 KiB in flight per block, and the other warps of the block
// are free for the next tile.  Sources and destinations may be local memory or a peer GPU's (CUDA-IPC mapping).
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* b, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem, const void* gmem, uint32_t bytes, unsigned long long* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem)), "l"(gmem), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gmem, const void* smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem), "r"(smem_u32(smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// The geometry of a tile, by value (a TileRec reference would force the caller's record into local memory).
struct BulkTile { int64_t ssy, ssz, dsy, dsz; uint32_t row0, nrows, ncols, ny; };
// Chunk c of a tile: `nr` rows starting at tile row r0, `bytes` bytes each, from byte `off` of the row.
struct BulkChunk { uint32_t r0, nr, off, bytes; };
__device__ __forceinline__ BulkChunk bulk_chunk(uint32_t nrows, uint32_t c, uint32_t rowb, uint32_t rows_per_chunk, uint32_t segs) {
    BulkChunk k;
    if (segs > 1) { k.r0 = c / segs; k.nr = 1; k.off = (c - k.r0 * segs) * kBulkStageBytes; k.bytes = min((uint32_t)kBulkStageBytes, rowb - k.off); }
    else { k.r0 = c * rows_per_chunk; k.nr = min(rows_per_chunk, nrows - k.r0); k.off = 0; k.bytes = rowb; }
    return k;
}
__device__ __forceinline__ int64_t bulk_row_off(uint32_t row, uint32_t ny, int64_t sy, int64_t sz) {
    uint32_t z = 0, y = row;
    if (row >= ny) { z = row / ny; y = row - z * ny; }
    return (int64_t)y * sy + (int64_t)z * sz;
}

// Moves one tile; called by ONE thread.  `phase`: parity bit per stage, carried from tile to tile (returned).  On
// return all stores of the tile are complete (their writes performed), so the caller may release the tile or let the
// block end.
__device__ __noinline__ uint32_t bulk_copy_tile(const BulkTile t, const char* __restrict__ src, char* __restrict__ dst,
                                                unsigned char* stages, unsigned long long* mbar, uint32_t phase) {
    const uint32_t rowb = t.ncols << 4;
    uint32_t rows_per_chunk = 1, segs = 1;
    if (rowb >= (uint32_t)kBulkStageBytes) segs = (rowb + kBulkStageBytes - 1) / kBulkStageBytes;
    else rows_per_chunk = kBulkStageBytes / rowb;
    const uint32_t nchunks = segs > 1 ? t.nrows * segs : (t.nrows + rows_per_chunk - 1) / rows_per_chunk;
    for (uint32_t c = 0; c < nchunks + (kBulkStages - 1); ++c) {
        if (c < nchunks) {
            const uint32_t s = c % kBulkStages;
            if (c >= (uint32_t)kBulkStages) bulk_wait_read0();      // the store that last used this stage has read it
            const BulkChunk k = bulk_chunk(t.nrows, c, rowb, rows_per_chunk, segs);
            mbar_expect_tx(mbar + s, k.nr * k.bytes);
            for (uint32_t q = 0; q < k.nr; ++q)
                bulk_g2s(stages + s * kBulkStageBytes + q * k.bytes, src + bulk_row_off(t.row0 + k.r0 + q, t.ny, t.ssy, t.ssz) + k.off, k.bytes, mbar + s);
        }
        if (c >= (uint32_t)(kBulkStages - 1)) {
            const uint32_t d = c - (kBulkStages - 1), s = d % kBulkStages;
            mbar_wait(mbar + s, (phase >> s) & 1u);
            phase ^= 1u << s;
            const BulkChunk k = bulk_chunk(t.nrows, d, rowb, rows_per_chunk, segs);
            for (uint32_t q = 0; q < k.nr; ++q)
                bulk_s2g(dst + bulk_row_off(t.row0 + k.r0 + q, t.ny, t.dsy, t.dsz) + k.off, stages + s * kBulkStageBytes + q * k.bytes, k.bytes);
            bulk_commit();
        }
    }
    bulk_wait0();
    return phase;
}

// ---- IGG_MODE_IPC_LL128: line-formatted staging (see igg_internal.h) -------------------------------------------
__device__ __forceinline__ void st_volatile_v2(char* p, unsigned long long lo, unsigned long long hi) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ void ld_volatile_v2(const char* p, unsigned long long& lo, unsigned long long& hi) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
}

// Position of payload byte `pb` of a box inside the field: merged rows of `rowb` bytes, ny rows per plane.  32-bit
// index arithmetic: a staging slot is smaller than 4 GiB (checked when the plan is built).
struct BoxCursor {
    int64_t off;          // byte offset from the box origin in the field
    uint32_t c, y;        // byte within the row, row within the plane
};
__device__ __forceinline__ BoxCursor box_seek(uint32_t pb, uint32_t rowb, uint32_t ny, int64_t sy, int64_t sz) {
    BoxCursor k;
    uint32_t r = 0, z = 0;
    k.c = pb;
    if (pb >= rowb) { r = pb / rowb; k.c = pb - r * rowb; }
    k.y = r;
    if (r >= ny) { z = r / ny; k.y = r - z * ny; }
    k.off = (int64_t)k.y * sy + (int64_t)z * sz + k.c;
    return k;
}
template <int W>
__device__ __forceinline__ void box_step(BoxCursor& k, uint32_t rowb, uint32_t ny, int64_t sy, int64_t sz) {
    k.c += W; k.off += W;
    if (k.c == rowb) {
        k.c = 0; k.off += sy - (int64_t)rowb; k.y += 1;
        if (k.y == ny) { k.y = 0; k.off += sz - (int64_t)ny * sy; }
    }
}
template <int W> struct WordOf;
template <> struct WordOf<1> { typedef unsigned char T; };
template <> struct WordOf<2> { typedef unsigned short T; };
template <> struct WordOf<4> { typedef unsigned int T; };
template <> struct WordOf<8> { typedef unsigned long long T; };

static_assert(kLinesPerTile == 2 * (kThreads / 8), "an LL128 tile is two passes of the block");

// PUSH: pack the lines [row0, row0 + nrows) of a slot and store them into the neighbour's staging, 8 lanes per line.
// Both passes of the tile load before either stores (four 8-byte loads in flight per lane).
template <int W>
__device__ __forceinline__ void push_ll_tile(const TileRec& t, const char* __restrict__ box, char* __restrict__ slot,
                                             unsigned long long epoch) {
    typedef typename WordOf<W>::T Wt;
    const uint32_t j = threadIdx.x & 7u, g0 = threadIdx.x >> 3;
    const uint32_t nbytes = (uint32_t)t.dst_sy;
    const bool narrow = t.ncols <= 16u;
    unsigned long long v[2][2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const uint32_t g = g0 + u * (kThreads / 8);
        v[u][0] = v[u][1] = 0ull;
        if (g < t.nrows) {
            const uint32_t pb = (t.row0 + g) * kLinePayload + 16u * j;
            int nb = j == 7 ? 8 : 16;
            if (pb + nb > nbytes) nb = nbytes > pb ? (int)(nbytes - pb) : 0;
            if (nb > 0) {
                BoxCursor k = box_seek(pb, t.ncols, t.ny, t.src_sy, t.src_sz);
#pragma unroll
                for (int b = 0; b < 16; b += W) {
                    if (b < nb) {
                        // strided x-face cells (the row is one word): ask L2 for 64 bytes, not the whole line
                        const unsigned long long w = narrow ? (unsigned long long)ld_strided(reinterpret_cast<const Wt*>(box + k.off))
                                                            : (unsigned long long)*reinterpret_cast<const Wt*>(box + k.off);
                        v[u][b >> 3] |= w << (8 * (b & 7));
                        box_step<W>(k, t.ncols, t.ny, t.src_sy, t.src_sz);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const uint32_t g = g0 + u * (kThreads / 8);
        if (g < t.nrows) {
            if (j == 7) v[u][1] = epoch;
            st_volatile_v2(slot + (int64_t)(t.row0 + g) * 128 + 16 * j, v[u][0], v[u][1]);
        }
    }
}

// UNPACK: wait for each line (its last 8 bytes equal this call's epoch), then scatter its payload into the field.
// A warp polls its 4 lines together; returns false on time-out.
template <int W>
__device__ __forceinline__ bool unpack_ll_tile(const TileRec& t, char* __restrict__ box, const char* __restrict__ slot,
                                               unsigned long long epoch, unsigned long long timeout_ns) {
    typedef typename WordOf<W>::T Wt;
    const uint32_t j = threadIdx.x & 7u, lane = threadIdx.x & 31u;
    const uint32_t nbytes = (uint32_t)t.dst_sy;
    const uint32_t passes = (t.nrows + kThreads / 8 - 1) / (kThreads / 8);
    for (uint32_t ps = 0; ps < passes; ++ps) {                 // every lane runs every pass (warp-wide votes below)
        const uint32_t g = ps * (kThreads / 8) + (threadIdx.x >> 3);
        const bool on = g < t.nrows;
        const uint32_t line = t.row0 + g;
        const char* src = slot + (int64_t)line * 128 + 16 * j;
        unsigned long long v[2] = {0ull, 0ull};
        unsigned long long t0 = 0;
        unsigned int spins = 0;
        while (true) {
            if (on) ld_volatile_v2(src, v[0], v[1]);
            const unsigned long long flag = __shfl_sync(0xffffffffu, v[1], lane | 7u);
            if (__all_sync(0xffffffffu, !on || flag == epoch)) break;
            if ((++spins & 0x3ffu) == 0) {
                unsigned long long now = globaltimer_ns();
                now = __shfl_sync(0xffffffffu, now, 0);        // one clock for the whole warp: the exit must be uniform
                if (t0 == 0) t0 = now;
                else if (now - t0 > timeout_ns) return false;
            }
        }
        if (on) {
            const uint32_t pb = line * kLinePayload + 16u * j;
            int nb = j == 7 ? 8 : 16;
            if (pb + nb > nbytes) nb = nbytes > pb ? (int)(nbytes - pb) : 0;
            if (nb > 0) {
                BoxCursor k = box_seek(pb, t.ncols, t.ny, t.src_sy, t.src_sz);
#pragma unroll
                for (int b = 0; b < 16; b += W) {
                    if (b < nb) {
                        *reinterpret_cast<Wt*>(box + k.off) = (Wt)(v[b >> 3] >> (8 * (b & 7)));
                        box_step<W>(k, t.ncols, t.ny, t.src_sy, t.src_sz);
                    }
                }
            }
        }
    }
    return true;
}

// One thread waits until *word >= want (acquire, system scope); a device-side time-out sets the status word.
__device__ __noinline__ void spin_until(const unsigned long long* word, unsigned long long want, const KParams& p) {
    unsigned long long t0 = 0;
    unsigned int spins = 0;
    while (ld_acquire_sys(word) < want) {
        if ((++spins & 0x3ff) == 0) {
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > p.timeout_ns) { *reinterpret_cast<volatile int*>(p.status) = IGG_ERR_TIMEOUT; break; }
        }
    }
}

template <bool LL>
__device__ __forceinline__ void halo_exchange_body(const KParams& p) {
    __shared__ unsigned long long s_epoch;            // flag value of this launch (exchange launches so far + 1)
    __shared__ char* s_push[IGG_NOFFSETS];            // staging half this launch stores into at the neighbour at offset o
    __shared__ const char* s_unpack[IGG_NOFFSETS];    // staging half of mine the neighbour at offset o stores into
    __shared__ unsigned long long* s_rflags[IGG_NOFFSETS];
    __shared__ int s_lflag_delta[IGG_NOFFSETS];
    __shared__ __align__(8) unsigned long long s_mbar[kBulkStages];
    extern __shared__ __align__(128) unsigned char s_stages[];      // kBulkStages * kBulkStageBytes when p.bulk
    uint32_t bulk_phase = 0;
    if (p.bulk && threadIdx.x == 0) {                 // only thread 0 ever touches the barriers and the stages
        for (int k = 0; k < kBulkStages; ++k) mbar_init(&s_mbar[k]);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Optional per-block timeline (igg_trace_enable): [0] globaltimer at start, then SM clock at [1] start,
    // [2] first tile record loaded, [3] first copy issued, [4] first push flagged, [5] first flag seen, [6] end.
#define IGG_TRACE(k) do { if (p.trace && threadIdx.x == 0) p.trace[(size_t)blockIdx.x * 8 + (k)] = (unsigned long long)clock64(); } while (0)
    if (p.trace && threadIdx.x == 0) p.trace[(size_t)blockIdx.x * 8] = globaltimer_ns();
    IGG_TRACE(1);

    int ti = blockIdx.x;
    TileRec t;
    if (ti < p.n_tiles) t = get_tile(p, ti);          // in flight together with the state / plan loads below
    else t.kind = 0xff;                               // a launch without tiles only advances the exchange state
    // Programmatic dependent launch: this grid may have been scheduled while the previous kernel of the stream was
    // still running (its launch latency is hidden); nothing that kernel wrote is touched before this point.
    // (The tile table and the plan constants are immutable.)  A no-op when the launch carries no programmatic dependency.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (p.state) {
        if (threadIdx.x < IGG_NOFFSETS) {
            const int o = threadIdx.x;
            const long long half = (long long)((__ldcg(&p.state->exchanges[o]) + 1u) & 1u);
            if (p.plan) {
                s_push[o] = p.plan->push_base[o] + half * p.plan->push_half[o];
                s_unpack[o] = p.plan->unpack_base[o] + half * p.plan->unpack_half;
                s_rflags[o] = p.plan->remote_flags[o];
                s_lflag_delta[o] = p.plan->local_flag_delta[o];
            }
        }
        if (threadIdx.x == 32) s_epoch = __ldcg(&p.state->call_epoch) + 1ull;
        __syncthreads();
        // Direct-to-halo: this launch has started, so every earlier kernel of my stream is done with my halos: tell
        // the neighbours that will store into them (their PUSH_DIRECT tiles wait for exactly this epoch).
        if (blockIdx.x == 0 && threadIdx.x < IGG_NOFFSETS && p.plan && ((p.plan->direct_mask >> threadIdx.x) & 1u))
            st_relaxed_sys(p.plan->remote_ready[threadIdx.x], s_epoch);
    }
    // Every block has now read the state it needs and says so; the last block to have said so advances the state for
    // the next launch (which is stream-ordered after this one) -- after its tiles, below: the count is issued here and
    // its result is only looked at there, so that no warp waits for the atomic's round trip before its loads.
    // (Launches that exchange with no other rank carry no state: nothing to read, no barrier, no count.)
    unsigned int seen = 0;
    if (p.state && threadIdx.x == 0) seen = atomicAdd(&p.state->blocks_done, 1u);
    // Let the next kernel of the stream begin its launch; it cannot touch memory before this grid has completed (it
    // starts with griddepcontrol.wait, or it is an ordinary launch that waits for completion anyway).  Launches that
    // exchange with no other rank say so at once; launches whose blocks wait for neighbours only after their tiles
    // (below), so that an early-scheduled grid does not sit on SM slots while this one is still pushing.
    if (!p.state) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (p.trace && threadIdx.x == 0)   // [7]: what the block's first tile is (kind, vector width, rows, columns)
        p.trace[(size_t)blockIdx.x * 8 + 7] = (unsigned long long)t.kind | ((unsigned long long)t.vec_log2 << 8) |
                                              ((unsigned long long)t.nrows << 16) | ((unsigned long long)t.ncols << 40);
    IGG_TRACE(2);

    bool first = true;
    while (ti < p.n_tiles) {
        if (t.kind == kOpWrapX) {
            wrapx_tile_dispatch(t, static_cast<char*>(p.field[t.field]));
            if (first) IGG_TRACE(3);
        } else if (LL && t.kind == kOpPushLL) {
            const char* box = static_cast<const char*>(p.field[t.field]) + t.src_off;
            char* slot = s_push[t.nbr] + t.dst_off;
            switch (t.vec_log2) {
                case 3: push_ll_tile<8>(t, box, slot, s_epoch); break;
                case 2: push_ll_tile<4>(t, box, slot, s_epoch); break;
                case 1: push_ll_tile<2>(t, box, slot, s_epoch); break;
                default: push_ll_tile<1>(t, box, slot, s_epoch); break;
            }
            if (first) IGG_TRACE(3);
        } else if (LL && t.kind == kOpUnpackLL) {
            char* box = static_cast<char*>(p.field[t.field]) + t.src_off;
            const char* slot = s_unpack[t.nbr] + t.dst_off;
            bool ok;
            switch (t.vec_log2) {
                case 3: ok = unpack_ll_tile<8>(t, box, slot, s_epoch, p.timeout_ns); break;
                case 2: ok = unpack_ll_tile<4>(t, box, slot, s_epoch, p.timeout_ns); break;
                case 1: ok = unpack_ll_tile<2>(t, box, slot, s_epoch, p.timeout_ns); break;
                default: ok = unpack_ll_tile<1>(t, box, slot, s_epoch, p.timeout_ns); break;
            }
            if (!ok && (threadIdx.x & 31u) == 0) *reinterpret_cast<volatile int*>(p.status) = IGG_ERR_TIMEOUT;
            if (first) IGG_TRACE(5);
        } else if (t.bulk && p.bulk && (!LL || t.kind == IGG_OP_LOCAL)) {
            // Long contiguous rows: thread 0 drives the bulk-copy pipeline, the other warps go on to the next tile.
            if (threadIdx.x == 0) {
                const BulkTile bt = {t.src_sy, t.src_sz, t.dst_sy, t.dst_sz, t.row0, t.nrows, t.ncols, t.ny};
                if (!LL && t.kind == IGG_OP_UNPACK) {
                    const unsigned long long* flag = p.local_flags + ((long long)t.flag_idx + s_lflag_delta[t.nbr]);
                    unsigned long long t0 = 0;
                    unsigned int spins = 0;
                    while (ld_acquire_sys(flag) < s_epoch) {
                        if ((++spins & 0x3ff) == 0) {
                            const unsigned long long now = globaltimer_ns();
                            if (t0 == 0) t0 = now;
                            else if (now - t0 > p.timeout_ns) { *reinterpret_cast<volatile int*>(p.status) = IGG_ERR_TIMEOUT; break; }
                        }
                    }
                    if (first) IGG_TRACE(5);
                    fence_proxy_async();      // what the acquire made visible to this thread, the bulk loads must see too
                    bulk_phase = bulk_copy_tile(bt, s_unpack[t.nbr] + t.src_off, static_cast<char*>(p.field[t.field]) + t.dst_off, s_stages, s_mbar, bulk_phase);
                } else {
                    const char* src = static_cast<const char*>(p.field[t.field]) + t.src_off;
                    char* dst = (!LL && t.kind == IGG_OP_PUSH) ? s_push[t.nbr] + t.dst_off : static_cast<char*>(p.field[t.field]) + t.dst_off;
                    bulk_phase = bulk_copy_tile(bt, src, dst, s_stages, s_mbar, bulk_phase);
                    if (!LL && t.kind == IGG_OP_PUSH) {
                        // the tile's bulk stores are complete: order them before the flag for the neighbour
                        fence_proxy_async();
                        fence_acq_rel_sys();
                        st_relaxed_sys(s_rflags[t.nbr] + t.flag_idx, s_epoch);
                        if (first) IGG_TRACE(4);
                    }
                }
                if (first) IGG_TRACE(3);
            }
        } else if (t.kind == kOpWait) {
            // Direct-to-halo, receiver's side: the neighbour stores into my halo itself; only its arrival flag is
            // waited for (so that the launch does not end before the halo is complete).
            if (threadIdx.x == 0) {
                spin_until(p.local_flags + ((long long)t.flag_idx + s_lflag_delta[t.nbr]), s_epoch, p);
                if (first) IGG_TRACE(5);
            }
        } else if (!LL && t.kind == IGG_OP_UNPACK) {
            // Wait until the neighbour has flagged this very tile (it flags each tile right after storing it).
            if (threadIdx.x == 0) {
                const unsigned long long* flag = p.local_flags + ((long long)t.flag_idx + s_lflag_delta[t.nbr]);
                const unsigned long long want = s_epoch;
                unsigned long long t0 = 0;
                unsigned int spins = 0;
                while (ld_acquire_sys(flag) < want) {
                    if ((++spins & 0x3ff) == 0) {
                        const unsigned long long now = globaltimer_ns();
                        if (t0 == 0) t0 = now;
                        else if (now - t0 > p.timeout_ns) { *reinterpret_cast<volatile int*>(p.status) = IGG_ERR_TIMEOUT; break; }
                    }
                }
                if (first) IGG_TRACE(5);
            }
            __syncthreads();
            copy_tile_dispatch<true>(t, s_unpack[t.nbr] + t.src_off, static_cast<char*>(p.field[t.field]) + t.dst_off);
            if (first) IGG_TRACE(3);
        } else {
            // PUSH: my field -> peer staging; LOCAL: my field -> my field; COPY (gather!, NCCL pack): field -> buffer;
            // COPY_IN (NCCL unpack): buffer filled by an earlier stream operation -> my field.
            // PUSH_DIRECT (direct-to-halo): my send box -> the receive box of the NEIGHBOUR's copy of the field, once
            // the neighbour's launch of this epoch has started (ready word: its halo may be overwritten); released
            // like a PUSH tile.
            const char* src = (t.kind == kOpCopyIn ? p.copy_src[t.nbr] : static_cast<const char*>(p.field[t.field])) + t.src_off;
            char* dst;
            if (t.kind == IGG_OP_LOCAL || t.kind == kOpCopyIn) dst = static_cast<char*>(p.field[t.field]) + t.dst_off;
            else if (!LL && t.kind == IGG_OP_PUSH) dst = s_push[t.nbr] + t.dst_off;
            else if (t.kind == kOpPushDirect) {
                dst = p.reg_table[(long long)p.field_reg[t.field] * p.nprocs + p.plan->peer_rank[t.nbr]] + p.field_reg_off[t.field] + t.dst_off;
                if (threadIdx.x == 0) spin_until(p.local_ready + t.nbr, s_epoch, p);
                __syncthreads();
            }
            else dst = p.copy_dst[t.nbr] + t.dst_off;
            copy_tile_dispatch<false>(t, src, dst);
            if (first) IGG_TRACE(3);
            if ((!LL && t.kind == IGG_OP_PUSH) || t.kind == kOpPushDirect) {
                // Release this tile to the neighbour: all stores of the block (ordered before thread 0 by the barrier)
                // become visible system-wide before the tile's arrival flag does.  Other warps move on to the next tile.
                __syncthreads();
                if (threadIdx.x == 0) {
                    fence_acq_rel_sys();
                    st_relaxed_sys(s_rflags[t.nbr] + t.flag_idx, s_epoch);
                    if (first) IGG_TRACE(4);
                }
            }
        }
        // Later waves simply stride by the grid.  (Measured alternative: handing the second wave first to the blocks
        // whose first tile was not a PUSH, so that no UNPACK tile queues behind a block's release fence -- no change
        // at 2 and 4 GPUs, slower at 8: 22.6 vs 20.0 us on cfg2.  The pushing blocks are the ones that are free
        // earliest; the blocks that began with an UNPACK are still waiting for their own flag.)
        first = false;
        ti += gridDim.x;
        if (ti >= p.n_tiles) break;
        t = get_tile(p, ti);
    }

    if (p.state) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (p.state && threadIdx.x == 0 && seen == gridDim.x - 1) {
        p.state->blocks_done = 0;
        p.state->call_epoch = s_epoch;
        const uint32_t mask = p.plan ? p.plan->active_mask : 0u;
        for (int o = 0; o < IGG_NOFFSETS; ++o)
            if ((mask >> o) & 1u) p.state->exchanges[o] += 1u;
    }

    // Host-synchronous calls: the last block to finish tells the host through a mapped word (cheaper than a stream
    // synchronisation).  Every store of the block happens-before thread 0's fence (barrier), every block's fence
    // before its count, and the last block's system fence before the word.
    if (p.notify) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned int fin = atomicAdd(p.fin_counter, 1u);
            if (fin == gridDim.x - 1) {
                *p.fin_counter = 0;
                __threadfence_system();
                *reinterpret_cast<volatile unsigned long long*>(p.notify) = p.notify_value;
            }
        }
    }
    IGG_TRACE(6);
#undef IGG_TRACE
}

// Two instantiations of the one body (a single kernel with both protocols' code spills at 64 registers): plans whose
// remote tiles are PUSH / UNPACK (fence protocol; also every plan without remote tiles, gather! and the NCCL path's
// pack / unpack) run the first, plans with PUSH_LL / UNPACK_LL tiles (LL128) the second.
__global__ void __launch_bounds__(kThreads, kBlocksPerSM) igg_halo_exchange_kernel(const __grid_constant__ KParams p) { halo_exchange_body<false>(p); }
__global__ void __launch_bounds__(kThreads, kBlocksPerSM) igg_halo_exchange_ll128_kernel(const __grid_constant__ KParams p) { halo_exchange_body<true>(p); }
}  // namespace

int halo_kernel_max_coresident_blocks(int device) {
    int per_sm = 0, per_sm_ll = 0, sms = 0;
    // the fence kernel may be launched with the bulk-copy stages as dynamic shared memory
    if (cudaFuncSetAttribute(igg_halo_exchange_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkStages * kBulkStageBytes) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(igg_halo_exchange_ll128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkStages * kBulkStageBytes) != cudaSuccess) return -1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, igg_halo_exchange_kernel, kThreads, kBulkStages * kBulkStageBytes) != cudaSuccess) return -1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_ll, igg_halo_exchange_ll128_kernel, kThreads, kBulkStages * kBulkStageBytes) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return -1;
    return (per_sm < per_sm_ll ? per_sm : per_sm_ll) * sms;      // a grid that is co-resident for either kernel
}

int launch_halo_kernel(const KParams& p, int grid, void* stream, bool ll128) {
    // Programmatic dependent launch: the next grid is scheduled while this one still runs and waits at
    // griddepcontrol.wait, which hides its launch latency.  Default: on for launches that exchange with no other rank
    // (no exchange state to read, nothing to spin on: cfg2 on one GPU 10.3 -> 8.1 us per stream-ordered call, 8.8 ->
    // 7.9 us replayed from a graph), off for launches whose blocks wait for a neighbour: measured slower per
    // stream-ordered call on 2 GPUs whether the dependents are released in the prologue (14.1 vs 13.0 us) or after
    // the tiles (14.4 vs 13.2 us; replayed from a graph 10.1 vs 10.5 us).  IGG_PDL=1 / 0 forces it on / off.
    static const int pdl_env = [] { const char* e = getenv("IGG_PDL"); return !e ? -1 : (e[0] == '1' ? 1 : 0); }();
    const bool pdl = pdl_env >= 0 ? pdl_env == 1 : p.state == nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = p.bulk ? kBulkStages * kBulkStageBytes : 0;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e = ll128 ? cudaLaunchKernelEx(&cfg, igg_halo_exchange_ll128_kernel, p)
                          : cudaLaunchKernelEx(&cfg, igg_halo_exchange_kernel, p);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(IGG_ERR_CUDA, "halo kernel launch failed: %s", cudaGetErrorString(e)); }
    return IGG_OK;
}

int launch_wrap_lines_kernel(const TileRec* d_tiles, int n_tiles, void* const* fields, int nfields, void* stream) {
    WrapFields f;
    std::memset(&f, 0, sizeof(f));
    for (int i = 0; i < nfields && i < IGG_MAX_FIELDS; ++i) f.field[i] = fields[i];
    igg_wrap_lines_kernel<<<n_tiles, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d_tiles, f);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "wrap-lines kernel launch failed: %s", cudaGetErrorString(e));
    return IGG_OK;
}

int launch_host_lines_kernel(const TileRec* d_tiles, int n_tiles, void* const* fields, int nfields, char* cache,
                             bool write_back, void* stream) {
    WrapFields f;
    std::memset(&f, 0, sizeof(f));
    for (int i = 0; i < nfields && i < IGG_MAX_FIELDS; ++i) f.field[i] = fields[i];
    if (write_back) igg_host_lines_kernel<true><<<n_tiles, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d_tiles, f, cache);
    else igg_host_lines_kernel<false><<<n_tiles, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d_tiles, f, cache);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "host-lines kernel launch failed: %s", cudaGetErrorString(e));
    return IGG_OK;
}

const char* halo_kernel_name() { return "igg_halo_exchange_kernel"; }

}  // namespace igg
