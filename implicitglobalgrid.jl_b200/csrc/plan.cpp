// Host-side exchange plan: which boxes of which field go where.  Pure C++ (no CUDA calls), so it can be
// exercised on a machine without a GPU through igg_plan_describe / igg_sendranges / igg_check_fields.
//
// Semantics follow the reference (paths relative to the reference root):
//   - per-field overlap and "has a halo" predicate      src/shared.jl:220, src/update_halo.jl:239,266,346,359,371
//   - send / receive boxes                              src/update_halo.jl:281-302
//   - who talks to whom                                 src/update_halo.jl:51-88 (+ MPI Cart_shift, src/create_global_grid.jl:98-100)
//   - argument checks                                   src/update_halo.jl:418-480
// The reference runs three dependent phases (z, x, y), each moving full faces.  Here all dimensions are exchanged
// in ONE phase with up to 26 neighbours: a face message excludes the cells that belong to another updated
// dimension's halo, and those cells arrive directly from the edge / corner neighbour instead.  Sources are never
// halo cells and destinations are always halo cells, so the exchange is hazard-free in place and bit-identical to
// the sequential scheme (checked against the oracle in tests/test_plan_vs_oracle.py).
#include <algorithm>
#include <array>
#include <cstdarg>
#include <cstdlib>

#include "igg_internal.h"

namespace igg {

static thread_local std::string g_last_error;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

const char* last_error_cstr() { return g_last_error.c_str(); }

static int64_t cart_rank(const int64_t c[3], const int64_t dims[3]) {
    return (c[0] * dims[1] + c[1]) * dims[2] + c[2];   // row-major (ref: src/gather.jl:40-46)
}

int check_grid(const igg_grid_desc& g) {
    int64_t np = 1;
    for (int d = 0; d < 3; ++d) {
        if (g.dims[d] < 1) return fail(IGG_ERR_GRID, "Invalid grid: dims[%d] = %lld < 1.", d, (long long)g.dims[d]);
        if (g.coords[d] < 0 || g.coords[d] >= g.dims[d])
            return fail(IGG_ERR_GRID, "Invalid grid: coords[%d] out of range.", d);
        if (g.nxyz[d] < 1) return fail(IGG_ERR_GRID, "Invalid arguments: nx, ny, and nz cannot be less than 1.");
        if (g.halowidths[d] < 1) return fail(IGG_ERR_GRID, "Invalid arguments: halowidths cannot be less than 1.");
        if (g.periods[d] != 0 && g.periods[d] != 1)
            return fail(IGG_ERR_GRID, "Invalid arguments: periodx, periody, and periodz must be either 0 or 1.");
        np *= g.dims[d];
    }
    if (np != g.nprocs) return fail(IGG_ERR_GRID, "Invalid grid: prod(dims) != nprocs.");
    if (cart_rank(g.coords, g.dims) != g.me)
        return fail(IGG_ERR_GRID, "Invalid grid: `me` is not the row-major Cartesian rank of `coords`.");
    for (int d = 0; d < 3; ++d) {
        const int64_t l = g.neighbors[d][0], r = g.neighbors[d][1];
        // ref: src/update_halo.jl:57,68,76
        const bool none_me = (l != g.me) && (r != g.me);
        const bool all_me = (l == g.me) && (r == g.me);
        if (!none_me && !all_me)
            return fail(IGG_ERR_GRID,
                        "Incoherent neighbors in dimension %d: either all neighbors must equal to me, or none.", d + 1);
        for (int s = 0; s < 2; ++s) {
            int64_t c[3] = {g.coords[0], g.coords[1], g.coords[2]};
            c[d] += s ? 1 : -1;
            int64_t want;
            if (g.periods[d]) {
                c[d] = (c[d] + g.dims[d]) % g.dims[d];
                want = cart_rank(c, g.dims);
            } else {
                want = (c[d] >= 0 && c[d] < g.dims[d]) ? cart_rank(c, g.dims) : IGG_PROC_NULL;
            }
            const int64_t got = g.neighbors[d][s] < 0 ? IGG_PROC_NULL : g.neighbors[d][s];
            if (got != want)
                return fail(IGG_ERR_UNSUPPORTED,
                            "neighbors[%d][%d] = %lld is not the rank at coords%+d (%lld): only disp=1 Cartesian "
                            "topologies are supported.", d, s, (long long)got, s ? 1 : -1, (long long)want);
        }
    }
    return IGG_OK;
}

static std::string join_positions(const std::vector<std::string>& v) {
    std::string s;
    for (size_t i = 0; i < v.size(); ++i) {
        if (i > 0) s += (i + 1 == v.size()) ? " and " : ", ";
        s += v[i];
    }
    return s;
}

// ref: src/update_halo.jl:418-480 (the checks that need Julia type information -- isbits, contiguity, array type --
// stay in the host-language wrapper).
int check_fields(const igg_grid_desc& g, int nfields, const igg_field* f) {
    if (nfields < 1) return fail(IGG_ERR_ARG, "update_halo needs at least one field.");
    if (nfields > IGG_MAX_FIELDS)
        return fail(IGG_ERR_ARG, "At most %d fields per call are supported (got %d).", IGG_MAX_FIELDS, nfields);
    for (int i = 0; i < nfields; ++i) {
        if (!f[i].ptr) return fail(IGG_ERR_ARG, "The field at position %d has a NULL pointer.", i + 1);
        if (f[i].ndims < 1 || f[i].ndims > 3) return fail(IGG_ERR_ARG, "The field at position %d has ndims outside 1..3.", i + 1);
        if (f[i].elsize < 1) return fail(IGG_ERR_ARG, "The field at position %d has elsize < 1.", i + 1);
        for (int d = 0; d < 3; ++d) {
            if (f[i].size[d] < 1) return fail(IGG_ERR_ARG, "The field at position %d has a size < 1.", i + 1);
            if (d >= f[i].ndims && f[i].size[d] != 1)
                return fail(IGG_ERR_ARG, "The field at position %d has size != 1 beyond its ndims.", i + 1);
        }
    }
    std::vector<std::string> pos;
    for (int i = 0; i < nfields; ++i)
        for (int d = 0; d < f[i].ndims; ++d)
            if (f[i].halowidths[d] < 1) { pos.push_back(std::to_string(i + 1)); break; }
    if (pos.size() > 1) return fail(IGG_ERR_FIELDS, "The fields at positions %s have a halowidth less than 1.", join_positions(pos).c_str());
    if (pos.size() == 1) return fail(IGG_ERR_FIELDS, "The field at position %s has a halowidth less than 1.", pos[0].c_str());

    pos.clear();
    for (int i = 0; i < nfields; ++i) {
        bool none = true;
        for (int d = 0; d < f[i].ndims; ++d)
            if (field_ol(g, f[i], d) >= 2 * f[i].halowidths[d]) none = false;
        if (none) pos.push_back(std::to_string(i + 1));
    }
    if (pos.size() > 1) return fail(IGG_ERR_FIELDS, "The fields at positions %s have no halo; remove them from the call.", join_positions(pos).c_str());
    if (pos.size() == 1) return fail(IGG_ERR_FIELDS, "The field at position %s has no halo; remove it from the call.", pos[0].c_str());

    std::vector<std::pair<int, int>> dup;
    for (int i = 0; i < nfields; ++i)
        for (int j = i + 1; j < nfields; ++j)
            if (f[i].ptr == f[j].ptr) dup.push_back({i + 1, j + 1});
    if (dup.size() > 2) {
        pos.clear();
        for (auto& p : dup) pos.push_back("[" + std::to_string(p.first) + ", " + std::to_string(p.second) + "]");
        return fail(IGG_ERR_FIELDS, "The pairs of fields with the positions %s are the same; remove any duplicates from the call.", join_positions(pos).c_str());
    }
    if (!dup.empty())
        return fail(IGG_ERR_FIELDS, "The field at position %d is a duplicate of the one at the position %d; remove the duplicate from the call.", dup[0].second, dup[0].first);

    pos.clear();
    for (int i = 1; i < nfields; ++i)
        if (f[i].elsize != f[0].elsize || f[i].ndims != f[0].ndims) pos.push_back(std::to_string(i + 1));
    if (pos.size() > 1) return fail(IGG_ERR_FIELDS, "The fields at positions %s are of different type than the first field; make sure that in a same call all fields are of the same type.", join_positions(pos).c_str());
    if (pos.size() == 1) return fail(IGG_ERR_FIELDS, "The field at position %s is of different type than the first field; make sure that in a same call all fields are of the same type.", pos[0].c_str());
    return IGG_OK;
}

// ref: src/update_halo.jl:281-302.  n = 1|2, dim = 1..3.
int ranges(const igg_grid_desc& g, const igg_field& f, int n, int dim, bool send, int64_t out[6]) {
    if (n < 1 || n > 2 || dim < 1 || dim > 3) return fail(IGG_ERR_ARG, "ranges: n must be 1|2 and dim 1..3.");
    const int d = dim - 1;
    const int64_t ol = field_ol(g, f, d), hw = f.halowidths[d];
    if (ol < 2 * hw) return fail(IGG_ERR_FIELDS, "Incoherent arguments: ol(A,dim)<2*halowidths[dim].");
    int64_t i;
    if (send) i = (n == 2) ? f.size[d] - (ol - 1) : 1 + (ol - hw);
    else      i = (n == 2) ? f.size[d] - (hw - 1) : 1;
    for (int k = 0; k < 3; ++k) { out[2 * k] = 1; out[2 * k + 1] = f.size[k]; }
    out[2 * d] = i;
    out[2 * d + 1] = i + hw - 1;
    return IGG_OK;
}

namespace {

struct Box { int64_t lo[3], hi[3]; };

int64_t round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

// Merges contiguous dimensions, picks the access width and cuts the copy into tiles.
// The copy moves `ext[1]*ext[2]` rows of `ext[0]*el` bytes; strides are in bytes.
struct Lowered { int64_t src_off, dst_off, s_sy, s_sz, d_sy, d_sz, row, ny, nz, v, ntiles, nbytes; };

Lowered emit_tiles(int64_t el, const int64_t ext[3], int64_t src_off, int64_t s_sy, int64_t s_sz, int64_t dst_off,
                   int64_t d_sy, int64_t d_sz, int64_t src_align, int64_t dst_align, int field_idx, int kind, int nbr,
                   std::vector<TileRec>& tiles, int64_t tile_bytes = kTileBytes) {
    int64_t row = ext[0] * el, ny = ext[1], nz = ext[2];
    if (ny == 1) { ny = nz; nz = 1; s_sy = s_sz; d_sy = d_sz; }
    while (ny > 1 && s_sy == row && d_sy == row) { row *= ny; ny = nz; nz = 1; s_sy = s_sz; d_sy = d_sz; }
    if (ny == 1) { s_sy = d_sy = 0; }
    if (nz == 1) { s_sz = d_sz = 0; }
    int64_t v = 16;
    auto bad = [&](int64_t vv) {
        if (row % vv) return true;
        if ((src_align + src_off) % vv) return true;
        if ((dst_align + dst_off) % vv) return true;
        if (s_sy % vv || s_sz % vv || d_sy % vv || d_sz % vv) return true;
        return false;
    };
    while (v > 1 && bad(v)) v /= 2;
    int vlog = 0;
    while ((1 << vlog) < v) ++vlog;

    const int64_t V = row / v, rows = ny * nz;
    const int64_t per_tile = tile_bytes / v;
    int64_t S, R;
    if (V >= per_tile) { S = per_tile; R = 1; } else { S = V; R = std::max<int64_t>(1, per_tile / V); }
    const int64_t segs = (V + S - 1) / S, rblocks = (rows + R - 1) / R;
    // Tiles with long 16-byte-aligned rows may go through the bulk-copy pipeline (IGG_BULK=0: never).
    static const bool bulk_on = !(getenv("IGG_BULK") && atoi(getenv("IGG_BULK")) == 0);
    const bool bulk = bulk_on && v == 16 && tile_bytes >= kBulkMinTileBytes && S * v >= kBulkMinRowBytes &&
                      (kind == IGG_OP_PUSH || kind == IGG_OP_LOCAL || kind == IGG_OP_UNPACK);       // (not the direct kinds)
    for (int64_t rb = 0; rb < rblocks; ++rb)
        for (int64_t sg = 0; sg < segs; ++sg) {
            TileRec t;
            std::memset(&t, 0, sizeof(t));
            t.src_off = src_off + sg * S * v;
            t.dst_off = dst_off + sg * S * v;
            t.src_sy = s_sy; t.src_sz = s_sz; t.dst_sy = d_sy; t.dst_sz = d_sz;
            t.row0 = (uint32_t)(rb * R);
            t.nrows = (uint32_t)std::min<int64_t>(R, rows - rb * R);
            t.ncols = (uint32_t)std::min<int64_t>(S, V - sg * S);
            t.ny = (uint32_t)ny;
            t.field = (uint16_t)field_idx;
            t.vec_log2 = (uint8_t)vlog;
            t.kind = (uint8_t)kind;
            t.nbr = (uint8_t)nbr;
            t.bulk = (uint8_t)((bulk && (int64_t)t.nrows * t.ncols * v >= kBulkMinTileBytes / 2) ? 1 : 0);
            tiles.push_back(t);
        }
    return Lowered{src_off, dst_off, s_sy, s_sz, d_sy, d_sz, row, ny, nz, v, segs * rblocks, row * rows};
}

// Fills the merged low-level form + tiles of one op.  `src_field` / `dst_field`: the box lives in the field array
// (strides of the field) or in a packed staging slot (strides of the box itself, origin = staging_off).
void lower_op(const igg_field& f, int field_idx, int kind, int nbr, const Box& sb, const Box& db, bool src_field,
              bool dst_field, int64_t staging_off, igg_plan_op& op, std::vector<TileRec>& tiles,
              int64_t tile_bytes = kTileBytes) {
    const int64_t el = f.elsize;
    int64_t b[3];
    for (int d = 0; d < 3; ++d) b[d] = sb.hi[d] - sb.lo[d] + 1;
    const int64_t fsy = f.size[0] * el, fsz = f.size[0] * f.size[1] * el;
    const int64_t psy = b[0] * el, psz = b[0] * b[1] * el;
    auto field_off = [&](const Box& x) { return (x.lo[0] - 1) * el + (x.lo[1] - 1) * fsy + (x.lo[2] - 1) * fsz; };
    const int64_t align = (int64_t)((uintptr_t)f.ptr % 16);
    const Lowered L = emit_tiles(el, b, src_field ? field_off(sb) : staging_off, src_field ? fsy : psy,
                                 src_field ? fsz : psz, dst_field ? field_off(db) : staging_off, dst_field ? fsy : psy,
                                 dst_field ? fsz : psz, src_field ? align : 0, dst_field ? align : 0, field_idx, kind,
                                 nbr, tiles, tile_bytes);
    op.src_off = L.src_off; op.dst_off = L.dst_off;
    op.src_sy = L.s_sy; op.src_sz = L.s_sz; op.dst_sy = L.d_sy; op.dst_sz = L.d_sz;
    op.row_bytes = L.row; op.ny = L.ny; op.nz = L.nz;
    op.vec_bytes = (int32_t)L.v;
    op.ntiles = (int32_t)L.ntiles;
    op.nbytes = L.nbytes;
}

}  // namespace

// Payload bytes per call above which kProtoAuto stages through the fence protocol (16-byte accesses, big tiles)
// instead of LL128 lines.  IGG_BULK_THRESHOLD (bytes) overrides it; it must be the same on every rank.
int64_t auto_bulk_threshold() {
    static const int64_t v = [] {
        const char* e = getenv("IGG_BULK_THRESHOLD");
        return (e && *e) ? (int64_t)atoll(e) : (int64_t)64 << 20;
    }();
    return v;
}

// Staged bytes per call from which fields in symmetric allocations are updated direct-to-halo.  The ready hand-shake
// and the release fence per tile make it slower than LL128 staging until the call is big enough for the saved unpack
// pass to matter.  Measured: 2 B200s in a row along z (profiles/r02/direct_ab_n2_r02.log) cfg2, 1 MiB staged: 16.0 vs
// 10.8 us; cfg4 F = 1 / 2 / 4 (4 / 8 / 16 MiB): 33.7 / 61.2 / 115.6 vs 35.5 / 71.1 / 143.6 us.  8 B200s, 2x2x2
// (profiles/r02/configs_r02_n8_sym.jsonl vs profiles/configs_r02_n8.jsonl): cfg3ii periodic (12 MiB) 63.5 vs 39.8 us,
// cfg4 F = 1 / 2 / 4 (12 / 24 / 48 MiB) 72.1 / 130.3 / 187.8 vs 43.5 / 97.9 / 209.4 us.  Hence 32 MiB.
// IGG_DIRECT_THRESHOLD overrides it (bytes; the same on every rank).
static int64_t direct_threshold() {
    static const int64_t v = [] {
        const char* e = getenv("IGG_DIRECT_THRESHOLD");
        return (e && *e) ? (int64_t)atoll(e) : (int64_t)32 << 20;
    }();
    return v;
}

int build_plan(const igg_grid_desc& g, int nfields, const igg_field* fields, int dims_mask, Plan& plan,
               const uint8_t* host_lines, int proto, const uint8_t* direct) {
    int rc = check_grid(g);
    if (rc) return rc;
    rc = check_fields(g, nfields, fields);
    if (rc) return rc;

    plan.key.clear();
    for (int i = 0; i < nfields; ++i) {
        FieldKey k;
        std::memset(&k, 0, sizeof(k));
        for (int d = 0; d < 3; ++d) { k.size[d] = fields[i].size[d]; k.hw[d] = fields[i].halowidths[d]; }
        k.elsize = fields[i].elsize; k.ndims = fields[i].ndims;
        k.align = (int32_t)((uintptr_t)fields[i].ptr % 16);
        k.lines = ((host_lines && host_lines[i]) ? 1 : 0) | ((direct && direct[i]) ? 2 : 0);
        plan.key.push_back(k);
    }
    plan.dims_mask = dims_mask;
    plan.ops.clear();
    plan.tiles.clear();
    plan.wrap_tiles.clear();
    plan.wx_ops.clear();
    plan.active_mask = 0;
    plan.direct_mask = 0;
    // Direct-to-halo for (offset, field): the field lives in a symmetric allocation and the box spans x with rows of
    // at least kDirectMinRowBytes -- decided from numbers that are the same on every rank.  Strided x-faces (and the
    // edges / corners with an x component) stay staged: they would cross NVLink as 8-byte stores.
    bool direct_on = false;          // set once the call's staged bytes are known (below)
    auto is_direct = [&](const int o[3], int i) {
        return direct_on && direct[i] && o[0] == 0 && fields[i].size[0] * (int64_t)fields[i].elsize >= kDirectMinRowBytes;
    };
    plan.bytes_sent = 0;

    bool exists[3][2];
    for (int d = 0; d < 3; ++d)
        for (int s = 0; s < 2; ++s) exists[d][s] = g.neighbors[d][s] >= 0;

    // has-halo predicate per field and dimension (ref: src/shared.jl:220; src/update_halo.jl:239)
    std::vector<std::array<bool, 3>> halo(nfields);
    for (int i = 0; i < nfields; ++i)
        for (int d = 0; d < 3; ++d)
            halo[i][d] = ((dims_mask >> d) & 1) && field_ol(g, fields[i], d) >= 2 * fields[i].halowidths[d] &&
                         (g.dims[d] > 1 || g.periods[d]);   // no rank has a neighbour in d otherwise (global fact)

    // Every rank must take part in a call or none (the exchange epoch is compared across ranks): whether a call
    // exchanges between ranks at all is decided from numbers that are the same everywhere.  A rank that happens to
    // have nothing to copy (arrays without interior cells, ref: src/update_halo.jl:239 accepts them) still launches.
    plan.global_active = false;
    for (int i = 0; i < nfields; ++i)
        for (int d = 0; d < 3; ++d) plan.global_active = plan.global_active || (halo[i][d] && g.dims[d] > 1);

    // Staging slots: one per (offset, field), identical on every rank (depends only on field sizes, halowidths and
    // the updated dimensions, never on which neighbours exist), so a sender can compute the receiver's layout.
    std::vector<int64_t> slot_off((size_t)IGG_NOFFSETS * nfields, -1);
    std::vector<int64_t> slot_n((size_t)IGG_NOFFSETS * nfields, 0);
    plan.remote_bytes = 0;
    plan.local_bytes = 0;
    for (int oi = 0; oi < IGG_NOFFSETS; ++oi) {
        int o[3];
        index_offset(oi, o);
        if (!o[0] && !o[1] && !o[2]) continue;
        for (int i = 0; i < nfields; ++i) {
            bool ok = true, remote = false;     // remote: the rank at this offset is another rank (on every rank alike)
            int64_t n = fields[i].elsize;
            for (int d = 0; d < 3; ++d) {
                if (o[d]) { ok = ok && halo[i][d]; remote = remote || g.dims[d] > 1; n *= fields[i].halowidths[d]; }
                else n *= fields[i].size[d];
            }
            if (ok && !remote) plan.local_bytes += n;         // (an upper bound: self-wraps of periodic one-rank dimensions)
            if (!ok || !remote) continue;       // self-wraps copy inside the array and need no staging slot
            slot_n[(size_t)oi * nfields + i] = n;
            plan.remote_bytes += n;
        }
    }
    direct_on = direct != nullptr && plan.remote_bytes >= direct_threshold();
    const bool ll128 = proto == kProtoLL128 || (proto == kProtoAuto && plan.remote_bytes < auto_bulk_threshold());
    plan.ll128 = ll128;
    // Fence protocol: tiles grow with the call so that a call has a few hundred PUSH tiles (one release fence each),
    // whatever its size: 8 KiB ... 128 KiB, the same number on every rank.
    int64_t tb = kTileBytes;
    while (tb < kMaxBulkTileBytes && plan.remote_bytes / tb > 256) tb *= 2;
    plan.tile_bytes = (int32_t)tb;
    // Self-wraps: 8-KiB tiles (the latency regime's tuning) until a call wraps more than 8 MiB, then growing alike.
    int64_t tbl = kTileBytes;
    if (plan.local_bytes > ((int64_t)8 << 20) && !plan.small_tiles)
        while (tbl < kMaxBulkTileBytes && plan.local_bytes / tbl > 512) tbl *= 2;
    // Arrival flags: one per tile.  A slot reserves an upper bound of the tiles any box that fits it can have
    // (every full tile covers more than kTileBytes/2 bytes), so that flag numbers too are identical on every rank.
    std::vector<uint32_t> slot_flag((size_t)IGG_NOFFSETS * nfields, 0), slot_nflag((size_t)IGG_NOFFSETS * nfields, 0);
    uint32_t nflags = 0;
    int64_t total = 0;
    for (int oi = 0; oi < IGG_NOFFSETS; ++oi) {
        plan.region_pack[oi] = total; plan.region_need[oi] = 0;
        plan.flag_pack[oi] = nflags;  plan.flag_need[oi] = 0;
    }
    for (int oi = 0; oi < IGG_NOFFSETS; ++oi) {
        int o[3];
        index_offset(oi, o);
        plan.region_pack[oi] = total;
        plan.flag_pack[oi] = nflags;
        // (the needs of offset oi are filled in after its fields, below)
        struct Close { Plan& p; int oi; int64_t& total; uint32_t& nflags;
                       ~Close() { p.region_need[oi] = total - p.region_pack[oi]; p.flag_need[oi] = nflags - p.flag_pack[oi]; } } close{plan, oi, total, nflags};
        if (!o[0] && !o[1] && !o[2]) continue;
        plan.peer[oi] = -1;
        for (int i = 0; i < nfields; ++i) {
            const int64_t n = slot_n[(size_t)oi * nfields + i];
            if (n == 0) continue;
            slot_off[(size_t)oi * nfields + i] = total;
            if (is_direct(o, i)) {              // no staging bytes, only arrival flags (one per tile)
                slot_flag[(size_t)oi * nfields + i] = nflags;
                slot_nflag[(size_t)oi * nfields + i] = (uint32_t)(2 * ((n + tb - 1) / tb) + 1);
                nflags += slot_nflag[(size_t)oi * nfields + i];
                continue;
            }
            if (ll128) {                        // lines of 120 payload bytes + 8 flag bytes; no separate flags
                if (n >= (int64_t)1 << 32)
                    return fail(IGG_ERR_ARG, "update_halo: a halo box of 4 GiB or more is not supported by the LL128 protocol.");
                total += (n + kLinePayload - 1) / kLinePayload * 128;
                continue;
            }
            total += round_up(n, 128);
            slot_flag[(size_t)oi * nfields + i] = nflags;
            slot_nflag[(size_t)oi * nfields + i] = (uint32_t)(2 * ((n + tb - 1) / tb) + 1);
            nflags += slot_nflag[(size_t)oi * nfields + i];
        }
    }
    plan.peer[13] = -1;
    plan.staging_bytes = total;
    plan.n_flags = nflags;

    std::vector<igg_plan_op> push, local, unpack;
    std::vector<TileRec> tpush, twrapx, tlocal, tunpack;

    // Host-resident fields whose x-faces go to another rank through cached whole lines (see the op loop): index into
    // plan.hostx or -1.  IGG_HOST_LINES_X=0 keeps the per-cell path.
    std::vector<int> hostx(nfields, -1);
    plan.hostx.clear();
    plan.hostx_cache_bytes = 0;
    static const bool hostx_on = !(getenv("IGG_HOST_LINES_X") && atoi(getenv("IGG_HOST_LINES_X")) == 0);
    for (int i = 0; i < nfields && hostx_on; ++i) {
        const igg_field& f = fields[i];
        const int64_t el = f.elsize, pitch = f.size[0] * el;
        if (host_lines && host_lines[i] && g.dims[0] > 1 && halo[i][0] && pitch % 64 == 0 && pitch >= 128 && 64 % el == 0 &&
            field_ol(g, f, 0) * el <= 64 && nfields + (int)plan.hostx.size() < IGG_MAX_FIELDS) {
            hostx[i] = (int)plan.hostx.size();
            HostXField hx;
            hx.field = i;
            plan.hostx.push_back(hx);
        }
    }

    for (int oi = 0; oi < IGG_NOFFSETS; ++oi) {
        int o[3];
        index_offset(oi, o);
        if (!o[0] && !o[1] && !o[2]) continue;
        // neighbour at coords + o
        bool nb_ok = true;
        int64_t c[3];
        for (int d = 0; d < 3; ++d) {
            c[d] = g.coords[d] + o[d];
            if (o[d]) {
                nb_ok = nb_ok && exists[d][o[d] > 0 ? 1 : 0];
                if (g.periods[d]) c[d] = (c[d] + g.dims[d]) % g.dims[d];
            }
        }
        if (!nb_ok) continue;
        const int64_t peer = cart_rank(c, g.dims);
        plan.peer[oi] = (int32_t)peer;

        for (int i = 0; i < nfields; ++i) {
            bool has = true;
            for (int d = 0; d < 3; ++d)
                if (o[d]) has = has && halo[i][d];
            if (!has) continue;
            if (peer != g.me && slot_off[(size_t)oi * nfields + i] < 0)
                return fail(IGG_ERR_STATE, "internal: exchange with another rank without a staging slot.");
            const igg_field& f = fields[i];
            Box snd, rcv_peer, rcv_mine;
            for (int d = 0; d < 3; ++d) {
                if (o[d]) {
                    int64_t r[6];
                    const int n = o[d] < 0 ? 1 : 2;
                    ranges(g, f, n, d + 1, true, r);      // my send box of side n ...
                    snd.lo[d] = r[2 * d]; snd.hi[d] = r[2 * d + 1];
                    ranges(g, f, 3 - n, d + 1, false, r); // ... lands in the receive box of side 3-n on the peer
                    rcv_peer.lo[d] = r[2 * d]; rcv_peer.hi[d] = r[2 * d + 1];
                    ranges(g, f, n, d + 1, false, r);     // what the neighbour at +o sends lands in my side n
                    rcv_mine.lo[d] = r[2 * d]; rcv_mine.hi[d] = r[2 * d + 1];
                } else {
                    // Cells of this dimension's own halo are delivered by the edge / corner neighbours instead.
                    const int64_t lo = (halo[i][d] && exists[d][0]) ? f.halowidths[d] + 1 : 1;
                    const int64_t hi = (halo[i][d] && exists[d][1]) ? f.size[d] - f.halowidths[d] : f.size[d];
                    snd.lo[d] = rcv_peer.lo[d] = rcv_mine.lo[d] = lo;
                    snd.hi[d] = rcv_peer.hi[d] = rcv_mine.hi[d] = hi;
                }
            }
            bool empty = false;
            for (int d = 0; d < 3; ++d) empty = empty || snd.hi[d] < snd.lo[d];
            if (empty) continue;

            igg_plan_op op;
            std::memset(&op, 0, sizeof(op));
            op.field = i;
            for (int d = 0; d < 3; ++d) {
                op.offset[d] = o[d];
                op.src_lo[d] = snd.lo[d]; op.src_hi[d] = snd.hi[d];
                op.dst_lo[d] = rcv_peer.lo[d]; op.dst_hi[d] = rcv_peer.hi[d];
            }
            op.peer = (int32_t)peer;
            if (peer == g.me) {
                op.kind = IGG_OP_LOCAL;
                op.staging_off = -1;
                op.flag_base = -1;
                // Host-resident field, pure x offset, the rank is its own x neighbour on both sides: the two
                // x-face wraps of a row go through whole 64-byte lines (kOpWrapLines).  The op description stays
                // the cell-level one; only the tiles differ.  Rows of this box are written by no other op, and the
                // interior cells of the two lines are rewritten with the values just read.
                const int64_t el = f.elsize, pitch = f.size[0] * el;
                const bool lines = host_lines && host_lines[i] && o[0] && !o[1] && !o[2] && exists[0][0] && exists[0][1] &&
                                   g.neighbors[0][0] == g.me && g.neighbors[0][1] == g.me && pitch % 64 == 0 &&
                                   pitch >= 128 && field_ol(g, f, 0) * el <= 64;
                if (lines) {
                    std::vector<TileRec> unused;
                    lower_op(f, i, IGG_OP_LOCAL, oi, snd, rcv_peer, true, true, 0, op, unused);
                    if (o[0] < 0) {       // once per field: the -x op emits the tiles for both sides
                        const int64_t ny = snd.hi[1] - snd.lo[1] + 1, nz = snd.hi[2] - snd.lo[2] + 1, rows = ny * nz;
                        for (int64_t r0 = 0; r0 < rows; r0 += kWrapRowsPerTile) {
                            TileRec t;
                            std::memset(&t, 0, sizeof(t));
                            t.src_off = (snd.lo[1] - 1) * pitch + (snd.lo[2] - 1) * pitch * f.size[1];
                            t.src_sy = pitch; t.src_sz = pitch * f.size[1];
                            t.dst_off = pitch; t.dst_sy = f.halowidths[0] * el; t.dst_sz = field_ol(g, f, 0) * el;
                            t.row0 = (uint32_t)r0;
                            t.nrows = (uint32_t)std::min<int64_t>(kWrapRowsPerTile, rows - r0);
                            t.ncols = 8; t.ny = (uint32_t)ny;
                            t.field = (uint16_t)i; t.vec_log2 = 4; t.kind = (uint8_t)kOpWrapLines; t.nbr = (uint8_t)oi;
                            plan.wrap_tiles.push_back(t);
                        }
                    }
                } else if (o[0] && !o[1] && !o[2] && exists[0][0] && exists[0][1]) {
                    // The two x-faces of a device-resident field wrap together (kOpWrapX): the -x op emits the tiles
                    // for both sides, one thread per row and word doing both directions.  The op descriptions stay
                    // the cell-level ones.  (The -x and +x boxes cover the same rows.)
                    std::vector<TileRec> unused;
                    lower_op(f, i, IGG_OP_LOCAL, oi, snd, rcv_peer, true, true, 0, op, unused);
                    if (o[0] < 0) {
                        const int64_t hwb = f.halowidths[0] * el, delta = (field_ol(g, f, 0) - f.halowidths[0]) * el;
                        const int64_t fsy = pitch, fsz = pitch * f.size[1];
                        const int64_t ny = snd.hi[1] - snd.lo[1] + 1, nz = snd.hi[2] - snd.lo[2] + 1, rows = ny * nz;
                        const int64_t row_org = (snd.lo[1] - 1) * fsy + (snd.lo[2] - 1) * fsz;
                        const int64_t lo_send = row_org + delta, hi_halo = row_org + (f.size[0] - f.halowidths[0]) * el;
                        const int64_t align = (int64_t)((uintptr_t)f.ptr % 16);
                        int64_t v = 16;
                        while (v > 1 && (hwb % v || delta % v || (align + lo_send) % v || (align + hi_halo) % v || fsy % v || fsz % v)) v /= 2;
                        int vlog = 0;
                        while ((1 << vlog) < v) ++vlog;
                        const int64_t ncols = hwb / v, per = std::max<int64_t>(1, kWrapXRowsPerTile / ncols);
                        WrapXOp w;
                        std::memset(&w, 0, sizeof(w));
                        w.src_off = lo_send; w.dst_off = hi_halo; w.sy = fsy; w.sz = fsz; w.delta = (uint32_t)delta;
                        w.ny = (uint32_t)ny; w.rows = (uint32_t)rows; w.ncols = (uint32_t)ncols; w.rows_per_tile = (uint32_t)per;
                        w.tile0 = (uint32_t)twrapx.size(); w.field = (uint16_t)i; w.vec_log2 = (uint8_t)vlog;
                        plan.wx_ops.push_back(w);
                        for (int64_t r0 = 0; r0 < rows; r0 += per) {
                            TileRec t;
                            std::memset(&t, 0, sizeof(t));
                            t.src_off = lo_send; t.dst_off = hi_halo;
                            t.src_sy = fsy; t.src_sz = fsz; t.dst_sy = delta;
                            t.row0 = (uint32_t)r0; t.nrows = (uint32_t)std::min<int64_t>(per, rows - r0);
                            t.ncols = (uint32_t)ncols; t.ny = (uint32_t)ny;
                            t.field = (uint16_t)i; t.vec_log2 = (uint8_t)vlog; t.kind = (uint8_t)kOpWrapX; t.nbr = (uint8_t)oi;
                            twrapx.push_back(t);
                        }
                    }
                } else {
                    lower_op(f, i, IGG_OP_LOCAL, oi, snd, rcv_peer, true, true, 0, op, tlocal, tbl);
                }
                local.push_back(op);
                plan.bytes_sent += op.nbytes;
            } else {
                op.kind = IGG_OP_PUSH;
                op.staging_off = slot_off[(size_t)opposite_index(oi) * nfields + i];
                if (is_direct(o, i)) {
                    // Straight into the neighbour's receive box (its copy of the field has my field's shape).  The op
                    // descriptions say so with staging_off = -1 on both sides; the receiver only waits.
                    const size_t first = tpush.size();
                    op.staging_off = -1;
                    lower_op(f, i, kOpPushDirect, oi, snd, rcv_peer, true, true, 0, op, tpush, tb);
                    if (tpush.size() - first > slot_nflag[(size_t)opposite_index(oi) * nfields + i])
                        return fail(IGG_ERR_STATE, "internal: a direct op has more tiles than its slot has arrival flags.");
                    for (size_t k = first; k < tpush.size(); ++k) {
                        tpush[k].flag_idx = slot_flag[(size_t)opposite_index(oi) * nfields + i] + (uint32_t)(k - first);
                        tpush[k].bulk = 0;
                    }
                    op.flag_base = (int32_t)slot_flag[(size_t)opposite_index(oi) * nfields + i];
                    push.push_back(op);
                    plan.active_mask |= 1u << oi;
                    plan.direct_mask |= 1u << oi;
                    plan.bytes_sent += op.nbytes;

                    igg_plan_op up;
                    std::memset(&up, 0, sizeof(up));
                    up.kind = IGG_OP_UNPACK;
                    up.field = i;
                    up.peer = (int32_t)peer;
                    for (int d = 0; d < 3; ++d) {
                        up.offset[d] = o[d];
                        up.dst_lo[d] = rcv_mine.lo[d]; up.dst_hi[d] = rcv_mine.hi[d];
                    }
                    up.staging_off = -1;
                    const size_t firstw = tunpack.size();
                    lower_op(f, i, kOpWait, oi, rcv_mine, rcv_mine, true, true, 0, up, tunpack, tb);
                    for (size_t k = firstw; k < tunpack.size(); ++k) {
                        tunpack[k].flag_idx = slot_flag[(size_t)oi * nfields + i] + (uint32_t)(k - firstw);
                        tunpack[k].bulk = 0;
                    }
                    up.flag_base = (int32_t)slot_flag[(size_t)oi * nfields + i];
                    up.nbytes = 0;
                    unpack.push_back(up);
                    continue;
                }
                // Page-locked HOST field, pure x offset to another rank: the strided cells of the x-face are not read
                // and written one by one over PCIe (a partial write costs the host a read-modify-write).  A kernel in
                // front of the exchange fetches the first and the last 64-byte line of every row of the box into a
                // device cache (kOpHostLinesX tiles), the exchange packs from / unpacks into that cache -- to the
                // tiles it is one more field: 128 bytes per row, the row's two lines side by side --, and a kernel
                // behind it writes the lines back whole.  No other op of the call touches these lines (the y/z faces
                // write y/z halo rows only; the edges and corners keep the per-cell path).
                const igg_field* kf = &f;
                igg_field fc;
                int kidx = i;
                Box ksnd = snd, krcv = rcv_mine;
                if (hostx[i] >= 0 && o[0] && !o[1] && !o[2]) {
                    HostXField& hx = plan.hostx[(size_t)hostx[i]];
                    const int64_t el = f.elsize, cpl = 128 / el;      // cells in the two cached lines of a row
                    if (hx.rows == 0) {                               // first x op of this field: lay its cache out
                        const int64_t pitch = f.size[0] * el;
                        hx.ny = snd.hi[1] - snd.lo[1] + 1;
                        hx.rows = hx.ny * (snd.hi[2] - snd.lo[2] + 1);
                        hx.cache_off = plan.hostx_cache_bytes;
                        plan.hostx_cache_bytes += hx.rows * 128;
                        hx.row_org = (snd.lo[1] - 1) * pitch + (snd.lo[2] - 1) * pitch * f.size[1];
                        hx.pitch = pitch;
                        hx.plane = pitch * f.size[1];
                    }
                    hx.sides |= o[0] < 0 ? 1u : 2u;
                    fc = f;
                    fc.ptr = nullptr;                                 // (the cache rows are 128-byte aligned)
                    fc.size[0] = cpl; fc.size[1] = hx.ny; fc.size[2] = hx.rows / hx.ny;
                    kf = &fc;
                    kidx = nfields + hostx[i];
                    const int64_t shift = o[0] < 0 ? 0 : cpl - f.size[0];     // the high end of the row lives in the second line
                    ksnd.lo[0] = snd.lo[0] + shift; ksnd.hi[0] = snd.hi[0] + shift;
                    krcv.lo[0] = rcv_mine.lo[0] + shift; krcv.hi[0] = rcv_mine.hi[0] + shift;
                    for (int d = 1; d < 3; ++d) {
                        ksnd.hi[d] = snd.hi[d] - snd.lo[d] + 1; ksnd.lo[d] = 1;
                        krcv.hi[d] = rcv_mine.hi[d] - rcv_mine.lo[d] + 1; krcv.lo[d] = 1;
                    }
                }
                if (ll128) {
                    // Same boxes, line-formatted staging: the op descriptions keep the field-side geometry (merged
                    // rows and strides); the tiles are runs of staging lines.
                    auto emit_ll = [&](int kind, int64_t foff, int64_t fsy, int64_t fsz, const igg_plan_op& d, std::vector<TileRec>& out) {
                        const int64_t nlines = (d.nbytes + kLinePayload - 1) / kLinePayload;
                        int wlog = 0;
                        while ((2 << wlog) <= std::min<int64_t>(8, d.vec_bytes)) ++wlog;
                        for (int64_t l0 = 0; l0 < nlines; l0 += kLinesPerTile) {
                            TileRec t;
                            std::memset(&t, 0, sizeof(t));
                            t.src_off = foff; t.src_sy = fsy; t.src_sz = fsz;
                            t.ny = (uint32_t)d.ny; t.ncols = (uint32_t)d.row_bytes;
                            t.row0 = (uint32_t)l0; t.nrows = (uint32_t)std::min<int64_t>(kLinesPerTile, nlines - l0);
                            t.dst_off = d.staging_off; t.dst_sy = d.nbytes;
                            t.field = (uint16_t)kidx; t.vec_log2 = (uint8_t)wlog; t.kind = (uint8_t)kind; t.nbr = (uint8_t)oi;
                            out.push_back(t);
                        }
                    };
                    std::vector<TileRec> unused;
                    lower_op(*kf, kidx, IGG_OP_PUSH, oi, ksnd, ksnd, true, false, op.staging_off, op, unused);
                    op.flag_base = -1;
                    emit_ll(kOpPushLL, op.src_off, op.src_sy, op.src_sz, op, tpush);
                    push.push_back(op);
                    plan.active_mask |= 1u << oi;
                    plan.bytes_sent += op.nbytes;

                    igg_plan_op up;
                    std::memset(&up, 0, sizeof(up));
                    up.kind = IGG_OP_UNPACK;
                    up.field = i;
                    up.peer = (int32_t)peer;
                    for (int d = 0; d < 3; ++d) {
                        up.offset[d] = o[d];
                        up.dst_lo[d] = rcv_mine.lo[d]; up.dst_hi[d] = rcv_mine.hi[d];
                    }
                    up.staging_off = slot_off[(size_t)oi * nfields + i];
                    up.flag_base = -1;
                    unused.clear();
                    lower_op(*kf, kidx, IGG_OP_UNPACK, oi, krcv, krcv, false, true, up.staging_off, up, unused);
                    emit_ll(kOpUnpackLL, up.dst_off, up.dst_sy, up.dst_sz, up, tunpack);
                    unpack.push_back(up);
                    continue;
                }
                const size_t first_push = tpush.size();
                lower_op(*kf, kidx, IGG_OP_PUSH, oi, ksnd, ksnd, true, false, op.staging_off, op, tpush, tb);
                if (tpush.size() - first_push > slot_nflag[(size_t)opposite_index(oi) * nfields + i])
                    return fail(IGG_ERR_STATE, "internal: an exchange op has more tiles than its slot has arrival flags.");
                for (size_t k = first_push; k < tpush.size(); ++k)      // flag numbers in the RECEIVER's slot table
                    tpush[k].flag_idx = slot_flag[(size_t)opposite_index(oi) * nfields + i] + (uint32_t)(k - first_push);
                op.flag_base = (int32_t)slot_flag[(size_t)opposite_index(oi) * nfields + i];
                push.push_back(op);
                plan.active_mask |= 1u << oi;
                plan.bytes_sent += op.nbytes;

                igg_plan_op up;
                std::memset(&up, 0, sizeof(up));
                up.kind = IGG_OP_UNPACK;
                up.field = i;
                up.peer = (int32_t)peer;
                for (int d = 0; d < 3; ++d) {
                    up.offset[d] = o[d];
                    up.dst_lo[d] = rcv_mine.lo[d]; up.dst_hi[d] = rcv_mine.hi[d];
                    up.src_lo[d] = up.src_hi[d] = 0;
                }
                up.staging_off = slot_off[(size_t)oi * nfields + i];
                const size_t first_unpack = tunpack.size();
                lower_op(*kf, kidx, IGG_OP_UNPACK, oi, krcv, krcv, false, true, up.staging_off, up, tunpack, tb);
                for (size_t k = first_unpack; k < tunpack.size(); ++k)
                    tunpack[k].flag_idx = slot_flag[(size_t)oi * nfields + i] + (uint32_t)(k - first_unpack);
                up.flag_base = (int32_t)slot_flag[(size_t)oi * nfields + i];
                unpack.push_back(up);
            }
        }
    }

    plan.n_push_tiles = (int32_t)tpush.size();
    plan.ops = push;
    plan.ops.insert(plan.ops.end(), local.begin(), local.end());
    plan.ops.insert(plan.ops.end(), unpack.begin(), unpack.end());
    // Order: PUSH, LOCAL, WRAPX, UNPACK.  Blocks are dispatched in index order, so the contiguous copies have their
    // tile records and their loads out before the strided x-face loads flood the memory system (measured: with WRAPX
    // first a few LOCAL blocks got their tile record 3-4.6 us late and ended the launch at 9.7 us instead of 7.5).
    plan.tiles = tpush;
    plan.tiles.insert(plan.tiles.end(), tlocal.begin(), tlocal.end());
    plan.wx_first = (int32_t)plan.tiles.size();
    plan.wx_count = (int32_t)twrapx.size();
    plan.tiles.insert(plan.tiles.end(), twrapx.begin(), twrapx.end());
    plan.tiles.insert(plan.tiles.end(), tunpack.begin(), tunpack.end());
    plan.hostx_tiles.clear();
    for (const HostXField& hx : plan.hostx)
        for (int64_t r0 = 0; r0 < hx.rows; r0 += kWrapRowsPerTile) {
            TileRec t;
            std::memset(&t, 0, sizeof(t));
            t.src_off = hx.row_org; t.src_sy = hx.pitch; t.src_sz = hx.plane;
            t.dst_off = hx.pitch; t.dst_sy = hx.cache_off; t.dst_sz = hx.sides;
            t.row0 = (uint32_t)r0; t.nrows = (uint32_t)std::min<int64_t>(kWrapRowsPerTile, hx.rows - r0);
            t.ncols = 8; t.ny = (uint32_t)hx.ny;
            t.field = (uint16_t)hx.field; t.vec_log2 = 4; t.kind = (uint8_t)kOpHostLinesX;
            plan.hostx_tiles.push_back(t);
        }
    plan.has_bulk = false;           // the LL128 kernel only takes the LOCAL ones through the pipeline
    for (const TileRec& t : plan.tiles) plan.has_bulk = plan.has_bulk || (t.bulk && (!ll128 || t.kind == IGG_OP_LOCAL));
    return IGG_OK;
}

// Tiles for copying a dense column-major block of `ext` elements into a box of a larger dense array whose first two
// extents are `dst_size` (used by gather!, ref: src/gather.jl:36-49, and for dense copies with dst_size == ext).
void build_block_copy_tiles(int64_t el, const int64_t ext[3], const int64_t dst_size[2], const int64_t dst_origin[3],
                            int64_t src_align, int64_t dst_align, std::vector<TileRec>& tiles) {
    const int64_t dsy = dst_size[0] * el, dsz = dst_size[0] * dst_size[1] * el;
    const int64_t dst_off = dst_origin[0] * el + dst_origin[1] * dsy + dst_origin[2] * dsz;
    emit_tiles(el, ext, 0, ext[0] * el, ext[0] * ext[1] * el, dst_off, dsy, dsz, src_align, dst_align, 0, 3 /*COPY*/, 0, tiles);
}

// One phase of the reference's scheme for dimension d (ref: src/update_halo.jl:51-88), used by the NCCL comparison
// path: PACK tiles copy the send box of side n (full extent in the other dimensions) of every field that has a halo
// in d into a contiguous per-side message (kind kOpCopy, base = push_base[n]); UNPACK tiles copy the per-side message
// into the receive box of side n (kind kOpCopyIn, base = unpack_base[n]).  bytes[n] = message size of side n.
int build_dim_phase(const igg_grid_desc& g, int nfields, const igg_field* fields, int d, std::vector<TileRec>& pack,
                    std::vector<TileRec>& unpack, int64_t bytes[2]) {
    bytes[0] = bytes[1] = 0;
    for (int n = 0; n < 2; ++n) {
        if (g.neighbors[d][n] < 0) continue;
        int64_t off = 0;
        for (int i = 0; i < nfields; ++i) {
            const igg_field& f = fields[i];
            if (field_ol(g, f, d) < 2 * f.halowidths[d]) continue;
            int64_t r[6], ext[3];
            const int64_t el = f.elsize, fsy = f.size[0] * el, fsz = f.size[0] * f.size[1] * el;
            const int64_t align = (int64_t)((uintptr_t)f.ptr % 16);
            ranges(g, f, n + 1, d + 1, true, r);
            for (int k = 0; k < 3; ++k) ext[k] = r[2 * k + 1] - r[2 * k] + 1;
            const int64_t foff_s = (r[0] - 1) * el + (r[2] - 1) * fsy + (r[4] - 1) * fsz;
            emit_tiles(el, ext, foff_s, fsy, fsz, off, ext[0] * el, ext[0] * ext[1] * el, align, 0, i, kOpCopy, n, pack);
            ranges(g, f, n + 1, d + 1, false, r);
            const int64_t foff_r = (r[0] - 1) * el + (r[2] - 1) * fsy + (r[4] - 1) * fsz;
            emit_tiles(el, ext, off, ext[0] * el, ext[0] * ext[1] * el, foff_r, fsy, fsz, 0, align, i, kOpCopyIn, n, unpack);
            off += round_up(ext[0] * ext[1] * ext[2] * el, 128);
        }
        bytes[n] = off;
    }
    return IGG_OK;
}

}  // namespace igg

// ---- C ABI: host-only helpers --------------------------------------------------------------------------------
namespace igg { const char* last_error_cstr(); }

extern "C" {

int igg_version(void) { return IGG_VERSION; }
const char* igg_last_error(void) { return igg::last_error_cstr(); }

int igg_sendranges(const igg_grid_desc* desc, const igg_field* f, int n, int dim, int64_t out[6]) {
    if (!desc || !f || !out) return igg::fail(IGG_ERR_ARG, "igg_sendranges: NULL argument.");
    return igg::ranges(*desc, *f, n, dim, true, out);
}

int igg_recvranges(const igg_grid_desc* desc, const igg_field* f, int n, int dim, int64_t out[6]) {
    if (!desc || !f || !out) return igg::fail(IGG_ERR_ARG, "igg_recvranges: NULL argument.");
    return igg::ranges(*desc, *f, n, dim, false, out);
}

int igg_check_fields(const igg_grid_desc* desc, int nfields, const igg_field* fields) {
    if (!desc || !fields) return igg::fail(IGG_ERR_ARG, "igg_check_fields: NULL argument.");
    return igg::check_fields(*desc, nfields, fields);
}

}  // extern "C"

int igg::dims_mask_from_order(const int* dims_order, int n, int* mask_out) {
    int mask = 0;
    if (!dims_order || n <= 0) { *mask_out = 7; return IGG_OK; }   // default dims=(3,1,2): all dimensions
    for (int i = 0; i < n; ++i) {
        if (dims_order[i] < 1 || dims_order[i] > 3) return igg::fail(IGG_ERR_ARG, "dims entries must be in 1..3.");
        mask |= 1 << (dims_order[i] - 1);
    }
    *mask_out = mask;
    return IGG_OK;
}

extern "C" {

int igg_plan_describe(const igg_grid_desc* desc, int nfields, const igg_field* fields, const int* dims_order,
                      int ndims_order, igg_plan_op* ops_out, int ops_cap, int* nops_out, int64_t* staging_bytes_out) {
    if (!desc || !fields || !nops_out) return igg::fail(IGG_ERR_ARG, "igg_plan_describe: NULL argument.");
    int mask;
    int rc = igg::dims_mask_from_order(dims_order, ndims_order, &mask);
    if (rc) return rc;
    igg::Plan plan;
    rc = igg::build_plan(*desc, nfields, fields, mask, plan);
    if (rc) return rc;
    *nops_out = (int)plan.ops.size();
    if (staging_bytes_out) *staging_bytes_out = plan.staging_bytes;
    if (ops_out) {
        if (ops_cap < (int)plan.ops.size()) return igg::fail(IGG_ERR_ARG, "igg_plan_describe: ops_cap too small (need %d).", (int)plan.ops.size());
        std::memcpy(ops_out, plan.ops.data(), plan.ops.size() * sizeof(igg_plan_op));
    }
    return IGG_OK;
}

int igg_plan_tiles(const igg_grid_desc* desc, int nfields, const igg_field* fields, const int* dims_order, int ndims_order,
                   int mode, void* tiles_out, int tiles_cap, int* ntiles_out, int64_t* staging_bytes_out,
                   int64_t* regions_out) {
    if (!desc || !fields || !ntiles_out) return igg::fail(IGG_ERR_ARG, "igg_plan_tiles: NULL argument.");
    const bool all_direct = (mode & 0x100) != 0;      // + 0x100: every field lives in a symmetric allocation
    const bool all_host = (mode & 0x200) != 0;        // + 0x200: every field is page-locked host memory on a 64-byte boundary
    mode &= 0xff;
    if (mode != IGG_MODE_IPC_1PHASE && mode != IGG_MODE_IPC_LL128 && mode != IGG_MODE_IPC_AUTO)
        return igg::fail(IGG_ERR_ARG, "igg_plan_tiles: mode must be an IPC mode.");
    uint8_t direct[IGG_MAX_FIELDS];
    std::memset(direct, all_direct ? 1 : 0, sizeof(direct));
    uint8_t host[IGG_MAX_FIELDS];                       // + 0x200 (before masking): every field is page-locked host memory
    std::memset(host, all_host ? 1 : 0, sizeof(host));
    int mask;
    int rc = igg::dims_mask_from_order(dims_order, ndims_order, &mask);
    if (rc) return rc;
    igg::Plan plan;
    rc = igg::build_plan(*desc, nfields, fields, mask, plan, all_host ? host : nullptr,
                         mode == IGG_MODE_IPC_LL128 ? igg::kProtoLL128 : mode == IGG_MODE_IPC_AUTO ? igg::kProtoAuto : igg::kProtoFence,
                         direct);
    if (rc) return rc;
    // host fields: the whole-line tiles (x self-wraps: kind 5; fetch / write-back of exchanged x-faces: kind 11) follow
    // the exchange kernel's table
    plan.tiles.insert(plan.tiles.end(), plan.wrap_tiles.begin(), plan.wrap_tiles.end());
    plan.tiles.insert(plan.tiles.end(), plan.hostx_tiles.begin(), plan.hostx_tiles.end());
    *ntiles_out = (int)plan.tiles.size();
    if (staging_bytes_out) *staging_bytes_out = plan.staging_bytes;
    if (regions_out)
        for (int o = 0; o < IGG_NOFFSETS; ++o) {
            regions_out[o] = plan.region_pack[o];
            regions_out[IGG_NOFFSETS + o] = plan.region_need[o];
            regions_out[2 * IGG_NOFFSETS + o] = plan.flag_pack[o];
            regions_out[3 * IGG_NOFFSETS + o] = plan.flag_need[o];
        }
    if (tiles_out) {
        if (tiles_cap < (int)plan.tiles.size()) return igg::fail(IGG_ERR_ARG, "igg_plan_tiles: tiles_cap too small (need %d).", (int)plan.tiles.size());
        std::memcpy(tiles_out, plan.tiles.data(), plan.tiles.size() * sizeof(igg::TileRec));
    }
    return IGG_OK;
}

}  // extern "C"
