/* cpu_ref.c -- C restatement of the reference's CPU path of update_halo!  --  TEST INFRASTRUCTURE ONLY.
 *
 * Used (a) as the timed CPU baseline of bench.py (`cpu_baseline`, `--impl reference`) and (b) as a second checker
 * beside oracle/igg_oracle.py.  Nothing under implicitglobalgrid.jl_b200/ links or loads it.
 * Parity pinning: the reference (Julia + MPI.jl) cannot run here; this file is validated against the numpy oracle,
 * which is pinned by the reference's closed-form fixtures (tests/test_oracle_reference_fixtures.py).
 *
 * What is restated (paths relative to the reference root):
 *   _update_halo!          src/update_halo.jl:44-89   three dependent phases, default order z, x, y
 *   iwrite_sendbufs!       src/update_halo.jl:236-243 -> write_h2h! :308-321   (pack, shape-specialised)
 *   irecv_halo!/isend_halo src/update_halo.jl:343-367 (MPI.Irecv!/Isend)       here: one memcpy into the neighbour
 *                                                     rank's receive buffer (what a shared-memory MPI does at best)
 *   sendrecv_halo_local    src/update_halo.jl:369-386 (rank is its own neighbour: recvbuf(3-n) <- sendbuf(n))
 *   iread_recvbufs!        src/update_halo.jl:263-270 -> read_h2h! :324-337    (unpack)
 *   memcopy!/memcopy_threads!  src/update_halo.jl:388-411: flat (x-/z-face, buffer) copies are split over the
 *                          threads when >= 32768 bytes (GG_THREADCOPY_THRESHOLD, src/shared.jl:32); y-faces and
 *                          hw>1 boxes are a single-threaded copyto! that runs along x.
 *   has-halo predicate     src/shared.jl:220, src/update_halo.jl:239
 * Ranks are OS threads (one per simulated MPI rank) that meet at barriers where MPI would wait.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GG_THREADCOPY_THRESHOLD 32768

typedef struct {
    int nranks, nfields, elsize, norder, copy_threads, iters;
    int64_t dims[3], periods[3], nxyz[3], overlaps[3];
    const int64_t* sizes; /* [field][3] */
    const int64_t* hws;   /* [field][3] */
    int order[3];
    char** fields;        /* [rank*nfields + field] */
    char** sendbuf;       /* [(rank*nfields + field)*2 + side] */
    char** recvbuf;
    pthread_barrier_t bar;
} job_t;

typedef struct { job_t* job; int rank; } arg_t;

static int64_t cart_rank(const int64_t c[3], const int64_t d[3]) { return (c[0] * d[1] + c[1]) * d[2] + c[2]; }

static int64_t neighbor(const job_t* j, int rank, int dim, int side) {
    int64_t c[3];
    int64_t r = rank;
    c[2] = r % j->dims[2]; r /= j->dims[2];
    c[1] = r % j->dims[1]; r /= j->dims[1];
    c[0] = r;
    c[dim] += side ? 1 : -1;
    if (j->periods[dim]) c[dim] = (c[dim] + j->dims[dim]) % j->dims[dim];
    else if (c[dim] < 0 || c[dim] >= j->dims[dim]) return -1;
    return cart_rank(c, j->dims);
}

/* memcopy_threads!: flat copy, threaded above the threshold */
static void flat_copy(char* dst, const char* src, int64_t bytes, int threads) {
#ifdef _OPENMP
    if (threads > 1 && bytes >= GG_THREADCOPY_THRESHOLD) {
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int t = 0; t < threads; ++t) {
            int64_t lo = bytes * t / threads, hi = bytes * (t + 1) / threads;
            memcpy(dst + lo, src + lo, (size_t)(hi - lo));
        }
        return;
    }
#endif
    (void)threads;
    memcpy(dst, src, (size_t)bytes);
}

/* strided element copy (the width-1 x-face view(A,ix,:,:) is a uniformly strided linear view) */
static void strided_copy(char* dst, int64_t dstride, const char* src, int64_t sstride, int64_t n, int el, int threads) {
#ifdef _OPENMP
    if (threads > 1 && n * el >= GG_THREADCOPY_THRESHOLD) {
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int64_t i = 0; i < n; ++i) memcpy(dst + i * dstride, src + i * sstride, (size_t)el);
        return;
    }
#endif
    (void)threads;
    if (el == 8) { for (int64_t i = 0; i < n; ++i) *(uint64_t*)(dst + i * dstride) = *(const uint64_t*)(src + i * sstride); }
    else if (el == 4) { for (int64_t i = 0; i < n; ++i) *(uint32_t*)(dst + i * dstride) = *(const uint32_t*)(src + i * sstride); }
    else for (int64_t i = 0; i < n; ++i) memcpy(dst + i * dstride, src + i * sstride, (size_t)el);
}

/* pack (to_buf=1) or unpack (to_buf=0) the box [lo,hi] (1-based, width hw in `dim`, full extent elsewhere) */
static void box_copy(char* A, const int64_t s[3], int el, int dim, int64_t lo, int64_t hw, char* buf, int to_buf, int threads) {
    const int64_t sx = s[0], sy = s[1], sz = s[2];
    if (dim == 2) {                       /* z-face(s): contiguous block of hw planes */
        char* a = A + (lo - 1) * sx * sy * el;
        int64_t bytes = hw * sx * sy * el;
        if (hw == 1) { if (to_buf) flat_copy(buf, a, bytes, threads); else flat_copy(a, buf, bytes, threads); }
        else { if (to_buf) memcpy(buf, a, (size_t)bytes); else memcpy(a, buf, (size_t)bytes); }
        return;
    }
    if (dim == 1) {                       /* y-face(s): rows of sx elements, copyto! along x, single thread */
        for (int64_t k = 0; k < sz; ++k)
            for (int64_t w = 0; w < hw; ++w) {
                char* a = A + ((lo - 1 + w) * sx + k * sx * sy) * el;
                char* b = buf + (w * sx + k * sx * hw) * el;
                if (to_buf) memcpy(b, a, (size_t)(sx * el)); else memcpy(a, b, (size_t)(sx * el));
            }
        return;
    }
    /* x-face(s) */
    if (hw == 1) {
        char* a = A + (lo - 1) * el;
        if (to_buf) strided_copy(buf, el, a, sx * el, sy * sz, el, threads);
        else strided_copy(a, sx * el, buf, el, sy * sz, el, threads);
        return;
    }
    for (int64_t r = 0; r < sy * sz; ++r) {   /* general case: hw contiguous elements per (y,z) row */
        char* a = A + (lo - 1 + r * sx) * el;
        char* b = buf + r * hw * el;
        if (to_buf) memcpy(b, a, (size_t)(hw * el)); else memcpy(a, b, (size_t)(hw * el));
    }
}

static void* rank_main(void* p) {
    arg_t* a = (arg_t*)p;
    job_t* j = a->job;
    const int me = a->rank, F = j->nfields, el = j->elsize;
    for (int it = 0; it < j->iters; ++it) {
        for (int q = 0; q < j->norder; ++q) {
            const int dim = j->order[q];
            const int64_t nb[2] = {neighbor(j, me, dim, 0), neighbor(j, me, dim, 1)};
            /* pack both sides, all fields (src/update_halo.jl:52-54) */
            for (int n = 0; n < 2; ++n) {
                if (nb[n] < 0) continue;
                for (int i = 0; i < F; ++i) {
                    const int64_t* s = j->sizes + 3 * i;
                    const int64_t hw = j->hws[3 * i + dim];
                    const int64_t ol = j->overlaps[dim] + (s[dim] - j->nxyz[dim]);
                    if (ol < 2 * hw) continue;
                    const int64_t lo = n ? s[dim] - (ol - 1) : 1 + (ol - hw);          /* sendranges :281-290 */
                    box_copy(j->fields[me * F + i], s, el, dim, lo, hw, j->sendbuf[(me * F + i) * 2 + n], 1, j->copy_threads);
                }
            }
            /* exchange: my send buffer of side n -> receive buffer of side 1-n on neighbor(n) (:57-74) */
            for (int n = 0; n < 2; ++n) {
                if (nb[n] < 0) continue;
                for (int i = 0; i < F; ++i) {
                    const int64_t* s = j->sizes + 3 * i;
                    const int64_t hw = j->hws[3 * i + dim];
                    const int64_t ol = j->overlaps[dim] + (s[dim] - j->nxyz[dim]);
                    if (ol < 2 * hw) continue;
                    int64_t bytes = hw * el;
                    for (int d = 0; d < 3; ++d) if (d != dim) bytes *= s[d];
                    flat_copy(j->recvbuf[((int)nb[n] * F + i) * 2 + (1 - n)], j->sendbuf[(me * F + i) * 2 + n], bytes,
                              nb[n] == me ? j->copy_threads : 1);
                }
            }
            pthread_barrier_wait(&j->bar);            /* MPI.Wait! on the receives (:79) */
            /* unpack (:78-81) */
            for (int n = 1; n >= 0; --n) {
                if (nb[n] < 0) continue;
                for (int i = 0; i < F; ++i) {
                    const int64_t* s = j->sizes + 3 * i;
                    const int64_t hw = j->hws[3 * i + dim];
                    const int64_t ol = j->overlaps[dim] + (s[dim] - j->nxyz[dim]);
                    if (ol < 2 * hw) continue;
                    const int64_t lo = n ? s[dim] - (hw - 1) : 1;                      /* recvranges :293-302 */
                    box_copy(j->fields[me * F + i], s, el, dim, lo, hw, j->recvbuf[(me * F + i) * 2 + n], 0, j->copy_threads);
                }
            }
            pthread_barrier_wait(&j->bar);            /* MPI.Waitall! on the sends (:85-87): buffers reusable */
        }
    }
    return NULL;
}

/* Runs `iters` update_halo! calls on `nranks` simulated ranks (row-major Cartesian ranks over dims), in place.
 * fields[rank*nfields + f] points to a dense column-major array of sizes[f].  dims_order holds 1..3.
 * Returns 0 and the elapsed wall time of the whole loop in *seconds_out. */
int igg_cpu_update_halo(int nranks, const int64_t dims[3], const int64_t periods[3], const int64_t nxyz[3],
                        const int64_t overlaps[3], int nfields, void** fields, const int64_t* sizes, const int64_t* hws,
                        int elsize, const int* dims_order, int norder, int copy_threads, int iters, double* seconds_out) {
    job_t j;
    memset(&j, 0, sizeof(j));
    if (nranks != dims[0] * dims[1] * dims[2] || norder < 1 || norder > 3) return -1;
    j.nranks = nranks; j.nfields = nfields; j.elsize = elsize; j.norder = norder; j.copy_threads = copy_threads; j.iters = iters;
    for (int d = 0; d < 3; ++d) { j.dims[d] = dims[d]; j.periods[d] = periods[d]; j.nxyz[d] = nxyz[d]; j.overlaps[d] = overlaps[d]; }
    for (int q = 0; q < norder; ++q) { if (dims_order[q] < 1 || dims_order[q] > 3) return -1; j.order[q] = dims_order[q] - 1; }
    j.sizes = sizes; j.hws = hws; j.fields = (char**)fields;
    const int nb = nranks * nfields * 2;
    j.sendbuf = (char**)calloc((size_t)nb, sizeof(char*));
    j.recvbuf = (char**)calloc((size_t)nb, sizeof(char*));
    for (int r = 0; r < nranks; ++r)
        for (int i = 0; i < nfields; ++i) {
            const int64_t* s = sizes + 3 * i;
            const int64_t* h = hws + 3 * i;
            int64_t m = s[0] * s[1] * h[2];                         /* max face elements (src/update_halo.jl:143) */
            if (s[0] * s[2] * h[1] > m) m = s[0] * s[2] * h[1];
            if (s[1] * s[2] * h[0] > m) m = s[1] * s[2] * h[0];
            m = (m + 31) / 32 * 32;                                 /* GG_ALLOC_GRANULARITY */
            for (int n = 0; n < 2; ++n) {
                j.sendbuf[(r * nfields + i) * 2 + n] = (char*)calloc((size_t)m, (size_t)elsize);
                j.recvbuf[(r * nfields + i) * 2 + n] = (char*)calloc((size_t)m, (size_t)elsize);
            }
        }
    pthread_barrier_init(&j.bar, NULL, (unsigned)nranks);
    pthread_t* th = (pthread_t*)calloc((size_t)nranks, sizeof(pthread_t));
    arg_t* args = (arg_t*)calloc((size_t)nranks, sizeof(arg_t));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < nranks; ++r) { args[r].job = &j; args[r].rank = r; pthread_create(&th[r], NULL, rank_main, &args[r]); }
    for (int r = 0; r < nranks; ++r) pthread_join(th[r], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (seconds_out) *seconds_out = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    pthread_barrier_destroy(&j.bar);
    for (int k = 0; k < nb; ++k) { free(j.sendbuf[k]); free(j.recvbuf[k]); }
    free(j.sendbuf); free(j.recvbuf); free(th); free(args);
    return 0;
}

int igg_cpu_ref_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
