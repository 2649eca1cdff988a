// xface_floor.cu -- what does the periodic x self-wrap of a 256^3 Float64 field cost on a B200, at best?
//
// cfg2's x-faces are 2 x 65 536 cells of 8 bytes, 2 KiB apart: cell 1 -> cell n-1 and cell n-2 -> cell 0 of every
// x-row.  This program times that data movement alone (no tile table, no plan, no exchange state) in several
// shapes, cold (ring of 24 arrays of 128 MiB, larger than the 126 MB L2) and hot (one array), each as one CUDA
// graph of K launches, so that the numbers can be set against the library's kernel:
//
//   two_pass   the library's round-1 structure: one pass copies column 1 -> column n-1, a second pass column
//              n-2 -> column 0 (two strided reads and two strided writes per row, in different passes)
//   row_owner  one thread owns a row: loads cell 1 and cell n-2 together, then stores both (R rows per thread)
//   row_owner16 the same with 16-byte accesses: reads {cell0,cell1} and {cell n-2,cell n-1}, writes both pairs back
//              (each pair shares a 32-byte sector anyway)
//   read_only / write_only   the two halves of row_owner on their own
//   empty      a kernel that does nothing (the launch floor inside a graph)
//
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o profiles/_bin/xface_floor profiles/xface_floor.cu
// Run:    profiles/_bin/xface_floor [K]          (prints one line per variant: us per launch cold / hot)
//         profiles/_bin/xface_floor K gran32     (the same after cudaLimitMaxL2FetchGranularity = 32)
//         ncu --cache-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
//             profiles/_bin/xface_floor 1 plain  (DRAM bytes per launch in the steady state of the cold ring)
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int N = 256;
constexpr long long ROWS = (long long)N * N;

__global__ void k_empty(double* a) {}

__global__ void k_column(double* __restrict__ a, int src, int dst) {   // one column copy, thread per row
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < ROWS) a[r * N + dst] = a[r * N + src];
}

template <int R>
__global__ void k_row_owner(double* __restrict__ a) {
    const long long r0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x);
    const long long stride = (long long)gridDim.x * blockDim.x;
    double lo[R], hi[R];
#pragma unroll
    for (int u = 0; u < R; ++u) {
        const long long r = r0 + u * stride;
        if (r < ROWS) { lo[u] = a[r * N + 1]; hi[u] = a[r * N + N - 2]; }
    }
#pragma unroll
    for (int u = 0; u < R; ++u) {
        const long long r = r0 + u * stride;
        if (r < ROWS) { a[r * N + N - 1] = lo[u]; a[r * N] = hi[u]; }
    }
}

template <int R>
__global__ void k_row_owner16(double* __restrict__ a) {
    const long long r0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x);
    const long long stride = (long long)gridDim.x * blockDim.x;
    double2 lo[R], hi[R];
#pragma unroll
    for (int u = 0; u < R; ++u) {
        const long long r = r0 + u * stride;
        if (r < ROWS) {
            lo[u] = *reinterpret_cast<const double2*>(a + r * N);
            hi[u] = *reinterpret_cast<const double2*>(a + r * N + N - 2);
        }
    }
#pragma unroll
    for (int u = 0; u < R; ++u) {
        const long long r = r0 + u * stride;
        if (r < ROWS) {
            const double2 v = make_double2(hi[u].x, lo[u].y);   // {cell n-2, cell 1}: the new content of BOTH pairs
            *reinterpret_cast<double2*>(a + r * N) = v;
            *reinterpret_cast<double2*>(a + r * N + N - 2) = v;
        }
    }
}

// Loads with an explicit L2 prefetch-size hint (PTX .L2::64B / .L2::128B): does the fetch granularity follow it?
template <int HINT>
__device__ __forceinline__ double ld_hint(const double* p) {
    double v;
    if (HINT == 64) asm volatile("ld.global.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 128) asm volatile("ld.global.L2::128B.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 1) asm volatile("ld.global.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 2) asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 3) asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 4) asm volatile("ld.global.cv.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 5) asm volatile("ld.global.nc.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 6) asm volatile("ld.global.L1::no_allocate.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (HINT == 7) asm volatile("ld.global.L1::evict_first.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else v = *p;
    return v;
}
template <int HINT>
__global__ void k_row_owner_hint(double* __restrict__ a) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < ROWS) {
        const double lo = ld_hint<HINT>(a + r * N + 1), hi = ld_hint<HINT>(a + r * N + N - 2);
        a[r * N + N - 1] = lo; a[r * N] = hi;
    }
}
// Stores that should not linger in L2 (.cs = evict-first) or go through (.wt).
template <int KIND>
__global__ void k_row_owner_st(double* __restrict__ a) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < ROWS) {
        const double lo = a[r * N + 1], hi = a[r * N + N - 2];
        if (KIND == 0) { __stcs(a + r * N + N - 1, lo); __stcs(a + r * N, hi); }
        else { __stwt(a + r * N + N - 1, lo); __stwt(a + r * N, hi); }
    }
}

__global__ void k_read_only(const double* __restrict__ a, double* sink) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < ROWS) {
        const double v = a[r * N + 1] + a[r * N + N - 2];
        if (v == 1.2345e300) *sink = v;
    }
}

__global__ void k_write_only(double* __restrict__ a) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < ROWS) { a[r * N + N - 1] = 1.0; a[r * N] = 2.0; }
}

struct Variant { const char* name; int id; int threads; int rows_per_thread; };

static void launch(const Variant& v, double* a, double* sink, cudaStream_t s) {
    const int t = v.threads;
    const long long thr = (ROWS + v.rows_per_thread - 1) / v.rows_per_thread;
    const int g = (int)((thr + t - 1) / t);
    switch (v.id) {
        case 0: k_empty<<<404, 256, 0, s>>>(a); break;
        case 1: k_column<<<g, t, 0, s>>>(a, 1, N - 1); k_column<<<g, t, 0, s>>>(a, N - 2, 0); break;
        case 2: if (v.rows_per_thread == 1) k_row_owner<1><<<g, t, 0, s>>>(a);
                else if (v.rows_per_thread == 2) k_row_owner<2><<<g, t, 0, s>>>(a);
                else k_row_owner<4><<<g, t, 0, s>>>(a);
                break;
        case 3: if (v.rows_per_thread == 1) k_row_owner16<1><<<g, t, 0, s>>>(a);
                else k_row_owner16<2><<<g, t, 0, s>>>(a);
                break;
        case 6: k_row_owner_hint<64><<<g, t, 0, s>>>(a); break;
        case 7: k_row_owner_hint<128><<<g, t, 0, s>>>(a); break;
        case 8: k_row_owner_hint<1><<<g, t, 0, s>>>(a); break;
        case 11: k_row_owner_hint<2><<<g, t, 0, s>>>(a); break;
        case 12: k_row_owner_hint<3><<<g, t, 0, s>>>(a); break;
        case 13: k_row_owner_hint<4><<<g, t, 0, s>>>(a); break;
        case 14: k_row_owner_hint<5><<<g, t, 0, s>>>(a); break;
        case 15: k_row_owner_hint<6><<<g, t, 0, s>>>(a); break;
        case 16: k_row_owner_hint<7><<<g, t, 0, s>>>(a); break;
        case 9: k_row_owner_st<0><<<g, t, 0, s>>>(a); break;
        case 10: k_row_owner_st<1><<<g, t, 0, s>>>(a); break;
        case 4: k_read_only<<<g, t, 0, s>>>(a, sink); break;
        case 5: k_write_only<<<g, t, 0, s>>>(a); break;
    }
}

static float time_graph(const Variant& v, std::vector<double*>& ring, int nring, double* sink, int K, cudaStream_t s) {
    cudaGraph_t graph;
    cudaGraphExec_t exec;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < K; ++i) launch(v, ring[i % nring], sink, s);
    CK(cudaStreamEndCapture(s, &graph));
    CK(cudaGraphInstantiate(&exec, graph, 0));
    CK(cudaGraphLaunch(exec, s));
    CK(cudaStreamSynchronize(s));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0, s));
        CK(cudaGraphLaunch(exec, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGraphExecDestroy(exec)); CK(cudaGraphDestroy(graph));
    CK(cudaEventDestroy(e0)); CK(cudaEventDestroy(e1));
    return best * 1e3f / K;
}

int main(int argc, char** argv) {
    const int K = argc > 1 ? atoi(argv[1]) : 480;
    const int NR = 24;
    std::vector<double*> ring(NR);
    const size_t bytes = (size_t)N * N * N * sizeof(double);
    for (int i = 0; i < NR; ++i) { CK(cudaMalloc(&ring[i], bytes)); CK(cudaMemset(ring[i], i + 1, bytes)); }
    double* sink;
    CK(cudaMalloc(&sink, 8));
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    const Variant vs[] = {
        {"empty_404x256", 0, 256, 1},
        {"two_pass_256", 1, 256, 1},
        {"row_owner_r1_t256", 2, 256, 1}, {"row_owner_r2_t256", 2, 256, 2}, {"row_owner_r4_t256", 2, 256, 4},
        {"row_owner_r1_t128", 2, 128, 1}, {"row_owner_r1_t64", 2, 64, 1}, {"row_owner_r2_t128", 2, 128, 2},
        {"row_owner16_r1_t256", 3, 256, 1}, {"row_owner16_r2_t256", 3, 256, 2}, {"row_owner16_r1_t128", 3, 128, 1},
        {"read_only_t256", 4, 256, 1}, {"write_only_t256", 5, 256, 1},
        {"row_owner_ldL2_64B", 6, 256, 1}, {"row_owner_ldL2_128B", 7, 256, 1}, {"row_owner_ld_noL1", 8, 256, 1},
        {"row_owner_st_cs", 9, 256, 1}, {"row_owner_st_wt", 10, 256, 1},
        {"row_owner_ld_volatile", 11, 256, 1}, {"row_owner_ld_relaxed_gpu", 12, 256, 1}, {"row_owner_ld_cv", 13, 256, 1},
        {"row_owner_ld_nc_L2_64B", 14, 256, 1}, {"row_owner_ld_noL1_L2_64B", 15, 256, 1}, {"row_owner_ld_L1ef_L2_64B", 16, 256, 1},
    };
    // `plain` mode (for ncu --cache-control none): two rounds over the ring, launch by launch, a few variants.
    if (argc > 2 && !strcmp(argv[2], "plain")) {
        const int ids[] = {2, 4, 5, 6, 9, 11, 12, 13, 15};
        for (int id : ids)
            for (const Variant& v : vs)
                if (v.id == id && v.threads == 256 && v.rows_per_thread == 1) {
                    for (int i = 0; i < 2 * NR; ++i) launch(v, ring[i % NR], sink, s);
                    CK(cudaStreamSynchronize(s));
                }
        printf("plain done\n");
        return 0;
    }
    if (argc > 2 && !strcmp(argv[2], "gran32")) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32));
    printf("variant,us_cold,us_hot\n");
    for (const Variant& v : vs) {
        const float cold = time_graph(v, ring, NR, sink, K, s);
        const float hot = time_graph(v, ring, 1, sink, K, s);
        printf("%s,%.3f,%.3f\n", v.name, cold, hot);
        fflush(stdout);
    }
    return 0;
}
