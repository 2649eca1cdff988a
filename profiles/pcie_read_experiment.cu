// What bounds a kernel that gathers one strided x-slab (65 536 cells of 8 B at 2 KiB pitch) out of page-locked HOST
// memory over PCIe?  (design input for bench.py's e2e number, where the fields live in pinned host memory)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/pcie_read_experiment profiles/pcie_read_experiment.cu
// Variants: load flavour (default / volatile / .nc / relaxed.sys), threads in flight, bytes fetched per cell (8, 16,
// 32: does a wider fetch of the same line cost more?), strided 8-byte stores to host, and a contiguous read for scale.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

constexpr int kN = 256;                       // 256^3 doubles
constexpr long long kCells = (long long)kN * kN;
constexpr long long kPitch = kN * 8;

template <int MODE>
__device__ __forceinline__ unsigned long long load8(const unsigned long long* p) {
    unsigned long long v;
    if (MODE == 0) v = *p;
    else if (MODE == 1) v = *reinterpret_cast<const volatile unsigned long long*>(p);
    else if (MODE == 2) v = __ldg(p);
    else asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// One cell per loop iteration and thread; `total` threads walk the 65 536 rows.
template <int MODE>
__global__ void gather8(const char* __restrict__ host, unsigned long long* __restrict__ out, long long x_off) {
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < kCells; r += total)
        out[r] = load8<MODE>(reinterpret_cast<const unsigned long long*>(host + r * kPitch + x_off));
}

// Fetch 16 or 32 bytes of the cell's line instead of 8 (keeps one cell).
template <int BYTES>
__global__ void gather_wide(const char* __restrict__ host, unsigned long long* __restrict__ out) {
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < kCells; r += total) {
        const uint4* p = reinterpret_cast<const uint4*>(host + r * kPitch);
        uint4 a = p[0];
        unsigned long long v = ((unsigned long long)a.w << 32) | a.z;      // cell x = 1
        if (BYTES == 32) { uint4 b = p[1]; v ^= b.x & 0u; }
        out[r] = v;
    }
}

__global__ void scatter8(char* __restrict__ host, const unsigned long long* __restrict__ in, long long x_off) {
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < kCells; r += total)
        *reinterpret_cast<unsigned long long*>(host + r * kPitch + x_off) = in[r];
}

// Write N16 x 16 bytes per row (16 = partial sector, 32 = one sector, 64 = a full host cache line).
template <int N16>
__global__ void scatter_wide(char* __restrict__ host, const unsigned long long* __restrict__ in) {
    // N16 neighbouring lanes write the N16 x 16 bytes of one row in one store instruction (one request per row)
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < kCells * N16; i += total) {
        const long long r = i / N16;
        const int k = (int)(i % N16);
        const unsigned long long v = in[r];
        uint4* p = reinterpret_cast<uint4*>(host + r * kPitch + 1024);      // mid-row: leaves the x-slabs alone
        p[k] = make_uint4((unsigned)v, (unsigned)(v >> 32), k, 0);
    }
}

// What a periodic single-rank update does with the x-slabs: host cell x=1 -> host cell x=255 and x=254 -> x=0.
__global__ void wrap_x(char* __restrict__ host) {
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < 2 * kCells; r += total) {
        const long long row = r >> 1;
        const long long so = (r & 1) ? 254 * 8 : 8, dofs = (r & 1) ? 0 : 255 * 8;
        *reinterpret_cast<unsigned long long*>(host + row * kPitch + dofs) =
            *reinterpret_cast<const unsigned long long*>(host + row * kPitch + so);
    }
}

// The same through full 64-byte lines: read the first and the last line of a row, patch the halo cells, write both
// lines back whole (no partial write reaches the host).
__global__ void wrap_x_lines(char* __restrict__ host) {
    // 8 neighbouring lanes per row: lanes 0-3 hold the first line (x = 0..7), lanes 4-7 the last one (x = 248..255);
    // a warp instruction touches 4 rows x 2 lines, each line one 64-byte request.  65 536 rows * 8 lanes is a
    // multiple of the grid stride for the grids used, so whole warps stay in the loop together.
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < kCells * 8; i += total) {
        const long long row = i >> 3;
        const int j = (int)(i & 7);
        uint4* p = reinterpret_cast<uint4*>(host + row * kPitch + (j < 4 ? 0 : kPitch - 128)) + j;
        uint4 v = *p;
        const int base = (threadIdx.x & 31) & ~7;
        const unsigned s1x = __shfl_sync(0xffffffffu, v.z, base), s1y = __shfl_sync(0xffffffffu, v.w, base);          // x = 1
        const unsigned s2x = __shfl_sync(0xffffffffu, v.x, base + 7), s2y = __shfl_sync(0xffffffffu, v.y, base + 7);  // x = 254
        if (j == 0) { v.x = s2x; v.y = s2y; }           // x = 0   <- x = 254
        if (j == 7) { v.z = s1x; v.w = s1y; }           // x = 255 <- x = 1
        *p = v;
    }
}

__global__ void contiguous(const uint4* __restrict__ host, uint4* __restrict__ out, long long n16) {
    const long long total = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += total) out[i] = host[i];
}

template <typename F>
static float timed(F launch, int reps = 10) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; ++i) launch();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms * 1000.f / reps;
}

int main() {
    char* h;
    const size_t bytes = (size_t)kN * kN * kN * 8;
    CK(cudaHostAlloc(&h, bytes, cudaHostAllocMapped));
    for (size_t i = 0; i < bytes / 8; ++i) reinterpret_cast<unsigned long long*>(h)[i] = i * 0x9E3779B97F4A7C15ull;
    char* hd;
    CK(cudaHostGetDevicePointer(&hd, h, 0));
    unsigned long long* out;
    CK(cudaMalloc(&out, bytes / kN));
    CK(cudaMemset(out, 0, bytes / kN));

    const int grids[] = {16, 64, 256, 592};
    for (int g : grids) {
        printf("gather 8 B x 65536 rows, %4d blocks x 256: default %.1f us, volatile %.1f us, ldg %.1f us, relaxed.sys %.1f us\n", g,
               timed([&] { gather8<0><<<g, 256>>>(hd, out, 8); }), timed([&] { gather8<1><<<g, 256>>>(hd, out, 8); }),
               timed([&] { gather8<2><<<g, 256>>>(hd, out, 8); }), timed([&] { gather8<3><<<g, 256>>>(hd, out, 8); }));
    }
    printf("gather 16 B per row (256 blocks): %.1f us;  32 B per row: %.1f us\n",
           timed([&] { gather_wide<16><<<256, 256>>>(hd, out); }), timed([&] { gather_wide<32><<<256, 256>>>(hd, out); }));
    printf("two x-slabs in one launch (x=1 and x=254 by two kernels on one stream): %.1f us\n",
           timed([&] { gather8<0><<<256, 256>>>(hd, out, 8); gather8<0><<<256, 256>>>(hd, out, 254 * 8); }));
    for (int g : {64, 256})
        printf("scatter 8 B x 65536 rows to host, %d blocks: %.1f us\n", g, timed([&] { scatter8<<<g, 256>>>(hd, out, 0); }));
    printf("read + write together (gather x=1, scatter x=0 on two streams would overlap; serial here): %.1f us\n",
           timed([&] { gather8<0><<<256, 256>>>(hd, out, 8); scatter8<<<256, 256>>>(hd, out, 0); }));
    {   // gather and scatter at the same time on two streams: do PCIe reads hide behind PCIe writes?
        cudaStream_t s1, s2;
        unsigned long long* out2;
        CK(cudaMalloc(&out2, bytes / kN));
        CK(cudaMemset(out2, 1, bytes / kN));
        CK(cudaStreamCreate(&s1));
        CK(cudaStreamCreate(&s2));
        cudaEvent_t a, b, c;
        CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); CK(cudaEventCreate(&c));
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(a, s1));
        CK(cudaStreamWaitEvent(s2, a, 0));
        for (int i = 0; i < 10; ++i) { gather8<0><<<128, 256, 0, s1>>>(hd, out, 8); scatter8<<<128, 256, 0, s2>>>(hd, out2, 0); }
        CK(cudaEventRecord(c, s2));
        CK(cudaStreamWaitEvent(s1, c, 0));
        CK(cudaEventRecord(b, s1));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        printf("gather (stream 1) and scatter (stream 2) concurrently, one slab each: %.1f us per pair\n", ms * 100.f);
    }
    printf("scatter to host, one request per row of 16 B: %.1f us, 32 B: %.1f us, 64 B: %.1f us\n",
           timed([&] { scatter_wide<1><<<256, 256>>>(hd, out); }), timed([&] { scatter_wide<2><<<256, 256>>>(hd, out); }),
           timed([&] { scatter_wide<4><<<256, 256>>>(hd, out); }));
    for (int g : {128, 512})
        printf("periodic x-wrap inside host memory, both slabs (8-byte cells), %d blocks: %.1f us\n", g,
               timed([&] { wrap_x<<<g, 256>>>(hd); }));
    for (int g : {128, 512})
        printf("the same through whole 64-byte lines (read 2 lines, write 2 lines per row), %d blocks: %.1f us\n", g,
               timed([&] { wrap_x_lines<<<g, 256>>>(hd); }));
    uint4* big;
    CK(cudaMalloc(&big, 1 << 20));
    for (int g : {32, 148})
        printf("contiguous 512 KiB host read, %d blocks: %.1f us;  1 MiB: %.1f us\n", g,
               timed([&] { contiguous<<<g, 256>>>(reinterpret_cast<const uint4*>(hd), big, (512 << 10) / 16); }),
               timed([&] { contiguous<<<g, 256>>>(reinterpret_cast<const uint4*>(hd), big, (1 << 20) / 16); }));
    printf("PCIE_EXPERIMENT_DONE\n");
    return 0;
}
