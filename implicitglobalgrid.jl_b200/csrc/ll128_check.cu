// LL128 self-test, run once at bootstrap between ring neighbours.
//
// What the LL128 protocol assumes and PTX does not promise: the 128-byte line that eight lanes of one warp store with
// one 16-byte instruction each (seven of payload, one of payload + epoch) reaches the peer GPU's L2 as ONE unit, and a
// warp that loads the line with the same instruction shape sees either the old line or the new one, never a mix.  It
// is what NCCL's LL128 relies on over NVLink (and switches off where the topology does not give it).  Here the
// assumption is checked on the very box and path the exchange will use, before the first exchange: every rank
// rewrites a few thousand lines in its right neighbour's control segment `rounds` times without any hand-shake while
// that neighbour reads them over and over, and every line image whose epoch word and payload do not belong to the
// same round counts as torn.  One torn line (or a neighbour that never shows up) makes the context fall back to
// the fence protocol for good (ctx.cu).
#include <cuda_runtime.h>

#include "igg_internal.h"

namespace igg {

namespace {

__device__ __forceinline__ unsigned long long mix(unsigned long long line, unsigned long long round, unsigned long long k) {
    unsigned long long x = (line << 32) ^ (round << 8) ^ k ^ 0x9e3779b97f4a7c15ull;
    x ^= x >> 29; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 32;
    return x | 1ull;
}

__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Blocks [0, gridDim/2) write into the right neighbour, blocks [gridDim/2, gridDim) read what the left neighbour
// writes here.  result[0] = torn line images, result[1] = time-outs, result[2] = line images checked.
__global__ void __launch_bounds__(256) igg_ll128_selftest_kernel(char* __restrict__ to_peer, const char* __restrict__ mine,
                                                                  int lines_per_block, int rounds, int corrupt,
                                                                  unsigned long long timeout_ns, unsigned long long* result) {
    const int half = gridDim.x / 2;
    const bool sender = (int)blockIdx.x < half;
    const unsigned j = threadIdx.x & 7u, lane = threadIdx.x & 31u;
    const unsigned long long line0 = (unsigned long long)(sender ? blockIdx.x : blockIdx.x - half) * lines_per_block;
    if (sender) {
        for (int r = 1; r <= rounds; ++r)
            for (int l = threadIdx.x >> 3; l < lines_per_block; l += 32) {
                const unsigned long long line = line0 + l;
                unsigned long long lo = mix(line, r, 2 * j), hi = j == 7 ? (unsigned long long)r : mix(line, r, 2 * j + 1);
                if (corrupt && j == 3 && r > 1) lo = mix(line, r - 1, 2 * j);     // IGG_LL128_SELFTEST_FAIL: a "torn" line
                asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(to_peer + line * 128 + 16 * j), "l"(lo), "l"(hi) : "memory");
            }
        return;
    }
    unsigned long long torn = 0, checked = 0;
    bool timed_out = false;
    const unsigned long long t0 = now_ns();
    for (int l = threadIdx.x >> 3; l < lines_per_block && !timed_out; l += 32) {       // uniform over the warp
        const unsigned long long line = line0 + l;
        unsigned spins = 0;
        while (true) {
            unsigned long long lo, hi;
            asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(mine + line * 128 + 16 * j) : "memory");
            const unsigned long long e = __shfl_sync(0xffffffffu, hi, lane | 7u);
            bool bad = false;
            if (e != 0) bad = lo != mix(line, e, 2 * j) || (j != 7 && hi != mix(line, e, 2 * j + 1));
            const unsigned badmask = __ballot_sync(0xffffffffu, bad);
            if (j == 7 && e != 0) { checked += 1; if ((badmask >> (lane & ~7u)) & 0xffu) torn += 1; }
            if (__all_sync(0xffffffffu, e == (unsigned long long)rounds)) break;
            if ((++spins & 0xffu) == 0) {
                unsigned long long t = now_ns();
                t = __shfl_sync(0xffffffffu, t, 0);
                if (t - t0 > timeout_ns) { timed_out = true; break; }
            }
        }
    }
    if (torn) atomicAdd(&result[0], torn);
    if (timed_out && lane == 0) atomicAdd(&result[1], 1ull);
    if (checked) atomicAdd(&result[2], checked);
}

}  // namespace

// Runs the test: `to_peer` = scratch in the right neighbour's control segment, `mine` = the same place in mine (both
// zeroed beforehand, `lines` lines of 128 bytes).  out[0..2] as above.  Synchronous.
int ll128_selftest(char* to_peer, const char* mine, int lines, int rounds, bool corrupt, unsigned long long timeout_ns,
                   unsigned long long out[3]) {
    unsigned long long* d_res = nullptr;
    cudaError_t e = cudaMalloc(&d_res, 3 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(d_res, 0, 3 * sizeof(unsigned long long));
    if (e != cudaSuccess) { if (d_res) cudaFree(d_res); return fail(IGG_ERR_CUDA, "LL128 self-test: %s", cudaGetErrorString(e)); }
    const int blocks = 16, lines_per_block = lines / blocks;
    igg_ll128_selftest_kernel<<<2 * blocks, 256>>>(to_peer, mine, lines_per_block, rounds, corrupt ? 1 : 0, timeout_ns, d_res);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(out, d_res, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaFree(d_res);
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "LL128 self-test: %s", cudaGetErrorString(e));
    return IGG_OK;
}

}  // namespace igg
