// The caller of the hot path: one time step of the reference's 3-D heat diffusion example
// (ref: examples/diffusion3D_multigpu_CuArrays_novis.jl:41-49), fused into one sm_100a kernel.
//
// The example does five broadcasts per step (qx, qy, qz, dTedt, T update; ~14 array passes) and then
// update_halo!(T).  Here T_new is written to a second array (the in-place update of the example reads only values
// of the previous step, so this is the same Jacobi step), every flux is evaluated with the example's operation
// order in IEEE double precision without fused multiply-add -- results are bit-identical to the example's
// arithmetic (checked against the numpy restatement in oracle/diffusion_oracle.py) -- and boundary cells are carried
// over unchanged.  `part` selects all cells, only the shell that the halo update sends (computed first so that the
// exchange can overlap the rest), or only the interior.
#include <cuda_runtime.h>

#include "igg_internal.h"

namespace {

__device__ __forceinline__ double flux(double lam_neg, double a, double b, double d) {
    // -lam .* (a - b) ./ d     (ref: examples/diffusion3D_multigpu_CuArrays_novis.jl:43-45)
    return __ddiv_rn(__dmul_rn(lam_neg, __dsub_rn(a, b)), d);
}

struct DiffParams {
    const double* T;
    const double* Cp;
    double* T2;
    long long nx, ny, nz;
    double lam_neg, dt, dx, dy, dz;
    long long lo[3], hi[3];   // 0-based inclusive box to process
    int fast;                 // 1: multiply by reciprocals and allow FMA (not bit-identical to the example, ~1e-15 rel.)
    double rdx, rdy, rdz;     // 1/dx, 1/dy, 1/dz for the fast variant
};

// One cell.  `t` = T at the cell, `tzm` / `tzp` = T at k-1 / k+1 (the callers keep the z column in registers).
__device__ __forceinline__ void update_cell(const DiffParams& p, long long i, long long j, long long k, long long c, double t, double tzm, double tzp) {
    const long long sx = 1, sy = p.nx;
    if (i == 0 || j == 0 || k == 0 || i == p.nx - 1 || j == p.ny - 1 || k == p.nz - 1) {
        p.T2[c] = t;                                       // boundary / halo cell: unchanged by the update (:47)
        return;
    }
    if (p.fast) {
        // Bandwidth-bound variant: the ten IEEE divisions per cell of the exact variant cost more than the memory
        // traffic on B200; here they become multiplications by precomputed reciprocals.
        const double lx = p.lam_neg * p.rdx * p.rdx, ly = p.lam_neg * p.rdy * p.rdy, lz = p.lam_neg * p.rdz * p.rdz;
        const double dqx = lx * ((p.T[c + sx] - t) - (t - p.T[c - sx]));
        const double dqy = ly * ((p.T[c + sy] - t) - (t - p.T[c - sy]));
        const double dqz = lz * ((tzp - t) - (t - tzm));
        p.T2[c] = t + p.dt * (-(dqx + dqy + dqz)) / p.Cp[c];
        return;
    }
    const double qx_hi = flux(p.lam_neg, p.T[c + sx], t, p.dx), qx_lo = flux(p.lam_neg, t, p.T[c - sx], p.dx);
    const double qy_hi = flux(p.lam_neg, p.T[c + sy], t, p.dy), qy_lo = flux(p.lam_neg, t, p.T[c - sy], p.dy);
    const double qz_hi = flux(p.lam_neg, tzp, t, p.dz), qz_lo = flux(p.lam_neg, t, tzm, p.dz);
    // dTedt = 1.0 ./ Cp .* (-d_xa(qx)./dx .- d_ya(qy)./dy .- d_za(qz)./dz)                                     (:46)
    const double ax = __ddiv_rn(-__dsub_rn(qx_hi, qx_lo), p.dx);
    const double ay = __ddiv_rn(__dsub_rn(qy_hi, qy_lo), p.dy);
    const double az = __ddiv_rn(__dsub_rn(qz_hi, qz_lo), p.dz);
    const double rhs = __dsub_rn(__dsub_rn(ax, ay), az);
    const double dTedt = __dmul_rn(__ddiv_rn(1.0, p.Cp[c]), rhs);
    p.T2[c] = __dadd_rn(t, __dmul_rn(p.dt, dTedt));       // T = T + dt*dTedt                                   (:47)
}

// Box kernel: a thread owns one (i, j) column of kZPerThread cells and marches along z with the three z-neighbours in
// registers.  Measured on B200, 256^3: 8 cells per thread 123 us (fast) / 274 us (exact) per step against 108.7 / 243
// with one cell per thread (fewer, longer threads hide less latency; the z-neighbours hit L2 anyway), so one it is.
constexpr int kZPerThread = 1;
__global__ void __launch_bounds__(256) igg_diffusion3d_kernel(const DiffParams p) {
    const long long i = p.lo[0] + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long j = p.lo[1] + (long long)blockIdx.y * blockDim.y + threadIdx.y;
    const long long k0 = p.lo[2] + (long long)blockIdx.z * kZPerThread;
    if (i > p.hi[0] || j > p.hi[1] || k0 > p.hi[2]) return;
    const long long sz = p.nx * p.ny;
    long long c = i + j * p.nx + k0 * sz;
    double tzm = k0 > 0 ? p.T[c - sz] : 0.0, t = p.T[c];
#pragma unroll
    for (int q = 0; q < kZPerThread; ++q) {
        const long long k = k0 + q;
        if (k > p.hi[2]) break;
        const double tzp = k < p.nz - 1 ? p.T[c + sz] : 0.0;
        update_cell(p, i, j, k, c, t, tzm, tzp);
        tzm = t; t = tzp; c += sz;
    }
}

// Shell kernel: the two outermost layers of the block (what update_halo! sends, plus the boundary copy) in ONE launch
// over the six slabs, cut so that no cell is visited twice: z-low, z-high (full planes), y-low, y-high (without the
// z slabs), x-low, x-high (without both).
__global__ void __launch_bounds__(256) igg_diffusion3d_shell_kernel(const DiffParams p) {
    const long long nx = p.nx, ny = p.ny, nz = p.nz;
    const long long nzs = 2 * nx * ny, nys = 2 * nx * (nz - 4), nxs = 2 * (ny - 4) * (nz - 4);
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long i, j, k;
    if (q < 2 * nzs) {
        const long long s = q / nzs, r = q - s * nzs;
        k = r / (nx * ny) + (s ? nz - 2 : 0);
        const long long rr = r % (nx * ny);
        j = rr / nx; i = rr - j * nx;
    } else if ((q -= 2 * nzs) < 2 * nys) {
        const long long s = q / nys, r = q - s * nys;
        k = 2 + r / (2 * nx);
        const long long rr = r % (2 * nx);
        j = rr / nx + (s ? ny - 2 : 0); i = rr % nx;
    } else if ((q -= 2 * nys) < 2 * nxs) {
        const long long s = q / nxs, r = q - s * nxs;
        k = 2 + r / (2 * (ny - 4));
        const long long rr = r % (2 * (ny - 4));
        j = 2 + rr / 2; i = (rr & 1) + (s ? nx - 2 : 0);
    } else {
        return;
    }
    const long long sz = nx * ny, c = i + j * nx + k * sz;
    update_cell(p, i, j, k, c, p.T[c], k > 0 ? p.T[c - sz] : 0.0, k < nz - 1 ? p.T[c + sz] : 0.0);
}

}  // namespace

// part_and_flags: bits 0-1 = part (0 = every cell, 1 = the two outermost layers -- boundary copy + the cells
// update_halo! sends --, 2 = the rest); bit 2 (value 4) = fast arithmetic (reciprocals + FMA, not bit-identical).
extern "C" int igg_diffusion3d_step(const double* T, const double* Cp, double* T2, const int64_t n[3], double lam,
                                    double dt, double dx, double dy, double dz, int part_and_flags, void* stream) {
    using namespace igg;
    if (!T || !Cp || !T2 || !n) return fail(IGG_ERR_ARG, "igg_diffusion3d_step: NULL argument.");
    if (n[0] < 3 || n[1] < 3 || n[2] < 3) return fail(IGG_ERR_ARG, "igg_diffusion3d_step: needs at least 3 cells per dimension.");
    const int part = part_and_flags & 3;
    if (part < 0 || part > 2) return fail(IGG_ERR_ARG, "igg_diffusion3d_step: part must be 0, 1 or 2.");
    DiffParams p;
    p.T = T; p.Cp = Cp; p.T2 = T2;
    p.nx = n[0]; p.ny = n[1]; p.nz = n[2];
    p.lam_neg = -lam; p.dt = dt; p.dx = dx; p.dy = dy; p.dz = dz;
    p.fast = (part_and_flags & 4) ? 1 : 0;
    p.rdx = 1.0 / dx; p.rdy = 1.0 / dy; p.rdz = 1.0 / dz;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    auto launch = [&](long long l0, long long l1, long long l2, long long h0, long long h1, long long h2) {
        if (h0 < l0 || h1 < l1 || h2 < l2) return;
        const dim3 block = (h0 - l0 + 1 <= 2) ? dim3(2, 128, 1) : dim3(128, 2, 1);   // thin x-slabs: threads along y
        p.lo[0] = l0; p.lo[1] = l1; p.lo[2] = l2;
        p.hi[0] = h0; p.hi[1] = h1; p.hi[2] = h2;
        const dim3 grid((unsigned)((h0 - l0 + 1 + block.x - 1) / block.x), (unsigned)((h1 - l1 + 1 + block.y - 1) / block.y),
                        (unsigned)((h2 - l2 + 1 + kZPerThread - 1) / kZPerThread));
        igg_diffusion3d_kernel<<<grid, block, 0, s>>>(p);
    };
    const long long X = n[0] - 1, Y = n[1] - 1, Z = n[2] - 1;
    if (part == 0) {
        launch(0, 0, 0, X, Y, Z);
    } else if (part == 2) {
        launch(2, 2, 2, X - 2, Y - 2, Z - 2);
    } else if (n[0] >= 5 && n[1] >= 5 && n[2] >= 5) {
        // the six slabs of thickness 2 in one launch (six launches cost ~9 us per step of the overlap schedule)
        const long long cells = 4 * (n[0] * n[1] + n[0] * (n[2] - 4) + (n[1] - 4) * (n[2] - 4));
        p.lo[0] = p.lo[1] = p.lo[2] = 0;
        p.hi[0] = X; p.hi[1] = Y; p.hi[2] = Z;
        igg_diffusion3d_shell_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, s>>>(p);
    } else {
        launch(0, 0, 0, X, Y, Z);              // blocks thinner than 5 cells are all shell
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(IGG_ERR_CUDA, "diffusion3d kernel launch failed: %s", cudaGetErrorString(e));
    return IGG_OK;
}
