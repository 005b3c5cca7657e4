// Coulomb (J) path on sm_100a — HBM-bandwidth-bound sweeps over the packed, tile-compacted B shard.
//
//   V[Q]   = sum_pst B[Q][pst] * dt[pst]          replaces the per-p gemv('N') loop  (ScreenedDF.jl:326-341)
//   Jt[pst] = 2 * sum_Q B[Q][pst] * V[Q]          replaces the per-p gemv('T', 2.0)  (ScreenedDF.jl:343-358)
//
// dt carries the reference's doubled off-diagonal density in tile space (kernels_k.cu: k_density_packed), so one pass
// over the q >= p tiles reproduces the sums over both triangles.
//
// Two implementations:
//   * two-pass (k_coulomb_v + k_coulomb_j): each streams the shard once, 128-bit loads, warp-shuffle reductions.
//   * fused single-pass (k_coulomb_fused): a persistent cooperative grid stages row chunks in shared memory with
//     bulk async copies; partial V are combined through a pipelined flag barrier in L2 and the J contribution of the
//     chunk is taken from shared memory, so HBM is read exactly once.
#include "dmma_common.cuh"
#include "jcdf_internal.h"

namespace jcdf {

__device__ __forceinline__ double2 ldg_stream(const double2* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

// ---------------------------------------------------------------------------------------------------------------------
// two-pass
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_coulomb_v(const double* __restrict__ Bp, const double* __restrict__ dt,
                                                    double* __restrict__ V, int64_t row_len, int q_len) {
    __shared__ double red[8];
    const int64_t n2 = row_len >> 1;
    const double2* d2 = reinterpret_cast<const double2*>(dt);
    for (int row = blockIdx.x; row < q_len; row += gridDim.x) {
        const double2* b2 = reinterpret_cast<const double2*>(Bp + (int64_t)row * row_len);
        double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
        int64_t i = threadIdx.x;
        for (; i + 3 * 256 < n2; i += 4 * 256) {
            double2 x0 = ldg_stream(b2 + i), x1 = ldg_stream(b2 + i + 256), x2 = ldg_stream(b2 + i + 512),
                    x3 = ldg_stream(b2 + i + 768);
            double2 y0 = __ldg(d2 + i), y1 = __ldg(d2 + i + 256), y2 = __ldg(d2 + i + 512), y3 = __ldg(d2 + i + 768);
            s0 = fma(x0.x, y0.x, fma(x0.y, y0.y, s0));
            s1 = fma(x1.x, y1.x, fma(x1.y, y1.y, s1));
            s2 = fma(x2.x, y2.x, fma(x2.y, y2.y, s2));
            s3 = fma(x3.x, y3.x, fma(x3.y, y3.y, s3));
        }
        for (; i < n2; i += 256) {
            double2 x = ldg_stream(b2 + i), y = __ldg(d2 + i);
            s0 = fma(x.x, y.x, fma(x.y, y.y, s0));
        }
        double s = warp_sum((s0 + s1) + (s2 + s3));
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x < 32) {
            double v = threadIdx.x < 8 ? red[threadIdx.x] : 0.0;
            v = warp_sum(v);
            if (threadIdx.x == 0) V[row] = v;
        }
        __syncthreads();
    }
}

// grid.x over column pairs, grid.y over row slabs; slabs > 1 accumulate with atomics into a zeroed Jt
__global__ void __launch_bounds__(256) k_coulomb_j(const double* __restrict__ Bp, const double* __restrict__ V,
                                                    double* __restrict__ Jt, int64_t row_len, int q_len,
                                                    int rows_per_slab, int use_atomic) {
    const int64_t n2 = row_len >> 1;
    const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (c >= n2) return;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(q_len, r0 + rows_per_slab);
    const double2* b2 = reinterpret_cast<const double2*>(Bp) + c;
    double ax = 0, ay = 0, bx = 0, by = 0;
    int r = r0;
    for (; r + 7 < r1; r += 8) {
        double2 x[8];
#pragma unroll
        for (int u = 0; u < 8; u++) x[u] = ldg_stream(b2 + (int64_t)(r + u) * n2);
#pragma unroll
        for (int u = 0; u < 8; u += 2) {
            const double v0 = __ldg(V + r + u), v1 = __ldg(V + r + u + 1);
            ax = fma(x[u].x, v0, ax);
            ay = fma(x[u].y, v0, ay);
            bx = fma(x[u + 1].x, v1, bx);
            by = fma(x[u + 1].y, v1, by);
        }
    }
    for (; r < r1; r++) {
        double2 x = ldg_stream(b2 + (int64_t)r * n2);
        const double v0 = __ldg(V + r);
        ax = fma(x.x, v0, ax);
        ay = fma(x.y, v0, ay);
    }
    const double jx = 2.0 * (ax + bx), jy = 2.0 * (ay + by);
    if (use_atomic) {
        atomicAdd(Jt + 2 * c, jx);
        atomicAdd(Jt + 2 * c + 1, jy);
    } else {
        reinterpret_cast<double2*>(Jt)[c] = make_double2(jx, jy);
    }
}

void launch_coulomb(const Layout& L, Shard& s, int sm_count) {
    if (s.q_len <= 0) return;
    int grid_v = (int)std::min<int64_t>(s.q_len, (int64_t)sm_count * 8);
    k_coulomb_v<<<grid_v, 256, 0, s.stream>>>(s.Bp, s.dt, s.V, L.row_len, (int)s.q_len);
    cudaEventRecord(s.ev[EV_V], s.stream);  // boundary V | J
    const int64_t n2 = L.row_len / 2;
    int gx = (int)((n2 + 255) / 256);
    // enough CTAs to fill the machine twice; deterministic (no atomics) whenever one slab suffices
    int slabs = 1;
    if (gx < 2 * sm_count) slabs = (int)std::min<int64_t>((2 * sm_count + gx - 1) / gx, std::max<int64_t>(1, s.q_len / 64));
    int rows_per_slab = (int)((s.q_len + slabs - 1) / slabs);
    if (slabs > 1) cudaMemsetAsync(s.Jt, 0, sizeof(double) * L.row_len, s.stream);
    dim3 grid(gx, slabs);
    k_coulomb_j<<<grid, 256, 0, s.stream>>>(s.Bp, s.V, s.Jt, L.row_len, (int)s.q_len, rows_per_slab, slabs > 1);
}

}  // namespace jcdf
