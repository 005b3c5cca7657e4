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

void launch_coulomb(const Layout& L, Shard& s, int sm_count, cudaStream_t st, bool have_v) {
    if (s.q_len <= 0) return;
    int grid_v = (int)std::min<int64_t>(s.q_len, (int64_t)sm_count * 8);
    // have_v: V already came out of the half-transform's accumulators (no density, no first sweep over B)
    if (!have_v) k_coulomb_v<<<grid_v, 256, 0, st>>>(s.Bp, s.dt, s.V, L.row_len, (int)s.q_len);
    record_event(s, s.ev[EV_V], st);  // boundary V | J
    const int64_t n2 = L.row_len / 2;
    int gx = (int)((n2 + 255) / 256);
    // enough CTAs to fill the machine twice; deterministic (no atomics) whenever one slab suffices
    int slabs = 1;
    if (gx < 2 * sm_count) slabs = (int)std::min<int64_t>((2 * sm_count + gx - 1) / gx, std::max<int64_t>(1, s.q_len / 64));
    int rows_per_slab = (int)((s.q_len + slabs - 1) / slabs);
    if (slabs > 1) cudaMemsetAsync(s.Jt, 0, sizeof(double) * L.row_len, st);
    dim3 grid(gx, slabs);
    k_coulomb_j<<<grid, 256, 0, st>>>(s.Bp, s.V, s.Jt, L.row_len, (int)s.q_len, rows_per_slab, slabs > 1);
}


// ---------------------------------------------------------------------------------------------------------------------
// fused single-pass Coulomb:  V = Bp dt  and  Jt = 2 Bp^T V  with ONE read of the packed shard from HBM.
//
// Cooperative persistent grid, one CTA of 384 threads per SM.  CTA c owns the column segment [c*seg, (c+1)*seg) of every
// packed row and keeps dt[segment] and the Jt[segment] accumulators in registers (EPT2 double2 each per thread).
// Rows are taken in chunks of R; every loop iteration i does
//   phase A (chunk i):        the chunk's row segments arrive FROM HBM through a ring of NSLOT shared-memory slots
//                             (1-D bulk async copies issued NSLOT chunks ahead; the lines also allocate in L2);
//                             partial V over the segment -> part[row][c] in L2, one release-increment of ready[i];
//   phase B (chunk i - LAG):  the same row segments again, now L2 HITS (LAG * R rows = 25-30 MB, a fraction of the
//                             126 MB L2), times the finished V[row] -> Jt accumulators.
// V[row] of chunk i - LAG + 1 is gathered one iteration early by warp 1 (fixed-order sum of the G partials: the result
// is deterministic) and the flag of chunk i - LAG + 2 is polled one iteration before that; both loads are issued at
// the top of an iteration and consumed at its end, so neither the flag latency nor the L2 round trip of the partials
// sits on the critical path, and there is a single __syncthreads per iteration.  HBM sees each element of Bp once; ncu's dram__bytes_read is the check.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int FJ_THREADS = 384;  // 12 warps = 3 per sub-partition: up to 168 registers/thread
constexpr int FJ_WARPS = FJ_THREADS / 32;
constexpr int FJ_LAG = 4;
constexpr int FJ_NSLOT = 3;
constexpr int FJ_PV = 5;  // partials per lane in the V gather: grid <= 160 CTAs
constexpr int FJ_MAXR = 8;

struct FJParams {
    const double* Bp;
    const double* dt;
    double* Jt;
    double* V;
    double* part;          // [q_len][G]
    unsigned int* ready;   // [n_chunks], zeroed before launch
    int64_t row_len;       // doubles per packed row
    int64_t seg;           // doubles per CTA segment (multiple of 2)
    int q_len, n_chunks;
};

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    JC_FUZZ(5);
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add_u32(unsigned int* p, unsigned int v) {
    JC_FUZZ(6);
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int EPT2, int R>
__global__ void __launch_bounds__(FJ_THREADS, 1) k_coulomb_fused(FJParams prm) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ double red[2][FJ_WARPS][R];
    __shared__ double Vs[2][R];
    __shared__ uint64_t full[FJ_NSLOT];
    const int G = gridDim.x, c = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t seg = prm.seg;
    const int seg2 = (int)(seg >> 1);
    const int64_t c0 = (int64_t)c * seg;
    const int64_t my_len = max((int64_t)0, min(seg, prm.row_len - c0));  // doubles, even
    const int my_d2 = (int)(my_len >> 1);
    const int64_t row2 = prm.row_len >> 1;
    const double2* B2 = reinterpret_cast<const double2*>(prm.Bp + c0);
    double* ring = reinterpret_cast<double*>(smem_raw);                  // [FJ_NSLOT][R][seg]
    double2* dts = reinterpret_cast<double2*>(ring + (size_t)FJ_NSLOT * R * seg);  // dt[segment]

    if (tid == 0) {
        for (int s = 0; s < FJ_NSLOT; s++) mbar_init(&full[s], 1);
        fence_barrier_init();
        fence_proxy_async();
    }
    for (int idx = tid; idx < seg2; idx += FJ_THREADS)
        dts[idx] = idx < my_d2 ? reinterpret_cast<const double2*>(prm.dt + c0)[idx] : make_double2(0.0, 0.0);
    __syncthreads();

    auto issue = [&](int chunk) {  // thread 0: R bulk copies HBM -> smem slot, completion on full[slot]
        const int slot = chunk % FJ_NSLOT;
        const int r0 = chunk * R, nr = min(R, prm.q_len - r0);
        if (my_len == 0) { mbar_arrive(&full[slot]); return; }
        mbar_arrive_expect_tx(&full[slot], (uint32_t)(nr * my_len * 8));
        for (int r = 0; r < nr; r++)
            bulk_g2s(ring + ((size_t)slot * R + r) * seg, prm.Bp + (int64_t)(r0 + r) * prm.row_len + c0,
                     (uint32_t)(my_len * 8), &full[slot]);
    };
    if (tid == 0)
        for (int i = 0; i < FJ_NSLOT && i < prm.n_chunks; i++) issue(i);

    double2 acc[EPT2], b[R][EPT2];
#pragma unroll
    for (int k = 0; k < EPT2; k++) acc[k] = make_double2(0.0, 0.0);
#pragma unroll
    for (int r = 0; r < R; r++)
#pragma unroll
        for (int k = 0; k < EPT2; k++) b[r][k] = make_double2(0.0, 0.0);

    for (int i = 0; i < prm.n_chunks + FJ_LAG; i++) {
        const bool doA = i < prm.n_chunks;
        const int j = i - FJ_LAG;
        const bool doB = j >= 0;
        // ---- warp 1, software-pipelined V gather: issue the loads now, use them at the end of the iteration ----
        const int jn = j + 1;   // chunk whose V is assembled in this iteration (its flag was checked last iteration)
        const int jp = j + 2;   // chunk whose flag is polled in this iteration (published LAG - 2 iterations ago)
        const bool gather = (warp == 1) && jn >= 0 && jn < prm.n_chunks;
        const bool poll = (warp == 1) && lane == 0 && jp >= 0 && jp < prm.n_chunks;
        unsigned int flag = (unsigned int)G;
        double pv[R][FJ_PV];
        if (poll) flag = ld_acquire_u32(prm.ready + jp);
        if (gather) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                const int row = jn * R + r;
#pragma unroll
                for (int m = 0; m < FJ_PV; m++) {
                    const int cc = lane + 32 * m;
                    pv[r][m] = (row < prm.q_len && cc < G) ? __ldcg(prm.part + (int64_t)row * G + cc) : 0.0;
                }
            }
        }
        // ---- phase B loads for the NEXT iteration (L2 hits): a whole iteration of latency budget ----
        double2 bn[R][EPT2];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int rowB = (j + 1) * R + r;
#pragma unroll
            for (int k = 0; k < EPT2; k++) {
                const int idx = tid + FJ_THREADS * k;
                bn[r][k] = (j + 1 >= 0 && rowB < prm.q_len && idx < my_d2) ? ldg_stream(B2 + (int64_t)rowB * row2 + idx)
                                                                           : make_double2(0.0, 0.0);
            }
        }
        // ---- phase A: partial V of chunk i over this segment, from the shared-memory slot ----
        if (doA) {
            const int slot = i % FJ_NSLOT;
            const int nr = min(R, prm.q_len - i * R);
            mbar_wait(&full[slot], (uint32_t)((i / FJ_NSLOT) & 1));
            const double2* sb = reinterpret_cast<const double2*>(ring + (size_t)slot * R * seg);
#pragma unroll
            for (int r = 0; r < R; r++) {
                double s0 = 0.0, s1 = 0.0;
                if (r < nr) {
#pragma unroll
                    for (int k = 0; k < EPT2; k++) {
                        const int idx = tid + FJ_THREADS * k;
                        if (idx < my_d2) {
                            const double2 v = sb[r * seg2 + idx], dd = dts[idx];
                            s0 = fma(v.x, dd.x, s0);
                            s1 = fma(v.y, dd.y, s1);
                        }
                    }
                }
                const double s = warp_sum(s0 + s1);
                if (lane == 0) red[i & 1][warp][r] = s;
            }
        }
        __syncthreads();  // red[i&1] complete, slot i free, Vs[j&1] (written by warp 1 last iteration) visible
        if (doA && warp == 0) {
            if (lane == 0 && i + FJ_NSLOT < prm.n_chunks) issue(i + FJ_NSLOT);
            if (lane < R && i * R + lane < prm.q_len) {
                double s = 0.0;
#pragma unroll
                for (int w = 0; w < FJ_WARPS; w++) s += red[i & 1][w][lane];
                __stcg(prm.part + (int64_t)(i * R + lane) * G + c, s);
            }
            __syncwarp();
            if (lane == 0) red_release_add_u32(prm.ready + i, 1u);
        }
        // ---- phase B: Jt += B[row, segment] * V[row] for chunk i - LAG ----
        if (doB) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                const double v = Vs[j & 1][r];
#pragma unroll
                for (int k = 0; k < EPT2; k++) {
                    acc[k].x = fma(b[r][k].x, v, acc[k].x);
                    acc[k].y = fma(b[r][k].y, v, acc[k].y);
                }
            }
        }
        // ---- warp 1: finish the gather (fixed-order sum -> deterministic) and make sure the polled flag is complete ----
        if (gather) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                double s = 0.0;
#pragma unroll
                for (int m = 0; m < FJ_PV; m++) s += pv[r][m];
                s = warp_sum(s);
                if (lane == 0) {
                    Vs[jn & 1][r] = s;
                    if (c == 0 && jn * R + r < prm.q_len) prm.V[jn * R + r] = s;
                }
            }
        }
        if (warp == 1) {
            if (poll)
                while (flag < (unsigned int)G) flag = ld_acquire_u32(prm.ready + jp);
            __syncwarp();
        }
#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int k = 0; k < EPT2; k++) b[r][k] = bn[r][k];
    }
#pragma unroll
    for (int k = 0; k < EPT2; k++) {
        const int idx = tid + FJ_THREADS * k;
        if (idx < my_d2) reinterpret_cast<double2*>(prm.Jt + c0)[idx] = make_double2(2.0 * acc[k].x, 2.0 * acc[k].y);
    }
}

static size_t fj_smem_bytes(int64_t seg, int R) { return (size_t)(FJ_NSLOT * R + 1) * seg * 8 + 128; }

template <int EPT2, int R>
static cudaError_t launch_fj(FJParams& prm, int grid, cudaStream_t st) {
    const size_t smem = fj_smem_bytes(prm.seg, R);
    cudaError_t e = cudaFuncSetAttribute(k_coulomb_fused<EPT2, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    void* args[] = {&prm};
    return cudaLaunchCooperativeKernel((const void*)k_coulomb_fused<EPT2, R>, dim3(grid), dim3(FJ_THREADS), args, smem, st);
}

// Returns false when the shard's rows are too long for the shared-memory ring + register segment (the caller falls
// back to the two-pass sweeps): (NSLOT * R + 1) * seg * 8 B must fit ~220 KB, i.e. packed rows up to ~8 MB.
bool coulomb_fused_plan(const Layout& L, const Shard& s, int sm_count, int64_t* seg_out, int* R_out, int* ept2_out) {
    if (s.q_len <= 0 || L.row_len <= 0 || sm_count > 32 * FJ_PV) return false;
    int64_t seg = (L.row_len + sm_count - 1) / sm_count;
    seg = (seg + 1) / 2 * 2;
    const int d2 = (int)(seg / 2);
    const int need = (d2 + FJ_THREADS - 1) / FJ_THREADS;
    if (need > 8) return false;
    const int ept2 = need <= 1 ? 1 : need <= 2 ? 2 : need <= 4 ? 4 : 8;
    const int R = ept2 == 1 ? 8 : ept2 == 2 ? 4 : ept2 == 4 ? 2 : 1;
    if (fj_smem_bytes(seg, R) > 220 * 1024) return false;
    *seg_out = seg;
    *ept2_out = ept2;
    *R_out = R;
    return true;
}

bool launch_coulomb_fused(const Layout& L, Shard& s, int sm_count, cudaStream_t st) {
    int64_t seg;
    int R, ept2;
    if (!coulomb_fused_plan(L, s, sm_count, &seg, &R, &ept2)) return false;
    FJParams prm;
    prm.Bp = s.Bp; prm.dt = s.dt; prm.Jt = s.Jt; prm.V = s.V; prm.part = s.Vpart; prm.ready = s.Vready;
    prm.row_len = L.row_len; prm.seg = seg; prm.q_len = (int)s.q_len;
    prm.n_chunks = (int)((s.q_len + R - 1) / R);
    cudaMemsetAsync(s.Vready, 0, sizeof(unsigned int) * (size_t)prm.n_chunks, st);
    cudaError_t e;
    switch (ept2) {
        case 1: e = launch_fj<1, 8>(prm, sm_count, st); break;
        case 2: e = launch_fj<2, 4>(prm, sm_count, st); break;
        case 4: e = launch_fj<4, 2>(prm, sm_count, st); break;
        default: e = launch_fj<8, 1>(prm, sm_count, st); break;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();  // clear; the caller falls back to the two-pass sweeps
        return false;
    }
    return true;
}

JC_DEFINE_FUZZ_SETTER(set_fuzz_seed_j)

}  // namespace jcdf
