// Per-function half-transform over a partner-compacted copy of B (sm_100a, FP64 DMMA m8n8k4 + 1-D bulk async copies).
//
// The reference's screened algorithm (calculate_W_screened, ScreenedDF.jl:242-289) does, for every basis function p, one
// GEMM over the K_p partners q the Schwarz mask kept:
//       W[:, :, p] = B[:, s_p : s_p + K_p - 1] * C_occ[kept q, :]          (M = Q, N = occ, K = K_p)
// so its flop count is exactly 2 Q occ P.  The tile kernel of kernels_k.cu pays for every 4-wide strip of a 32 x 32
// tile that holds one surviving pair (w50: 35 % of the dense k-steps for 12.9 % surviving pairs).  This file restates the
// reference's per-p GEMM for the tensor cores instead:
//
//   Bc [chunk][row][128]   row = (p, k-th kept partner of p), partner lists padded to multiples of 4 rows; the 128
//                          auxiliary rows of a chunk are the fastest index (swizzled by row & 3), so a stage of 32
//                          partner rows is ONE contiguous 32 KiB bulk copy and lands conflict-free as the A operand;
//   Ci [q][ldci]           C_occ rows in internal order, row-major: the gather "kept rows of C" is one bulk copy per
//                          partner row, issued by the 32 lanes of a producer warp (row stride NI + 4 doubles in smem);
//   W  [kc][p64][64][32]   a work item (p, chunk, n tile) owns whole 256-byte runs of the k-fastest W tiles.
//
// Executed flops = 2 Q occ_tile sum_p ceil4(K_p): 1.02 x the algorithmic count at w50.
#include <algorithm>

#include "dmma_common.cuh"
#include "jcdf_internal.h"

namespace jcdf {

constexpr int PP_QC = 128;       // auxiliary rows per chunk (CTA tile M)
constexpr int PP_KS = 32;        // partner rows per pipeline stage
constexpr int PP_NSTAGE = 3;
constexpr int PP_CONSUMERS = 8;  // 2 (M halves of 64 rows) x 4 (N quarters of 8 * NT orbitals)
constexpr int PP_THREADS = (PP_CONSUMERS + 1) * 32;

// ---------------------------------------------------------------------------------------------------------------------
// Bp (tile-packed, one auxiliary row contiguous) -> Bc (partner-compacted, 128 auxiliary rows contiguous).  One-off per
// B formation; a 32-row x 128-column transpose through shared memory.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pp_pack(const double* __restrict__ Bp, int64_t row_len,
                                                  const int32_t* __restrict__ src, double* __restrict__ Bc, int64_t rows,
                                                  int q_len) {
    __shared__ double tile[32][PP_QC + 1];
    const int64_t r0 = (int64_t)blockIdx.x * 32;
    const int ch = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sr = (r0 + lane < rows) ? src[r0 + lane] : -1;
    for (int m = warp; m < PP_QC; m += 8) {
        const int64_t Q = (int64_t)ch * PP_QC + m;
        tile[lane][m] = (sr >= 0 && Q < q_len) ? Bp[Q * row_len + sr] : 0.0;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < 32 * PP_QC; idx += 256) {
        const int r = idx >> 7, m = idx & (PP_QC - 1);
        if (r0 + r < rows) Bc[((int64_t)ch * rows + r0 + r) * PP_QC + (m ^ ((r & 3) << 2))] = tile[r][m];
    }
}

// Ci[q][i] = C_occ[inv[q], i], zero beyond N / occ; row n_rows - 1 is the all-zero row the padding partners point to.
__global__ void k_pp_prep(const double* __restrict__ Craw, int64_t ldc, double* __restrict__ Ci, int N, int occ,
                          int n_rows, int ldci, const int32_t* __restrict__ inv) {
    const int64_t total = (int64_t)n_rows * ldci;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        // q fastest among the threads of a warp: coalesced reads of the column-major upload
        const int q = (int)(idx % n_rows);
        const int i = (int)(idx / n_rows);
        Ci[(int64_t)q * ldci + i] = (q < N && i < occ) ? Craw[inv[q] + ldc * i] : 0.0;
    }
}

struct PPParams {
    const double* Bc;
    const double* Ci;
    double* W;
    const int32_t* off;
    const int32_t* q;
    int64_t rows;
    int ldci, n_fn, n_nt, chunks, q_len, npt64, occ_pad;
    int64_t n_items;
};

template <int NT, bool PARTIAL>
__device__ __forceinline__ void pp_kstep(double (&acc)[8][NT][2], const double* __restrict__ sA,
                                         const double* __restrict__ sC, int m0, int n0, int g, int t, int ks, int mt_valid,
                                         int nt_valid) {
    constexpr int LDC = 32 * NT + 4;
    const int kr = 4 * ks + t;
    double a[8], b[NT];
#pragma unroll
    for (int jm = 0; jm < 8; jm++) a[jm] = sA[kr * PP_QC + ((m0 + 8 * jm + g) ^ (t << 2))];
#pragma unroll
    for (int jn = 0; jn < NT; jn++) b[jn] = sC[kr * LDC + n0 + 8 * jn + g];
#pragma unroll
    for (int jm = 0; jm < 8; jm++) {
        if (PARTIAL && jm >= mt_valid) continue;
#pragma unroll
        for (int jn = 0; jn < NT; jn++) {
            if (PARTIAL && jn >= nt_valid) continue;
            dmma(acc[jm][jn][0], acc[jm][jn][1], a[jm], b[jn]);
        }
    }
}

// One work item = (function p, chunk of 128 auxiliary rows, tile of NI = 32 NT occupied orbitals):
//     out[Q, i] = sum_k Bc[chunk][off_p + k][Q] * Ci[q_k(p)][i]        (M = Q, N = i, K = padded partner count of p)
// Warp 8 is the producer: lane 0 issues the A copy, lane j the copy of partner row j; warps 0-7 own 64 x (8 NT) blocks.
template <int NT>
__global__ void __launch_bounds__(PP_THREADS, 1) k_pp_half_transform(PPParams prm) {
    constexpr int NI = 32 * NT;
    constexpr int LDC = NI + 4;  // = 4 mod 16 doubles: the b fragment rows t = 0..3 fall on distinct bank quads
    constexpr int STAGE_DOUBLES = PP_KS * PP_QC + PP_KS * LDC;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stages = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + (size_t)PP_NSTAGE * STAGE_DOUBLES);
    uint64_t* empty = full + PP_NSTAGE;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < PP_NSTAGE; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], PP_CONSUMERS);
        }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    const int chunks = prm.chunks, n_nt = prm.n_nt;

    if (warp == PP_CONSUMERS) {
        // ------------------------------------------------ producer warp ------------------------------------------------
        PipeState pp;
        for (int64_t item = blockIdx.x; item < prm.n_items; item += gridDim.x) {
            const int ch = (int)(item % chunks);
            const int64_t r1 = item / chunks;
            const int nt = (int)(r1 % n_nt);
            const int p = (int)(r1 / n_nt);
            const int r_begin = prm.off[p];
            const int k4 = prm.off[p + 1] - r_begin;
            const double* Abase = prm.Bc + ((int64_t)ch * prm.rows + r_begin) * PP_QC;
            const double* Cbase = prm.Ci + (int64_t)nt * NI;
            for (int k0 = 0; k0 < k4; k0 += PP_KS) {
                const int nr = min(PP_KS, k4 - k0);
                const int qrow = (lane < nr) ? prm.q[r_begin + k0 + lane] : 0;
                mbar_wait(&empty[pp.stage], pp.phase ^ 1u);
                double* st = stages + (size_t)pp.stage * STAGE_DOUBLES;
                if (lane == 0) {
                    mbar_arrive_expect_tx(&full[pp.stage], (uint32_t)nr * (uint32_t)((PP_QC + NI) * 8));
                    bulk_g2s(st, Abase + (int64_t)k0 * PP_QC, (uint32_t)nr * PP_QC * 8u, &full[pp.stage]);
                }
                __syncwarp();
                if (lane < nr)
                    bulk_g2s(st + PP_KS * PP_QC + lane * LDC, Cbase + (int64_t)qrow * prm.ldci, NI * 8u, &full[pp.stage]);
                pp.advance<PP_NSTAGE>();
            }
        }
        return;
    }

    // --------------------------------------------------- consumers ---------------------------------------------------
    PipeState ps;
    const int g = lane >> 2, t = lane & 3;
    const int m0 = (warp >> 2) * 64;
    const int n0 = (warp & 3) * 8 * NT;
    for (int64_t item = blockIdx.x; item < prm.n_items; item += gridDim.x) {
        const int ch = (int)(item % chunks);
        const int64_t r1 = item / chunks;
        const int nt = (int)(r1 % n_nt);
        const int p = (int)(r1 / n_nt);
        const int k4 = prm.off[p + 1] - prm.off[p];
        const int rows_valid = min(PP_QC, prm.q_len - ch * PP_QC);
        const int mt_valid = max(0, min(8, (rows_valid - m0 + 7) >> 3));
        const int nt_valid = max(0, min(NT, (prm.occ_pad - nt * NI - n0 + 7) >> 3));
        const bool any = mt_valid > 0 && nt_valid > 0;
        const bool whole = mt_valid == 8 && nt_valid == NT;
        double acc[8][NT][2];
#pragma unroll
        for (int jm = 0; jm < 8; jm++)
#pragma unroll
            for (int jn = 0; jn < NT; jn++) acc[jm][jn][0] = acc[jm][jn][1] = 0.0;

        for (int k0 = 0; k0 < k4; k0 += PP_KS) {
            const int nks = min(PP_KS, k4 - k0) >> 2;
            mbar_wait(&full[ps.stage], ps.phase);
            const double* sA = stages + (size_t)ps.stage * STAGE_DOUBLES;
            const double* sC = sA + PP_KS * PP_QC;
            if (whole) {
                if (nks == PP_KS / 4) {
#pragma unroll 2
                    for (int ks = 0; ks < PP_KS / 4; ks++) pp_kstep<NT, false>(acc, sA, sC, m0, n0, g, t, ks, 8, NT);
                } else {
                    for (int ks = 0; ks < nks; ks++) pp_kstep<NT, false>(acc, sA, sC, m0, n0, g, t, ks, 8, NT);
                }
            } else if (any) {
                for (int ks = 0; ks < nks; ks++) pp_kstep<NT, true>(acc, sA, sC, m0, n0, g, t, ks, mt_valid, nt_valid);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[ps.stage]);
            ps.advance<PP_NSTAGE>();
        }
        if (!any) continue;
        // epilogue: the thread's (i, i + 1) pair is adjacent in the k-fastest W tile -> 16-byte stores, 4 lanes = 64 bytes
        const int p64 = p >> 6, pr = p & 63;
        const int sw = (pr & 3) << 2;
#pragma unroll
        for (int jm = 0; jm < 8; jm++) {
            const int64_t Qrow = (int64_t)ch * PP_QC + m0 + 8 * jm + g;
            if (Qrow >= prm.q_len) continue;
#pragma unroll
            for (int jn = 0; jn < NT; jn++) {
                const int i = nt * NI + n0 + 8 * jn + 2 * t;
                if (i >= prm.occ_pad) continue;
                const int64_t krow = Qrow * prm.occ_pad + i;
                const int64_t kc = krow >> 5;
                const int kr = (int)(krow & 31);
                *reinterpret_cast<double2*>(prm.W + ((kc * prm.npt64 + p64) * 64 + pr) * 32 + (kr ^ sw)) =
                    make_double2(acc[jm][jn][0], acc[jm][jn][1]);
            }
        }
    }
}

static size_t pp_smem_bytes(int nt) {
    return (size_t)PP_NSTAGE * (PP_KS * PP_QC + PP_KS * (32 * nt + 4)) * 8 + 2 * PP_NSTAGE * 8 + 16;
}

template <int NT>
static void launch_pp_t(const PPParams& prm, int grid, cudaStream_t st) {
    const size_t smem = pp_smem_bytes(NT);
    cudaFuncSetAttribute(k_pp_half_transform<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_pp_half_transform<NT><<<grid, PP_THREADS, smem, st>>>(prm);
}

void launch_pp_pack(const Layout& L, Shard& s) {
    if (s.pp_chunks <= 0 || L.pp_rows <= 0) return;
    dim3 grid((unsigned)((L.pp_rows + 31) / 32), (unsigned)s.pp_chunks);
    k_pp_pack<<<grid, 256, 0, s.stream>>>(s.Bp, L.row_len, s.d_pp_src, s.Bc, L.pp_rows, (int)s.q_len);
}

void launch_pp_prep(const Layout& L, Shard& s, int64_t ldc) {
    const int ldci = L.pp_n_nt * 32 * L.pp_nt;
    const int64_t total = (int64_t)(L.n_pad + 1) * ldci;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    k_pp_prep<<<blocks, 256, 0, s.stream>>>(s.Craw, ldc, s.Ci, (int)L.N, (int)L.occ, L.n_pad + 1, ldci, s.d_inv);
}

void launch_pp_half_transform(const Layout& L, Shard& s, int sm_count) {
    PPParams prm;
    prm.Bc = s.Bc;
    prm.Ci = s.Ci;
    prm.W = s.W;
    prm.off = s.d_pp_off;
    prm.q = s.d_pp_q;
    prm.rows = L.pp_rows;
    prm.ldci = L.pp_n_nt * 32 * L.pp_nt;
    prm.n_fn = (int)L.N;
    prm.n_nt = L.pp_n_nt;
    prm.chunks = s.pp_chunks;
    prm.q_len = (int)s.q_len;
    prm.npt64 = L.npt64;
    prm.occ_pad = L.occ_pad;
    prm.n_items = (int64_t)L.N * L.pp_n_nt * s.pp_chunks;
    const int grid = (int)std::min<int64_t>(prm.n_items, sm_count);
    if (grid <= 0) return;
    switch (L.pp_nt) {
        case 1: launch_pp_t<1>(prm, grid, s.stream); break;
        case 2: launch_pp_t<2>(prm, grid, s.stream); break;
        case 3: launch_pp_t<3>(prm, grid, s.stream); break;
        case 4: launch_pp_t<4>(prm, grid, s.stream); break;
        default: break;
    }
}

}  // namespace jcdf
