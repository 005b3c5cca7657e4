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
#include <utility>
#include <vector>

#include "dmma_common.cuh"
#include "jcdf_internal.h"

namespace jcdf {

constexpr int PP_QC = 128;       // auxiliary rows per chunk (CTA tile M)
constexpr int PP_KS = 32;        // partner rows per pipeline stage
constexpr int PP_NSTAGE = 3;
constexpr int PP_CONSUMERS = 8;  // 2 (M halves of 64 rows) x 4 (N quarters of 8 * NT orbitals)
constexpr int PP_THREADS = (PP_CONSUMERS + 1) * 32;
// Warp-specialised register split (setmaxnreg): the CTA is launched with 12 warps = 3 warpgroups at 168 registers per thread;
// the producer's warpgroup (warp 8 + three warps that exit at once) gives registers back (-> 40) and the two consumer
// warpgroups take them (-> 232), so ptxas can keep a whole k-step of fragments in flight (it uses 186) instead of loading
// every a fragment right before its first DMMA.  128 x 40 + 256 x 232 = 64512 <= 65536.
constexpr int PP_THREADS_BIG = (PP_CONSUMERS + 4) * 32;
// Layout 2 (two CTAs per SM): one consumer warpgroup (2 x 2 warps) + the producer's warpgroup = 256 threads, launched at the
// 128-register cap of two resident CTAs; the producer's warpgroup drops to 40, the consumers rise to 216
// (2 x (128 x 40 + 128 x 216) = 65536); two stages of 49 KiB per CTA (4 in flight per SM).
constexpr int PP2_CONSUMERS = 4;
constexpr int PP2_THREADS = (PP2_CONSUMERS + 4) * 32;
constexpr int PP2_NSTAGE = 2;
__host__ __device__ constexpr int pp_layout_orbitals(int layout, int nt) { return (layout == 1 ? 8 : layout == 2 ? 16 : 32) * nt; }
__host__ __device__ constexpr int pp_layout_slots(int layout) { return layout == 1 ? 1 : layout == 2 ? 2 : 4; }

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
        JC_CHECK(r0 + r >= rows || ((int64_t)ch * rows + r0 + r) * PP_QC + (m ^ ((r & 3) << 2)) < (int64_t)gridDim.y * rows * PP_QC,
                 "pack: write outside Bc");
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
    int ldci, n_fn, n_nt, chunks, q_len, npt64, occ_w;
    // items of CTA c: item_rec[item_off[c] .. item_off[c + 1])  (cost-balanced on the host).  One 16-byte record per item,
    // { p, chunk | orbital tile << 24, first partner row of p, padded partner count of p }: no divisions and no dependent
    // loads at an item start
    const int4* item_rec;
    const int32_t* item_off;
    double* vitem;  // [CTA][chunk][warp column][128] partial Coulomb V of the CTA's items of a chunk (nullptr: not wanted)
    int64_t w_elems, bc_elems, ci_elems, vitem_elems;  // buffer extents in doubles (read by the bounds-checked build only)
};

// One pipeline stage (nks k-steps of 4 partner rows) of a warp block of MTV x NTV DMMA tiles inside the warp's 8 x NT
// accumulator block.  MTV / NTV are compile-time: a predicated-off DMMA still takes its slot in the tensor pipe, so the
// ragged last chunk and the ragged last orbital tile get their own instantiations instead of predicates.
template <int MTW, int NT, int LDC, int MTV, int NTV>
__device__ __forceinline__ void pp_slab(double (&acc)[MTW][NT][2], const double* __restrict__ sA,
                                        const double* __restrict__ sC, int m0, int n0, int g, int t, int nks) {
    // (m0 + 8 jm + g) ^ (t << 2): the swizzle touches bits 2-3, 8 jm bit 3 and up -> one base for even jm, one for odd jm
    const double* pa0 = sA + t * PP_QC + ((m0 + g) ^ (t << 2));
    const double* pa1 = sA + t * PP_QC + ((m0 + 8 + g) ^ (t << 2)) - 8;
    const double* pc = sC + t * LDC + n0 + g;
#pragma unroll 2
    for (int ks = 0; ks < nks; ks++) {
        double a[MTV], b[NTV];
#pragma unroll
        for (int jm = 0; jm < MTV; jm++) a[jm] = ((jm & 1) ? pa1 : pa0)[ks * 4 * PP_QC + 8 * jm];
#pragma unroll
        for (int jn = 0; jn < NTV; jn++) b[jn] = pc[ks * 4 * LDC + 8 * jn];
#pragma unroll
        for (int jm = 0; jm < MTV; jm++)
#pragma unroll
            for (int jn = 0; jn < NTV; jn++) dmma(acc[jm][jn][0], acc[jm][jn][1], a[jm], b[jn]);
    }
}

template <int MTW, int NT, int LDC, int NTV>
__device__ __forceinline__ void pp_slab_m(double (&acc)[MTW][NT][2], const double* sA, const double* sC, int m0, int n0,
                                          int g, int t, int nks, int mtv) {
    switch (mtv) {
        case 1: pp_slab<MTW, NT, LDC, 1, NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 2: pp_slab<MTW, NT, LDC, (MTW >= 2 ? 2 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 3: if (MTW >= 3) pp_slab<MTW, NT, LDC, (MTW >= 3 ? 3 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 4: if (MTW >= 4) pp_slab<MTW, NT, LDC, (MTW >= 4 ? 4 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 5: if (MTW >= 5) pp_slab<MTW, NT, LDC, (MTW >= 5 ? 5 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 6: if (MTW >= 6) pp_slab<MTW, NT, LDC, (MTW >= 6 ? 6 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 7: if (MTW >= 7) pp_slab<MTW, NT, LDC, (MTW >= 7 ? 7 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        case 8: if (MTW >= 8) pp_slab<MTW, NT, LDC, (MTW >= 8 ? 8 : 1), NTV>(acc, sA, sC, m0, n0, g, t, nks); break;
        default: break;
    }
}

template <int MTW, int NT, int LDC>
__device__ __forceinline__ void pp_slab_mn(double (&acc)[MTW][NT][2], const double* sA, const double* sC, int m0, int n0,
                                           int g, int t, int nks, int mtv, int ntv) {
    if (ntv == NT) pp_slab_m<MTW, NT, LDC, NT>(acc, sA, sC, m0, n0, g, t, nks, mtv);
    else if (NT > 1 && ntv == 1) pp_slab_m<MTW, NT, LDC, 1>(acc, sA, sC, m0, n0, g, t, nks, mtv);
    else if (NT > 2 && ntv == 2) pp_slab_m<MTW, NT, LDC, (NT > 2 ? 2 : 1)>(acc, sA, sC, m0, n0, g, t, nks, mtv);
    else if (NT > 3 && ntv == 3) pp_slab_m<MTW, NT, LDC, (NT > 3 ? 3 : 1)>(acc, sA, sC, m0, n0, g, t, nks, mtv);
    else if (ntv > 0) pp_slab_m<MTW, NT, LDC, NT>(acc, sA, sC, m0, n0, g, t, nks, mtv);  // layout 1 tail: whole width (zero C columns)
}

// One work item = (function p, chunk of 128 auxiliary rows, tile of NI = 32 NT occupied orbitals):
//     out[Q, i] = sum_k Bc[chunk][off_p + k][Q] * Ci[q_k(p)][i]        (M = Q, N = i, K = padded partner count of p)
// Warp 8 is the producer: lane 0 issues the A copy, lane j the copy of partner row j; warps 0-7 own 64 x (8 NT) blocks.
// LAYOUT 0: consumer warps 2 (M halves of 64 rows) x 4 (N quarters of 8 NT orbitals), NI = 32 NT — the least shared
//           memory traffic per DMMA, but the orbital count is padded to a multiple of 32 per item;
// LAYOUT 1: consumer warps 8 x 1, every warp 16 rows x ALL NI = 8 NT orbitals (NT <= 12) — no padding beyond the 8-wide
//           DMMA tile and the same DMMA count on every sub-partition; chosen when it executes fewer tiles (occ = 81:
//           11 tiles instead of 12 on the busiest sub-partition).
// WITH_V: the epilogue also leaves the item's share of the Coulomb vector V (a separate instantiation: the block costs
//         registers the NT = 4 kernel does not have to spare).
// LAYOUT 2: TWO CTAs per SM, each 2 x 2 consumer warps of 64 x (8 NT) blocks, NI = 16 NT, two stages per CTA: the CTAs of
//           an SM run out of phase, so the item epilogue (W stores, V dot products) and the pipeline refill of one sit under
//           the DMMAs of the other; costs twice the passes over Bc of layout 0 (L2 hits: the items of a (p, chunk) are adjacent).
template <int NT, int LAYOUT, bool WITH_V, bool BIGREG>
// 9 warps are allocated like 12 (register file granularity of 4 warps): 168 registers per thread is the cap at launch
__global__ void __launch_bounds__(LAYOUT == 2 ? PP2_THREADS : BIGREG ? PP_THREADS_BIG : PP_THREADS, LAYOUT == 2 ? 2 : 1)
k_pp_half_transform(PPParams prm) {
    static_assert(LAYOUT != 2 || BIGREG, "layout 2 exists with the register split only");
    constexpr int MTW = LAYOUT == 1 ? 2 : 8;        // 8-row DMMA tiles per warp
    constexpr int NI = pp_layout_orbitals(LAYOUT, NT);   // orbitals per item
    constexpr int LDC = NI + 4;  // = 4 (or 12) mod 16 doubles: the b fragment rows t = 0..3 fall on distinct bank quads
    constexpr int STAGE_DOUBLES = PP_KS * PP_QC + PP_KS * LDC;
    constexpr int NSTG = LAYOUT == 2 ? PP2_NSTAGE : PP_NSTAGE;
    constexpr int NCONS = LAYOUT == 2 ? PP2_CONSUMERS : PP_CONSUMERS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stages = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + (size_t)NSTG * STAGE_DOUBLES);
    uint64_t* empty = full + NSTG;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTG; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NCONS);
        }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    const int chunks = prm.chunks, n_nt = prm.n_nt;

    if (warp >= NCONS) {
        if constexpr (BIGREG) {
            asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");  // all four warps of the producer's warpgroup
            if (warp > NCONS) return;
        }
        // ------------------------------------------------ producer warp ------------------------------------------------
        PipeState pp;
        for (int j = prm.item_off[blockIdx.x]; j < prm.item_off[blockIdx.x + 1]; j++) {
            const int4 rec = prm.item_rec[j];
            const int ch = rec.y & 0xFFFFFF, nt = rec.y >> 24;
            const int r_begin = rec.z;
            const int k4 = rec.w;
            const double* Abase = prm.Bc + ((int64_t)ch * prm.rows + r_begin) * PP_QC;
            const double* Cbase = prm.Ci + (int64_t)nt * NI;
            for (int k0 = 0; k0 < k4; k0 += PP_KS) {
                const int nr = min(PP_KS, k4 - k0);
                const int qrow = (lane < nr) ? prm.q[r_begin + k0 + lane] : 0;
                mbar_wait(&empty[pp.stage], pp.phase ^ 1u);
                double* st = stages + (size_t)pp.stage * STAGE_DOUBLES;
                if (lane == 0) {
                    JC_CHECK(Abase + (int64_t)k0 * PP_QC >= prm.Bc && Abase + (int64_t)(k0 + nr) * PP_QC <= prm.Bc + prm.bc_elems,
                             "half-transform: A copy outside Bc");
                    mbar_arrive_expect_tx(&full[pp.stage], (uint32_t)nr * (uint32_t)((PP_QC + NI) * 8));
                    bulk_g2s(st, Abase + (int64_t)k0 * PP_QC, (uint32_t)nr * PP_QC * 8u, &full[pp.stage]);
                }
                __syncwarp();
                if (lane < nr)
                    JC_CHECK(qrow >= 0 && Cbase + (int64_t)qrow * prm.ldci + NI <= prm.Ci + prm.ci_elems, "half-transform: C row outside Ci");
                if (lane < nr)
                    bulk_g2s(st + PP_KS * PP_QC + lane * LDC, Cbase + (int64_t)qrow * prm.ldci, NI * 8u, &full[pp.stage]);
                pp.advance<NSTG>();
            }
        }
        return;
    }

    // --------------------------------------------------- consumers ---------------------------------------------------
    if constexpr (LAYOUT == 2) asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    else if constexpr (BIGREG) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    PipeState ps;
    const int g = lane >> 2, t = lane & 3;
    const int m0 = LAYOUT == 1 ? warp * 16 : LAYOUT == 2 ? (warp >> 1) * 64 : (warp >> 2) * 64;
    const int n0 = LAYOUT == 1 ? 0 : LAYOUT == 2 ? (warp & 1) * 8 * NT : (warp & 3) * 8 * NT;
    // with the register split the record of the NEXT item is fetched while the current one computes (four registers the
    // 168-register variant does not have: it reads its record at the item start)
    const int j_end = prm.item_off[blockIdx.x + 1];
    int4 rec_next = make_int4(0, 0, 0, 0);
    if constexpr (BIGREG) {
        if (prm.item_off[blockIdx.x] < j_end) rec_next = prm.item_rec[prm.item_off[blockIdx.x]];
    }
    for (int j = prm.item_off[blockIdx.x]; j < j_end; j++) {
        int4 rec;
        if constexpr (BIGREG) {
            rec = rec_next;
            if (j + 1 < j_end) rec_next = prm.item_rec[j + 1];
        } else {
            rec = prm.item_rec[j];
        }
        const int p = rec.x, ch = rec.y & 0xFFFFFF, nt = rec.y >> 24;
        const int k4 = rec.w;
        const int rows_valid = min(PP_QC, prm.q_len - ch * PP_QC);
        const int mt_valid = max(0, min(MTW, (rows_valid - m0 + 7) >> 3));
        const int nt_valid = max(0, min(NT, (prm.occ_w - nt * NI - n0 + 7) >> 3));
        const bool any = mt_valid > 0 && nt_valid > 0;
        const bool whole = mt_valid == MTW && nt_valid == NT;
        double acc[MTW][NT][2];
#pragma unroll
        for (int jm = 0; jm < MTW; jm++)
#pragma unroll
            for (int jn = 0; jn < NT; jn++) acc[jm][jn][0] = acc[jm][jn][1] = 0.0;
        if constexpr (WITH_V) {  // the epilogue reads the warp's 8 NT orbitals of row p of C: have them in L1 by then
            if (lane < NT) asm volatile("prefetch.global.L1 [%0];" ::"l"(prm.Ci + (int64_t)p * prm.ldci + nt * NI + n0 + 8 * lane));
        }

        for (int k0 = 0; k0 < k4; k0 += PP_KS) {
            const int nks = min(PP_KS, k4 - k0) >> 2;
            mbar_wait(&full[ps.stage], ps.phase);
            const double* sA = stages + (size_t)ps.stage * STAGE_DOUBLES;
            const double* sC = sA + PP_KS * PP_QC;
            if (whole) pp_slab<MTW, NT, LDC, MTW, NT>(acc, sA, sC, m0, n0, g, t, nks);
            else if (any) pp_slab_mn<MTW, NT, LDC>(acc, sA, sC, m0, n0, g, t, nks, mt_valid, nt_valid);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[ps.stage]);
            ps.advance<NSTG>();
        }
        if (!any) continue;
        // epilogue: the thread's (i, i + 1) pair is adjacent in the k-fastest W tile -> 16-byte stores, 4 lanes = 64 bytes.
        // 32-bit index arithmetic up to the tile number (k rows and tiles are far below 2^31), one wide multiply per store.
        const int pr = p & 63;
        const int sw = (pr & 3) << 2;
        double* const wbase = prm.W + ((int64_t)(p >> 6) * 64 + pr) * 32;
        const int tile_stride = prm.npt64;  // W tiles per k chunk
        const int i0 = nt * NI + n0 + 2 * t;
        const int q0 = ch * PP_QC + m0 + g;
        // every row and orbital of the warp's block exists (a last 8-row tile may be partly beyond q_len: guarded stores)
        const bool all_in = (rows_valid - m0 >= 8 * MTW) && (nt * NI + n0 + 8 * NT <= prm.occ_w);
        // (Tried: leave the block in the accumulators and issue its 32 stores between the DMMAs of the next item's first k-step,
        // which then starts from C = 0 without a clearing pass — parity green, but 15.02 ms against 14.73 at w50: the stores
        // and their address arithmetic inside the DMMA stream cost more than the idle gap they remove.)
        if (BIGREG && all_in) {  // the common case: no bounds to check (a second copy of the store loop: only where registers allow)
#pragma unroll
            for (int jm = 0; jm < MTW; jm++) {
                const int krow0 = (q0 + 8 * jm) * prm.occ_w + i0;
#pragma unroll
                for (int jn = 0; jn < NT; jn++) {
                    const int krow = krow0 + 8 * jn;
                    double* dst = wbase + (int64_t)((krow >> 5) * tile_stride) * 2048 + ((krow & 31) ^ sw);
                    JC_CHECK(dst >= prm.W && dst + 2 <= prm.W + prm.w_elems, "half-transform: W store outside W");
                    *reinterpret_cast<double2*>(dst) = make_double2(acc[jm][jn][0], acc[jm][jn][1]);
                }
            }
        } else {
#pragma unroll
            for (int jm = 0; jm < MTW; jm++) {
                const int Qrow = q0 + 8 * jm;
                if (!all_in && Qrow >= prm.q_len) continue;
                const int krow0 = Qrow * prm.occ_w + i0;
#pragma unroll
                for (int jn = 0; jn < NT; jn++) {
                    if (!all_in && i0 + 8 * jn >= prm.occ_w) continue;  // i and occ_w are even: the pair is in or out together
                    const int krow = krow0 + 8 * jn;
                    double* dst = wbase + (int64_t)((krow >> 5) * tile_stride) * 2048 + ((krow & 31) ^ sw);
                    JC_CHECK(dst >= prm.W && dst + 2 <= prm.W + prm.w_elems, "half-transform: W store outside W");
                    *reinterpret_cast<double2*>(dst) = make_double2(acc[jm][jn][0], acc[jm][jn][1]);
                }
            }
        }
        // Coulomb V from the accumulators: D[p,q] = sum_i C[p,i] C[q,i], so
        //     V[Q] = sum_pq B[Q,p,q] D[p,q] = sum_p sum_i C[p,i] W[Q,i,p]
        // and the item's share is a dot product of each accumulator row with row p of C: MTW * NT * 2 DFMA per thread
        // and ITEM, no density pass and no first sweep over B.
        if constexpr (WITH_V) {
            // opaque copies: nothing of this block may be hoisted above the k loop, where every register is taken
            int pe = p, nte = nt, lne = lane;
            asm volatile("" : "+r"(pe), "+r"(nte), "+r"(lne));
            const int te = lne & 3, ge = lne >> 2;
            const double* cp = prm.Ci + (int64_t)pe * prm.ldci + nte * NI + n0 + 2 * te;
            // after the W stores only the accumulators are live: row p of C is loaded whole (one latency), then MTW
            // independent dot-product chains
            double2 c[NT];
#pragma unroll
            for (int jn = 0; jn < NT; jn++) c[jn] = *reinterpret_cast<const double2*>(cp + 8 * jn);
            double v[MTW];
#pragma unroll
            for (int jm = 0; jm < MTW; jm++) v[jm] = 0.0;
#pragma unroll
            for (int jn = 0; jn < NT; jn++)
#pragma unroll
                for (int jm = 0; jm < MTW; jm++) v[jm] = fma(acc[jm][jn][0], c[jn].x, fma(acc[jm][jn][1], c[jn].y, v[jm]));
#pragma unroll
            for (int jm = 0; jm < MTW; jm++) v[jm] += __shfl_xor_sync(0xffffffffu, v[jm], 1);
#pragma unroll
            for (int jm = 0; jm < MTW; jm++) v[jm] += __shfl_xor_sync(0xffffffffu, v[jm], 2);
            if (te == 0) {
                // The shares of all the CTA's items of one chunk add up in the warp's own slot [CTA][chunk][warp column][128]:
                // one thread per address and program order, so the fire-and-forget reductions are bit-reproducible.  (Rows
                // beyond q_len / orbitals beyond occ add exact zeros: zero rows of Bc, zero columns of C.  Summing in
                // shared memory and flushing once per chunk made ptxas spill accumulators inside the k loop.)
                int che = ch, we = warp;
                asm volatile("" : "+r"(che), "+r"(we));
                constexpr int NWC = pp_layout_slots(LAYOUT);
                double* vdst = prm.vitem + ((((int64_t)blockIdx.x * chunks + che) * NWC + (we & (NWC - 1))) * PP_QC) + m0 + ge;
                JC_CHECK(vdst >= prm.vitem && vdst + 8 * (MTW - 1) < prm.vitem + prm.vitem_elems, "half-transform: V slot outside Vitem");
#pragma unroll
                for (int jm = 0; jm < MTW; jm++) asm volatile("red.global.add.f64 [%0], %1;" ::"l"(vdst + 8 * jm), "d"(v[jm]) : "memory");
            }
        }
    }
}

// =====================================================================================================================
// Coulomb over the partner-compacted copy: the rows with q >= p (a contiguous tail of every partner list, like the column
// tails calculate_coulomb_screened walks, ScreenedDF.jl:326-358) are streamed twice, 1 KiB per row and chunk:
//     V[Q]  = sum_rows Bc[chunk][row][Q] * dvec[row]          dvec = D[p,q] (x 2 off the diagonal, :305-316)
//     J[row] = 2 * sum_Q Bc[chunk][row][Q] * V[Q]
// Every partial sum is written to its own slot and reduced in a fixed order (no atomics: bit-reproducible).
// =====================================================================================================================
constexpr int PPJ_SLAB = 512;  // rows per block

__device__ __forceinline__ double2 ld_stream2(const double* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

// one warp per row: dvec[row] = w * sum_i Ci[p][i] * Ci[q][i]
__global__ void __launch_bounds__(256) k_pp_density(const double* __restrict__ Ci, int ldci, int occ,
                                                     const int32_t* __restrict__ rp, const int32_t* __restrict__ rq,
                                                     const int32_t* __restrict__ jdst, double* __restrict__ dvec,
                                                     int64_t rows) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp0; r < rows; r += nwarps) {
        if (jdst[r] < 0) {
            if (lane == 0) dvec[r] = 0.0;
            continue;
        }
        const int p = rp[r], q = rq[r];
        const double* cp = Ci + (int64_t)p * ldci;
        const double* cq = Ci + (int64_t)q * ldci;
        double acc = 0.0;
        for (int i = lane; i < occ; i += 32) acc = fma(cp[i], cq[i], acc);
        acc = warp_sum(acc);
        if (lane == 0) dvec[r] = (p == q) ? acc : 2.0 * acc;
    }
}

__global__ void __launch_bounds__(256) k_pp_coulomb_v(const double* __restrict__ Bc, const double* __restrict__ dvec,
                                                       double* __restrict__ Vpart, int64_t rows, int qpad) {
    __shared__ double part[8][PP_QC];
    const int ch = blockIdx.y;
    const int64_t r0 = (int64_t)blockIdx.x * PPJ_SLAB;
    const int64_t r1 = min(rows, r0 + PPJ_SLAB);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const double* base = Bc + (int64_t)ch * rows * PP_QC + lane * 4;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int64_t r = r0 + w; r < r1; r += 32) {  // rows r, r+8, r+16, r+24: all = w (mod 8), so (row & 3) = w & 3
        double d[4];
        double2 x[4], y[4];
#pragma unroll
        for (int u = 0; u < 4; u++) d[u] = (r + 8 * u < r1) ? dvec[r + 8 * u] : 0.0;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            x[u] = y[u] = make_double2(0.0, 0.0);
            if (d[u] != 0.0) {
                const double* src = base + (r + 8 * u) * PP_QC;
                x[u] = ld_stream2(src);
                y[u] = ld_stream2(src + 2);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            acc[0] = fma(d[u], x[u].x, acc[0]);
            acc[1] = fma(d[u], x[u].y, acc[1]);
            acc[2] = fma(d[u], y[u].x, acc[2]);
            acc[3] = fma(d[u], y[u].y, acc[3]);
        }
    }
    const int lg = (lane * 4) ^ ((w & 3) << 2);  // un-swizzle: the auxiliary rows this thread's four columns hold
#pragma unroll
    for (int j = 0; j < 4; j++) part[w][lg + j] = acc[j];
    __syncthreads();
    if (threadIdx.x < PP_QC) {
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 8; k++) v += part[k][threadIdx.x];
        Vpart[(int64_t)blockIdx.x * qpad + ch * PP_QC + threadIdx.x] = v;
    }
}

__global__ void k_pp_vreduce(const double* __restrict__ Vpart, double* __restrict__ V, int n_slab, int qpad, int q_len) {
    const int Q = blockIdx.x * blockDim.x + threadIdx.x;
    if (Q >= qpad) return;
    double v = 0.0;
    for (int s = 0; s < n_slab; s++) v += Vpart[(int64_t)s * qpad + Q];
    if (Q < q_len) V[Q] = v;
}

// V[Q] = sum of the per-(CTA, chunk) partials the half-transform left, in a fixed order.  One block per 32 auxiliary
// functions: 32 groups of 32 threads take the CTAs k (mod 32) (n_fn = number of CTAs), then one sum over the 32 groups.
__global__ void __launch_bounds__(1024) k_pp_vitem_reduce(const double* __restrict__ vitem, double* __restrict__ V,
                                                          int n_fn, int chunks, int slots, int q_len) {
    __shared__ double part[32][33];
    const int ch = blockIdx.x >> 2, c = (blockIdx.x & 3) * 32 + (threadIdx.x & 31), k = threadIdx.x >> 5;
    double a = 0.0;
    for (int p = k; p < n_fn; p += 32) {
        const double* src = vitem + ((int64_t)p * chunks + ch) * slots * PP_QC + c;
        double b = 0.0;
        for (int sl = 0; sl < slots; sl++) b += __ldcs(src + sl * PP_QC);
        a += b;
    }
    part[k][threadIdx.x & 31] = a;
    __syncthreads();
    if (k == 0) {
        double v = 0.0;
#pragma unroll
        for (int kk = 0; kk < 32; kk++) v += part[kk][threadIdx.x];
        const int Q = ch * PP_QC + c;
        if (Q < q_len) V[Q] = v;
    }
}

__global__ void __launch_bounds__(256) k_pp_coulomb_j(const double* __restrict__ Bc, const double* __restrict__ V,
                                                       const int32_t* __restrict__ jdst, double* __restrict__ Jpart,
                                                       int64_t rows, int q_len) {
    const int ch = blockIdx.y;
    const int64_t r0 = (int64_t)blockIdx.x * PPJ_SLAB;
    const int64_t r1 = min(rows, r0 + PPJ_SLAB);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lg = (lane * 4) ^ ((w & 3) << 2);
    double v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int Q = ch * PP_QC + lg + j;
        v[j] = Q < q_len ? V[Q] : 0.0;
    }
    const double* base = Bc + (int64_t)ch * rows * PP_QC + lane * 4;
    double* out = Jpart + (int64_t)ch * rows;
    for (int64_t r = r0 + w; r < r1; r += 32) {
        bool on[4];
        double2 x[4], y[4];
#pragma unroll
        for (int u = 0; u < 4; u++) on[u] = (r + 8 * u < r1) && jdst[r + 8 * u] >= 0;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            x[u] = y[u] = make_double2(0.0, 0.0);
            if (on[u]) {
                const double* src = base + (r + 8 * u) * PP_QC;
                x[u] = ld_stream2(src);
                y[u] = ld_stream2(src + 2);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (!on[u]) continue;  // warp-uniform
            double dot = fma(x[u].x, v[0], fma(x[u].y, v[1], fma(y[u].x, v[2], y[u].y * v[3])));
            dot = warp_sum(dot);
            if (lane == 0) out[r + 8 * u] = dot;
        }
    }
}

// Jt[tile-space offset of (hi, lo)] = 2 * sum_chunks Jpart[chunk][row]   (the epilogue reads Jt as before)
__global__ void k_pp_jreduce(const double* __restrict__ Jpart, const int32_t* __restrict__ jdst, double* __restrict__ Jt,
                             int64_t rows, int chunks) {
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const int32_t dst = jdst[r];
    if (dst < 0) return;
    double s = 0.0;
    for (int c = 0; c < chunks; c++) s += Jpart[(int64_t)c * rows + r];
    Jt[dst] = 2.0 * s;
}

// ---------------------------------------------------------------------------------------------------------------------
// Single HBM pass: one cooperative persistent grid.  CTA c owns a fixed range of the q >= p rows and walks the auxiliary
// index in quarters of a chunk (32 functions = 256 B of every row, 26 MB per quarter at w50):
//   pass 1 of quarter s      partial V over the CTA's rows -> Vp[s][c][32] in L2, one release-increment of flag[s]
//   pass 2 of quarter s-LAG  once all CTAs have contributed: V = sum_c Vp (fixed order), the SAME 256-byte pieces are
//                            read again — now L2 hits, LAG + 1 quarters stay resident — and J[row] += B . V accumulates
//                            in shared memory; J leaves the SM once, at the end.
// HBM sees every q >= p element of Bc once; no atomics on data, every sum in a fixed order (bit-reproducible).
// Measured (real w50, 150 quarters of 26 MB): 2.1 ms against 1.3 ms for the two streaming passes — every quarter costs a
// grid-wide flag round plus a 148-way gather of partial V (about 14 us of serial latency for 4 us of traffic), and the
// L2 cannot hold a larger step between the passes.  Kept as an option (coulomb_pp_fused = 1), off by default.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int PJF_THREADS = 512;
constexpr int PJF_LAG = 2;
constexpr int PJF_UNROLL = 8;    // rows in flight per thread (16 bytes each)
constexpr int PJF_MAX_ROWS = 8192;  // q >= p rows per CTA that fit the shared-memory row cache

struct PJFParams {
    const double* Bc;
    const double* dvec;
    const int32_t* trirow;  // [n_tri] indices of the q >= p rows, ascending
    const int32_t* jdst;
    double* Vp;             // [quarters][gridDim.x][32]
    unsigned int* flag;     // [quarters], zeroed before launch
    double* Jt;
    double* V;
    int64_t rows;
    int n_tri, chunks, q_len;
};

__device__ __forceinline__ unsigned int pjf_ld_acquire(const unsigned int* p) {
    JC_FUZZ(7);
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void pjf_red_release(unsigned int* p) {
    JC_FUZZ(8);
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ double2 pjf_ld2(const double* p) {  // plain (L2-allocating) 16-byte load
    double2 v;
    asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(PJF_THREADS, 1) k_pp_coulomb_fused(PJFParams prm) {
    extern __shared__ __align__(16) unsigned char pjf_smem[];
    const int G = gridDim.x, c = blockIdx.x, tid = threadIdx.x;
    const int t0 = (int)((int64_t)prm.n_tri * c / G), t1 = (int)((int64_t)prm.n_tri * (c + 1) / G);
    const int nloc = t1 - t0;
    double* Jacc = reinterpret_cast<double*>(pjf_smem);             // [nloc]
    double* sd = Jacc + nloc;                                       // [nloc] weighted density of the row
    double* red = sd + nloc;                                        // [32][32]
    double* sV = red + 32 * 32;                                     // [32]
    int32_t* soff = reinterpret_cast<int32_t*>(sV + 32);            // [nloc] row index (its low 2 bits = swizzle phase)
    for (int i = tid; i < nloc; i += PJF_THREADS) {
        const int r = prm.trirow[t0 + i];
        soff[i] = r;
        sd[i] = prm.dvec[r];
        Jacc[i] = 0.0;
    }
    __syncthreads();
    // half-warp = one row piece of 16 x 16 bytes; l = logical pair of auxiliary rows; warp wp handles rows 2 wp + h (+ 32 k)
    const int hw = tid >> 4, l = tid & 15, wp = tid >> 5, h = (tid >> 4) & 1;
    const int nq = prm.chunks * 4;
    for (int step = 0; step < nq + PJF_LAG; step++) {
        if (step < nq) {
            const double* base = prm.Bc + (int64_t)(step >> 2) * prm.rows * PP_QC + (step & 3) * 32;
            double ax = 0.0, ay = 0.0;
            int ib = 2 * wp;  // rows ib + h: the trip count is uniform per warp (pass 2 shuffles inside the same loops)
            for (; ib + 32 * (PJF_UNROLL - 1) < nloc; ib += 32 * PJF_UNROLL) {  // PJF_UNROLL rows in flight per thread
                double2 v[PJF_UNROLL];
                double d[PJF_UNROLL];
#pragma unroll
                for (int u = 0; u < PJF_UNROLL; u++) {
                    const int i = ib + 32 * u + h;
                    const bool ok = i < nloc;
                    const int r = ok ? soff[i] : 0;
                    d[u] = ok ? sd[i] : 0.0;
                    v[u] = ok ? pjf_ld2(base + (int64_t)r * PP_QC + ((l ^ ((r & 3) << 1)) << 1)) : make_double2(0.0, 0.0);
                }
#pragma unroll
                for (int u = 0; u < PJF_UNROLL; u++) {
                    ax = fma(d[u], v[u].x, ax);
                    ay = fma(d[u], v[u].y, ay);
                }
            }
            for (; ib < nloc; ib += 32) {
                const int i = ib + h;
                if (i < nloc) {
                    const int r = soff[i];
                    const double2 v = pjf_ld2(base + (int64_t)r * PP_QC + ((l ^ ((r & 3) << 1)) << 1));
                    ax = fma(sd[i], v.x, ax);
                    ay = fma(sd[i], v.y, ay);
                }
            }
            red[hw * 32 + 2 * l] = ax;
            red[hw * 32 + 2 * l + 1] = ay;
            __syncthreads();
            if (tid < 32) {
                double v = 0.0;
#pragma unroll 8
                for (int k = 0; k < 32; k++) v += red[k * 32 + tid];
                prm.Vp[((int64_t)step * G + c) * 32 + tid] = v;
                __threadfence();
            }
            __syncthreads();
            if (tid == 0) pjf_red_release(prm.flag + step);
        }
        const int j = step - PJF_LAG;
        if (j >= 0) {
            if (tid == 0)
                while (pjf_ld_acquire(prm.flag + j) < (unsigned int)G) {}
            __syncthreads();
            {   // V of quarter j: 16 groups of threads sum interleaved CTAs, then one fixed-order sum of the 16 partials
                const int k = tid >> 5, m = tid & 31;
                double v = 0.0;
                for (int cc = k; cc < G; cc += 16) v += __ldcg(prm.Vp + ((int64_t)j * G + cc) * 32 + m);
                red[k * 32 + m] = v;
            }
            __syncthreads();
            if (tid < 32) {
                double v = 0.0;
#pragma unroll
                for (int k = 0; k < 16; k++) v += red[k * 32 + tid];
                sV[tid] = v;
                const int Q = (j >> 2) * PP_QC + (j & 3) * 32 + tid;
                if (c == 0 && Q < prm.q_len) prm.V[Q] = v;
            }
            __syncthreads();
            const double vx = sV[2 * l], vy = sV[2 * l + 1];
            const double* base = prm.Bc + (int64_t)(j >> 2) * prm.rows * PP_QC + (j & 3) * 32;
            int ib = 2 * wp;
            for (; ib + 32 * (PJF_UNROLL - 1) < nloc; ib += 32 * PJF_UNROLL) {
                double2 v[PJF_UNROLL];
#pragma unroll
                for (int u = 0; u < PJF_UNROLL; u++) {
                    const int i = ib + 32 * u + h;
                    const int r = i < nloc ? soff[i] : 0;
                    v[u] = i < nloc ? pjf_ld2(base + (int64_t)r * PP_QC + ((l ^ ((r & 3) << 1)) << 1)) : make_double2(0.0, 0.0);
                }
#pragma unroll
                for (int u = 0; u < PJF_UNROLL; u++) {
                    const int i = ib + 32 * u + h;
                    double dot = fma(v[u].x, vx, v[u].y * vy);
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
                    if (l == 0 && i < nloc) Jacc[i] += dot;
                }
            }
            for (; ib < nloc; ib += 32) {
                const int i = ib + h;
                const int r = i < nloc ? soff[i] : 0;
                const double2 v = i < nloc ? pjf_ld2(base + (int64_t)r * PP_QC + ((l ^ ((r & 3) << 1)) << 1)) : make_double2(0.0, 0.0);
                double dot = fma(v.x, vx, v.y * vy);
#pragma unroll
                for (int o = 8; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
                if (l == 0 && i < nloc) Jacc[i] += dot;
            }
            __syncthreads();  // red / sV are reused by the next step
        }
    }
    for (int i = tid; i < nloc; i += PJF_THREADS) prm.Jt[prm.jdst[soff[i]]] = 2.0 * Jacc[i];
}

static size_t pjf_smem_bytes(int nloc) { return (size_t)nloc * (8 + 8 + 4) + (32 * 32 + 32) * 8 + 64; }

// false: the shard does not fit (rows per CTA) or the cooperative launch failed -> the caller runs the two passes
bool launch_pp_coulomb_fused(const Layout& L, Shard& s, int sm_count, cudaStream_t st) {
    if (s.pp_chunks <= 0 || L.pp_n_tri <= 0 || s.ppj_flag == nullptr) return false;
    const int grid = sm_count;
    const int nloc_max = (int)((L.pp_n_tri + grid - 1) / grid) + 1;
    if (nloc_max > PJF_MAX_ROWS) return false;
    PJFParams prm;
    prm.Bc = s.Bc; prm.dvec = s.dvec; prm.trirow = s.d_pp_trirow; prm.jdst = s.d_pp_jdst;
    prm.Vp = s.ppj_vp; prm.flag = s.ppj_flag; prm.Jt = s.Jt; prm.V = s.V;
    prm.rows = L.pp_rows; prm.n_tri = (int)L.pp_n_tri; prm.chunks = s.pp_chunks; prm.q_len = (int)s.q_len;
    const size_t smem = pjf_smem_bytes(nloc_max);
    if (cudaFuncSetAttribute(k_pp_coulomb_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    cudaMemsetAsync(s.ppj_flag, 0, sizeof(unsigned int) * (size_t)s.pp_chunks * 4, st);
    void* args[] = {&prm};
    if (cudaLaunchCooperativeKernel((const void*)k_pp_coulomb_fused, dim3(grid), dim3(PJF_THREADS), args, smem, st) !=
        cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return true;
}

void launch_pp_density(const Layout& L, Shard& s) {
    const int ldci = L.pp_n_nt * pp_item_orbitals(L);
    const int blocks = (int)std::min<int64_t>((L.pp_rows + 7) / 8, 148 * 16);
    if (blocks <= 0) return;
    k_pp_density<<<blocks, 256, 0, s.stream>>>(s.Ci, ldci, (int)L.occ, s.d_pp_p, s.d_pp_q, s.d_pp_jdst, s.dvec, L.pp_rows);
}

// two streaming passes; v_in_ht: the half-transform already left the partial V sums, only their reduction remains
void launch_pp_coulomb(const Layout& L, Shard& s, cudaStream_t st, cudaEvent_t ev_v, bool v_in_ht) {
    if (s.pp_chunks <= 0 || L.pp_rows <= 0) return;
    const int n_slab = (int)((L.pp_rows + PPJ_SLAB - 1) / PPJ_SLAB);
    const int qpad = s.pp_chunks * PP_QC;
    dim3 grid((unsigned)n_slab, (unsigned)s.pp_chunks);
    if (v_in_ht) {
        k_pp_vitem_reduce<<<s.pp_chunks * 4, 1024, 0, st>>>(s.Vitem, s.V, s.pp_ctas, s.pp_chunks, pp_v_slots(L), (int)s.q_len);
    } else {
        k_pp_coulomb_v<<<grid, 256, 0, st>>>(s.Bc, s.dvec, s.VpartP, L.pp_rows, qpad);
        k_pp_vreduce<<<(qpad + 127) / 128, 128, 0, st>>>(s.VpartP, s.V, n_slab, qpad, (int)s.q_len);
    }
    if (ev_v) record_event(s, ev_v, st);
    k_pp_coulomb_j<<<grid, 256, 0, st>>>(s.Bc, s.V, s.d_pp_jdst, s.Jpart, L.pp_rows, (int)s.q_len);
    k_pp_jreduce<<<(unsigned)((L.pp_rows + 255) / 256), 256, 0, st>>>(s.Jpart, s.d_pp_jdst, s.Jt, L.pp_rows, s.pp_chunks);
}

int64_t pp_coulomb_slabs(const Layout& L) { return (L.pp_rows + PPJ_SLAB - 1) / PPJ_SLAB; }

static size_t pp_smem_bytes(int ni, int nstage) {
    return (size_t)nstage * (PP_KS * PP_QC + PP_KS * (ni + 4)) * 8 + 2 * nstage * 8 + 16;
}

template <int NT, int LAYOUT, bool WITH_V, bool BIGREG>
static void launch_pp_k(const PPParams& prm, int grid, size_t smem, cudaStream_t st) {
    cudaFuncSetAttribute(k_pp_half_transform<NT, LAYOUT, WITH_V, BIGREG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_pp_half_transform<NT, LAYOUT, WITH_V, BIGREG><<<grid, LAYOUT == 2 ? PP2_THREADS : BIGREG ? PP_THREADS_BIG : PP_THREADS, smem, st>>>(prm);
}

template <int NT, int LAYOUT>
static void launch_pp_t(const PPParams& prm, int grid, bool bigreg, cudaStream_t st) {
    const size_t smem = pp_smem_bytes(pp_layout_orbitals(LAYOUT, NT), LAYOUT == 2 ? PP2_NSTAGE : PP_NSTAGE);
    if (prm.vitem != nullptr) {
        cudaMemsetAsync(prm.vitem, 0, sizeof(double) * (size_t)grid * prm.chunks * pp_layout_slots(LAYOUT) * PP_QC, st);  // the slots are summed into
        if (bigreg || LAYOUT == 2) launch_pp_k<NT, LAYOUT, true, true>(prm, grid, smem, st);
        else launch_pp_k<NT, LAYOUT, true, LAYOUT == 2>(prm, grid, smem, st);
    } else {
        if (bigreg || LAYOUT == 2) launch_pp_k<NT, LAYOUT, false, true>(prm, grid, smem, st);
        else launch_pp_k<NT, LAYOUT, false, LAYOUT == 2>(prm, grid, smem, st);
    }
}

// Work lists of the per-function half-transform.  Items differ in cost — the partner count of p, the ragged last
// 128-row chunk, the ragged last orbital tile — and a static round-robin aliases with the chunk count (gly18 shard,
// 10 chunks of which the last holds 15 rows: CTAs of one parity got all the cheap items and the others 11 % more than
// the mean).  Items are dealt longest-first onto the least-loaded CTA from the DMMA count the busiest sub-partition
// issues; each CTA then walks its list in ascending (p-major) order so co-running CTAs share C rows in L2.
void plan_pp_items(const Layout& L, Shard& s, int sm_count, bool p_major, std::vector<int32_t>& list, std::vector<int32_t>& off) {
    // list: four int32 per item (the int4 records the kernel reads)
    const int chunks = s.pp_chunks, n_nt = L.pp_n_nt, NT = L.pp_nt;
    const int64_t n_items = (int64_t)L.N * n_nt * chunks;
    const int n_ctas = (int)std::max<int64_t>(1, std::min<int64_t>(n_items, (int64_t)sm_count * pp_ctas_per_sm(L)));
    s.pp_ctas = n_items > 0 ? n_ctas : 0;
    list.clear();
    off.assign((size_t)n_ctas + 1, 0);
    if (n_items <= 0 || n_items >= ((int64_t)1 << 31)) { s.pp_ctas = 0; return; }
    // row units of a chunk (both 64-row halves of the CTA tile) and orbital tiles of the busiest warp column
    std::vector<int> m_units((size_t)chunks), n_units((size_t)n_nt);
    const int NI = pp_item_orbitals(L);
    for (int ch = 0; ch < chunks; ch++) {
        const int rows = (int)std::min<int64_t>(PP_QC, s.q_len - (int64_t)ch * PP_QC);
        if (L.pp_layout != 1) {
            const int m0 = std::max(0, std::min(8, (rows + 7) >> 3)), m1 = std::max(0, std::min(8, (rows - 64 + 7) >> 3));
            m_units[(size_t)ch] = m0 + m1;
        } else {  // warps w and w + 4 share a sub-partition, 16 rows each
            int best = 0;
            for (int w = 0; w < 4; w++) {
                const int a = std::max(0, std::min(2, (rows - 16 * w + 7) >> 3));
                const int b = std::max(0, std::min(2, (rows - 16 * (w + 4) + 7) >> 3));
                best = std::max(best, a + b);
            }
            m_units[(size_t)ch] = best;
        }
    }
    for (int nt = 0; nt < n_nt; nt++) {
        int u = std::max(0, std::min(NT, (L.occ_w - nt * NI + 7) >> 3));
        if (L.pp_layout == 1 && u > 3 && u < NT) u = NT;  // the ragged last item of layout 1 runs the whole width
        n_units[(size_t)nt] = u;
    }
    std::vector<int64_t> cost((size_t)n_items);
    std::vector<int32_t> order((size_t)n_items);
    for (int64_t it = 0; it < n_items; it++) {
        const int ch = (int)(it % chunks);
        const int nt = (int)((it / chunks) % n_nt);
        const int p = (int)(it / chunks / n_nt);
        const int ksteps = (L.pp_off[(size_t)p + 1] - L.pp_off[(size_t)p]) / 4;
        // 16 pipe cycles per DMMA; ~1500 cycles of epilogue and ramp per item
        cost[(size_t)it] = (int64_t)ksteps * m_units[(size_t)ch] * n_units[(size_t)nt] * 16 + 1500;
        order[(size_t)it] = (int32_t)it;
    }
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return cost[(size_t)a] > cost[(size_t)b]; });
    // least-loaded CTA through a binary heap of (load, cta)
    std::vector<std::pair<int64_t, int>> heap;
    for (int c = 0; c < n_ctas; c++) heap.emplace_back(0, c);
    auto cmp = [](const std::pair<int64_t, int>& a, const std::pair<int64_t, int>& b) { return a > b; };
    std::make_heap(heap.begin(), heap.end(), cmp);
    std::vector<std::vector<int32_t>> bins((size_t)n_ctas);
    for (int32_t it : order) {
        std::pop_heap(heap.begin(), heap.end(), cmp);
        auto& top = heap.back();
        top.first += cost[(size_t)it];
        bins[(size_t)top.second].push_back(it);
        std::push_heap(heap.begin(), heap.end(), cmp);
    }
    for (int c = 0; c < n_ctas; c++) {
        // p-major (default: co-running CTAs work on the same functions and share their gathered rows of C in L2; w50 14.9 ms)
        // or chunk-major (the rows of Bc a CTA reads form ascending runs inside one chunk; 15.4 ms)
        std::sort(bins[(size_t)c].begin(), bins[(size_t)c].end(), [&](int32_t a, int32_t b) {
            const int ca = a % chunks, cb = b % chunks;
            return (!p_major && ca != cb) ? ca < cb : a < b;
        });
        off[(size_t)c] = (int32_t)(list.size() / 4);
        for (int32_t it : bins[(size_t)c]) {
            const int ch = it % chunks, nt = (it / chunks) % n_nt, p = it / chunks / n_nt;
            list.push_back(p);
            list.push_back(ch | (nt << 24));
            list.push_back(L.pp_off[(size_t)p]);
            list.push_back(L.pp_off[(size_t)p + 1] - L.pp_off[(size_t)p]);
        }
    }
    off[(size_t)n_ctas] = (int32_t)(list.size() / 4);
}

void launch_pp_pack(const Layout& L, Shard& s) {
    if (s.pp_chunks <= 0 || L.pp_rows <= 0) return;
    dim3 grid((unsigned)((L.pp_rows + 31) / 32), (unsigned)s.pp_chunks);
    k_pp_pack<<<grid, 256, 0, s.stream>>>(s.Bp, L.row_len, s.d_pp_src, s.Bc, L.pp_rows, (int)s.q_len);
}

void launch_pp_prep(const Layout& L, Shard& s, int64_t ldc) {
    const int ldci = L.pp_n_nt * pp_item_orbitals(L);
    const int64_t total = (int64_t)(L.n_pad + 1) * ldci;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    k_pp_prep<<<blocks, 256, 0, s.stream>>>(s.Craw, ldc, s.Ci, (int)L.N, (int)L.occ, L.n_pad + 1, ldci, s.d_inv);
}

void launch_pp_half_transform(const Layout& L, Shard& s, int sm_count, bool v_in_ht, bool bigreg) {
    PPParams prm;
    prm.Bc = s.Bc;
    prm.Ci = s.Ci;
    prm.W = s.W;
    prm.off = s.d_pp_off;
    prm.q = s.d_pp_q;
    prm.rows = L.pp_rows;
    prm.ldci = L.pp_n_nt * pp_item_orbitals(L);
    prm.n_fn = (int)L.N;
    prm.n_nt = L.pp_n_nt;
    prm.chunks = s.pp_chunks;
    prm.q_len = (int)s.q_len;
    prm.npt64 = L.npt64;
    prm.occ_w = L.occ_w;
    prm.item_rec = reinterpret_cast<const int4*>(s.d_ppitems);
    prm.item_off = s.d_ppoff;
    prm.vitem = v_in_ht ? s.Vitem : nullptr;
    prm.w_elems = (int64_t)s.xplan.chunks * L.npt64 * 2048;
    prm.bc_elems = (int64_t)s.pp_chunks * L.pp_rows * PP_QC;
    prm.ci_elems = (int64_t)(L.n_pad + 1) * prm.ldci;
    prm.vitem_elems = (int64_t)std::max(1, s.pp_ctas) * s.pp_chunks * pp_v_slots(L) * PP_QC;
    const int grid = s.pp_ctas;
    (void)sm_count;
    if (grid <= 0) return;
    if (L.pp_layout == 2) {
        switch (L.pp_nt) {
            case 1: launch_pp_t<1, 2>(prm, grid, bigreg, s.stream); break;
            case 2: launch_pp_t<2, 2>(prm, grid, bigreg, s.stream); break;
            case 3: launch_pp_t<3, 2>(prm, grid, bigreg, s.stream); break;
            case 4: launch_pp_t<4, 2>(prm, grid, bigreg, s.stream); break;
            default: break;
        }
    } else if (L.pp_layout == 0) {
        switch (L.pp_nt) {
            case 1: launch_pp_t<1, 0>(prm, grid, bigreg, s.stream); break;
            case 2: launch_pp_t<2, 0>(prm, grid, bigreg, s.stream); break;
            case 3: launch_pp_t<3, 0>(prm, grid, bigreg, s.stream); break;
            case 4: launch_pp_t<4, 0>(prm, grid, bigreg, s.stream); break;
            default: break;
        }
    } else {
        switch (L.pp_nt) {
            case 1: launch_pp_t<1, 1>(prm, grid, bigreg, s.stream); break;
            case 2: launch_pp_t<2, 1>(prm, grid, bigreg, s.stream); break;
            case 3: launch_pp_t<3, 1>(prm, grid, bigreg, s.stream); break;
            case 4: launch_pp_t<4, 1>(prm, grid, bigreg, s.stream); break;
            case 5: launch_pp_t<5, 1>(prm, grid, bigreg, s.stream); break;
            case 6: launch_pp_t<6, 1>(prm, grid, bigreg, s.stream); break;
            case 7: launch_pp_t<7, 1>(prm, grid, bigreg, s.stream); break;
            case 8: launch_pp_t<8, 1>(prm, grid, bigreg, s.stream); break;
            case 9: launch_pp_t<9, 1>(prm, grid, bigreg, s.stream); break;
            case 10: launch_pp_t<10, 1>(prm, grid, bigreg, s.stream); break;
            case 11: launch_pp_t<11, 1>(prm, grid, bigreg, s.stream); break;
            case 12: launch_pp_t<12, 1>(prm, grid, bigreg, s.stream); break;
            default: break;
        }
    }
}

JC_DEFINE_FUZZ_SETTER(set_fuzz_seed_pp)

}  // namespace jcdf
