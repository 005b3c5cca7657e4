// Exchange (K) path of the DF-RHF Fock build on sm_100a: FP64 tensor-core (DMMA m8n8k4) tiles fed by 1-D bulk
// async copies (TMA unit) through an mbarrier ring.
//
//   half-transform   W[(Q,i), p] = sum_q C_occ[q,i] * B[Q,p,q]        replaces calculate_W_screened
//                                                                      (ScreenedDF.jl:242-289; dense :216 of DensityFitting.jl)
//   exchange         K[p,q]     = sum_(Q,i) W[(Q,i),p] * W[(Q,i),q]    replaces calculate_K_lower_diagonal_block_no_screen
//                                                                      (ScreenedDF.jl:548-641; dense DensityFitting.jl:219)
//
// (the per-function half-transform used on sparse masks lives in kernels_pp.cu; W is stored as [chunk of 32 k][64-wide
// p tile][64 p][32 k], k = Q_local * occ_w + i, so both half-transforms write and the exchange reads the same tiles)
// Both kernels are persistent (one CTA per SM; the exchange walks cost-balanced work lists, the tile half-transform a
// static round-robin); 8 warps issue DMMA and thread 0 is
// also the copy producer, running NSTAGE - 1 slabs ahead through a full/empty mbarrier ring.  All operands are stored
// in HBM pre-swizzled (dmma_common.cuh: swz) so a dense bulk copy lands bank-conflict-free in shared memory.
#include <algorithm>

#include "dmma_common.cuh"
#include "jcdf_internal.h"

namespace jcdf {

constexpr int NSTAGE = 3;
constexpr int CONSUMER_WARPS = 8;
// 8 warps = 2 per SM sub-partition: every warp issues DMMA, thread 0 additionally issues the bulk copies.  (A ninth,
// dedicated producer warp would put 3 warps on one sub-partition, which caps the kernel at 168 registers/thread and
// fills that sub-partition's register file, so no Coulomb CTA could co-reside.)
constexpr int CTA_THREADS = CONSUMER_WARPS * 32;
constexpr int HT_THREADS = CTA_THREADS + 32;  // half-transform: plus one dedicated producer warp (168 registers/thread)
// register split with setmaxnreg (see kernels_pp.cu): 12 warps, the producer's warpgroup drops to 40 registers per thread and
// the two consumer warpgroups rise to 232
constexpr int HT_THREADS_BIG = CTA_THREADS + 128;
constexpr int MAX_REGS = 208;

// =====================================================================================================================
// C_occ re-layout and packed density
// =====================================================================================================================
// Cs[kb][mt][kr][swz(kr, m)] = C_occ[inv[q = 32 kb + kr], i = mt*bm + m]   (zero beyond N / occ)
__global__ void k_prep_cs(const double* __restrict__ Craw, int64_t ldc, double* __restrict__ Cs, int N, int occ, int nb,
                          int n_mt, int bm, const int32_t* __restrict__ inv) {
    int64_t total = (int64_t)nb * n_mt * 32 * bm;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        int kr = (int)(idx % 32);
        int64_t r1 = idx / 32;
        int m = (int)(r1 % bm);
        int64_t r2 = r1 / bm;
        int mt = (int)(r2 % n_mt);
        int kb = (int)(r2 / n_mt);
        int q = kb * 32 + kr, i = mt * bm + m;
        double v = (q < N && i < occ) ? Craw[inv[q] + ldc * i] : 0.0;  // q: internal order, inv[q]: the caller's row
        Cs[(((int64_t)kb * n_mt + mt) * 32 + kr) * bm + swz(kr, m)] = v;
    }
}

// dt[tile][r][swz(r,c)] = w * sum_i C[p,i] C[q,i],  p = 32 a + r, q = 32 b + c,  w = 1 on diagonal tiles (full
// square stored) and 2 on off-diagonal tiles — the doubled off-diagonal density of copy_screened_density_to_array
// (ScreenedDF.jl:305-316) expressed in tile space.
__global__ void __launch_bounds__(256) k_density_packed(const double* __restrict__ Craw, int64_t ldc,
                                                         double* __restrict__ dt, const int32_t* __restrict__ tile_a,
                                                         const int32_t* __restrict__ tile_b, int N, int occ,
                                                         const int32_t* __restrict__ inv) {
    constexpr int IC = 64;                      // occupied orbitals staged per pass
    __shared__ double sa[IC][33], sb[IC][33];   // [i][p] / [i][q], padded against bank conflicts
    const int tile = blockIdx.x;
    const int a = tile_a[tile], b = tile_b[tile];
    const double w = (a == b) ? 1.0 : 2.0;
    const int c = threadIdx.x & 31, r0 = threadIdx.x >> 5;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int i0 = 0; i0 < occ; i0 += IC) {
        for (int ii = r0; ii < IC; ii += 8) {
            const int i = i0 + ii, p = a * 32 + c, q = b * 32 + c;
            sa[ii][c] = (i < occ && p < N) ? Craw[inv[p] + ldc * i] : 0.0;
            sb[ii][c] = (i < occ && q < N) ? Craw[inv[q] + ldc * i] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int ii = 0; ii < IC; ii++) {
            const double vb = sb[ii][c];
#pragma unroll
            for (int k = 0; k < 4; k++) acc[k] = fma(sa[ii][r0 + 8 * k], vb, acc[k]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int r = r0 + 8 * k;
        dt[(int64_t)tile * TILE_ELEMS + r * 32 + swz(r, c)] = w * acc[k];
    }
}

void launch_prep(const Layout& L, Shard& s, int64_t ldc, bool with_density) {
    int64_t total = (int64_t)L.nb * L.n_mt * 32 * L.bm;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    if (!(L.pp_enabled && s.Bc != nullptr))  // the per-function half-transform reads Ci (launch_pp_prep) instead
        k_prep_cs<<<blocks, 256, 0, s.stream>>>(s.Craw, ldc, s.Cs, (int)L.N, (int)L.occ, L.nb, L.n_mt, L.bm, s.d_inv);
    if (with_density)
        k_density_packed<<<(int)L.tiles_kept, 256, 0, s.stream>>>(s.Craw, ldc, s.dt, s.d_tile_a, s.d_tile_b, (int)L.N,
                                                               (int)L.occ, s.d_inv);
}

// =====================================================================================================================
// Half-transform
// =====================================================================================================================
struct HTParams {
    const double* Bp;
    int64_t row_len;
    const int32_t* tile_index;
    const int32_t* kb_list;   // [nb][nb]: the q blocks kb for which tile (max(pb,kb), min(pb,kb)) is stored, ascending
    const int32_t* kb_count;  // [nb]
    const int32_t* tile_strips;  // [tiles_kept]: bits 0-7 non-empty 4-row strips, bits 8-15 non-empty 4-column strips
    const double* Cs;
    double* W;
    int q_len, nb, n_mt, npt64, occ_w, n_valid_last;
    int64_t n_items;
    int64_t w_elems, bp_elems, cs_elems;  // buffer extents in doubles (read by the bounds-checked build only)
    double* vt;   // [CTA][M half][q_len] partial Coulomb vector from the accumulators (WITH_V; nullptr otherwise)
};

// One work item = (quad of auxiliary rows Q0..Q0+3, 32-wide p block pb, M tile of BM = 8*MT occupied orbitals):
//   out[Qj][i, p] = sum_q Cs[q][i] * Bsym_Qj[q][p]      (M = i, N = p, K = q)
// The k loop walks only the q blocks kb whose tile (max(pb,kb), min(pb,kb)) the mask kept — the same list for every
// auxiliary row, so the four rows of a quad do exactly the same work and a stage (one C slab + the four rows' copies
// of one tile) never idles a warp: the cost of the half-transform is proportional to the number of stored tiles.
// A stored tile is read as [k][n] when kb >= pb and transposed as [n][k] when kb < pb, so it serves both triangles.
// Each consumer warp owns a (BM/2) x 32 block: wm = warp >> 2 selects the M half, j = warp & 3 the auxiliary row of
// the quad, so a k-step of 4 costs MH + 4 fragment loads for MH * 4 DMMAs.
template <int MT, bool TRANSPOSED>
__device__ __forceinline__ void ht_kstep(double (&acc)[MT / 2][4][2], const double* __restrict__ sA,
                                         const double* __restrict__ sB, int m0, int g, int t, int ks) {
    constexpr int BM = 8 * MT;
    constexpr int MH = MT / 2;
    const int kr = 4 * ks + t;
    double a[MH], b[4];
#pragma unroll
    for (int jm = 0; jm < MH; jm++) a[jm] = sA[kr * BM + ((m0 + 8 * jm + g) ^ (t << 2))];
#pragma unroll
    for (int jn = 0; jn < 4; jn++) {
        const int n = 8 * jn + g;
        b[jn] = TRANSPOSED ? sB[n * 32 + (kr ^ ((n & 3) << 2))] : sB[kr * 32 + (n ^ (t << 2))];
    }
#pragma unroll
    for (int jm = 0; jm < MH; jm++)
#pragma unroll
        for (int jn = 0; jn < 4; jn++) dmma(acc[jm][jn][0], acc[jm][jn][1], a[jm], b[jn]);
}

// strips: bit ks set = the four k rows 4 ks .. 4 ks + 3 of the tile hold an unscreened pair (empty strips are exact
// zeros in HBM, so skipping their k-steps changes nothing); 0xFF takes the unrolled path
template <int MT, bool TRANSPOSED>
__device__ __forceinline__ void ht_slab(double (&acc)[MT / 2][4][2], const double* __restrict__ sA,
                                        const double* __restrict__ sB, int m0, int g, int t, uint32_t strips) {
    if (strips == 0xFFu) {
#pragma unroll 2
        for (int ks = 0; ks < KSLAB / 4; ks++) ht_kstep<MT, TRANSPOSED>(acc, sA, sB, m0, g, t, ks);
    } else {
        for (uint32_t m = strips; m; m &= m - 1) ht_kstep<MT, TRANSPOSED>(acc, sA, sB, m0, g, t, __ffs(m) - 1);
    }
}

// Launched with HT_THREADS = 288: warp 8 is a dedicated copy producer (its lane 0 runs the whole produce() sequence, as
// far ahead as the ring allows) and warps 0-7 only compute.
// WITH_V: the epilogue also adds the item's share of the Coulomb vector, V[Q] = sum_p sum_i C[p,i] W[Q,i,p], to
// the slot of (CTA, M half, Q) — one warp, hence one thread, per address: the reductions are bit-reproducible.
template <int MT, bool BIGREG, bool WITH_V>
__global__ void __launch_bounds__(BIGREG ? HT_THREADS_BIG : HT_THREADS, 1) k_half_transform(HTParams prm) {
    constexpr int BM = 8 * MT;
    constexpr int MH = MT / 2;
    constexpr int STAGE_DOUBLES = 32 * BM + 4 * TILE_ELEMS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stages = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + (size_t)NSTAGE * STAGE_DOUBLES);
    uint64_t* empty = full + NSTAGE;
    int* stage_kb = reinterpret_cast<int*>(empty + NSTAGE);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], CONSUMER_WARPS);
        }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    const int nb = prm.nb, n_mt = prm.n_mt;

    // ---- producer warp: walks the same (item, stored tile) sequence ahead of the consumers ----
    if (warp >= CONSUMER_WARPS) {
        if constexpr (BIGREG) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");  // all four warps of this warpgroup
        if (warp > CONSUMER_WARPS || lane != 0) return;
        PipeState pp;
        for (int64_t item = blockIdx.x; item < prm.n_items; item += gridDim.x) {
            const int mt = (int)(item % n_mt);
            const int64_t r1 = item / n_mt;
            const int pb = (int)(r1 % nb);
            const int64_t q0 = (r1 / nb) * 4;
            const int nq = (int)min((int64_t)4, (int64_t)prm.q_len - q0);
            const double* Brow = prm.Bp + q0 * prm.row_len;
            const int nk = prm.kb_count[pb];
            const int32_t* kl = prm.kb_list + (int64_t)pb * nb;
            for (int s = 0; s < nk; s++) {
                const int kb = kl[s];
                const int32_t tix = prm.tile_index[max(pb, kb) * nb + min(pb, kb)];
                mbar_wait(&empty[pp.stage], pp.phase ^ 1u);
                // k runs over the tile's rows when kb >= pb (tile (kb, pb) read straight), over its columns otherwise
                const uint32_t sm = (uint32_t)prm.tile_strips[tix];
                stage_kb[pp.stage] = kb | (int)(((kb >= pb) ? (sm & 0xFFu) : ((sm >> 8) & 0xFFu)) << 16);
                double* st = stages + (size_t)pp.stage * STAGE_DOUBLES;
                JC_CHECK(tix >= 0 && ((int64_t)kb * n_mt + mt + 1) * (32 * BM) <= prm.cs_elems &&
                             (q0 + nq - 1) * prm.row_len + ((int64_t)tix + 1) * TILE_ELEMS <= prm.bp_elems,
                         "tile half-transform: copy outside Cs / Bp");
                mbar_arrive_expect_tx(&full[pp.stage], 32u * BM * 8u + (uint32_t)nq * (TILE_ELEMS * 8u));
                bulk_g2s(st, prm.Cs + ((int64_t)kb * n_mt + mt) * (32 * BM), 32u * BM * 8u, &full[pp.stage]);
                for (int j = 0; j < nq; j++)
                    bulk_g2s(st + 32 * BM + j * TILE_ELEMS, Brow + j * prm.row_len + (int64_t)tix * TILE_ELEMS, TILE_ELEMS * 8u,
                             &full[pp.stage]);
                pp.advance<NSTAGE>();
            }
        }
        return;
    }

    // ---------------------------------------------------- consumers ----------------------------------------------------
    if constexpr (BIGREG) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    PipeState ps;
    const int g = lane >> 2, t = lane & 3;
    const int j = warp & 3;                 // auxiliary row inside the quad
    const int m0 = (warp >> 2) * (BM / 2);  // first occupied-orbital row of this warp's half
    for (int64_t item = blockIdx.x; item < prm.n_items; item += gridDim.x) {
        const int mt = (int)(item % n_mt);
        const int64_t r1 = item / n_mt;
        const int pb = (int)(r1 % nb);
        const int64_t qrow = (r1 / nb) * 4 + j;
        const bool valid = qrow < prm.q_len;
        const int nk = prm.kb_count[pb];
        double acc[MH][4][2];
#pragma unroll
        for (int jm = 0; jm < MH; jm++)
#pragma unroll
            for (int jn = 0; jn < 4; jn++) acc[jm][jn][0] = acc[jm][jn][1] = 0.0;

        for (int s = 0; s < nk; s++) {
            mbar_wait(&full[ps.stage], ps.phase);
            if (valid) {
                const int info = stage_kb[ps.stage];
                const int kb = info & 0xFFFF;
                const uint32_t strips = (uint32_t)info >> 16;
                const double* sA = stages + (size_t)ps.stage * STAGE_DOUBLES;
                const double* sB = sA + 32 * BM + j * TILE_ELEMS;
                if (kb < pb) ht_slab<MT, true>(acc, sA, sB, m0, g, t, strips);
                else ht_slab<MT, false>(acc, sA, sB, m0, g, t, strips);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[ps.stage]);
            ps.advance<NSTAGE>();
        }
        // epilogue: W tile [64 p][32 k] (k = (Q,i) fastest, swizzled); 8 lanes of equal t fill 64 contiguous bytes
        if (valid) {
#pragma unroll
            for (int jm = 0; jm < MH; jm++) {
                const int i = mt * BM + m0 + 8 * jm + g;
                if (i < prm.occ_w) {
                    const int64_t krow = qrow * prm.occ_w + i;
                    const int64_t kc = krow >> 5;
                    const int kr = (int)(krow & 31);
#pragma unroll
                    for (int jn = 0; jn < 4; jn++) {
                        const int p = 32 * pb + 8 * jn + 2 * t;  // even: p and p + 1 share the 64-wide tile
                        const int p64 = p >> 6, pr = p & 63;
                        double* w = prm.W + ((kc * prm.npt64 + p64) * 64 + pr) * 32;
                        JC_CHECK(w >= prm.W && w + 64 <= prm.W + prm.w_elems, "tile half-transform: W store outside W");
                        w[swz(pr, kr)] = acc[jm][jn][0];
                        w[32 + swz(pr + 1, kr)] = acc[jm][jn][1];
                    }
                }
            }
            if constexpr (WITH_V) {
                // C[p, i] of the warp's block comes from the same Cs slab the k loop used for q block pb (zero beyond N / occ)
                const double* cs = prm.Cs + ((int64_t)pb * n_mt + mt) * (32 * BM);
                double v = 0.0;
#pragma unroll
                for (int jm = 0; jm < MH; jm++) {
                    const int m = m0 + 8 * jm + g;
#pragma unroll
                    for (int jn = 0; jn < 4; jn++) {
                        const int kr0 = 8 * jn + 2 * t;
                        v = fma(acc[jm][jn][0], cs[kr0 * BM + swz(kr0, m)], v);
                        v = fma(acc[jm][jn][1], cs[(kr0 + 1) * BM + swz(kr0 + 1, m)], v);
                    }
                }
                v = warp_sum(v);
                if (lane == 0)
                    asm volatile("red.global.add.f64 [%0], %1;" ::"l"(prm.vt + ((int64_t)blockIdx.x * 2 + (warp >> 2)) * prm.q_len + qrow), "d"(v)
                                 : "memory");
            }
        }
    }
}

// V[Q] = sum over (CTA, M half) of the partial Coulomb vectors the tile half-transform left, in a fixed order
__global__ void k_vt_reduce(const double* __restrict__ vt, double* __restrict__ V, int n_parts, int q_len) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= q_len) return;
    double v = 0.0;
    for (int c = 0; c < n_parts; c++) v += vt[(int64_t)c * q_len + q];
    V[q] = v;
}

size_t half_transform_smem_bytes(int bm) {
    return (size_t)NSTAGE * (32 * bm + 4 * TILE_ELEMS) * 8 + 2 * NSTAGE * 8 + NSTAGE * 4 + 16;
}

template <int MT, bool BIGREG, bool WITH_V>
static void launch_ht_k(const HTParams& prm, int grid, size_t smem, cudaStream_t st) {
    cudaFuncSetAttribute(k_half_transform<MT, BIGREG, WITH_V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_half_transform<MT, BIGREG, WITH_V><<<grid, BIGREG ? HT_THREADS_BIG : HT_THREADS, smem, st>>>(prm);
}

template <int MT>
static void launch_ht_t(const HTParams& prm, int grid, bool bigreg, cudaStream_t st) {
    size_t smem = half_transform_smem_bytes(8 * MT);
    if (prm.vt != nullptr) {
        if (bigreg) launch_ht_k<MT, true, true>(prm, grid, smem, st);
        else launch_ht_k<MT, false, true>(prm, grid, smem, st);
    } else {
        if (bigreg) launch_ht_k<MT, true, false>(prm, grid, smem, st);
        else launch_ht_k<MT, false, false>(prm, grid, smem, st);
    }
}

void launch_half_transform(const Layout& L, Shard& s, int sm_count, bool bigreg, bool v_in_ht) {
    HTParams prm;
    prm.Bp = s.Bp;
    prm.row_len = L.row_len;
    prm.tile_index = s.d_tile_index;
    prm.kb_list = s.d_kb_list;
    prm.kb_count = s.d_kb_count;
    prm.tile_strips = s.d_tile_strips;
    prm.Cs = s.Cs;
    prm.W = s.W;
    prm.q_len = (int)s.q_len;
    prm.nb = L.nb;
    prm.n_mt = L.n_mt;
    prm.npt64 = L.npt64;
    prm.occ_w = L.occ_w;
    prm.n_valid_last = (int)(L.N - 32 * (L.nb - 1));
    prm.n_items = ((int64_t)s.q_len + 3) / 4 * L.nb * L.n_mt;
    prm.w_elems = (int64_t)s.xplan.chunks * L.npt64 * 2048;
    prm.bp_elems = (int64_t)s.q_len * L.row_len;
    prm.cs_elems = (int64_t)L.nb * L.n_mt * 32 * L.bm;
    int grid = (int)std::min<int64_t>(prm.n_items, sm_count);
    if (grid <= 0) return;
    prm.vt = (v_in_ht && s.Vt != nullptr) ? s.Vt : nullptr;
    if (prm.vt) cudaMemsetAsync(s.Vt, 0, sizeof(double) * (size_t)grid * 2 * (size_t)s.q_len, s.stream);  // the slots are summed into
    switch (L.bm / 8) {
        case 2: launch_ht_t<2>(prm, grid, bigreg, s.stream); break;
        case 4: launch_ht_t<4>(prm, grid, bigreg, s.stream); break;
        case 6: launch_ht_t<6>(prm, grid, bigreg, s.stream); break;
        case 8: launch_ht_t<8>(prm, grid, bigreg, s.stream); break;
        case 10: launch_ht_t<10>(prm, grid, bigreg, s.stream); break;
        case 12: launch_ht_t<12>(prm, grid, bigreg, s.stream); break;
        case 14: launch_ht_t<14>(prm, grid, bigreg, s.stream); break;
        case 16: launch_ht_t<16>(prm, grid, bigreg, s.stream); break;
        default: break;
    }
    if (prm.vt) k_vt_reduce<<<(unsigned)((s.q_len + 127) / 128), 128, 0, s.stream>>>(s.Vt, s.V, grid * 2, (int)s.q_len);
}

// =====================================================================================================================
// Exchange (SYRK-like):  Kpart[item][m][n] = sum_{k in slice} W[k][128 ta + m] * W[k][128 tb + n]
// =====================================================================================================================
struct XParams {
    const double* W;
    double* Kpart;
    const int32_t* item_list;  // items of CTA c: item_list[item_off[c] .. item_off[c + 1])  (cost-balanced on the host)
    const int32_t* item_off;
    long long* cta_cycles;     // [gridDim.x] busy cycles of this launch (profiling aid)
    const uint8_t* half_dead;  // block-screened exchange: per tile, bit h = row half h lies wholly in dropped K blocks (or nullptr)
    int npt64, n_tiles, n_slices, chunks, chunks_per_slice, N;
    int64_t w_elems, kpart_elems;  // buffer extents in doubles (read by the bounds-checked build only)
};

constexpr int X_STAGE_DOUBLES = 4 * 32 * 64;  // A0 A1 B0 B1, each a [64 p][32 k] W tile

// Which block of the 128 x 128 output tile (ta, tb) a warp owns, in units of 8 rows / columns: rows 8 r0 .. 8 (r0 + mt),
// columns 8 c0 .. 8 (c0 + nt), mt <= 8, nt <= 4.
//   * off-diagonal tiles: the plain 2 x 4 arrangement of 64 x 32 blocks;
//   * diagonal tiles: only the lower triangle is needed (the reference computes its diagonal blocks whole,
//     ScreenedDF.jl:596-601) — it is covered by eight rectangles chosen so that the two warps of every SM sub-partition
//     (w and w + 4) together issue 6 + 4 = 10 of the 16 row units a full tile costs;
//   * rows / columns beyond N (the last tile row) are clipped.
// Shared by the kernel and the host-side cost model of plan_exchange.
__host__ __device__ inline void exchange_warp_block(int ta, int tb, int warp, int N, int& r0, int& c0, int& mt, int& nt) {
    if (ta == tb) {
        // Eight rectangles over the lower triangle of the 16 x 16 grid of 8 x 8 DMMA tiles, column strips of 4:
        //   cols 0-3: rows 0-5, 6-11, 12-15   cols 4-7: rows 4-9, 10-15   cols 8-11: rows 8-11, 12-15   cols 12-15: rows 12-15
        // paired (w, w + 4) as 6 + 4 row units on every sub-partition (an 8 + 2 pairing of the same cover measured the
        // same to 0.3 %: the pipe total per sub-partition is what counts).
        //   r0 = {0, 6, 4, 10, 12, 8, 12, 12}   c0 = {0, 0, 4, 4, 0, 8, 8, 12}   mt = {6, 6, 6, 6, 4, 4, 4, 4}
        r0 = (int)((0xCC8CA460u >> (4 * warp)) & 15u);
        c0 = (int)((0xC8804400u >> (4 * warp)) & 15u);
        mt = (int)((0x44446666u >> (4 * warp)) & 15u);
        nt = 4;
    } else {
        // column strip warp & 3; the valid rows of the tile (16 units, fewer in the last tile row) split evenly between
        // the two warps of the sub-partition
        const int rows16 = min(16, (N - 128 * ta + 7) >> 3);
        const int half = (rows16 + 1) >> 1;
        r0 = (warp >> 2) ? half : 0;
        mt = (warp >> 2) ? rows16 - half : half;
        c0 = 4 * (warp & 3); nt = 4;
    }
    const int rows8 = (N - 128 * ta + 7) >> 3, cols8 = (N - 128 * tb + 7) >> 3;
    mt = max(0, min(mt, rows8 - r0));
    nt = max(0, min(nt, cols8 - c0));
    if (mt == 0 || nt == 0) mt = nt = 0;
}

// A rectangle that starts ON the diagonal of a diagonal tile (r0 == c0) only needs its lower staircase: row unit jm needs
// the column units <= jm.  All four diagonal-crossing rectangles of the cover start on it.
__host__ __device__ inline bool exchange_block_is_stair(int ta, int tb, int r0, int c0) { return ta == tb && r0 == c0; }

// DMMA tiles per k-step of a warp block
__host__ __device__ inline int exchange_block_tiles(int ta, int tb, int r0, int c0, int mt, int nt) {
    if (!exchange_block_is_stair(ta, tb, r0, c0)) return mt * nt;
    int n = 0;
    for (int jm = 0; jm < mt; jm++) n += min(nt, jm + 1);
    return n;
}

__host__ __device__ inline void exchange_decode_tile(int tile, int& ta, int& tb) {
    int a = 0;
    while ((a + 1) * (a + 2) / 2 <= tile) a++;
    ta = a;
    tb = tile - a * (a + 1) / 2;
}

// DMMA row units (x 4 column tiles) the busiest sub-partition issues per k-step of 4 for this tile: the cost model
// half_dead (block-screened exchange): bit h = the warps of row half h of an off-diagonal tile have nothing to compute
int exchange_tile_cost(int tile, int N, int half_dead) {
    int ta, tb, best = 0;
    exchange_decode_tile(tile, ta, tb);
    for (int j = 0; j < 4; j++) {
        int c = 0;
        for (int w = j; w < 8; w += 4) {
            int r0, c0, mt, nt;
            exchange_warp_block(ta, tb, w, N, r0, c0, mt, nt);
            if (ta != tb && ((half_dead >> (w >> 2)) & 1)) continue;
            c += exchange_block_tiles(ta, tb, r0, c0, mt, nt);
        }
        best = std::max(best, c);
    }
    return best;
}

// Rows [row_begin, row_end) of the 128-row tile ta (tile-local) that the four warps of row half `half` own in an off-diagonal
// tile (ta, tb): what the block-screened planner tests against the dropped K blocks.  Empty for a diagonal tile.
void exchange_half_rows(int ta, int tb, int half, int N, int& row_begin, int& row_end) {
    row_begin = row_end = 0;
    if (ta == tb) return;
    int r0, c0, mt, nt;
    exchange_warp_block(ta, tb, 4 * half, N, r0, c0, mt, nt);
    row_begin = 8 * r0;
    row_end = min(8 * (r0 + mt), N - 128 * ta);
    if (row_end < row_begin) row_end = row_begin;
}

// One pipeline stage (32 k rows) of a warp block of MT x NT DMMA tiles.  MT and NT are compile-time so a clipped or
// diagonal block issues exactly MT * NT DMMAs per k-step (a predicated-off DMMA still takes its slot in the tensor pipe:
// measured — the predicated version of this loop left the pipe cycles of the diagonal tiles unchanged).
template <int MT, int NT>
__device__ __forceinline__ void x_slab(double (&acc)[8][4][2], const double* __restrict__ sA, const double* __restrict__ sB,
                                       int g, int t) {
    const int sw = (g & 3) << 2;
#pragma unroll 2
    for (int k0 = 0; k0 < KSLAB; k0 += 4) {
        const int kr = (k0 + t) ^ sw;
        double a[MT], b[NT];
#pragma unroll
        for (int jm = 0; jm < MT; jm++) a[jm] = sA[(8 * jm + g) * 32 + kr];
#pragma unroll
        for (int jn = 0; jn < NT; jn++) b[jn] = sB[(8 * jn + g) * 32 + kr];
#pragma unroll
        for (int jm = 0; jm < MT; jm++)
#pragma unroll
            for (int jn = 0; jn < NT; jn++) dmma(acc[jm][jn][0], acc[jm][jn][1], a[jm], b[jn]);
    }
}

// Staircase block of a rectangle that starts on the diagonal: row unit jm times column units 0 .. min(3, jm)
template <int MT>
__device__ __forceinline__ void x_slab_stair(double (&acc)[8][4][2], const double* __restrict__ sA,
                                             const double* __restrict__ sB, int g, int t) {
    const int sw = (g & 3) << 2;
#pragma unroll 2
    for (int k0 = 0; k0 < KSLAB; k0 += 4) {
        const int kr = (k0 + t) ^ sw;
        double a[MT], b[4];
#pragma unroll
        for (int jm = 0; jm < MT; jm++) a[jm] = sA[(8 * jm + g) * 32 + kr];
#pragma unroll
        for (int jn = 0; jn < 4; jn++)
            if (jn < MT) b[jn] = sB[(8 * jn + g) * 32 + kr];
#pragma unroll
        for (int jm = 0; jm < MT; jm++)
#pragma unroll
            for (int jn = 0; jn < 4; jn++)
                if (jn <= jm) dmma(acc[jm][jn][0], acc[jm][jn][1], a[jm], b[jn]);  // compile-time after unrolling
    }
}

__device__ __forceinline__ void x_slab_stair_mt(double (&acc)[8][4][2], const double* sA, const double* sB, int g, int t,
                                                int mt) {
    switch (mt) {
        case 1: x_slab_stair<1>(acc, sA, sB, g, t); break;
        case 2: x_slab_stair<2>(acc, sA, sB, g, t); break;
        case 3: x_slab_stair<3>(acc, sA, sB, g, t); break;
        case 4: x_slab_stair<4>(acc, sA, sB, g, t); break;
        case 5: x_slab_stair<5>(acc, sA, sB, g, t); break;
        case 6: x_slab_stair<6>(acc, sA, sB, g, t); break;
        default: break;
    }
}

template <int NT>
__device__ __forceinline__ void x_slab_mt(double (&acc)[8][4][2], const double* __restrict__ sA,
                                          const double* __restrict__ sB, int g, int t, int mt) {
    switch (mt) {
        case 1: x_slab<1, NT>(acc, sA, sB, g, t); break;
        case 2: x_slab<2, NT>(acc, sA, sB, g, t); break;
        case 3: x_slab<3, NT>(acc, sA, sB, g, t); break;
        case 4: x_slab<4, NT>(acc, sA, sB, g, t); break;
        case 5: x_slab<5, NT>(acc, sA, sB, g, t); break;
        case 6: x_slab<6, NT>(acc, sA, sB, g, t); break;
        case 7: x_slab<7, NT>(acc, sA, sB, g, t); break;
        case 8: x_slab<8, NT>(acc, sA, sB, g, t); break;
        default: break;
    }
}

// DEDICATED = false: 8 warps, thread 0 doubles as the copy producer (it refills a slot only after ALL warps released it, so
//   warp 0 meets the slowest warp at every stage).  DEDICATED = true: 12 warps with the setmaxnreg register split of the
//   half-transform kernels — warp 8 only produces, the eight DMMA warps never look at the empty barriers.
template <bool DEDICATED>
__device__ __forceinline__ void exchange_body(const XParams& prm) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stages = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + (size_t)NSTAGE * X_STAGE_DOUBLES);
    uint64_t* empty = full + NSTAGE;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], CONSUMER_WARPS);
        }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    const long long t_begin = clock64();
    const int npt64 = prm.npt64;
    const bool is_producer = DEDICATED ? (threadIdx.x == CTA_THREADS) : (threadIdx.x == 0);
    const int j_begin = prm.item_off[blockIdx.x], j_end = prm.item_off[blockIdx.x + 1];

    auto decode = [&](int item, int& ta, int& tb, int& kc0, int& kc1) {
        exchange_decode_tile(item % prm.n_tiles, ta, tb);
        const int slice = item / prm.n_tiles;
        kc0 = slice * prm.chunks_per_slice;
        kc1 = min(prm.chunks, kc0 + prm.chunks_per_slice);
    };

    // ---- copy issue (thread 0 only), NSTAGE - 1 chunks ahead of the consumers ----
    PipeState pp;
    int p_j = j_begin;
    int p_kc = -1, p_kc1 = 0, p_col[4] = {0, 0, 0, 0}, p_nvalid = 0;
    auto produce = [&]() {
        while (p_kc < 0 || p_kc >= p_kc1) {  // next item (skips empty slices)
            if (p_kc >= 0) p_j++;
            if (p_j >= j_end) return;
            int ta, tb, kc0, kc1;
            decode(prm.item_list[p_j], ta, tb, kc0, kc1);
            p_col[0] = 2 * ta; p_col[1] = 2 * ta + 1; p_col[2] = 2 * tb; p_col[3] = 2 * tb + 1;
            if (ta == tb) p_col[2] = p_col[3] = npt64;  // diagonal tile: B0 B1 = A0 A1, read from the A slots
            p_nvalid = 0;
#pragma unroll
            for (int x = 0; x < 4; x++) p_nvalid += (p_col[x] < npt64);
            p_kc = kc0;
            p_kc1 = kc1;
        }
        mbar_wait(&empty[pp.stage], pp.phase ^ 1u);
        double* st = stages + (size_t)pp.stage * X_STAGE_DOUBLES;
        mbar_arrive_expect_tx(&full[pp.stage], (uint32_t)p_nvalid * 32u * 64u * 8u);
#pragma unroll
        for (int x = 0; x < 4; x++)
            JC_CHECK(p_col[x] >= npt64 || ((int64_t)p_kc * npt64 + p_col[x] + 1) * 2048 <= prm.w_elems, "exchange: copy outside W");
#pragma unroll
        for (int x = 0; x < 4; x++)
            if (p_col[x] < npt64)
                bulk_g2s(st + x * 2048, prm.W + ((int64_t)p_kc * npt64 + p_col[x]) * 2048, 2048u * 8u, &full[pp.stage]);
        pp.advance<NSTAGE>();
        p_kc++;
    };
    if constexpr (DEDICATED) {
        if (warp >= CONSUMER_WARPS) {
            asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");  // all four warps of the producer's warpgroup
            if (is_producer)
                while (p_j < j_end) produce();
            return;
        }
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    } else {
        if (is_producer)
            for (int s = 0; s < NSTAGE - 1; s++) produce();
    }

    PipeState ps;
    const int g = lane >> 2, t = lane & 3;
    for (int j = j_begin; j < j_end; j++) {
        const int item = prm.item_list[j];
        int ta, tb, kc0, kc1;
        decode(item, ta, tb, kc0, kc1);
        int r0, c0, mt, nt;
        exchange_warp_block(ta, tb, warp, prm.N, r0, c0, mt, nt);
        // block-screened exchange: a row half whose every element lies in dropped K blocks is not computed (its warps only keep
        // the ring moving); both warps of every sub-partition sit in different halves, so the saving is the same on all four
        if (prm.half_dead != nullptr && ta != tb && ((prm.half_dead[item % prm.n_tiles] >> (warp >> 2)) & 1)) mt = nt = 0;
        const bool active = mt > 0;
        const bool whole = (mt == 8) && (nt == 4);
        const bool stair = exchange_block_is_stair(ta, tb, r0, c0) && mt > 0;  // mt <= 6 on these rectangles, nt >= min(4, mt)
        // A0 | A1 and B0 | B1 are adjacent [64][32] tiles: 128 rows of one stride, so a block may straddle the pair
        const int a_off = r0 * 8 * 32;
        const int b_off = (ta == tb ? 0 : 2 * 2048) + c0 * 8 * 32;
        double acc[8][4][2];
#pragma unroll
        for (int jm = 0; jm < 8; jm++)
#pragma unroll
            for (int jn = 0; jn < 4; jn++) acc[jm][jn][0] = acc[jm][jn][1] = 0.0;

        for (int kc = kc0; kc < kc1; kc++) {
            mbar_wait(&full[ps.stage], ps.phase);
            // refill the slot every warp released one stage ago BEFORE computing this one: the copy then has two full
            // stage times to land.  (Also tried: register double-buffering of the fragments across k-steps and stage
            // boundaries — 244 registers, 3 % slower; the two warps per sub-partition already cover each other's bubbles.)
            // Also tried: cp.async.bulk.prefetch.L2 of the chunk four stages beyond the ring — 6 % slower.
            if constexpr (!DEDICATED) {
                if (is_producer) produce();
                __syncwarp();
            }
            const double* sA = stages + (size_t)ps.stage * X_STAGE_DOUBLES + a_off;
            const double* sB = stages + (size_t)ps.stage * X_STAGE_DOUBLES + b_off;
            if (whole) x_slab<8, 4>(acc, sA, sB, g, t);
            else if (stair) x_slab_stair_mt(acc, sA, sB, g, t, mt);
            else if (nt == 4) x_slab_mt<4>(acc, sA, sB, g, t, mt);
            else if (nt == 3) x_slab_mt<3>(acc, sA, sB, g, t, mt);
            else if (nt == 2) x_slab_mt<2>(acc, sA, sB, g, t, mt);
            else if (nt == 1) x_slab_mt<1>(acc, sA, sB, g, t, mt);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[ps.stage]);
            ps.advance<NSTAGE>();
        }
        if (!active) continue;
        double* out = prm.Kpart + (int64_t)item * (128 * 128);
#pragma unroll
        for (int jm = 0; jm < 8; jm++) {
            if (jm >= mt) continue;
            const int m = 8 * (r0 + jm) + g;
#pragma unroll
            for (int jn = 0; jn < 4; jn++) {
                if (jn >= nt) continue;
                const int n = 8 * (c0 + jn) + 2 * t;
                JC_CHECK(m < 128 && n + 1 < 128 && (int64_t)(item + 1) * (128 * 128) <= prm.kpart_elems, "exchange: store outside Kpart");
                *reinterpret_cast<double2*>(out + m * 128 + n) = make_double2(acc[jm][jn][0], acc[jm][jn][1]);
            }
        }
    }
    if (threadIdx.x == 0) prm.cta_cycles[blockIdx.x] = clock64() - t_begin;
}

__global__ void __maxnreg__(MAX_REGS) k_exchange(XParams prm) { exchange_body<false>(prm); }
__global__ void __launch_bounds__(HT_THREADS_BIG, 1) k_exchange_ws(XParams prm) { exchange_body<true>(prm); }

size_t exchange_smem_bytes() { return (size_t)NSTAGE * X_STAGE_DOUBLES * 8 + 2 * NSTAGE * 8 + 16; }

void launch_exchange(const Layout& L, Shard& s, int sm_count, bool dedicated_producer) {
    XParams prm;
    prm.W = s.W;
    prm.Kpart = s.Kpart;
    prm.item_list = s.d_xitems;
    prm.item_off = s.d_xoff;
    prm.cta_cycles = s.d_xcycles;
    prm.half_dead = (L.xs_halves_skipped > 0) ? s.d_xs_half_dead : nullptr;
    prm.npt64 = L.npt64;
    prm.n_tiles = s.xplan.n_tiles;
    prm.n_slices = s.xplan.n_slices;
    prm.chunks = s.xplan.chunks;
    prm.chunks_per_slice = s.xplan.chunks_per_slice;
    prm.N = (int)L.N;
    prm.w_elems = (int64_t)s.xplan.chunks * L.npt64 * 2048;
    prm.kpart_elems = (int64_t)s.xplan.n_tiles * s.xplan.n_slices * 128 * 128;
    const int grid = s.xplan.n_ctas;
    if (grid <= 0) return;
    size_t smem = exchange_smem_bytes();
    if (dedicated_producer) {
        cudaFuncSetAttribute(k_exchange_ws, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_exchange_ws<<<grid, HT_THREADS_BIG, smem, s.stream>>>(prm);
    } else {
        cudaFuncSetAttribute(k_exchange, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_exchange<<<grid, CTA_THREADS, smem, s.stream>>>(prm);
    }
}

// =====================================================================================================================
// Fock epilogue:  F = [H] + Jt (scatter from tile space) - sum_slices Kpart, mirrored to both triangles
// (copy_screened_coulomb_to_fock!, ScreenedDF.jl:367-378; H add DensityFitting.jl:62-65; block mirror :598-601)
// =====================================================================================================================
__global__ void k_fock_epilogue(double* __restrict__ F, const double* __restrict__ H, const double* __restrict__ Jt,
                                const double* __restrict__ Kpart, const int32_t* __restrict__ tile_index,
                                const int32_t* __restrict__ perm, int N, int nb, int n_tiles, int n_slices,
                                const double* __restrict__ Fprev, const int32_t* __restrict__ xs_blk_of,
                                const uint8_t* __restrict__ xs_keep, int xs_nblk) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int p = blockIdx.y;
    if (q > p || p >= N) return;
    if (Fprev != nullptr) {
        // block-screened exchange (calculate_K_lower_diagonal_block, ScreenedDF.jl:494-512): a dropped block is not
        // written by the K GEMM and not touched by the Coulomb copy (:367-378: all its pairs are screened), so it keeps
        // the content of two_electron_fock — the caller's buffer as it arrived — plus H on rank 0
        const int bp = xs_blk_of[p], bq = xs_blk_of[q];
        if (bp >= 0 && bq >= 0 && !xs_keep[bp * xs_nblk + bq]) {
            double vpq = Fprev[p + (int64_t)N * q], vqp = Fprev[q + (int64_t)N * p];
            if (H != nullptr) {
                vpq += H[p + (int64_t)N * q];
                vqp += H[q + (int64_t)N * p];
            }
            F[p + (int64_t)N * q] = vpq;
            F[q + (int64_t)N * p] = vqp;
            return;
        }
    }
    // (p, q) in the caller's order -> internal order; K and J are symmetric, so read the (hi, lo) element
    const int pi = perm[p], qi = perm[q];
    const int hi = max(pi, qi), lo = min(pi, qi);
    const int ta = hi >> 7, tb = lo >> 7;
    const int tile = ta * (ta + 1) / 2 + tb;
    const int64_t off = (int64_t)(hi & 127) * 128 + (lo & 127);
    double k = 0.0;
    for (int s = 0; s < n_slices; s++) k += Kpart[((int64_t)s * n_tiles + tile) * (128 * 128) + off];
    double v = -k;
    const int ti = tile_index[(hi >> 5) * nb + (lo >> 5)];
    if (ti >= 0) v += Jt[(int64_t)ti * TILE_ELEMS + (hi & 31) * 32 + swz(hi & 31, lo & 31)];
    double vpq = v, vqp = v;
    if (H != nullptr) {
        vpq += H[p + (int64_t)N * q];
        vqp += H[q + (int64_t)N * p];
    }
    F[p + (int64_t)N * q] = vpq;
    F[q + (int64_t)N * p] = vqp;
}

// After the all-reduce: NCCL may sum F[p,q] and F[q,p] over the ranks in different orders (ring chunks), so the two can
// differ in the last bit; copy the lower triangle over the upper one to keep F exactly symmetric on every rank
// (the reference mirrors explicitly too: ScreenedDF.jl:375, :596-601).
__global__ void k_mirror_lower(double* __restrict__ F, int N) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int p = blockIdx.y;
    if (q < p && p < N) F[q + (int64_t)N * p] = F[p + (int64_t)N * q];
}

void launch_mirror_lower(const Layout& L, Shard& s) {
    dim3 block(128);
    dim3 grid((unsigned)((L.N + 127) / 128), (unsigned)L.N);
    k_mirror_lower<<<grid, block, 0, s.stream>>>(s.F, (int)L.N);
}

// Ksum[tile][m][n] = sum_slices Kpart[slice][tile][m][n], slices in ascending order (deterministic).  The slices are
// read as coalesced streams here, so the permuted gather of the epilogue touches one L2-resident copy only.
__global__ void k_exchange_reduce(const double* __restrict__ Kpart, double* __restrict__ Ksum, int64_t tile_elems_total,
                                  int n_slices) {
    const int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * 2;
    if (i >= tile_elems_total) return;
    double2 acc = make_double2(0.0, 0.0);
    for (int s = 0; s < n_slices; s++) {
        const double2 v = *reinterpret_cast<const double2*>(Kpart + (int64_t)s * tile_elems_total + i);
        acc.x += v.x;
        acc.y += v.y;
    }
    *reinterpret_cast<double2*>(Ksum + i) = acc;
}

void launch_epilogue(const Layout& L, Shard& s, bool add_H) {
    const double* K = s.Kpart;
    int n_slices = s.xplan.n_slices;
    if (n_slices > 1) {
        const int64_t total = (int64_t)s.xplan.n_tiles * 128 * 128;
        k_exchange_reduce<<<(unsigned)((total / 2 + 255) / 256), 256, 0, s.stream>>>(s.Kpart, s.Ksum, total, n_slices);
        K = s.Ksum;
        n_slices = 1;
    }
    dim3 block(128);
    dim3 grid((unsigned)((L.N + 127) / 128), (unsigned)L.N);
    k_fock_epilogue<<<grid, block, 0, s.stream>>>(s.F, add_H ? s.Hdev : nullptr, s.Jt, K, s.d_tile_index, s.d_perm,
                                                  (int)L.N, L.nb, s.xplan.n_tiles, n_slices,
                                                  L.xs_any_dropped ? s.Fprev : nullptr, s.d_xs_blk_of, s.d_xs_keep, L.xs_nblk);
}

JC_DEFINE_FUZZ_SETTER(set_fuzz_seed_k)

}  // namespace jcdf

extern "C" int32_t jcdf_debug_exchange_block(int32_t ta, int32_t tb, int32_t warp, int64_t N, int32_t* out4 /* 5 values */) {
    if (!out4 || ta < tb || tb < 0 || warp < 0 || warp > 7 || N <= 128 * (int64_t)ta) {
        jcdf::set_error("jcdf_debug_exchange_block: need out4, 0 <= tb <= ta, 0 <= warp < 8, 128 ta < N");
        return JCDF_ERR_INVALID;
    }
    int r0, c0, mt, nt;
    jcdf::exchange_warp_block(ta, tb, warp, (int)N, r0, c0, mt, nt);
    out4[0] = r0; out4[1] = c0; out4[2] = mt; out4[3] = nt;
    out4[4] = (mt > 0 && jcdf::exchange_block_is_stair(ta, tb, r0, c0)) ? 1 : 0;
    return JCDF_OK;
}
