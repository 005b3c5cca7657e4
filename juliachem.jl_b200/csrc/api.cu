// C-ABI of libjcdf (include/jcdf.h): host-side state, layout construction, uploads, metric factorisation, the
// per-iteration Fock build sequence and the NCCL reduce.  All arithmetic of the hot path lives in kernels_*.cu.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <cublas_v2.h>
#include <cusolverDn.h>

#include "jcdf_internal.h"
#include "jcdf_macros.h"
#include "nccl_dyn.h"

namespace jcdf {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }

struct EighCache {
    int dev = 0;
    cublasHandle_t blas = nullptr;
    cusolverDnHandle_t sol = nullptr;
    double *dF = nullptr, *dX = nullptr, *dT = nullptr, *dW = nullptr, *work = nullptr;
    int lwork = 0;
    int* dinfo = nullptr;
    ~EighCache() {
        cudaSetDevice(dev);
        void* ptrs[] = {dF, dX, dT, dW, work, dinfo};
        for (void* p : ptrs)
            if (p) cudaFree(p);
        if (blas) cublasDestroy(blas);
        if (sol) cusolverDnDestroy(sol);
    }
};

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

template <typename T>
static int dev_alloc(T** p, size_t count) {
    *p = nullptr;
    if (count == 0) count = 1;
    JC_CUDA(cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T)));
    return JCDF_OK;
}

// keep_metric: the two-centre metric / its inverse factor do not depend on the mask (jcdf_set_mask keeps them)
static void free_shard_buffers(Shard& s, bool keep_metric = false) {
    cudaSetDevice(s.dev);
    if (s.gexec) { cudaGraphExecDestroy(s.gexec); s.gexec = nullptr; }  // the graph holds the old buffer addresses
    s.gkey = 0;
    s.graph_broken = false;
    void* ptrs[] = {s.d_pp_trirow, s.ppj_flag, s.ppj_vp, s.d_ppitems, s.d_ppoff, s.d_xcycles, s.Ksum, s.d_xitems, s.d_xoff, s.Bc, s.Ci, s.d_pp_off, s.d_pp_q, s.d_pp_src, s.d_pp_p, s.d_pp_jdst, s.dvec, s.VpartP, s.Vitem, s.Vt, s.Jpart, s.Vpart, s.Vready, s.Bp, s.W, s.Cs, s.Craw, s.dt, s.Jt, s.V, s.Kpart, s.F, s.Hdev, s.Traw, s.d_tile_index, s.d_tile_a,
                    s.d_tile_b, s.d_dest_of_col, s.d_fn_mask, s.d_kb_list, s.d_kb_count, s.d_perm, s.d_inv, s.d_tile_strips, s.Fprev, s.d_xs_blk_of, s.d_xs_keep, s.d_xs_half_dead};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (!keep_metric) {
        if (s.Linv) cudaFree(s.Linv);
        if (s.Metric) cudaFree(s.Metric);
        s.Linv = s.Metric = nullptr;
    }
    s.Bp = s.W = s.Cs = s.Craw = s.dt = s.Jt = s.V = s.Kpart = s.F = s.Hdev = s.Traw = nullptr;
    s.d_tile_index = s.d_tile_a = s.d_tile_b = s.d_kb_list = s.d_kb_count = s.d_perm = s.d_inv = s.d_tile_strips = nullptr;
    s.Vpart = nullptr;
    s.Vready = nullptr;
    s.Fprev = nullptr;
    s.d_xs_blk_of = nullptr;
    s.d_xs_keep = nullptr;
    s.d_xs_half_dead = nullptr;
    s.Bc = s.Ci = nullptr;
    s.d_pp_off = s.d_pp_q = s.d_pp_src = s.d_pp_p = s.d_pp_jdst = nullptr;
    s.dvec = s.VpartP = s.Vitem = s.Vt = s.Jpart = nullptr;
    s.d_xitems = s.d_xoff = nullptr;
    s.d_xcycles = nullptr;
    s.d_ppitems = s.d_ppoff = nullptr;
    s.d_pp_trirow = nullptr;
    s.ppj_flag = nullptr;
    s.ppj_vp = nullptr;
    s.pp_ctas = 0;
    s.Ksum = nullptr;
    s.pp_chunks = 0;
    s.bc_valid = false;
    s.d_dest_of_col = nullptr;
    s.d_fn_mask = nullptr;
    s.traw_capacity = 0;
    s.b_ready = s.t_uploaded = s.has_H = false;
}

// df_exchange_screen = true (option "exchange_block_screen" = the requested df_exchange_n_blocks, already defaulted by the
// caller): the reference's K block grid and block_screen_matrix, calculate_exchange_block_screen_matrix
// (ScreenedDF.jl:385-457) — below 100 functions one block of width N (:392-394), else width = N div n_blocks (:396); a
// lower-triangular block is kept iff its sub-block of basis_function_screen_matrix holds any pair (:433-441); the strip
// beyond width * n_blocks is never screened (:519-544).  The exchange tiles of this library live in the INTERNAL function
// order, so a 128 x 128 tile is left out of the work lists only when every element of it maps into dropped blocks; the
// epilogue decides element by element in the caller's order, which is what makes the result the reference's.
static void plan_exchange_block_screen(jcdf_handle* h, const uint8_t* fn_mask) {
    Layout& L = h->lay;
    const int64_t N = L.N;
    L.xs_nblk = L.xs_width = 0;
    L.xs_blocks_total = L.xs_blocks_kept = L.xs_tiles_skipped = 0;
    L.xs_any_dropped = false;
    L.xs_blk_of.clear();
    L.xs_keep.clear();
    L.xs_tile_skip.clear();
    L.xs_half_dead.clear();
    L.xs_halves_skipped = 0;
    // the dense statement of the contraction (NULL mask) has no K blocks: DensityFitting.jl:185-224
    if (h->opt_exchange_block_screen <= 0 || fn_mask == nullptr) return;
    int nblk = (int)std::min<int64_t>(h->opt_exchange_block_screen, N);
    int width;
    if (N < 100) { width = (int)N; nblk = 1; }
    else width = (int)(N / nblk);
    L.xs_nblk = nblk;
    L.xs_width = width;
    L.xs_blocks_total = (int64_t)nblk * (nblk + 1) / 2;
    L.xs_blk_of.assign((size_t)N, -1);
    for (int64_t p = 0; p < (int64_t)width * nblk; p++) L.xs_blk_of[(size_t)p] = (int32_t)(p / width);
    L.xs_keep.assign((size_t)nblk * nblk, 0);
    for (int bp = 0; bp < nblk; bp++)
        for (int bq = 0; bq <= bp; bq++) {
            bool any = false;
            for (int64_t c = (int64_t)bq * width; c < (int64_t)(bq + 1) * width && !any; c++)
                for (int64_t r = (int64_t)bp * width; r < (int64_t)(bp + 1) * width; r++)
                    if (fn_mask[r + N * c]) { any = true; break; }
            L.xs_keep[(size_t)bp * nblk + bq] = L.xs_keep[(size_t)bq * nblk + bp] = any ? 1 : 0;
            L.xs_blocks_kept += any;
        }
    L.xs_any_dropped = L.xs_blocks_kept < L.xs_blocks_total;
    // exchange tiles (internal order) whose every in-range element lies in a dropped block
    const int n128 = L.n128;
    L.xs_tile_skip.assign((size_t)n128 * (n128 + 1) / 2, 0);
    if (!L.xs_any_dropped) return;
    L.xs_half_dead.assign((size_t)n128 * (n128 + 1) / 2, 0);
    // distinct K blocks of a run of internal rows / columns (caller-order block of every function in it); strip = a function
    // of the never-screened remainder, which keeps everything it touches alive
    auto blocks_in = [&](int64_t i0, int64_t i1, bool& strip) {
        std::vector<int32_t> out;
        std::vector<uint8_t> seen((size_t)nblk, 0);
        strip = false;
        for (int64_t i = i0; i < std::min<int64_t>(N, i1); i++) {
            const int32_t b = L.xs_blk_of[(size_t)L.inv[(size_t)i]];
            if (b < 0) strip = true;
            else if (!seen[(size_t)b]) { seen[(size_t)b] = 1; out.push_back(b); }
        }
        return out;
    };
    auto all_dropped = [&](const std::vector<int32_t>& rows, bool rstrip, const std::vector<int32_t>& cols, bool cstrip) {
        if (rstrip || cstrip) return false;
        for (int32_t bp : rows)
            for (int32_t bq : cols)
                if (L.xs_keep[(size_t)bp * nblk + bq]) return false;
        return true;
    };
    for (int ta = 0; ta < n128; ta++)
        for (int tb = 0; tb <= ta; tb++) {
            const size_t tile = (size_t)ta * (ta + 1) / 2 + tb;
            bool rs, cs;
            const std::vector<int32_t> rows = blocks_in(128 * (int64_t)ta, 128 * (int64_t)(ta + 1), rs);
            const std::vector<int32_t> cols = blocks_in(128 * (int64_t)tb, 128 * (int64_t)(tb + 1), cs);
            if (all_dropped(rows, rs, cols, cs)) {
                L.xs_tile_skip[tile] = 1;
                L.xs_tiles_skipped++;
                continue;
            }
            // the two row halves of an off-diagonal tile (the rows the warps w < 4 / w >= 4 own, kernels_k.cu)
            for (int half = 0; half < 2 && ta != tb; half++) {
                int r_begin, r_end;
                exchange_half_rows(ta, tb, half, (int)N, r_begin, r_end);
                if (r_end <= r_begin) continue;
                bool hs;
                const std::vector<int32_t> hrows = blocks_in(128 * (int64_t)ta + r_begin, 128 * (int64_t)ta + r_end, hs);
                if (all_dropped(hrows, hs, cols, cs)) {
                    L.xs_half_dead[tile] |= (uint8_t)(1u << half);
                    L.xs_halves_skipped++;
                }
            }
        }
}

// Split of the exchange contraction index: items = tiles x slices dealt round-robin to one CTA per SM.
static void plan_exchange(const Layout& L, Shard& s, int sm_count, bool balance, int64_t forced_slices) {
    ExchangePlan& x = s.xplan;
    x.chunks = (int)((s.q_len * L.occ_w + 31) / 32);
    x.n_tiles = L.n128 * (L.n128 + 1) / 2;
    int max_slices = std::max(1, x.chunks / 8);
    const int live_tiles = std::max(1, x.n_tiles - (int)L.xs_tiles_skipped);  // block-screened exchange: tiles with work
    int want = std::max(1, (12 * sm_count + live_tiles - 1) / live_tiles);
    int s0 = std::min(want, max_slices);
    int best = s0;
    double best_eff = -1;
    for (int c = s0; c <= std::min(max_slices, s0 + 48); c++) {
        int64_t items = (int64_t)live_tiles * c;
        int64_t waves = (items + sm_count - 1) / sm_count;
        double eff = (double)items / (double)(waves * sm_count);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = c; }
    }
    if (forced_slices > 0) best = (int)std::min<int64_t>(forced_slices, std::max(1, x.chunks));  // option "exchange_slices"
    x.chunks_per_slice = (x.chunks + best - 1) / best;
    if (x.chunks_per_slice < 1) x.chunks_per_slice = 1;
    x.n_slices = std::max(1, (x.chunks + x.chunks_per_slice - 1) / x.chunks_per_slice);
    // Items cost differently (diagonal tiles issue 10/16 of a full tile's DMMAs, the last tile row is clipped to N, the
    // last slice may be short): deal them to the persistent CTAs longest-first onto the least-loaded CTA.
    const int n_items = x.n_tiles * x.n_slices;
    auto skipped = [&](int it) { return !L.xs_tile_skip.empty() && L.xs_tile_skip[(size_t)(it % x.n_tiles)] != 0; };
    x.n_ctas = std::max(1, std::min(live_tiles * x.n_slices, sm_count));
    std::vector<int64_t> cost((size_t)n_items);
    std::vector<int> tile_cost((size_t)x.n_tiles);
    for (int t = 0; t < x.n_tiles; t++)
        tile_cost[(size_t)t] = exchange_tile_cost(t, (int)L.N, L.xs_half_dead.empty() ? 0 : L.xs_half_dead[(size_t)t]);
    for (int it = 0; it < n_items; it++) {
        const int slice = it / x.n_tiles;
        const int kc0 = slice * x.chunks_per_slice, kc1 = std::min(x.chunks, kc0 + x.chunks_per_slice);
        cost[(size_t)it] = (int64_t)tile_cost[(size_t)(it % x.n_tiles)] * std::max(0, kc1 - kc0) + 6 * 32;  // + epilogue
    }
    std::vector<int32_t> order;
    order.reserve((size_t)n_items);
    for (int it = 0; it < n_items; it++)
        if (!skipped(it)) order.push_back(it);  // a skipped tile's Kpart stays at its initial zeros and is never read
    std::vector<int64_t> load((size_t)x.n_ctas, 0);
    std::vector<std::vector<int32_t>> bins((size_t)x.n_ctas);
    if (!balance) {  // plain round-robin in item order
        for (size_t k = 0; k < order.size(); k++) bins[k % (size_t)x.n_ctas].push_back(order[k]);
        order.clear();
    }
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return cost[(size_t)a] > cost[(size_t)b]; });
    for (int32_t it : order) {
        int best_c = 0;
        for (int c = 1; c < x.n_ctas; c++)
            if (load[(size_t)c] < load[(size_t)best_c]) best_c = c;
        load[(size_t)best_c] += cost[(size_t)it];
        bins[(size_t)best_c].push_back(it);
    }
    x.item_list.clear();
    x.item_off.assign((size_t)x.n_ctas + 1, 0);
    for (int c = 0; c < x.n_ctas; c++) {
        // ascending item id inside a CTA: consecutive slices of a tile keep its W columns warm in L2
        std::sort(bins[(size_t)c].begin(), bins[(size_t)c].end());
        x.item_off[(size_t)c] = (int32_t)x.item_list.size();
        x.item_list.insert(x.item_list.end(), bins[(size_t)c].begin(), bins[(size_t)c].end());
    }
    x.item_off[(size_t)x.n_ctas] = (int32_t)x.item_list.size();
}

static int alloc_shard_buffers(jcdf_handle* h, Shard& s) {
    Layout& L = h->lay;
    JC_CUDA(cudaSetDevice(s.dev));
    plan_exchange(L, s, h->sm_count, h->opt_exchange_balance != 0, h->opt_exchange_slices);
    int rc;
    const size_t nB = (size_t)s.q_len * (size_t)L.row_len;
    if ((rc = dev_alloc(&s.Bp, nB))) return rc;
    JC_CUDA(cudaMemsetAsync(s.Bp, 0, std::max<size_t>(nB, 1) * sizeof(double), s.stream));
    const size_t nW = (size_t)s.xplan.chunks * L.npt64 * 2048;
    if ((rc = dev_alloc(&s.W, nW))) return rc;
    JC_CUDA(cudaMemsetAsync(s.W, 0, std::max<size_t>(nW, 1) * sizeof(double), s.stream));
    if ((rc = dev_alloc(&s.Cs, (size_t)L.nb * L.n_mt * 32 * L.bm))) return rc;
    if ((rc = dev_alloc(&s.Craw, (size_t)L.N * L.occ))) return rc;
    if ((rc = dev_alloc(&s.dt, (size_t)L.row_len))) return rc;
    if ((rc = dev_alloc(&s.Jt, (size_t)L.row_len))) return rc;
    JC_CUDA(cudaMemsetAsync(s.Jt, 0, std::max<size_t>((size_t)L.row_len, 1) * sizeof(double), s.stream));
    if ((rc = dev_alloc(&s.V, (size_t)std::max<int64_t>(s.q_len, 1)))) return rc;
    if ((rc = dev_alloc(&s.Vpart, (size_t)std::max<int64_t>(s.q_len, 1) * h->sm_count))) return rc;
    if ((rc = dev_alloc(&s.Vready, (size_t)std::max<int64_t>(s.q_len, 1)))) return rc;
    if ((rc = dev_alloc(&s.Vt, (size_t)std::max<int64_t>(s.q_len, 1) * h->sm_count * 2))) return rc;
    if ((rc = dev_alloc(&s.Kpart, (size_t)s.xplan.n_tiles * s.xplan.n_slices * 128 * 128))) return rc;
    // Kpart is zeroed once: blocks no warp owns (above the diagonal, beyond N) are never written but are summed by
    // k_exchange_reduce, and uninitialised memory could hold NaN patterns
    JC_CUDA(cudaMemsetAsync(s.Kpart, 0, (size_t)s.xplan.n_tiles * s.xplan.n_slices * 128 * 128 * sizeof(double), s.stream));
    if ((rc = dev_alloc(&s.Ksum, (size_t)s.xplan.n_tiles * 128 * 128))) return rc;
    if ((rc = dev_alloc(&s.d_xitems, s.xplan.item_list.size()))) return rc;
    if ((rc = dev_alloc(&s.d_xoff, s.xplan.item_off.size()))) return rc;
    if ((rc = dev_alloc(&s.d_xcycles, (size_t)std::max(1, s.xplan.n_ctas)))) return rc;
    JC_CUDA(cudaMemcpyAsync(s.d_xitems, s.xplan.item_list.data(), sizeof(int32_t) * s.xplan.item_list.size(), cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_xoff, s.xplan.item_off.data(), sizeof(int32_t) * s.xplan.item_off.size(), cudaMemcpyHostToDevice, s.stream));
    if ((rc = dev_alloc(&s.F, (size_t)L.N * L.N))) return rc;
    if ((rc = dev_alloc(&s.Hdev, (size_t)L.N * L.N))) return rc;
    if (L.xs_any_dropped) {  // block-screened exchange: the caller's F_out as it arrived + the block tables of the epilogue
        if ((rc = dev_alloc(&s.Fprev, (size_t)L.N * L.N))) return rc;
        if ((rc = dev_alloc(&s.d_xs_blk_of, L.xs_blk_of.size()))) return rc;
        if ((rc = dev_alloc(&s.d_xs_keep, L.xs_keep.size()))) return rc;
        if ((rc = dev_alloc(&s.d_xs_half_dead, L.xs_half_dead.size()))) return rc;
        JC_CUDA(cudaMemcpyAsync(s.d_xs_half_dead, L.xs_half_dead.data(), L.xs_half_dead.size(), cudaMemcpyHostToDevice, s.stream));
        JC_CUDA(cudaMemsetAsync(s.Fprev, 0, sizeof(double) * L.N * L.N, s.stream));
        JC_CUDA(cudaMemcpyAsync(s.d_xs_blk_of, L.xs_blk_of.data(), sizeof(int32_t) * L.xs_blk_of.size(), cudaMemcpyHostToDevice, s.stream));
        JC_CUDA(cudaMemcpyAsync(s.d_xs_keep, L.xs_keep.data(), L.xs_keep.size(), cudaMemcpyHostToDevice, s.stream));
    }
    if ((rc = dev_alloc(&s.d_tile_index, (size_t)L.nb * L.nb))) return rc;
    if ((rc = dev_alloc(&s.d_tile_a, (size_t)L.tiles_kept))) return rc;
    if ((rc = dev_alloc(&s.d_tile_b, (size_t)L.tiles_kept))) return rc;
    if ((rc = dev_alloc(&s.d_dest_of_col, (size_t)L.pairs_kept))) return rc;
    if ((rc = dev_alloc(&s.d_perm, (size_t)L.n_pad))) return rc;
    if ((rc = dev_alloc(&s.d_inv, (size_t)L.n_pad))) return rc;
    if ((rc = dev_alloc(&s.d_tile_strips, (size_t)L.tiles_kept))) return rc;
    JC_CUDA(cudaMemcpyAsync(s.d_perm, L.perm.data(), sizeof(int32_t) * L.perm.size(), cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_inv, L.inv.data(), sizeof(int32_t) * L.inv.size(), cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_tile_strips, L.tile_strips.data(), sizeof(int32_t) * L.tile_strips.size(), cudaMemcpyHostToDevice, s.stream));
    if ((rc = dev_alloc(&s.d_kb_list, (size_t)L.nb * L.nb))) return rc;
    if ((rc = dev_alloc(&s.d_kb_count, (size_t)L.nb))) return rc;
    JC_CUDA(cudaMemcpyAsync(s.d_kb_list, L.kb_list.data(), sizeof(int32_t) * L.kb_list.size(), cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_kb_count, L.kb_count.data(), sizeof(int32_t) * L.kb_count.size(), cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_tile_index, L.tile_index.data(), sizeof(int32_t) * L.tile_index.size(),
                            cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_tile_a, L.tile_a.data(), sizeof(int32_t) * L.tile_a.size(), cudaMemcpyHostToDevice,
                            s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_tile_b, L.tile_b.data(), sizeof(int32_t) * L.tile_b.size(), cudaMemcpyHostToDevice,
                            s.stream));
    JC_CUDA(cudaMemcpyAsync(s.d_dest_of_col, L.dest_of_col.data(), sizeof(int64_t) * L.dest_of_col.size(),
                            cudaMemcpyHostToDevice, s.stream));
    if (L.pp_enabled) {
        // partner-compacted copy of the shard for the per-function half-transform (kernels_pp.cu)
        s.pp_chunks = (int)((s.q_len + 127) / 128);
        const size_t nBc = (size_t)s.pp_chunks * (size_t)L.pp_rows * 128;
        cudaError_t e = cudaErrorMemoryAllocation;
        if (h->opt_pp_max_bytes <= 0 || (int64_t)(nBc * sizeof(double)) <= h->opt_pp_max_bytes)
            e = cudaMalloc(reinterpret_cast<void**>(&s.Bc), std::max<size_t>(nBc, 1) * sizeof(double));
        if (e != cudaSuccess) {
            cudaGetLastError();
            s.Bc = nullptr;
            if (h->opt_ht_mode == 2) {
                set_error("out of device memory for the partner-compacted copy of B (ht_mode = 2)");
                return JCDF_ERR_CUDA;
            }
            L.pp_enabled = false;  // automatic choice: fall back to the tile kernel on every shard
        } else {
            const int ldci = L.pp_n_nt * pp_item_orbitals(L);
            if ((rc = dev_alloc(&s.Ci, (size_t)(L.n_pad + 1) * ldci))) return rc;
            if ((rc = dev_alloc(&s.d_pp_off, L.pp_off.size()))) return rc;
            if ((rc = dev_alloc(&s.d_pp_q, L.pp_q.size()))) return rc;
            if ((rc = dev_alloc(&s.d_pp_src, L.pp_src.size()))) return rc;
            JC_CUDA(cudaMemcpyAsync(s.d_pp_off, L.pp_off.data(), sizeof(int32_t) * L.pp_off.size(), cudaMemcpyHostToDevice, s.stream));
            JC_CUDA(cudaMemcpyAsync(s.d_pp_q, L.pp_q.data(), sizeof(int32_t) * L.pp_q.size(), cudaMemcpyHostToDevice, s.stream));
            JC_CUDA(cudaMemcpyAsync(s.d_pp_src, L.pp_src.data(), sizeof(int32_t) * L.pp_src.size(), cudaMemcpyHostToDevice, s.stream));
            if ((rc = dev_alloc(&s.d_pp_p, L.pp_p.size()))) return rc;
            if ((rc = dev_alloc(&s.d_pp_jdst, L.pp_jdst.size()))) return rc;
            JC_CUDA(cudaMemcpyAsync(s.d_pp_p, L.pp_p.data(), sizeof(int32_t) * L.pp_p.size(), cudaMemcpyHostToDevice, s.stream));
            JC_CUDA(cudaMemcpyAsync(s.d_pp_jdst, L.pp_jdst.data(), sizeof(int32_t) * L.pp_jdst.size(), cudaMemcpyHostToDevice, s.stream));
            {
                std::vector<int32_t> list, off;
                plan_pp_items(L, s, h->sm_count, h->opt_pp_order == 1, list, off);
                if ((rc = dev_alloc(&s.d_ppitems, list.size()))) return rc;
                if ((rc = dev_alloc(&s.d_ppoff, off.size()))) return rc;
                JC_CUDA(cudaMemcpy(s.d_ppitems, list.data(), sizeof(int32_t) * list.size(), cudaMemcpyHostToDevice));
                JC_CUDA(cudaMemcpy(s.d_ppoff, off.data(), sizeof(int32_t) * off.size(), cudaMemcpyHostToDevice));
            }
            if ((rc = dev_alloc(&s.dvec, (size_t)L.pp_rows))) return rc;
            if ((rc = dev_alloc(&s.d_pp_trirow, L.pp_trirow.size()))) return rc;
            JC_CUDA(cudaMemcpyAsync(s.d_pp_trirow, L.pp_trirow.data(), sizeof(int32_t) * L.pp_trirow.size(), cudaMemcpyHostToDevice, s.stream));
            if ((rc = dev_alloc(&s.ppj_flag, (size_t)s.pp_chunks * 4))) return rc;
            if ((rc = dev_alloc(&s.ppj_vp, (size_t)s.pp_chunks * 4 * h->sm_count * 32))) return rc;
            if ((rc = dev_alloc(&s.VpartP, (size_t)pp_coulomb_slabs(L) * s.pp_chunks * 128))) return rc;
            if ((rc = dev_alloc(&s.Jpart, (size_t)s.pp_chunks * L.pp_rows))) return rc;
            {   // partial V per (CTA, chunk, warp column); pairs a CTA's list never visits are never written: cleared once
                const size_t nv = (size_t)std::max(1, s.pp_ctas) * s.pp_chunks * pp_v_slots(L) * 128;
                if ((rc = dev_alloc(&s.Vitem, nv))) return rc;
                JC_CUDA(cudaMemsetAsync(s.Vitem, 0, nv * sizeof(double), s.stream));
            }
        }
    }
    s.bc_valid = false;
    JC_CUDA(cudaStreamSynchronize(s.stream));
    return JCDF_OK;
}

static int default_shards(jcdf_handle* h) {
    const int nd = (int)h->shards.size();
    const int64_t Q = h->lay.Q;
    for (int d = 0; d < nd; d++) {
        h->shards[d].q_begin = Q * d / nd;
        h->shards[d].q_len = Q * (d + 1) / nd - Q * d / nd;
    }
    return JCDF_OK;
}

static int ensure_comm(jcdf_handle* h) {
    if (h->comm_ready) return JCDF_OK;
    const int nd = (int)h->shards.size();
    if (h->n_ranks == 1 && nd == 1) { h->comm_ready = true; return JCDF_OK; }
    const NcclApi& n = nccl();
    if (!n.ok) { set_error(std::string("NCCL unavailable: ") + n.why); return JCDF_ERR_NCCL; }
    if (h->n_ranks == 1) {
        std::vector<int> devs(nd);
        std::vector<void*> comms(nd, nullptr);
        for (int d = 0; d < nd; d++) devs[d] = h->shards[d].dev;
        JC_NCCL(n.CommInitAll(comms.data(), nd, devs.data()));
        for (int d = 0; d < nd; d++) h->shards[d].nccl_comm = comms[d];
        h->comm_ready = true;
        return JCDF_OK;
    }
    set_error("multi-rank communicator not initialised: call jcdf_comm_init_rank first");
    return JCDF_ERR_INVALID;
}

// potrf!('L') of s.Metric (Q x Q, lower triangle read) in place and s.Linv = L^-1 by a triangular solve against the
// identity (same result as trtri!('L','N'), DensityFitting.jl:138-139) on the shard's device and stream.
int factor_metric_on_device(Shard& s, int64_t Q) {
    struct Guard {  // released on every exit path
        cusolverDnHandle_t sol = nullptr;
        cublasHandle_t blas = nullptr;
        double* work = nullptr;
        int* dinfo = nullptr;
        ~Guard() {
            if (work) cudaFree(work);
            if (dinfo) cudaFree(dinfo);
            if (sol) cusolverDnDestroy(sol);
            if (blas) cublasDestroy(blas);
        }
    } g;
    JC_CUSOLVER(cusolverDnCreate(&g.sol));
    JC_CUSOLVER(cusolverDnSetStream(g.sol, s.stream));
    JC_CUBLAS(cublasCreate(&g.blas));
    JC_CUBLAS(cublasSetStream(g.blas, s.stream));
    int lwork = 0, rc;
    JC_CUSOLVER(cusolverDnDpotrf_bufferSize(g.sol, CUBLAS_FILL_MODE_LOWER, (int)Q, s.Metric, (int)Q, &lwork));
    if ((rc = dev_alloc(&g.work, (size_t)lwork))) return rc;
    if ((rc = dev_alloc(&g.dinfo, 1))) return rc;
    JC_CUSOLVER(cusolverDnDpotrf(g.sol, CUBLAS_FILL_MODE_LOWER, (int)Q, s.Metric, (int)Q, g.work, lwork, g.dinfo));
    int info = 0;
    JC_CUDA(cudaMemcpyAsync(&info, g.dinfo, sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    JC_CUDA(cudaStreamSynchronize(s.stream));
    if (info != 0) {
        set_error("two-centre metric is not positive definite (potrf info = " + std::to_string(info) + ")");
        return JCDF_ERR_NUMERIC;
    }
    launch_set_identity(s.Linv, Q, s.stream);
    const double one = 1.0;
    JC_CUBLAS(cublasDtrsm(g.blas, CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, CUBLAS_DIAG_NON_UNIT, (int)Q, (int)Q,
                          &one, s.Metric, (int)Q, s.Linv, (int)Q));
    JC_CUDA(cudaStreamSynchronize(s.stream));
    return JCDF_OK;
}

// Mask-driven regrouping of the basis functions into 32-wide blocks (internal order only; every C-ABI argument stays
// in the caller's order).  The Schwarz mask of a real molecule is sparse at the function-pair level, but in the input
// order (atoms in file order, tight and diffuse shells interleaved) nearly every 32 x 32 tile holds some surviving pair.
// Functions with similar partner sets — the same neighbourhood and a similar extent — are put in the same block:
// a block is grown from the unassigned function of smallest degree by repeatedly adding the function that enlarges the
// union of the block's partner set least (ties: largest overlap).  Rows are bitsets; O(N^3 / 64) word operations.
static void regroup_functions(const uint8_t* fn_mask, int64_t N, std::vector<int32_t>& inv_out) {
    const int64_t nw = (N + 63) / 64;
    std::vector<uint64_t> rows((size_t)(N * nw), 0);
    std::vector<int32_t> deg((size_t)N, 0);
    for (int64_t c = 0; c < N; c++)
        for (int64_t r = 0; r < N; r++)
            if (fn_mask[r + N * c]) {
                rows[(size_t)(r * nw + (c >> 6))] |= 1ull << (c & 63);
                deg[(size_t)r]++;
            }
    std::vector<uint8_t> used((size_t)N, 0);
    std::vector<uint64_t> uni((size_t)nw);
    inv_out.clear();
    inv_out.reserve((size_t)N);
    int64_t remaining = N;
    while (remaining > 0) {
        int64_t seed = -1;
        for (int64_t i = 0; i < N; i++)
            if (!used[(size_t)i] && (seed < 0 || deg[(size_t)i] < deg[(size_t)seed])) seed = i;
        used[(size_t)seed] = 1;
        remaining--;
        inv_out.push_back((int32_t)seed);
        for (int64_t w = 0; w < nw; w++) uni[(size_t)w] = rows[(size_t)(seed * nw + w)];
        for (int k = 1; k < 32 && remaining > 0; k++) {
            int64_t best = -1;
            int best_add = 0, best_ov = 0;
            for (int64_t i = 0; i < N; i++) {
                if (used[(size_t)i]) continue;
                int add = 0, ov = 0;
                const uint64_t* ri = &rows[(size_t)(i * nw)];
                for (int64_t w = 0; w < nw; w++) {
                    add += __builtin_popcountll(ri[w] & ~uni[(size_t)w]);
                    ov += __builtin_popcountll(ri[w] & uni[(size_t)w]);
                }
                if (best < 0 || add < best_add || (add == best_add && ov > best_ov)) { best = i; best_add = add; best_ov = ov; }
            }
            used[(size_t)best] = 1;
            remaining--;
            inv_out.push_back((int32_t)best);
            for (int64_t w = 0; w < nw; w++) uni[(size_t)w] |= rows[(size_t)(best * nw + w)];
        }
    }
}

// Partner lists of the per-function half-transform (kernels_pp.cu): for every function p, in internal order, the kept
// partners q ascending, padded to a multiple of four rows, with the offset of B[Q, p, q] inside a tile-packed row.
// Chooses between the two half-transform kernels from the DMMA k-steps each would execute on this mask.
template <typename Kept>
static void build_partner_lists(jcdf_handle* h, Kept kept) {
    Layout& L = h->lay;
    const int64_t N = L.N;
    const int nb = L.nb;
    L.pp_enabled = false;
    L.pp_off.clear(); L.pp_q.clear(); L.pp_src.clear(); L.pp_p.clear(); L.pp_jdst.clear(); L.pp_trirow.clear(); L.pp_n_tri = 0;
    L.pp_rows = 0;
    // occupied-orbital tiling.  Layout 0: 4 consumer warps across n, each 8 * nt wide (nt <= 4) — the orbital count of
    // an item is padded to a multiple of 32.  Layout 1: every warp takes all 8 * nt orbitals of the item (nt <= 12) — no
    // padding beyond the 8-wide DMMA tile; taken when it executes fewer tiles on the busiest sub-partition.
    const int n8 = L.occ_pad / 8;
    const int n_nt0 = (n8 + 15) / 16, nt0 = ((n8 + n_nt0 - 1) / n_nt0 + 3) / 4;
    const int n_nt1 = (n8 + 11) / 12, nt1 = (n8 + n_nt1 - 1) / n_nt1;
    const bool layout1 = h->opt_pp_layout == 1 || (h->opt_pp_layout < 0 && n_nt1 * nt1 * 100 < n_nt0 * 4 * nt0 * 97);
    L.pp_layout = layout1 ? 1 : 0;
    L.pp_n_nt = layout1 ? n_nt1 : n_nt0;
    L.pp_nt = layout1 ? nt1 : nt0;
    if (h->opt_pp_layout == 2) {
        // Layout 2: two CTAs per SM, each 2 (M halves) x 2 (N halves of 8 nt orbitals, nt <= 4) consumer warps: items of
        // 16 nt orbitals; the two CTAs of an SM run out of phase, so one's item epilogue / pipeline fill sits under the other's DMMAs
        const int n_nt2 = (n8 + 7) / 8;
        L.pp_layout = 2;
        L.pp_n_nt = n_nt2;
        L.pp_nt = ((n8 + n_nt2 - 1) / n_nt2 + 1) / 2;
    }
    if (h->opt_ht_mode == 1 || L.row_len >= ((int64_t)1 << 31)) return;
    // executed multiply-adds per auxiliary row of either kernel
    double cost_tiles = 0.0;
    for (int pb = 0; pb < nb; pb++)
        for (int x = 0; x < L.kb_count[pb]; x++) {
            const int kb = L.kb_list[(size_t)pb * nb + x];
            const uint32_t sm = (uint32_t)L.tile_strips[(size_t)L.tile_index[(size_t)std::max(pb, kb) * nb + std::min(pb, kb)]];
            cost_tiles += __builtin_popcount(kb >= pb ? (sm & 0xFFu) : ((sm >> 8) & 0xFFu));
        }
    cost_tiles *= 4.0 * 32.0 * (double)L.bm * L.n_mt;
    int64_t rows = 0;
    std::vector<int32_t> cnt((size_t)N, 0);
    for (int64_t p = 0; p < N; p++) {
        int c = 0;
        for (int64_t q = 0; q < N; q++) c += kept(p, q) ? 1 : 0;
        cnt[(size_t)p] = c;
        rows += (c + 3) / 4 * 4;
    }
    const double cost_pp = (double)rows * L.pp_n_nt * (double)pp_item_orbitals(L);
    if (rows >= ((int64_t)1 << 31)) return;
    // measured on the C20H42 shape (52.8 % of the pairs kept, cost ratio 0.93): per-function 18.2 ms + compact Coulomb
    // 4.4 ms against tiles 19.3 ms + 5.3 ms — the compact family wins as soon as it executes fewer DMMAs at all
    if (h->opt_ht_mode != 2 && !(cost_pp < 0.97 * cost_tiles)) return;
    L.pp_rows = rows;
    L.pp_off.resize((size_t)N + 1);
    L.pp_q.reserve((size_t)rows);
    L.pp_src.reserve((size_t)rows);
    L.pp_p.reserve((size_t)rows);
    L.pp_jdst.reserve((size_t)rows);
    for (int64_t p = 0; p < N; p++) {
        L.pp_off[(size_t)p] = (int32_t)L.pp_q.size();
        for (int64_t q = 0; q < N; q++) {
            if (!kept(p, q)) continue;
            const int ap = (int)(p / 32), aq = (int)(q / 32);
            int64_t off;
            if (ap >= aq) {
                const int rr = (int)(p % 32), cc = (int)(q % 32);
                off = (int64_t)L.tile_index[(size_t)ap * nb + aq] * 1024 + rr * 32 + (cc ^ ((rr & 3) << 2));
            } else {
                const int rr = (int)(q % 32), cc = (int)(p % 32);
                off = (int64_t)L.tile_index[(size_t)aq * nb + ap] * 1024 + rr * 32 + (cc ^ ((rr & 3) << 2));
            }
            L.pp_q.push_back((int32_t)q);
            L.pp_src.push_back((int32_t)off);
            L.pp_p.push_back((int32_t)p);
            // Coulomb half: rows with q >= p; the epilogue reads J at (row hi = q, column lo = p) of tile (aq, ap)
            int64_t jd = -1;
            if (q >= p) {
                const int rr = (int)(q % 32), cc = (int)(p % 32);
                jd = (int64_t)L.tile_index[(size_t)aq * nb + ap] * 1024 + rr * 32 + (cc ^ ((rr & 3) << 2));
            }
            if (jd >= 0) L.pp_trirow.push_back((int32_t)L.pp_jdst.size());
            L.pp_jdst.push_back((int32_t)jd);
        }
        while (L.pp_q.size() % 4) {
            L.pp_q.push_back((int32_t)L.n_pad);
            L.pp_src.push_back(-1);
            L.pp_p.push_back((int32_t)p);
            L.pp_jdst.push_back(-1);
        }
    }
    L.pp_off[(size_t)N] = (int32_t)L.pp_q.size();
    L.pp_n_tri = (int64_t)L.pp_trirow.size();
    L.pp_enabled = true;
}

}  // namespace jcdf

using namespace jcdf;

extern "C" {

const char* jcdf_last_error(void) { return g_err.c_str(); }

int32_t jcdf_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        set_error(std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
        return JCDF_ERR_CUDA;
    }
    return n;
}

int32_t jcdf_create(int64_t N, int64_t Q, int64_t occ, int32_t ndev, const int32_t* dev_ids, jcdf_handle** out) {
    JC_REQUIRE(out != nullptr, "jcdf_create: out is NULL");
    *out = nullptr;
    JC_REQUIRE(N > 0 && Q > 0 && occ > 0 && occ <= N, "jcdf_create: need N > 0, Q > 0, 0 < occ <= N");
    JC_REQUIRE(N < (1 << 21) && Q < (1 << 21), "jcdf_create: N and Q must be below 2^21");
    JC_REQUIRE(ndev >= 1, "jcdf_create: ndev must be >= 1");
    int avail = 0;
    JC_CUDA(cudaGetDeviceCount(&avail));
    // mirrors error("Number of devices requested is greater than the number of devices available") GPUDF.jl:39-41
    JC_REQUIRE(ndev <= avail, "Number of devices requested is greater than the number of devices available");
    jcdf_handle* h = new (std::nothrow) jcdf_handle();
    JC_REQUIRE(h != nullptr, "jcdf_create: out of host memory");
    Layout& L = h->lay;
    L.N = N; L.Q = Q; L.occ = occ;
    L.nb = (int)((N + 31) / 32);
    L.n_pad = 32 * L.nb;
    L.occ_pad = (int)round_up(occ, 8);
    L.occ_w = (int)round_up(occ, 2);
    L.n_mt = (L.occ_pad + 127) / 128;
    L.bm = (int)round_up((L.occ_pad + L.n_mt - 1) / L.n_mt, 16);
    L.npt64 = (L.n_pad + 63) / 64;
    L.n_pt128 = (L.nb + 3) / 4;
    L.n128 = (L.npt64 + 1) / 2;
    h->shards.resize(ndev);
    for (int d = 0; d < ndev; d++) {
        Shard& s = h->shards[d];
        s.dev = dev_ids ? dev_ids[d] : d;
        if (s.dev < 0 || s.dev >= avail) {
            delete h;
            set_error("jcdf_create: device id out of range");
            return JCDF_ERR_INVALID;
        }
        cudaError_t e = cudaSetDevice(s.dev);
        int prio_lo = 0, prio_hi = 0;
        if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
        if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, prio_hi);
        if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&s.stream_j, cudaStreamNonBlocking, prio_lo);
        for (int i = 0; i < EV_COUNT && e == cudaSuccess; i++) e = cudaEventCreate(&s.ev[i]);
        for (int i = 0; i < 3 && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&s.ev_dep[i], cudaEventDisableTiming);
        if (e != cudaSuccess) {
            set_error(std::string("jcdf_create: ") + cudaGetErrorString(e));
            delete h;
            return JCDF_ERR_CUDA;
        }
    }
    cudaDeviceProp prop;
    JC_CUDA(cudaGetDeviceProperties(&prop, h->shards[0].dev));
    h->sm_count = prop.multiProcessorCount;
    if (prop.major < 10) {
        delete h;
        set_error("libjcdf is built for sm_100a (Blackwell B200) only; device compute capability is " +
                  std::to_string(prop.major) + "." + std::to_string(prop.minor));
        return JCDF_ERR_CUDA;
    }
    default_shards(h);
    *out = h;
    return JCDF_OK;
}

int32_t jcdf_destroy(jcdf_handle* h) {
    if (!h) return JCDF_OK;
    scf_destroy(h);
    integrals_destroy(h);
    delete h->eigh;
    h->eigh = nullptr;
    if (h->fout_registered) {
        cudaHostUnregister(h->fout_registered);
        cudaGetLastError();
        h->fout_registered = nullptr;
    }
    for (Shard& s : h->shards) {
        cudaSetDevice(s.dev);
        if (s.stream) cudaStreamSynchronize(s.stream);
        if (s.nccl_comm && nccl().ok) nccl().CommDestroy(s.nccl_comm);
        free_shard_buffers(s);
        for (int i = 0; i < EV_COUNT; i++)
            if (s.ev[i]) cudaEventDestroy(s.ev[i]);
        for (int i = 0; i < 3; i++)
            if (s.ev_dep[i]) cudaEventDestroy(s.ev_dep[i]);
        if (s.stream_j) cudaStreamDestroy(s.stream_j);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    delete h;
    return JCDF_OK;
}

int32_t jcdf_set_shards(jcdf_handle* h, const int64_t* q_begin, const int64_t* q_len) {
    JC_REQUIRE(h && q_begin && q_len, "jcdf_set_shards: NULL argument");
    JC_REQUIRE(!h->lay.mask_set, "jcdf_set_shards must precede jcdf_set_mask");
    for (size_t d = 0; d < h->shards.size(); d++) {
        JC_REQUIRE(q_begin[d] >= 0 && q_len[d] >= 0 && q_begin[d] + q_len[d] <= h->lay.Q,
                   "jcdf_set_shards: shard outside [0, Q)");
        for (size_t e = 0; e < d; e++)  // shards of one process must not overlap (each row would be contracted twice)
            JC_REQUIRE(q_len[d] == 0 || q_len[e] == 0 || q_begin[d] >= q_begin[e] + q_len[e] || q_begin[e] >= q_begin[d] + q_len[d],
                       "jcdf_set_shards: shards overlap");
    }
    for (size_t d = 0; d < h->shards.size(); d++) {
        h->shards[d].q_begin = q_begin[d];
        h->shards[d].q_len = q_len[d];
    }
    h->shards_set = true;
    return JCDF_OK;
}

int32_t jcdf_set_mask(jcdf_handle* h, const uint8_t* fn_mask, const int64_t* map) {
    JC_REQUIRE(h != nullptr, "jcdf_set_mask: NULL handle");
    JC_REQUIRE((fn_mask == nullptr) == (map == nullptr), "jcdf_set_mask: pass both fn_mask and sparse_pq_index_map or neither");
    Layout& L = h->lay;
    const int64_t N = L.N;
    const int nb = L.nb;
    for (Shard& s : h->shards) free_shard_buffers(s, true);
    auto kept_ref = [&](int64_t r, int64_t c) -> bool { return fn_mask ? fn_mask[r + N * c] != 0 : true; };
    int64_t P = 0, Ptri = 0;
    for (int64_t c = 0; c < N; c++)
        for (int64_t r = 0; r < N; r++)
            if (kept_ref(r, c)) {
                P++;
                if (r >= c) Ptri++;
                if (fn_mask && !fn_mask[c + N * r]) {
                    set_error("jcdf_set_mask: basis_function_screen_matrix is not symmetric");
                    return JCDF_ERR_INVALID;
                }
            }
    for (int64_t r = 0; r < N; r++) JC_REQUIRE(kept_ref(r, r), "jcdf_set_mask: a diagonal pair (p,p) is screened");
    // internal function order: identity, or the mask-driven regrouping when it stores fewer tiles
    L.perm.resize((size_t)L.n_pad);
    L.inv.resize((size_t)L.n_pad);
    for (int i = 0; i < L.n_pad; i++) L.perm[i] = L.inv[i] = i;
    auto count_tiles = [&](const std::vector<int32_t>& inv) {
        int64_t n = 0;
        for (int a = 0; a < nb; a++)
            for (int b = 0; b <= a; b++) {
                bool any = false;
                const int r1 = (int)std::min<int64_t>(N, 32 * (a + 1)), c1 = (int)std::min<int64_t>(N, 32 * (b + 1));
                for (int c = 32 * b; c < c1 && !any; c++)
                    for (int r = 32 * a; r < r1; r++)
                        if (kept_ref(inv[r], inv[c])) { any = true; break; }
                n += any;
            }
        return n;
    };
    L.regrouped = false;
    if (fn_mask && P < N * N && h->opt_regroup) {
        std::vector<int32_t> inv2;
        regroup_functions(fn_mask, N, inv2);
        inv2.resize((size_t)L.n_pad);
        for (int i = (int)N; i < L.n_pad; i++) inv2[i] = i;
        if (count_tiles(inv2) < count_tiles(L.inv)) {
            L.inv = inv2;
            for (int i = 0; i < L.n_pad; i++) L.perm[L.inv[i]] = i;
            L.regrouped = true;
        }
    }
    auto kept = [&](int64_t ri, int64_t ci) -> bool { return kept_ref(L.inv[ri], L.inv[ci]); };  // internal coordinates
    // tile table: a tile is stored iff it holds at least one unscreened pair
    L.tile_index.assign((size_t)nb * nb, -1);
    L.tile_a.clear();
    L.tile_b.clear();
    L.tiles_total = (int64_t)nb * (nb + 1) / 2;
    for (int a = 0; a < nb; a++)
        for (int b = 0; b <= a; b++) {
            bool any = false;
            const int64_t r1 = std::min<int64_t>(N, 32 * (a + 1)), c1 = std::min<int64_t>(N, 32 * (b + 1));
            for (int64_t c = 32 * b; c < c1 && !any; c++)
                for (int64_t r = 32 * a; r < r1; r++)
                    if (kept(r, c)) { any = true; break; }
            if (any) {
                L.tile_index[(size_t)a * nb + b] = (int32_t)L.tile_a.size();
                L.tile_a.push_back(a);
                L.tile_b.push_back(b);
            }
        }
    L.tiles_kept = (int64_t)L.tile_a.size();
    L.kb_list.assign((size_t)nb * nb, 0);
    L.kb_count.assign((size_t)nb, 0);
    for (int pb = 0; pb < nb; pb++)
        for (int kb = 0; kb < nb; kb++)
            if (L.tile_index[(size_t)std::max(pb, kb) * nb + std::min(pb, kb)] >= 0) L.kb_list[(size_t)pb * nb + L.kb_count[pb]++] = kb;
    // strips of four rows / four columns of every stored tile that hold an unscreened pair: bits 0-7 row strips,
    // bits 8-15 column strips (the half-transform skips the k-steps of empty strips)
    L.tile_strips.assign((size_t)L.tiles_kept, 0);
    for (int64_t ti = 0; ti < L.tiles_kept; ti++) {
        const int a = L.tile_a[(size_t)ti], b = L.tile_b[(size_t)ti];
        uint32_t bits = 0;
        for (int r = 0; r < 32; r++)
            for (int c = 0; c < 32; c++) {
                const int64_t ri = 32 * a + r, ci = 32 * b + c;
                if (ri < N && ci < N && kept(ri, ci)) bits |= (1u << (r >> 2)) | (1u << (8 + (c >> 2)));
            }
        L.tile_strips[(size_t)ti] = (int32_t)bits;
    }
    L.row_len = L.tiles_kept * 1024;
    L.pairs_kept = P;
    L.pairs_tri = Ptri;
    // reference column -> packed offset.  Column of pair (row r, col c) is map[r + N*c] - 1 (SchwarzScreening.jl:72-81).
    L.dest_of_col.assign((size_t)P, -1);
    L.src_of_col.assign((size_t)P, 0);
    for (int64_t c = 0; c < N; c++)
        for (int64_t r = 0; r < N; r++) {
            if (!kept_ref(r, c)) {
                if (map) JC_REQUIRE(map[r + N * c] == 0, "jcdf_set_mask: index map non-zero on a screened pair");
                continue;
            }
            const int64_t col = map ? map[r + N * c] - 1 : r + N * c;
            JC_REQUIRE(col >= 0 && col < P, "jcdf_set_mask: sparse_pq_index_map entry out of range");
            const int ri = L.perm[r], ci = L.perm[c];
            const int a = ri / 32, b = ci / 32;
            if (a < b) {  // mirror copy; the (c, r) column fills tile (b, a)
                const int32_t tm = L.tile_index[(size_t)b * nb + a];
                const int rm = ci % 32, cm = ri % 32;
                L.src_of_col[(size_t)col] = (int64_t)tm * 1024 + rm * 32 + (cm ^ ((rm & 3) << 2));
                continue;
            }
            const int32_t ti = L.tile_index[(size_t)a * nb + b];
            const int rr = ri % 32, cc = ci % 32;
            L.dest_of_col[(size_t)col] = (int64_t)ti * 1024 + rr * 32 + (cc ^ ((rr & 3) << 2));
            L.src_of_col[(size_t)col] = L.dest_of_col[(size_t)col];
        }
    build_partner_lists(h, [&](int64_t ri, int64_t ci) { return kept(ri, ci); });
    plan_exchange_block_screen(h, fn_mask);
    const bool pp_wanted = L.pp_enabled;
    L.mask_set = true;
    integrals_on_set_mask(h);
    for (Shard& s : h->shards) {
        int rc = alloc_shard_buffers(h, s);
        if (rc) return rc;
        if (fn_mask) {
            JC_CUDA(cudaSetDevice(s.dev));
            if ((rc = dev_alloc(&s.d_fn_mask, (size_t)N * N))) return rc;
            JC_CUDA(cudaMemcpy(s.d_fn_mask, fn_mask, (size_t)N * N, cudaMemcpyHostToDevice));
        }
    }
    if (pp_wanted && !L.pp_enabled) {
        // the compact copy did not fit on some shard and every shard falls back to the tile family: give back what the
        // shards allocated before that one (up to 7.9 GB each at w50)
        for (Shard& s : h->shards) {
            cudaSetDevice(s.dev);
            void** ptrs[] = {(void**)&s.Bc, (void**)&s.Ci, (void**)&s.Vitem, (void**)&s.Jpart, (void**)&s.VpartP, (void**)&s.dvec,
                             (void**)&s.ppj_vp, (void**)&s.d_ppitems, (void**)&s.d_ppoff};
            for (void** q : ptrs)
                if (*q) { cudaFree(*q); *q = nullptr; }
            s.pp_ctas = 0;
        }
    }
    return JCDF_OK;
}

static int upload_window(jcdf_handle* h, int32_t dev, const double* T_dev, int64_t ld, int64_t col_begin,
                         int64_t col_count) {
    Shard& s = h->shards[dev];
    launch_scatter_pairs(h->lay, s, T_dev, ld, col_begin, col_count);
    JC_CUDA(cudaGetLastError());
    JC_CUDA(cudaStreamSynchronize(s.stream));
    return JCDF_OK;
}

static int upload_host(jcdf_handle* h, int32_t dev, const double* T, int64_t ld) {
    JC_REQUIRE(h && T, "upload: NULL argument");
    JC_REQUIRE(h->lay.mask_set, "upload: call jcdf_set_mask first");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "upload: device index out of range");
    Shard& s = h->shards[dev];
    JC_REQUIRE(ld >= s.q_len, "upload: ld smaller than the shard's row count");
    JC_CUDA(cudaSetDevice(s.dev));
    if (s.q_len == 0) return JCDF_OK;
    const int64_t P = h->lay.pairs_kept;
    // staging window of at most 1 GiB
    int64_t win = std::max<int64_t>(1, std::min<int64_t>(P, ((int64_t)1 << 27) / std::max<int64_t>(s.q_len, 1)));
    if (s.traw_capacity < win * s.q_len) {
        if (s.Traw) cudaFree(s.Traw);
        s.Traw = nullptr;
        int rc = dev_alloc(&s.Traw, (size_t)(win * s.q_len));
        if (rc) return rc;
        s.traw_capacity = win * s.q_len;
    }
    for (int64_t c0 = 0; c0 < P; c0 += win) {
        const int64_t nc = std::min(win, P - c0);
        JC_CUDA(cudaMemcpy2DAsync(s.Traw, sizeof(double) * s.q_len, T + ld * c0, sizeof(double) * ld,
                                  sizeof(double) * s.q_len, (size_t)nc, cudaMemcpyHostToDevice, s.stream));
        int rc = upload_window(h, dev, s.Traw, s.q_len, c0, nc);
        if (rc) return rc;
    }
    cudaFree(s.Traw);
    s.Traw = nullptr;
    s.traw_capacity = 0;
    return JCDF_OK;
}

int32_t jcdf_upload_three_center(jcdf_handle* h, int32_t dev, const double* T, int64_t ld) {
    int rc = upload_host(h, dev, T, ld);
    if (rc) return rc;
    h->shards[dev].t_uploaded = true;
    h->shards[dev].b_ready = false;
    h->shards[dev].bc_valid = false;
    return JCDF_OK;
}

int32_t jcdf_upload_three_center_device(jcdf_handle* h, int32_t dev, const double* T_dev, int64_t ld, int64_t col_begin,
                                        int64_t col_count) {
    JC_REQUIRE(h && T_dev, "jcdf_upload_three_center_device: NULL argument");
    JC_REQUIRE(h->lay.mask_set, "jcdf_upload_three_center_device: call jcdf_set_mask first");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "device index out of range");
    JC_REQUIRE(col_begin >= 0 && col_count >= 0 && col_begin + col_count <= h->lay.pairs_kept, "column window out of range");
    JC_CUDA(cudaSetDevice(h->shards[dev].dev));
    JC_CUDA(cudaDeviceSynchronize());  // the caller's producer stream is unknown
    int rc = upload_window(h, dev, T_dev, ld, col_begin, col_count);
    if (rc) return rc;
    h->shards[dev].t_uploaded = true;
    h->shards[dev].b_ready = false;
    h->shards[dev].bc_valid = false;
    return JCDF_OK;
}

int32_t jcdf_set_B(jcdf_handle* h, int32_t dev, const double* B, int64_t ld) {
    int rc = upload_host(h, dev, B, ld);
    if (rc) return rc;
    h->shards[dev].b_ready = true;
    h->shards[dev].bc_valid = false;
    return JCDF_OK;
}

int32_t jcdf_set_B_device(jcdf_handle* h, int32_t dev, const double* B_dev, int64_t ld, int64_t col_begin,
                          int64_t col_count) {
    int rc = jcdf_upload_three_center_device(h, dev, B_dev, ld, col_begin, col_count);
    if (rc) return rc;
    h->shards[dev].b_ready = true;
    h->shards[dev].bc_valid = false;
    return JCDF_OK;
}

int32_t jcdf_get_B_rows(jcdf_handle* h, int32_t dev, int64_t row_begin, int64_t row_count, double* out, int64_t ld) {
    JC_REQUIRE(h && out, "jcdf_get_B_rows: NULL argument");
    JC_REQUIRE(h->lay.mask_set, "jcdf_get_B_rows: call jcdf_set_mask first");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "jcdf_get_B_rows: device index out of range");
    Shard& s = h->shards[dev];
    JC_REQUIRE(row_begin >= 0 && row_count >= 0 && row_begin + row_count <= s.q_len, "jcdf_get_B_rows: rows outside the shard");
    JC_REQUIRE(ld >= row_count, "jcdf_get_B_rows: ld smaller than the row count");
    if (row_count == 0) return JCDF_OK;
    const Layout& L = h->lay;
    const int64_t P = L.pairs_kept;
    JC_CUDA(cudaSetDevice(s.dev));
    // scope guard: the two scratch buffers are released on every exit path
    struct Scratch {
        int64_t* d_src = nullptr;
        double* win_buf = nullptr;
        ~Scratch() { if (d_src) cudaFree(d_src); if (win_buf) cudaFree(win_buf); }
    } sc;
    int rc;
    if ((rc = dev_alloc(&sc.d_src, (size_t)P))) return rc;
    JC_CUDA(cudaMemcpyAsync(sc.d_src, L.src_of_col.data(), sizeof(int64_t) * P, cudaMemcpyHostToDevice, s.stream));
    const int64_t win = std::max<int64_t>(1, std::min<int64_t>(P, ((int64_t)1 << 26) / std::max<int64_t>(row_count, 1)));
    if ((rc = dev_alloc(&sc.win_buf, (size_t)(win * row_count)))) return rc;
    for (int64_t c0 = 0; c0 < P; c0 += win) {
        const int64_t nc = std::min(win, P - c0);
        launch_gather_pairs(L, s, sc.win_buf, row_count, c0, nc, sc.d_src, row_begin, row_count);
        JC_CUDA(cudaGetLastError());
        JC_CUDA(cudaMemcpy2DAsync(out + ld * c0, sizeof(double) * ld, sc.win_buf, sizeof(double) * row_count,
                                  sizeof(double) * row_count, (size_t)nc, cudaMemcpyDeviceToHost, s.stream));
        JC_CUDA(cudaStreamSynchronize(s.stream));
    }
    return JCDF_OK;
}

int32_t jcdf_get_B(jcdf_handle* h, int32_t dev, double* out, int64_t ld) {
    JC_REQUIRE(h && out, "jcdf_get_B: NULL argument");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "jcdf_get_B: device index out of range");
    JC_REQUIRE(ld >= h->shards[dev].q_len, "jcdf_get_B: ld smaller than the shard's row count");
    return jcdf_get_B_rows(h, dev, 0, h->shards[dev].q_len, out, ld);
}

int32_t jcdf_upload_metric(jcdf_handle* h, const double* J_PQ) {
    JC_REQUIRE(h && J_PQ, "jcdf_upload_metric: NULL argument");
    const int64_t Q = h->lay.Q;
    h->host_metric.assign(J_PQ, J_PQ + Q * Q);
    h->metric_on_device = false;
    h->max_PP = 0.0;
    for (int64_t i = 0; i < Q; i++) h->max_PP = std::max(h->max_PP, std::fabs(J_PQ[i + Q * i]));
    return JCDF_OK;
}

// potrf!('L') + trtri!('L','N') on the device, then B_g = sum_{h<=g} Linv[Q_g, Q_h] T_h per packed column window.
int32_t jcdf_factor_and_form_B(jcdf_handle* h) {
    JC_REQUIRE(h != nullptr, "jcdf_factor_and_form_B: NULL handle");
    JC_REQUIRE(h->lay.mask_set, "jcdf_factor_and_form_B: call jcdf_set_mask first");
    JC_REQUIRE(!h->host_metric.empty() || h->metric_on_device,
               "jcdf_factor_and_form_B: call jcdf_upload_metric or jcdf_compute_two_center first");
    JC_REQUIRE(h->n_ranks == 1, "jcdf_factor_and_form_B: multi-process B formation is not implemented; form B per rank and call jcdf_set_B");
    const Layout& L = h->lay;
    const int64_t Q = L.Q;
    const int nd = (int)h->shards.size();
    int64_t covered = 0;
    for (int d = 0; d < nd; d++) {
        JC_REQUIRE(h->shards[d].t_uploaded || h->shards[d].q_len == 0, "jcdf_factor_and_form_B: a shard has no three-centre integrals");
        JC_REQUIRE(h->shards[d].q_begin == covered, "jcdf_factor_and_form_B: shards must tile [0, Q) contiguously in device order");
        covered += h->shards[d].q_len;
    }
    JC_REQUIRE(covered == Q, "jcdf_factor_and_form_B: shards do not cover all auxiliary functions");

    // --- factorisation on the first device -----------------------------------------------------------------------
    Shard& s0 = h->shards[0];
    JC_CUDA(cudaSetDevice(s0.dev));
    int rc;
    if (!s0.Metric && (rc = dev_alloc(&s0.Metric, (size_t)Q * Q))) return rc;
    if (!s0.Linv && (rc = dev_alloc(&s0.Linv, (size_t)Q * Q))) return rc;
    if (!h->metric_on_device)
        JC_CUDA(cudaMemcpyAsync(s0.Metric, h->host_metric.data(), sizeof(double) * Q * Q, cudaMemcpyHostToDevice, s0.stream));
    h->metric_on_device = false;  // potrf overwrites it with L
    if ((rc = factor_metric_on_device(s0, Q))) return rc;
    const double one = 1.0, zero = 0.0;

    // --- B formation, packed column windows ---------------------------------------------------------------------------
    // Row-major packed rows [q][row_len] viewed column-major are (row_len x q) with ld = row_len, so
    //   R_g^T (w x Qg) = sum_h  T_h^T (w x Qh) * Linv[Qg, Qh]^T .
    std::vector<double*> linv_d(nd, nullptr), tmpT(nd, nullptr), R(nd, nullptr);
    std::vector<cublasHandle_t> bl(nd, nullptr);
    int64_t max_q = 0;
    for (int d = 0; d < nd; d++) max_q = std::max(max_q, h->shards[d].q_len);
    const int64_t win = std::max<int64_t>(1024, std::min<int64_t>(L.row_len, (((int64_t)1 << 26) / std::max<int64_t>(max_q, 1)) / 1024 * 1024));
    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        if (d == 0) linv_d[d] = s0.Linv;
        else {
            if ((rc = dev_alloc(&linv_d[d], (size_t)Q * Q))) return rc;
            JC_CUDA(cudaMemcpyPeer(linv_d[d], s.dev, s0.Linv, s0.dev, sizeof(double) * Q * Q));
        }
        if (nd > 1 && (rc = dev_alloc(&tmpT[d], (size_t)(win * max_q)))) return rc;
        if ((rc = dev_alloc(&R[d], (size_t)(win * std::max<int64_t>(s.q_len, 1))))) return rc;
        JC_CUBLAS(cublasCreate(&bl[d]));
        JC_CUBLAS(cublasSetStream(bl[d], s.stream));
    }
    for (int64_t c0 = 0; c0 < L.row_len; c0 += win) {
        const int64_t w = std::min(win, L.row_len - c0);
        for (int g = 0; g < nd; g++) {
            Shard& sg = h->shards[g];
            if (sg.q_len == 0) continue;
            JC_CUDA(cudaSetDevice(sg.dev));
            for (int hh = 0; hh <= g; hh++) {
                Shard& sh = h->shards[hh];
                if (sh.q_len == 0) continue;
                const double* Tsrc;
                int64_t ldT;
                if (hh == g) { Tsrc = sg.Bp + c0; ldT = L.row_len; }
                else {
                    JC_CUDA(cudaMemcpy2DAsync(tmpT[g], sizeof(double) * w, sh.Bp + c0, sizeof(double) * L.row_len,
                                              sizeof(double) * w, (size_t)sh.q_len, cudaMemcpyDefault, sg.stream));
                    Tsrc = tmpT[g];
                    ldT = w;
                }
                const double* Lblk = linv_d[g] + sg.q_begin + Q * sh.q_begin;
                JC_CUBLAS(cublasDgemm(bl[g], CUBLAS_OP_N, CUBLAS_OP_T, (int)w, (int)sg.q_len, (int)sh.q_len, &one, Tsrc,
                                      (int)ldT, Lblk, (int)Q, hh == 0 ? &zero : &one, R[g], (int)w));
            }
        }
        for (int g = 0; g < nd; g++) { JC_CUDA(cudaSetDevice(h->shards[g].dev)); JC_CUDA(cudaStreamSynchronize(h->shards[g].stream)); }
        for (int g = 0; g < nd; g++) {
            Shard& sg = h->shards[g];
            if (sg.q_len == 0) continue;
            JC_CUDA(cudaSetDevice(sg.dev));
            JC_CUDA(cudaMemcpy2DAsync(sg.Bp + c0, sizeof(double) * L.row_len, R[g], sizeof(double) * w, sizeof(double) * w,
                                      (size_t)sg.q_len, cudaMemcpyDeviceToDevice, sg.stream));
        }
        for (int g = 0; g < nd; g++) { JC_CUDA(cudaSetDevice(h->shards[g].dev)); JC_CUDA(cudaStreamSynchronize(h->shards[g].stream)); }
    }
    for (int d = 0; d < nd; d++) {
        JC_CUDA(cudaSetDevice(h->shards[d].dev));
        if (d > 0 && linv_d[d]) cudaFree(linv_d[d]);
        if (tmpT[d]) cudaFree(tmpT[d]);
        if (R[d]) cudaFree(R[d]);
        if (bl[d]) cublasDestroy(bl[d]);
        h->shards[d].b_ready = true;
        h->shards[d].bc_valid = false;
    }
    return JCDF_OK;
}

int32_t jcdf_get_metric_inverse(jcdf_handle* h, double* Linv_out) {
    JC_REQUIRE(h && Linv_out, "jcdf_get_metric_inverse: NULL argument");
    JC_REQUIRE(h->shards[0].Linv != nullptr, "jcdf_get_metric_inverse: no factorisation yet");
    JC_CUDA(cudaSetDevice(h->shards[0].dev));
    JC_CUDA(cudaMemcpy(Linv_out, h->shards[0].Linv, sizeof(double) * h->lay.Q * h->lay.Q, cudaMemcpyDeviceToHost));
    return JCDF_OK;
}

int32_t jcdf_set_core_hamiltonian(jcdf_handle* h, const double* H) {
    JC_REQUIRE(h && H, "jcdf_set_core_hamiltonian: NULL argument");
    JC_REQUIRE(h->lay.mask_set, "jcdf_set_core_hamiltonian: call jcdf_set_mask first");
    Shard& s = h->shards[0];
    JC_CUDA(cudaSetDevice(s.dev));
    JC_CUDA(cudaMemcpy(s.Hdev, H, sizeof(double) * h->lay.N * h->lay.N, cudaMemcpyHostToDevice));
    s.has_H = true;
    return JCDF_OK;
}

// What the per-build launch sequence of one shard depends on (everything else is fixed once jcdf_set_mask has run).
struct BuildSchedule {
    bool use_pp, pp_j, v_in_ht, overlap, add_H;
    bool tile_v;  // tile family: V from the half-transform's accumulators, then the J sweep alone
    uint64_t key(const jcdf_handle* h) const {
        return 1ull | (uint64_t)use_pp << 1 | (uint64_t)pp_j << 2 | (uint64_t)v_in_ht << 3 | (uint64_t)overlap << 4 |
               (uint64_t)add_H << 5 | (uint64_t)(h->opt_coulomb_fused & 3) << 6 | (uint64_t)(h->opt_coulomb_pp_fused & 1) << 8 |
               (uint64_t)(h->opt_pp_bigreg & 1) << 9 | (uint64_t)(h->opt_exchange_producer & 1) << 10 | (uint64_t)tile_v << 11;
    }
};

// Launches prep .. epilogue of one shard on its stream (between the H2D copies and the all-reduce).  Called eagerly, or
// once under stream capture to build the graph the later builds replay.
static int issue_build_kernels(jcdf_handle* h, Shard& s, const BuildSchedule& b) {
    const Layout& L = h->lay;
    if (!b.pp_j) launch_prep(L, s, L.N, !b.tile_v);
    if (b.use_pp) launch_pp_prep(L, s, L.N);
    // V = sum_p sum_i C[p,i] W[Q,i,p] comes out of the half-transform's accumulators: no density, no V sweep
    if (b.pp_j && !b.v_in_ht) launch_pp_density(L, s);
    JC_CUDA(record_event(s, s.ev[EV_PREP], s.stream));
    if (b.overlap) JC_CUDA(cudaEventRecord(s.ev_dep[0], s.stream));
    if (b.use_pp) launch_pp_half_transform(L, s, h->sm_count, b.v_in_ht, h->opt_pp_bigreg != 0);
    else launch_half_transform(L, s, h->sm_count, h->opt_pp_bigreg != 0, b.tile_v);
    JC_CUDA(cudaPeekAtLastError());  // a launch that did not fit fails the build here, not as a wrong Fock matrix
    JC_CUDA(record_event(s, s.ev[EV_W], s.stream));
    if (b.overlap) JC_CUDA(cudaEventRecord(s.ev_dep[1], s.stream));
    launch_exchange(L, s, h->sm_count, h->opt_exchange_producer != 0);
    JC_CUDA(cudaPeekAtLastError());
    JC_CUDA(record_event(s, s.ev[EV_K], s.stream));
    // Coulomb sweeps (HBM-bound, 32-40 registers, no dynamic smem).  With coulomb_overlap they go to a second,
    // lower-priority stream: their CTAs co-reside with the persistent tensor-core CTAs (one per SM, compute-bound,
    // issue slots ~16 % busy) and soak up the idle HBM bandwidth instead of queueing behind them.
    // (fork / join through dedicated dependency events: plain records, which a capture turns into graph edges)
    cudaStream_t sj = b.overlap ? s.stream_j : s.stream;
    if (b.overlap) {
        // the fork point is after the half-transform when it leaves the Coulomb partials, after the prep kernels otherwise:
        // recorded here, i.e. after the exchange launch, it would serialise the two streams, so the events are recorded
        // at their places above (dep[0] after prep, dep[1] after the half-transform)
        JC_CUDA(cudaStreamWaitEvent(sj, s.ev_dep[((b.pp_j && b.v_in_ht) || b.tile_v) ? 1 : 0], 0));
    }
    JC_CUDA(record_event(s, s.ev[EV_JBEGIN], sj));
    JC_CUDA(record_event(s, s.ev[EV_V], sj));  // two-pass: overwritten at the V | J boundary; fused: V time reads 0
    if (b.pp_j) {
        s.last_coulomb_fused = false;
        s.last_ppj_fused = h->opt_coulomb_pp_fused && launch_pp_coulomb_fused(L, s, h->sm_count, sj);
        if (!s.last_ppj_fused) {
            launch_pp_coulomb(L, s, sj, s.ev[EV_V], b.v_in_ht);
        }
    } else {
        // fused single pass pays off on long column segments (w50 dense: 3.8 TB/s effective against <= 3.05 for two
        // passes at the copy bandwidth); on short ones its per-chunk grid-wide flags dominate and two streaming
        // passes are faster (measured: benzene dimer 1.01 -> 0.91 ms, C20H42 5.54 -> 5.26 ms)
        const bool want_fused = !b.tile_v && (h->opt_coulomb_fused == 1 ||
                                              (h->opt_coulomb_fused == 2 && L.row_len / std::max(1, h->sm_count) >= 4500));
        s.last_coulomb_fused = want_fused && launch_coulomb_fused(L, s, h->sm_count, sj);
        if (!s.last_coulomb_fused) launch_coulomb(L, s, h->sm_count, sj, b.tile_v);  // tile_v: V is there already, J sweep only
    }
    s.last_pp_j = b.pp_j;
    s.last_tile_v = b.tile_v;
    s.last_v_in_ht = b.v_in_ht && !s.last_ppj_fused;
    JC_CUDA(record_event(s, s.ev[EV_J], sj));
    if (b.overlap) {
        JC_CUDA(cudaEventRecord(s.ev_dep[2], sj));
        JC_CUDA(cudaStreamWaitEvent(s.stream, s.ev_dep[2], 0));
    }
    launch_epilogue(L, s, b.add_H);
    JC_CUDA(record_event(s, s.ev[EV_EPI], s.stream));
    JC_CUDA(cudaGetLastError());
    return JCDF_OK;
}

// Replays the shard's graph of issue_build_kernels, capturing it first when there is none for this schedule.  Any failure
// of capture / instantiation marks the shard and the build is launched eagerly instead (same kernels, same order).
static int run_build_kernels(jcdf_handle* h, Shard& s, const BuildSchedule& b) {
    const bool want_graph = h->opt_graph && !s.graph_broken;
    if (!want_graph) return issue_build_kernels(h, s, b);
    const uint64_t key = b.key(h);
    if (s.gexec == nullptr || s.gkey != key) {
        if (s.gexec) { cudaGraphExecDestroy(s.gexec); s.gexec = nullptr; }
        cudaGraph_t g = nullptr;
        int rc = JCDF_OK;
        if (cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();
            s.graph_broken = true;
            return issue_build_kernels(h, s, b);
        }
        s.capturing = true;
        rc = issue_build_kernels(h, s, b);
        s.capturing = false;
        const cudaError_t ec = cudaStreamEndCapture(s.stream, &g);
        if (rc != JCDF_OK || ec != cudaSuccess || g == nullptr ||
            cudaGraphInstantiate(&s.gexec, g, 0) != cudaSuccess) {
            cudaGetLastError();
            if (g) cudaGraphDestroy(g);
            s.gexec = nullptr;
            s.graph_broken = true;
            return issue_build_kernels(h, s, b);
        }
        cudaGraphDestroy(g);
        s.gkey = key;
    }
    JC_CUDA(cudaGraphLaunch(s.gexec, s.stream));
    s.graph_replays++;
    return JCDF_OK;
}

// Once per communicator: the q_len of all shards of all ranks must add up to Q (a rank left on the default shards, or
// ranges that overlap across ranks, would otherwise return a wrong Fock matrix without any error).
static int check_global_coverage(jcdf_handle* h) {
    if (h->coverage_checked) return JCDF_OK;
    const int nd = (int)h->shards.size();
    const NcclApi& n = nccl();
    std::vector<double> got((size_t)nd, 0.0);
    JC_NCCL(n.GroupStart());
    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        const double mine = (double)s.q_len;
        JC_CUDA(cudaMemcpyAsync(s.V, &mine, sizeof(double), cudaMemcpyHostToDevice, s.stream));  // V is rebuilt every build
        JC_CUDA(cudaStreamSynchronize(s.stream));
        JC_NCCL(n.AllReduce(s.V, s.V, 1, NCCL_DOUBLE, NCCL_SUM, s.nccl_comm, s.stream));
    }
    JC_NCCL(n.GroupEnd());
    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        JC_CUDA(cudaMemcpyAsync(&got[(size_t)d], s.V, sizeof(double), cudaMemcpyDeviceToHost, s.stream));
        JC_CUDA(cudaStreamSynchronize(s.stream));
    }
    for (int d = 0; d < nd; d++)
        if (got[(size_t)d] != (double)h->lay.Q) {
            set_error("jcdf_fock_build: the shards of all ranks hold " + std::to_string((int64_t)got[(size_t)d]) +
                      " auxiliary rows in total, expected Q = " + std::to_string(h->lay.Q) + " (jcdf_set_shards on every rank)");
            return JCDF_ERR_INVALID;
        }
    h->coverage_checked = true;
    return JCDF_OK;
}

int fock_build_impl(jcdf_handle* h, const double* C_occ, int64_t ldc, const double* H, double* F_out,
                           double* stage_seconds, bool device_ptrs) {
    JC_REQUIRE(h && C_occ && F_out, "jcdf_fock_build: NULL argument");
    const Layout& L = h->lay;
    JC_REQUIRE(L.mask_set, "jcdf_fock_build: call jcdf_set_mask first");
    JC_REQUIRE(ldc >= L.N, "jcdf_fock_build: ldc < N");
    const int nd = (int)h->shards.size();
    for (int d = 0; d < nd; d++) JC_REQUIRE(h->shards[d].b_ready, "jcdf_fock_build: B has not been formed on every device");
    int rc = ensure_comm(h);
    if (rc) return rc;
    const bool owns_rank0 = (h->rank == 0);
    const bool reduce = (h->n_ranks * nd > 1);
    if (reduce && (rc = check_global_coverage(h))) return rc;
    const size_t nn = (size_t)L.N * L.N;
    if (!device_ptrs && h->opt_register_output) {
        // The reference hands over scf_data.two_electron_fock, a pageable Julia Array that lives as long as the SCF
        // (DensityFitting.jl:45): a device-to-host copy into pageable memory is staged by the driver at a fraction of the
        // PCIe rate.  Once the same pointer arrives a second time it is page-locked in place (cudaHostRegister) and stays so
        // until another buffer shows up or the handle is destroyed.  Already pinned or managed memory is left alone;
        // a refusal is not an error.  (C_occ is a fresh slice every iteration, DensityFitting.jl:49 — not worth registering.)
        if (h->fout_registered && h->fout_registered != (void*)F_out) {
            cudaHostUnregister(h->fout_registered);
            cudaGetLastError();
            h->fout_registered = nullptr;
        }
        if (!h->fout_registered && h->fout_seen == (const void*)F_out) {
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, F_out) == cudaSuccess && at.type == cudaMemoryTypeUnregistered &&
                cudaHostRegister(F_out, sizeof(double) * nn, cudaHostRegisterPortable) == cudaSuccess)
                h->fout_registered = F_out;
            cudaGetLastError();
        }
        h->fout_seen = F_out;
    }
    auto t0 = std::chrono::steady_clock::now();

    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        JC_CUDA(cudaEventRecord(s.ev[EV_START], s.stream));
        // C_occ -> device, compacted to ld = N
        JC_CUDA(cudaMemcpy2DAsync(s.Craw, sizeof(double) * L.N, C_occ, sizeof(double) * ldc, sizeof(double) * L.N,
                                  (size_t)L.occ, device_ptrs ? cudaMemcpyDefault : cudaMemcpyHostToDevice, s.stream));
        BuildSchedule b;
        b.add_H = false;
        if (d == 0 && owns_rank0) {
            if (H != nullptr) {
                JC_CUDA(cudaMemcpyAsync(s.Hdev, H, sizeof(double) * nn, device_ptrs ? cudaMemcpyDefault : cudaMemcpyHostToDevice,
                                        s.stream));
                s.has_H = true;
            }
            b.add_H = s.has_H;
        }
        if (L.xs_any_dropped)
            // block-screened exchange: F_out is IN/OUT.  The reference's dropped K blocks are neither computed nor written
            // (ScreenedDF.jl:494-510), the Coulomb copy skips them too (all their pairs are screened, :367-378), so they keep
            // what scf_data.two_electron_fock held — the caller's buffer — and receive H on rank 0 (DensityFitting.jl:62-65).
            JC_CUDA(cudaMemcpyAsync(s.Fprev, F_out, sizeof(double) * nn, device_ptrs ? cudaMemcpyDefault : cudaMemcpyHostToDevice,
                                    s.stream));
        JC_CUDA(cudaEventRecord(s.ev[EV_H2D], s.stream));
        b.use_pp = L.pp_enabled && s.Bc != nullptr;
        if (b.use_pp && !s.bc_valid) {  // first build after B changed: refresh the partner-compacted copy
            launch_pp_pack(L, s);
            s.bc_valid = true;
        }
        b.pp_j = b.use_pp && h->opt_coulomb_pp;  // Coulomb over the compact copy as well
        b.v_in_ht = b.pp_j && h->opt_coulomb_v_in_ht && !h->opt_coulomb_pp_fused && s.Vitem != nullptr;
        // tile family with the automatic Coulomb choice: V from the accumulators of the tile half-transform, then one J sweep
        // (an explicit coulomb_fused = 0 / 1 keeps the two-pass / single-pass kernels selectable)
        b.tile_v = !b.use_pp && h->opt_coulomb_v_in_ht && h->opt_coulomb_fused == 2 && s.Vt != nullptr;
        b.overlap = h->opt_coulomb_overlap && (b.pp_j || !h->opt_coulomb_fused || b.tile_v);
        if ((rc = run_build_kernels(h, s, b))) return rc;
    }
    auto t_launched = std::chrono::steady_clock::now();
    if (reduce) {
        const NcclApi& n = nccl();
        JC_NCCL(n.GroupStart());
        for (int d = 0; d < nd; d++) {
            Shard& s = h->shards[d];
            JC_NCCL(n.AllReduce(s.F, s.F, nn, NCCL_DOUBLE, NCCL_SUM, s.nccl_comm, s.stream));
        }
        JC_NCCL(n.GroupEnd());
        for (int d = 0; d < nd; d++) {
            JC_CUDA(cudaSetDevice(h->shards[d].dev));
            launch_mirror_lower(L, h->shards[d]);
        }
    }
    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        JC_CUDA(cudaEventRecord(s.ev[EV_AR], s.stream));
    }
    {
        Shard& s = h->shards[0];
        JC_CUDA(cudaSetDevice(s.dev));
        JC_CUDA(cudaMemcpyAsync(F_out, s.F, sizeof(double) * nn, device_ptrs ? cudaMemcpyDefault : cudaMemcpyDeviceToHost,
                                s.stream));
        JC_CUDA(cudaEventRecord(s.ev[EV_D2H], s.stream));
    }
    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        JC_CUDA(cudaStreamSynchronize(s.stream));
    }
    auto t1 = std::chrono::steady_clock::now();
    double mx[JCDF_N_STAGES] = {};
    for (int d = 0; d < nd; d++) {
        Shard& s = h->shards[d];
        JC_CUDA(cudaSetDevice(s.dev));
        auto el = [&](int a, int b) { float ms = 0; cudaEventElapsedTime(&ms, s.ev[a], s.ev[b]); return (double)ms; };
        s.last_ms[JCDF_T_H2D] = el(EV_START, EV_H2D);
        s.last_ms[JCDF_T_DENSITY] = el(EV_H2D, EV_PREP);
        s.last_ms[JCDF_T_W] = el(EV_PREP, EV_W);
        s.last_ms[JCDF_T_K] = el(EV_W, EV_K);
        s.last_ms[JCDF_T_V] = el(EV_JBEGIN, EV_V);
        s.last_ms[JCDF_T_J] = el(EV_V, EV_J);
        s.last_ms[JCDF_T_EPILOGUE] = (h->opt_coulomb_overlap && (s.last_pp_j || !h->opt_coulomb_fused || s.last_tile_v)) ? el(EV_K, EV_EPI) : el(EV_J, EV_EPI);
        s.last_ms[JCDF_T_ALLREDUCE] = el(EV_EPI, EV_AR);
        s.last_ms[JCDF_T_D2H] = d == 0 ? el(EV_AR, EV_D2H) : 0.0;
        s.last_ms[JCDF_T_TOTAL] = std::chrono::duration<double, std::milli>(t1 - t0).count();
        for (int i = 0; i < JCDF_N_STAGES; i++) mx[i] = std::max(mx[i], s.last_ms[i]);
    }
    h->last_host_launch_ms = std::chrono::duration<double, std::milli>(t_launched - t0).count();
    if (stage_seconds)
        for (int i = 0; i < JCDF_N_STAGES; i++) stage_seconds[i] = mx[i] * 1e-3;
    return JCDF_OK;
}

int32_t jcdf_fock_build(jcdf_handle* h, const double* C_occ, int64_t ldc, const double* H, double* F_out,
                        double* stage_seconds) {
    return fock_build_impl(h, C_occ, ldc, H, F_out, stage_seconds, false);
}

int32_t jcdf_fock_build_device(jcdf_handle* h, const double* C_occ_dev, int64_t ldc, const double* H_dev,
                               double* F_out_dev, double* stage_seconds) {
    return fock_build_impl(h, C_occ_dev, ldc, H_dev, F_out_dev, stage_seconds, true);
}

int32_t jcdf_eigh(jcdf_handle* h, const double* F, const double* X, double* evals, double* C) {
    JC_REQUIRE(h && F && X && evals && C, "jcdf_eigh: NULL argument");
    const int64_t N = h->lay.N;
    Shard& s = h->shards[0];
    JC_CUDA(cudaSetDevice(s.dev));
    // called once per SCF iteration from iteration(): handles and workspaces live in the handle
    if (!h->eigh) h->eigh = new (std::nothrow) EighCache();
    JC_REQUIRE(h->eigh != nullptr, "jcdf_eigh: out of host memory");
    EighCache& e = *h->eigh;
    e.dev = s.dev;
    int rc;
    if (!e.blas) {
        JC_CUBLAS(cublasCreate(&e.blas));
        JC_CUBLAS(cublasSetStream(e.blas, s.stream));
    }
    if (!e.sol) {
        JC_CUSOLVER(cusolverDnCreate(&e.sol));
        JC_CUSOLVER(cusolverDnSetStream(e.sol, s.stream));
    }
    if (!e.dF && (rc = dev_alloc(&e.dF, (size_t)N * N))) return rc;
    if (!e.dX && (rc = dev_alloc(&e.dX, (size_t)N * N))) return rc;
    if (!e.dT && (rc = dev_alloc(&e.dT, (size_t)N * N))) return rc;
    if (!e.dW && (rc = dev_alloc(&e.dW, (size_t)N))) return rc;
    if (!e.dinfo && (rc = dev_alloc(&e.dinfo, 1))) return rc;
    if (!e.work) {
        JC_CUSOLVER(cusolverDnDsyevd_bufferSize(e.sol, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, (int)N, e.dF, (int)N, e.dW, &e.lwork));
        if ((rc = dev_alloc(&e.work, (size_t)e.lwork))) return rc;
    }
    JC_CUDA(cudaMemcpyAsync(e.dF, F, sizeof(double) * N * N, cudaMemcpyHostToDevice, s.stream));
    JC_CUDA(cudaMemcpyAsync(e.dX, X, sizeof(double) * N * N, cudaMemcpyHostToDevice, s.stream));
    const double one = 1.0, zero = 0.0;
    // workspace_a = X^T F ; Fp = workspace_a X   (symm!/gemm!, SCF.jl:1083-1084)
    JC_CUBLAS(cublasDgemm(e.blas, CUBLAS_OP_T, CUBLAS_OP_N, (int)N, (int)N, (int)N, &one, e.dX, (int)N, e.dF, (int)N, &zero, e.dT, (int)N));
    JC_CUBLAS(cublasDgemm(e.blas, CUBLAS_OP_N, CUBLAS_OP_N, (int)N, (int)N, (int)N, &one, e.dT, (int)N, e.dX, (int)N, &zero, e.dF, (int)N));
    JC_CUSOLVER(cusolverDnDsyevd(e.sol, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_UPPER, (int)N, e.dF, (int)N, e.dW, e.work, e.lwork, e.dinfo));
    // C = X U  (symm!, SCF.jl:1095)
    JC_CUBLAS(cublasDgemm(e.blas, CUBLAS_OP_N, CUBLAS_OP_N, (int)N, (int)N, (int)N, &one, e.dX, (int)N, e.dF, (int)N, &zero, e.dT, (int)N));
    int info = 0;
    JC_CUDA(cudaMemcpyAsync(&info, e.dinfo, sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    JC_CUDA(cudaMemcpyAsync(evals, e.dW, sizeof(double) * N, cudaMemcpyDeviceToHost, s.stream));
    JC_CUDA(cudaMemcpyAsync(C, e.dT, sizeof(double) * N * N, cudaMemcpyDeviceToHost, s.stream));
    JC_CUDA(cudaStreamSynchronize(s.stream));
    if (info != 0) {
        set_error("syevd did not converge (info = " + std::to_string(info) + ")");
        return JCDF_ERR_NUMERIC;
    }
    return JCDF_OK;
}

int32_t jcdf_comm_unique_id(uint8_t id_out[128]) {
    JC_REQUIRE(id_out != nullptr, "jcdf_comm_unique_id: NULL argument");
    const NcclApi& n = nccl();
    if (!n.ok) { set_error(std::string("NCCL unavailable: ") + n.why); return JCDF_ERR_NCCL; }
    NcclApi::UniqueId id;
    JC_NCCL(n.GetUniqueId(&id));
    memcpy(id_out, id.internal, 128);
    return JCDF_OK;
}

int32_t jcdf_comm_init_rank(jcdf_handle* h, int32_t n_ranks, int32_t rank, const uint8_t id_bytes[128]) {
    JC_REQUIRE(h && id_bytes, "jcdf_comm_init_rank: NULL argument");
    JC_REQUIRE(n_ranks >= 1 && rank >= 0 && rank < n_ranks, "jcdf_comm_init_rank: bad rank");
    JC_REQUIRE(!h->comm_ready, "jcdf_comm_init_rank: communicator already initialised");
    // the default shards split ALL of Q over this process's devices: with several ranks every rank would contract every
    // auxiliary row and the all-reduce would return H + n_ranks (2J - K)
    JC_REQUIRE(n_ranks == 1 || h->shards_set, "jcdf_comm_init_rank: call jcdf_set_shards with this rank's auxiliary rows first");
    const NcclApi& n = nccl();
    if (!n.ok) { set_error(std::string("NCCL unavailable: ") + n.why); return JCDF_ERR_NCCL; }
    const int nd = (int)h->shards.size();
    h->n_ranks = n_ranks;
    h->rank = rank;
    NcclApi::UniqueId id;
    memcpy(id.internal, id_bytes, 128);
    JC_NCCL(n.GroupStart());
    for (int d = 0; d < nd; d++) {
        JC_CUDA(cudaSetDevice(h->shards[d].dev));
        JC_NCCL(n.CommInitRank(&h->shards[d].nccl_comm, n_ranks * nd, id, rank * nd + d));
    }
    JC_NCCL(n.GroupEnd());
    h->comm_ready = true;
    return JCDF_OK;
}

int32_t jcdf_get_layout(jcdf_handle* h, jcdf_layout_info* o) {
    JC_REQUIRE(h && o, "jcdf_get_layout: NULL argument");
    const Layout& L = h->lay;
    memset(o, 0, sizeof(*o));
    o->N = L.N; o->Q = L.Q; o->occ = L.occ;
    o->n_pad = L.n_pad; o->occ_pad = L.occ_pad;
    o->tiles_total = L.tiles_total; o->tiles_kept = L.tiles_kept;
    o->pairs_kept = L.pairs_kept; o->pairs_tri = L.pairs_tri;
    o->packed_row_len = L.row_len;
    o->exchange_blocks_total = L.xs_blocks_total;
    o->exchange_blocks_kept = L.xs_blocks_kept;
    o->exchange_block_width = L.xs_width;
    o->exchange_tiles_total = (int64_t)L.n128 * (L.n128 + 1) / 2;
    o->exchange_tiles_skipped = L.xs_tiles_skipped;
    o->exchange_half_tiles_skipped = L.xs_halves_skipped;
    for (const Shard& s : h->shards) {
        o->q_local += s.q_len;
        o->bytes_B += s.q_len * L.row_len * 8;
        if (L.pp_enabled && s.Bc) o->bytes_B += (int64_t)s.pp_chunks * L.pp_rows * 128 * 8;
        o->bytes_W += (int64_t)s.xplan.chunks * L.npt64 * 2048 * 8;
        o->graph_replays += s.graph_replays;
    }
    // prep_cs, density, half_transform, exchange, coulomb (fused: 1, two-pass: 2), epilogue
    // per-function path with its own Coulomb: prep, density, half_transform, exchange, V, V reduce, J, J reduce, epilogue
    const bool ppj = !h->shards.empty() && L.pp_enabled && h->shards[0].Bc != nullptr && h->opt_coulomb_pp;
    const int kred = (!h->shards.empty() && h->shards[0].xplan.n_slices > 1) ? 1 : 0;  // k_exchange_reduce
    const bool ppjf = ppj && h->shards[0].last_ppj_fused;  // one cooperative Coulomb kernel instead of V, V reduce, J, J reduce
    const bool vht = ppj && !ppjf && h->opt_coulomb_v_in_ht && !h->opt_coulomb_pp_fused;  // density, V pass, V reduce -> one reduce
    // tile family with V from the half-transform: prep, half_transform, V reduce, exchange, J sweep, epilogue
    const bool tile_v = !ppj && !h->shards.empty() && h->shards[0].last_tile_v;
    o->kernel_launches_per_build = kred + (ppjf ? 6 : ppj ? (vht ? 7 : 9) : (tile_v || (!h->shards.empty() && h->shards[0].last_coulomb_fused) ? 6 : 7)) +
                                   ((h->n_ranks * (int)h->shards.size() > 1) ? 1 : 0);  // + mirror after the all-reduce
    return JCDF_OK;
}

int32_t jcdf_fill_synthetic_B(jcdf_handle* h, uint64_t seed, double decay_length) {
    JC_REQUIRE(h != nullptr, "jcdf_fill_synthetic_B: NULL handle");
    JC_REQUIRE(h->lay.mask_set, "jcdf_fill_synthetic_B: call jcdf_set_mask first");
    for (Shard& s : h->shards) {
        JC_CUDA(cudaSetDevice(s.dev));
        launch_fill_synthetic(h->lay, s, seed, decay_length);
        JC_CUDA(cudaGetLastError());
        s.b_ready = true;
        s.bc_valid = false;
    }
    return JCDF_OK;
}

int32_t jcdf_set_option(jcdf_handle* h, const char* name, int64_t value) {
    JC_REQUIRE(h && name, "jcdf_set_option: NULL argument");
    const std::string n(name);
    struct Opt { const char* name; int64_t* slot; bool before_mask; int64_t lo, hi; };
    const int64_t big = (int64_t)1 << 62;
    const Opt opts[] = {
        {"coulomb_fused", &h->opt_coulomb_fused, false, 0, 2},
        {"coulomb_overlap", &h->opt_coulomb_overlap, false, 0, 1},
        {"pp_max_bytes", &h->opt_pp_max_bytes, true, 0, big},
        {"pp_layout", &h->opt_pp_layout, true, -1, 2},
        {"coulomb_pp", &h->opt_coulomb_pp, false, 0, 1},
        {"coulomb_v_in_ht", &h->opt_coulomb_v_in_ht, false, 0, 1},
        {"coulomb_pp_fused", &h->opt_coulomb_pp_fused, false, 0, 1},
        {"exchange_balance", &h->opt_exchange_balance, true, 0, 1},
        {"ht_mode", &h->opt_ht_mode, true, 0, 2},
        {"regroup", &h->opt_regroup, true, 0, 1},
        {"exchange_slices", &h->opt_exchange_slices, true, 0, 1 << 20},
        {"exchange_block_screen", &h->opt_exchange_block_screen, true, 0, 1 << 20},
        {"pp_order", &h->opt_pp_order, true, 0, 1},
        {"pp_bigreg", &h->opt_pp_bigreg, false, 0, 1},
        {"exchange_producer", &h->opt_exchange_producer, false, 0, 1},
        {"graph", &h->opt_graph, false, 0, 1},
        {"register_output", &h->opt_register_output, false, 0, 1},
    };
    for (const Opt& o : opts) {
        if (n != o.name) continue;
        // validate BEFORE storing: a refused call leaves the handle as it was
        if (value < o.lo || value > o.hi) {
            set_error("jcdf_set_option: value " + std::to_string(value) + " out of range for '" + n + "'");
            return JCDF_ERR_INVALID;
        }
        if (o.before_mask)
            for (Shard& s : h->shards)
                if (s.Bp != nullptr) {
                    set_error("jcdf_set_option: '" + n + "' must be set before jcdf_set_mask");
                    return JCDF_ERR_INVALID;
                }
        *o.slot = value;
        if (!o.before_mask) h->graph_dirty = true;  // the captured launch sequence depends on the schedule options
        return JCDF_OK;
    }
    set_error("jcdf_set_option: unknown option '" + n + "'");
    return JCDF_ERR_INVALID;
}

int32_t jcdf_last_kernel_ms(jcdf_handle* h, int32_t dev, double* out) {
    JC_REQUIRE(h && out, "jcdf_last_kernel_ms: NULL argument");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "jcdf_last_kernel_ms: device index out of range");
    for (int i = 0; i < JCDF_N_STAGES; i++) out[i] = h->shards[dev].last_ms[i];
    return JCDF_OK;
}

int32_t jcdf_exchange_cta_cycles(jcdf_handle* h, int32_t dev, int64_t* out, int32_t n) {
    JC_REQUIRE(h && out, "jcdf_exchange_cta_cycles: NULL argument");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "jcdf_exchange_cta_cycles: device index out of range");
    Shard& s = h->shards[dev];
    JC_REQUIRE(s.d_xcycles != nullptr, "jcdf_exchange_cta_cycles: call jcdf_set_mask first");
    const int m = std::min<int>(n, s.xplan.n_ctas);
    JC_CUDA(cudaSetDevice(s.dev));
    JC_CUDA(cudaStreamSynchronize(s.stream));
    static_assert(sizeof(long long) == sizeof(int64_t), "clock64 width");
    JC_CUDA(cudaMemcpy(out, s.d_xcycles, sizeof(int64_t) * (size_t)std::max(m, 0), cudaMemcpyDeviceToHost));
    return s.xplan.n_ctas;
}

int32_t jcdf_checked_set_fuzz(jcdf_handle* h, uint32_t seed) {
    JC_REQUIRE(h != nullptr, "jcdf_checked_set_fuzz: NULL handle");
    for (Shard& s : h->shards) {
        JC_CUDA(cudaSetDevice(s.dev));
        JC_CUDA(cudaStreamSynchronize(s.stream));
        set_fuzz_seed_k(seed);
        set_fuzz_seed_pp(seed);
        set_fuzz_seed_j(seed);
        JC_CUDA(cudaGetLastError());
    }
    return JCDF_OK;
}

int32_t jcdf_get_stream(jcdf_handle* h, int32_t dev, void** stream_out) {
    JC_REQUIRE(h && stream_out, "jcdf_get_stream: NULL argument");
    JC_REQUIRE(dev >= 0 && dev < (int)h->shards.size(), "jcdf_get_stream: device index out of range");
    *stream_out = reinterpret_cast<void*>(h->shards[dev].stream);
    return JCDF_OK;
}

}  // extern "C"
