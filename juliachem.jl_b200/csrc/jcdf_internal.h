// Internal types shared by the libjcdf translation units (not part of the C ABI).
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <cuda_runtime.h>

#include "../../include/jcdf.h"

namespace jcdf {

// Geometry of the tile-compacted layout; replicated on every device.
struct Layout {
    int64_t N = 0, Q = 0, occ = 0;
    int nb = 0;         // 32-wide blocks per dimension
    int n_pad = 0;      // 32 * nb
    int occ_pad = 0;    // occ rounded up to 8
    int occ_w = 0;      // occ rounded up to 2: row stride of the flattened contraction index k = Q_local * occ_w + i of W
    int bm = 0;         // half-transform M tile (rows i per work item), multiple of 16, <= 128
    int n_mt = 0;       // M tiles: bm * n_mt >= occ_pad
    int npt64 = 0;      // 64-wide W column tiles
    int n_pt128 = 0;    // 128-wide half-transform p tiles = ceil(nb / 4)
    int n128 = 0;       // 128-wide exchange tiles per dimension
    int64_t pairs_kept = 0, pairs_tri = 0;
    int64_t tiles_total = 0, tiles_kept = 0;
    int64_t row_len = 0;  // tiles_kept * 1024 doubles per auxiliary row
    bool mask_set = false;
    std::vector<int32_t> tile_index;  // [a * nb + b], a >= b: index of the stored tile or -1
    std::vector<int32_t> tile_a, tile_b;
    std::vector<int32_t> perm, inv;           // caller's function index -> internal index and back (n_pad entries)
    std::vector<int32_t> tile_strips;         // per stored tile: bits 0-7 non-empty 4-row strips, bits 8-15 4-column strips
    bool regrouped = false;
    std::vector<int32_t> kb_list, kb_count;  // per 32-wide block pb: the blocks kb with a stored tile (either order)
    std::vector<int64_t> dest_of_col;  // reference pair column -> offset inside a packed row, or -1
    std::vector<int64_t> src_of_col;   // reference pair column -> packed offset holding its value (mirror columns included)
    // ---- partner-compacted layout of the per-function half-transform (kernels_pp.cu) ----
    // For every function p (internal order) the kept partners q, ascending, padded to a multiple of 4 rows.
    bool pp_enabled = false;           // the per-function path is selected for this mask
    int pp_layout = 0;                 // consumer warp arrangement: 0 = 2 x 4 (NI = 32 pp_nt orbitals per item), 1 = 8 x 1 (NI = 8 pp_nt),
                                       // 2 = TWO CTAs per SM of 2 x 2 consumer warps each (NI = 16 pp_nt)
    int pp_nt = 0;                     // 8-wide occupied-orbital tiles per consumer warp
    int pp_n_nt = 0;                   // items along the occupied orbitals: pp_n_nt * NI >= occ_w
    int64_t pp_rows = 0;               // sum over p of the padded partner counts
    std::vector<int32_t> pp_off;       // [N + 1] first row of p
    std::vector<int32_t> pp_q;         // [pp_rows] partner (internal index), n_pad on padding rows
    std::vector<int32_t> pp_src;       // [pp_rows] offset of B[Q, p, q] inside a packed row of Bp, -1 on padding rows
    std::vector<int32_t> pp_p;         // [pp_rows] the function p of the row
    std::vector<int32_t> pp_trirow;    // indices of the q >= p rows (pp_jdst >= 0), ascending
    int64_t pp_n_tri = 0;
    std::vector<int32_t> pp_jdst;      // [pp_rows] q >= p rows (the Coulomb half): offset of (q, p) inside Jt; -1 elsewhere
    // ---- block-screened exchange (option "exchange_block_screen" = df_exchange_screen, ScreenedDF.jl:385-457) ----
    int xs_nblk = 0;                   // the reference's df_exchange_n_blocks after its own adjustments (0 = option off)
    int xs_width = 0;                  // K_block_width
    int64_t xs_blocks_total = 0, xs_blocks_kept = 0;  // lower-triangular blocks / blocks_to_calculate
    bool xs_any_dropped = false;
    std::vector<int32_t> xs_blk_of;    // [N] block of a function in the CALLER's order, -1 in the never-screened remainder strip
    std::vector<uint8_t> xs_keep;      // [xs_nblk * xs_nblk] symmetric block_screen_matrix
    std::vector<uint8_t> xs_tile_skip;  // [n128 (n128 + 1) / 2] 128 x 128 exchange tiles lying wholly inside dropped blocks
    int64_t xs_tiles_skipped = 0;
    std::vector<uint8_t> xs_half_dead;  // per exchange tile: bit h set = the rows the warps of row half h own lie wholly in dropped blocks
    int64_t xs_halves_skipped = 0;      // such halves in tiles that are NOT skipped as a whole
};

inline int pp_item_orbitals(const Layout& L) { return (L.pp_layout == 1 ? 8 : L.pp_layout == 2 ? 16 : 32) * L.pp_nt; }
inline int pp_v_slots(const Layout& L) { return L.pp_layout == 1 ? 1 : L.pp_layout == 2 ? 2 : 4; }  // partial-V slots (warp columns) per (CTA, chunk)
inline int pp_ctas_per_sm(const Layout& L) { return L.pp_layout == 2 ? 2 : 1; }

// event points recorded on the shard's stream during one build, in launch order
enum EvPoint { EV_START = 0, EV_H2D, EV_PREP, EV_W, EV_K, EV_JBEGIN, EV_V, EV_J, EV_EPI, EV_AR, EV_D2H, EV_COUNT };

struct ExchangePlan {
    int n_tiles = 0;     // lower-triangular 128x128 output tiles
    int n_slices = 0;    // split of the contraction index per tile
    int chunks = 0;      // 32-row k chunks held by the shard
    int chunks_per_slice = 0;
    int n_ctas = 0;      // persistent CTAs (<= SM count)
    std::vector<int32_t> item_list, item_off;  // cost-balanced item lists per CTA (longest-processing-time first)
};

struct Shard {
    int dev = 0;
    int64_t q_begin = 0, q_len = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream_j = nullptr;  // Coulomb sweeps, concurrent with the tensor-core kernels when coulomb_overlap = 1
    // HBM-resident operands
    double* Bp = nullptr;      // [q_len][row_len] packed, swizzled tiles
    double* Bc = nullptr;      // [pp_chunks][pp_rows][128] partner-compacted copy of Bp (128 auxiliary rows fastest)
    double* Ci = nullptr;      // [n_pad + 1][ldci] C_occ rows in internal order, row-major, zero padded (last row = 0)
    int32_t* d_pp_off = nullptr;
    int32_t* d_pp_q = nullptr;
    int32_t* d_pp_src = nullptr;
    int32_t* d_pp_p = nullptr;
    int32_t* d_pp_jdst = nullptr;
    int32_t* d_pp_trirow = nullptr;
    unsigned int* ppj_flag = nullptr;  // [pp_chunks * 4] quarter flags of the single-pass compact Coulomb kernel
    double* ppj_vp = nullptr;          // [pp_chunks * 4][sm_count][32] per-CTA partial V
    bool last_ppj_fused = false;
    bool last_v_in_ht = false;
    bool last_tile_v = false;  // the last build took V from the TILE half-transform's accumulators
    int32_t* d_ppitems = nullptr;  // cost-balanced work lists of the per-function half-transform
    int32_t* d_ppoff = nullptr;
    int pp_ctas = 0;
    double* dvec = nullptr;    // [pp_rows] weighted density in partner-row space
    double* VpartP = nullptr;  // [slabs][pp_chunks * 128] partial V of the compact Coulomb pass
    double* Vt = nullptr;      // [sm_count][2][q_len] partial V left by the tile half-transform (coulomb_v_in_ht)
    double* Vitem = nullptr;   // [pp_ctas][pp_chunks][pp_v_slots][128] partial V left by the per-function half-transform
    double* Jpart = nullptr;   // [pp_chunks][pp_rows] partial J per chunk
    int pp_chunks = 0;         // ceil(q_len / 128)
    bool bc_valid = false;     // Bc mirrors the current Bp
    double* W = nullptr;       // [chunks][npt64][64 p][32 k] half-transformed tensor, k = (Q_local, i) flattened
    double* Cs = nullptr;      // [nb][n_mt][32][bm] C_occ slabs for the half-transform
    double* Craw = nullptr;    // N x occ column-major upload
    double* dt = nullptr;      // [row_len] packed, weighted density
    double* Jt = nullptr;      // [row_len] packed Coulomb vector
    double* V = nullptr;       // [q_len]
    double* Vpart = nullptr;   // [q_len][sm_count] per-CTA partial V of the fused Coulomb kernel
    unsigned int* Vready = nullptr;  // [q_len] chunk flags of the fused Coulomb kernel
    double* Kpart = nullptr;   // [n_tiles * n_slices][128*128]
    double* Ksum = nullptr;    // [n_tiles][128*128] slices of Kpart summed (k_exchange_reduce)
    int32_t* d_xitems = nullptr;  // exchange item lists (ExchangePlan)
    int32_t* d_xoff = nullptr;
    long long* d_xcycles = nullptr;  // [n_ctas] busy cycles of the last exchange launch
    double* F = nullptr;       // N x N column-major partial / reduced Fock
    double* Fprev = nullptr;   // N x N: the caller's F_out as it arrived (block-screened exchange: dropped blocks keep it)
    int32_t* d_xs_blk_of = nullptr;  // [N]
    uint8_t* d_xs_keep = nullptr;    // [xs_nblk * xs_nblk]
    uint8_t* d_xs_half_dead = nullptr;  // [n_tiles]
    double* Hdev = nullptr;    // N x N (only where H is added)
    double* Traw = nullptr;    // staging window for uploads
    int64_t traw_capacity = 0;
    int32_t* d_tile_index = nullptr;
    int32_t* d_tile_a = nullptr;
    int32_t* d_tile_b = nullptr;
    int32_t* d_perm = nullptr;      // [n_pad]
    int32_t* d_inv = nullptr;       // [n_pad]
    int32_t* d_tile_strips = nullptr;  // [tiles_kept]
    int32_t* d_kb_list = nullptr;   // [nb][nb]
    int32_t* d_kb_count = nullptr;  // [nb]
    int64_t* d_dest_of_col = nullptr;
    uint8_t* d_fn_mask = nullptr;   // N x N or nullptr when unscreened
    double* Linv = nullptr;    // Q x Q on the first device only
    double* Metric = nullptr;
    bool has_H = false;
    bool last_coulomb_fused = false;
    bool last_pp_j = false;    // the last build ran the Coulomb passes over the partner-compacted copy
    bool b_ready = false;
    bool t_uploaded = false;
    ExchangePlan xplan;
    cudaEvent_t ev[EV_COUNT] = {};
    cudaEvent_t ev_dep[3] = {};    // fork (after prep / after the half-transform) and join of the Coulomb stream (coulomb_overlap)
    // CUDA graph of the per-build kernel sequence (prep .. epilogue): captured on the first build, replayed afterwards
    cudaGraphExec_t gexec = nullptr;
    uint64_t gkey = 0;             // schedule the graph was captured for (family, Coulomb variant, H add)
    bool capturing = false;        // the launchers record timing events as external event-record nodes while set
    bool graph_broken = false;     // capture or instantiation failed once: this shard launches eagerly from then on
    int64_t graph_replays = 0;
    double last_ms[JCDF_N_STAGES] = {};
    void* nccl_comm = nullptr;
};

struct ScfState;        // scf.cu
struct IntegralsState;  // integrals.cu
struct EighCache;       // api.cu: handles + workspaces of jcdf_eigh, created on first use, released by jcdf_destroy

}  // namespace jcdf

struct jcdf_handle {
    jcdf::Layout lay;
    std::vector<jcdf::Shard> shards;
    int n_ranks = 1, rank = 0;   // process-level (one process per GPU mode)
    bool comm_ready = false;
    bool shards_set = false;
    int sm_count = 148;
    std::vector<double> host_metric;
    int64_t opt_coulomb_fused = 2;    // tile-family Coulomb: 2 = single-pass cooperative kernel on long column segments,
                                      // two streaming passes on short ones; 1 = single pass whenever it fits; 0 = two passes
    int64_t opt_coulomb_overlap = 0;  // two-pass sweeps on a second stream, concurrent with the tensor-core kernels
    int64_t opt_exchange_slices = 0;   // > 0: that many slices of the exchange contraction index per tile (0 = chosen from the sizes)
    int64_t opt_pp_bigreg = 1;         // per-function kernel: setmaxnreg register split between producer and consumer warpgroups
    int64_t opt_exchange_producer = 1;  // exchange kernel: dedicated producer warp + setmaxnreg split (0: thread 0 of a DMMA warp produces)
    int64_t opt_pp_order = 1;          // per-function work lists: 1 = p-major inside a CTA (measured faster: 14.9 vs 15.4 ms at w50), 0 = chunk-major
    int64_t opt_exchange_balance = 1;  // cost-balanced exchange item lists (0: round-robin in item order)
    int64_t opt_regroup = 1;          // mask-driven regrouping of the basis functions into 32-wide blocks
    int64_t opt_coulomb_v_in_ht = 1;   // compact Coulomb: V from the per-function half-transform's accumulators (no density, no V sweep)
    int64_t opt_coulomb_pp_fused = 0;  // compact Coulomb as ONE HBM pass (cooperative kernel): measured slower than the two
                                       // streaming passes (2.1 vs 1.3 ms at w50: 150 grid-wide flag rounds), so off by default
    int64_t opt_pp_layout = -1;        // per-function half-transform warp arrangement: -1 = chosen from occ, 0 = 2 x 4, 1 = 8 x 1,
                                       // 2 = two CTAs per SM of 2 x 2 consumer warps (one CTA's item epilogue under the other's DMMAs)
    int64_t opt_pp_max_bytes = 0;      // cap on the partner-compacted copy per shard (0 = whatever cudaMalloc grants)
    int64_t opt_coulomb_pp = 1;       // Coulomb over the partner-compacted copy whenever the per-function path is active
    int64_t opt_ht_mode = 0;          // half-transform: 0 = choose from the mask, 1 = tile kernel, 2 = per-function kernel
    int64_t opt_exchange_block_screen = 0;  // > 0: df_exchange_screen = true with that many K blocks (ScreenedDF.jl:433-449, 459-545)
    int64_t opt_graph = 1;            // replay the per-build launch sequence as a CUDA graph (device-pointer builds)
    bool graph_dirty = true;          // options / operands changed since the capture
    int64_t opt_register_output = 0;  // 1: page-lock the caller's F_out once it has been seen twice (a Julia Array is pageable);
                                      // opt-in, because the caller must keep that buffer alive until jcdf_destroy
    const void* fout_seen = nullptr;  // F_out of the previous host-pointer build
    void* fout_registered = nullptr;  // range this handle registered with cudaHostRegister (released by jcdf_destroy)
    double last_host_launch_ms = 0.0;  // host time the last build spent issuing work before the collective
    bool coverage_checked = false;    // the shards of all ranks were seen to cover Q exactly once
    jcdf::ScfState* scf = nullptr;    // on-device SCF driver state (jcdf_scf_*), owned
    jcdf::IntegralsState* ints = nullptr;  // basis sets + integral scratch (jcdf_set_basis / jcdf_compute_*), owned
    jcdf::EighCache* eigh = nullptr;       // jcdf_eigh state, owned
    bool metric_on_device = false;    // shards[0].Metric holds (P|Q) (jcdf_compute_two_center) — consumed by the factorisation
    double max_PP = 0.0;              // max_P |(P|P)| of the last metric seen (get_max_P_P, SchwarzScreening.jl:86-95)
};

namespace jcdf {

// ---- kernel launchers (kernels_k.cu / kernels_j.cu / kernels_misc.cu) ----------------------------------------------
void launch_prep(const Layout& L, Shard& s, int64_t ldc, bool with_density);
void launch_half_transform(const Layout& L, Shard& s, int sm_count, bool bigreg, bool v_in_ht);
void launch_pp_pack(const Layout& L, Shard& s);                        // kernels_pp.cu: Bp -> Bc
void launch_pp_prep(const Layout& L, Shard& s, int64_t ldc);           // C_occ -> Ci
void launch_pp_half_transform(const Layout& L, Shard& s, int sm_count, bool v_in_ht, bool bigreg);
void launch_pp_density(const Layout& L, Shard& s);
void plan_pp_items(const Layout& L, Shard& s, int sm_count, bool p_major, std::vector<int32_t>& list, std::vector<int32_t>& off);
void launch_pp_coulomb(const Layout& L, Shard& s, cudaStream_t st, cudaEvent_t ev_v, bool v_in_ht);
int64_t pp_coulomb_slabs(const Layout& L);
bool launch_pp_coulomb_fused(const Layout& L, Shard& s, int sm_count, cudaStream_t st);
void launch_exchange(const Layout& L, Shard& s, int sm_count, bool dedicated_producer);
void launch_coulomb(const Layout& L, Shard& s, int sm_count, cudaStream_t st, bool have_v);
bool launch_coulomb_fused(const Layout& L, Shard& s, int sm_count, cudaStream_t st);
void launch_epilogue(const Layout& L, Shard& s, bool add_H);
void launch_mirror_lower(const Layout& L, Shard& s);
void launch_scatter_pairs(const Layout& L, Shard& s, const double* T_dev, int64_t ld, int64_t col_begin,
                          int64_t col_count);
void launch_gather_pairs(const Layout& L, Shard& s, double* T_dev, int64_t ld, int64_t col_begin, int64_t col_count,
                         const int64_t* d_src_of_col, int64_t row_begin, int64_t row_count);
void launch_fill_synthetic(const Layout& L, Shard& s, uint64_t seed, double decay_length);
size_t half_transform_smem_bytes(int bm);
size_t exchange_smem_bytes();
int exchange_tile_cost(int tile, int N, int half_dead = 0);  // kernels_k.cu: DMMA row units per k-step on the busiest sub-partition
void exchange_half_rows(int ta, int tb, int half, int N, int& row_begin, int& row_end);  // kernels_k.cu: rows of tile ta the warps of a row half own
int kernels_per_build();

void set_fuzz_seed_k(unsigned int seed);   // checked build: schedule-fuzzing seed of each kernel translation unit
void set_fuzz_seed_pp(unsigned int seed);
void set_fuzz_seed_j(unsigned int seed);
void set_error(const std::string& msg);
// timing event on the shard's stream: a plain record when launching eagerly, an external event-record NODE while the
// launch sequence is being captured into a graph (so every replay still timestamps the stage boundaries)
inline cudaError_t record_event(const Shard& s, cudaEvent_t ev, cudaStream_t st) {
    return s.capturing ? cudaEventRecordWithFlags(ev, st, cudaEventRecordExternal) : cudaEventRecord(ev, st);
}
void scf_destroy(jcdf_handle* h);        // scf.cu
void integrals_destroy(jcdf_handle* h);  // integrals.cu
void integrals_on_set_mask(jcdf_handle* h);
int factor_metric_on_device(Shard& s, int64_t Q);  // api.cu
void launch_set_identity(double* A, int64_t n, cudaStream_t st);  // kernels_misc.cu

}  // namespace jcdf

// One Fock build over host (device_ptrs = false) or device-resident operands; defined in api.cu, also driven by scf.cu.
extern "C" __attribute__((visibility("hidden"))) int fock_build_impl(jcdf_handle* h, const double* C_occ, int64_t ldc,
                                                                      const double* H, double* F_out,
                                                                      double* stage_seconds, bool device_ptrs);
