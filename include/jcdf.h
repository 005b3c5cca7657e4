/*
 * libjcdf — B200-native (sm_100a) density-fitted RHF Fock build behind JuliaChem.jl's DF-RHF seam.
 *
 * C-ABI boundary (plain pointers and sizes, no C++/torch types).  Every entry point names the reference interface it
 * replaces; file:line are relative to the reference checkout (j-hayes/JuliaChem.jl @ 2064729).  The Julia side that
 * binds these with `ccall` is shown in INTEGRATION.md and shipped as julia/B200DF.jl; the in-repo precedent for a
 * pointer-level C call from this module is `call_gemm!` (src/rhf/energy/DensityFitting/ScreenedDF.jl:645-673).
 *
 * Conventions
 *   - all matrices are column-major FP64 exactly as Julia passes them; index maps are Int64, 1-based, 0 = screened;
 *     Bool arrays are 1 byte per element (Julia `Array{Bool}`).
 *   - every function returns 0 on success and a negative jcdf_status on failure; jcdf_last_error() then returns a
 *     thread-local, NUL-terminated description.  No C++ exception crosses this boundary.
 *   - host pointers are borrowed for the duration of the call only; all device memory is owned by the handle.
 *   - calls on one handle are serial (one caller thread at a time); the library selects its devices itself on
 *     entry, so the caller may migrate between OS threads (Julia tasks).
 */
#ifndef JCDF_H
#define JCDF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct jcdf_handle jcdf_handle;

typedef enum jcdf_status {
    JCDF_OK = 0,
    JCDF_ERR_INVALID = -1,   /* bad argument / call order */
    JCDF_ERR_CUDA = -2,      /* CUDA runtime / cuSOLVER / cuBLAS failure */
    JCDF_ERR_NCCL = -3,      /* NCCL failure or libnccl not loadable */
    JCDF_ERR_NOMEM = -4,     /* device allocation failed */
    JCDF_ERR_NUMERIC = -5    /* metric not positive definite, eigensolver did not converge */
} jcdf_status;

/* stage_seconds[] slots filled by jcdf_fock_build — the reference's per-iteration JCTiming keys
 * (src/shared/JCTiming.jl:58-65; filled at ScreenedDF.jl:238-239, 362-364, DensityFitting.jl:64,70, GPUDF.jl:281-301) */
enum {
    JCDF_T_W = 0,         /* W_time      half-transform                           */
    JCDF_T_K = 1,         /* K_time      exchange                                 */
    JCDF_T_V = 2,         /* V_time      Coulomb intermediate                     */
    JCDF_T_J = 3,         /* J_time      Coulomb matrix                           */
    JCDF_T_ALLREDUCE = 4, /* fock_MPI_time  (NCCL all-reduce)                     */
    JCDF_T_DENSITY = 5,   /* density_time   (C_occ re-layout + packed density)    */
    JCDF_T_H2D = 6,       /* C_occ upload                                         */
    JCDF_T_D2H = 7,       /* Fock download (fock_gpu_cpu_copy_reduce_time)        */
    JCDF_T_TOTAL = 8,     /* fock_time   whole call, host clock                   */
    JCDF_T_EPILOGUE = 9,  /* H add + J scatter + symmetrise (H_add_time, gpu_copy_J_time, gpu_copy_sym_time) */
    JCDF_N_STAGES = 10
};

const char* jcdf_last_error(void);
/* number of visible CUDA devices, or a negative status (replaces length(CUDA.devices()), GPUDF.jl:39) */
int32_t jcdf_device_count(void);

/* Create the per-SCF state the reference keeps in scf_data / scf_data.gpu_data from iteration 1
 * (DensityFitting.jl:34-47; GPUDF.jl:79-155; types src/shared/SCFData.jl:19-37, GPUData_cuda.jl:4-38).
 *   N    primary basis functions (scf_data.mu), Q auxiliary functions (scf_data.A), occ = nels / 2
 *   ndev devices of THIS process (scf_options.num_devices, Constants.jl:67-68); dev_ids NULL = 0..ndev-1          */
int32_t jcdf_create(int64_t N, int64_t Q, int64_t occ, int32_t ndev, const int32_t* dev_ids, jcdf_handle** out);
int32_t jcdf_destroy(jcdf_handle* h);

/* Q sharding.  q_begin/q_len (0-based first auxiliary function, count) for each of this process's devices:
 * the ranges static_load_rank_indicies returns (DynamicLoad.jl:160-203; per device GPUDF.jl:1026-1056).
 * Default when never called: one process owning all of Q, split evenly by function count over its devices.      */
int32_t jcdf_set_shards(jcdf_handle* h, const int64_t* q_begin, const int64_t* q_len);

/* Screening mask and pair compaction (schwarz_screen_itegrals_df, SchwarzScreening.jl:9-83):
 *   fn_mask               N x N Bool, basis_function_screen_matrix (symmetric, diagonal kept)
 *   sparse_pq_index_map   N x N Int64, [q,p] -> 1-based column of T/B in the reference pair order, 0 = screened
 * Both NULL = unscreened ("dense") with column (q,p) -> q + N*p + 1 (setup_unscreened_screening_matricies :97-111
 * modulo its transposed fill, which is the same map for a symmetric tensor).
 * Builds the tile-compacted HBM layout.  Must precede any upload.                                */
int32_t jcdf_set_mask(jcdf_handle* h, const uint8_t* fn_mask, const int64_t* sparse_pq_index_map);

/* Rows of the raw three-centre integrals (P|mn) of device `dev`'s shard: T is (q_len[dev] x P) column-major with
 * leading dimension ld, columns in the reference pair order — exactly three_center_integrals[device_id] of
 * GPUDF.jl:51-57 / calculate_three_center_integrals (ThreeCenterIntegrals.jl:9).  Stored until
 * jcdf_factor_and_form_B.                                                                                        */
int32_t jcdf_upload_three_center(jcdf_handle* h, int32_t dev, const double* T, int64_t ld);
/* Same, for a column window [col_begin, col_begin+col_count) that already lives in device memory on device `dev`
 * (leading dimension ld).  Windows may arrive in any order and may be freed by the caller after the call.        */
int32_t jcdf_upload_three_center_device(jcdf_handle* h, int32_t dev, const double* T_dev, int64_t ld,
                                        int64_t col_begin, int64_t col_count);
/* Two-centre metric (P|Q), Q x Q column-major, lower triangle valid (calculate_two_center_intgrals,
 * TwoCenterIntegrals.jl:7; only the lower triangle is filled :39-41).                                            */
int32_t jcdf_upload_metric(jcdf_handle* h, const double* J_PQ);
/* On-device potrf!('L') + trtri!('L','N') + B = Linv * T (calculate_B!, DensityFitting.jl:128-183;
 * ScreenedDF.jl:92-103; calculate_B_GPU!, GPUDF.jl:828-1008 which does the factorisation on the CPU :890-891).
 * No CPU fallback.  Requires every device's T and the metric.                                                    */
int32_t jcdf_factor_and_form_B(jcdf_handle* h);
/* Alternative to the three calls above when the caller already holds B = L^-1 (P|mn) rows of the shard
 * (scf_data.D of ScreenedDF.jl:103 / gpu_data.device_B): same layout as jcdf_upload_three_center.                 */
int32_t jcdf_set_B(jcdf_handle* h, int32_t dev, const double* B, int64_t ld);
int32_t jcdf_set_B_device(jcdf_handle* h, int32_t dev, const double* B_dev, int64_t ld, int64_t col_begin,
                          int64_t col_count);
/* Copy device `dev`'s packed shard back in the reference layout (q_len[dev] x P column-major, leading dimension ld,
 * columns in the reference pair order): T after jcdf_upload_three_center / jcdf_compute_three_center, B after
 * jcdf_factor_and_form_B / jcdf_set_B — scf_data.D of ScreenedDF.jl:103.  For tests and debugging.                */
int32_t jcdf_get_B(jcdf_handle* h, int32_t dev, double* out, int64_t ld);
/* Same for the row window [row_begin, row_begin + row_count) of the shard (local row numbers): out is
 * (row_count x P) column-major with leading dimension ld.  Lets a checker pull a shard too large for one host array
 * slab by slab (the Fock build is a plain sum over auxiliary rows).                                                */
int32_t jcdf_get_B_rows(jcdf_handle* h, int32_t dev, int64_t row_begin, int64_t row_count, double* out, int64_t ld);
/* Copy the inverse Cholesky factor Linv (Q x Q, lower, column-major) back — for tests (J_AB_invt, ScreenedDF.jl:95) */
int32_t jcdf_get_metric_inverse(jcdf_handle* h, double* Linv_out);
/* Keep H resident on the shard-0 device so later jcdf_fock_build calls may pass H = NULL (H is constant over an SCF:
 * scf_data-level state in the reference, SCF.jl:141-181).                                                        */
int32_t jcdf_set_core_hamiltonian(jcdf_handle* h, const double* H);

/* One Fock build: replaces the body of df_rhf_fock_build! after `occupied_orbital_coefficients = coefficients[:,1:occ]`
 * (DensityFitting.jl:49-71) for contraction_mode == "B200":
 *     F = [H on the shard-0 device] + sum over devices of ( 2 J[Dm] - K[Dm] ),  Dm = C_occ C_occ^T,
 * i.e. calculate_exchange_screened! + calculate_coulomb_screened (ScreenedDF.jl:227-378), `+= H` on rank 0
 * (DensityFitting.jl:62-65) and the reduce over devices/ranks (GPUDF.jl:267-277, DensityFitting.jl:68-71) done as one
 * NCCL all-reduce on the devices.
 *   C_occ  N x occ column-major, leading dimension ldc (host)
 *   H      N x N column-major or NULL (host); only added when this process owns global shard 0
 *   F_out  N x N column-major (host), fully overwritten, exactly symmetric; not retained.  With the option
 *          "exchange_block_screen" and a mask that drops K blocks it is IN/OUT like scf_data.two_electron_fock in the
 *          reference: the elements of dropped blocks come back as F_out_in + H (rank 0), summed over ranks — see the option
 *   stage_seconds  NULL or double[JCDF_N_STAGES]                                                                 */
int32_t jcdf_fock_build(jcdf_handle* h, const double* C_occ, int64_t ldc, const double* H, double* F_out,
                        double* stage_seconds);
/* Device-resident variant: C_occ, H and F_out are device pointers on this process's first device; no host copies.
 * Used to measure the contraction with inputs resident in HBM.                                                   */
int32_t jcdf_fock_build_device(jcdf_handle* h, const double* C_occ_dev, int64_t ldc, const double* H_dev,
                               double* F_out_dev, double* stage_seconds);

/* On-device replacement of iteration()'s linear algebra (src/rhf/energy/SCF.jl:1072-1099):
 *   Fp = X^T F X ; (evals, U) = eigen(Hermitian(Fp)) ascending ; C = X U.
 * F, X: N x N column-major (host).  evals: N.  C: N x N.                                                         */
int32_t jcdf_eigh(jcdf_handle* h, const double* F, const double* X, double* evals, double* C);

/* ---- integral front-end on the GPU -------------------------------------------------------------------
 * Replaces, for contraction_mode == "B200", the host integral drivers of iteration 1 and the H2D copy of their
 * results (GPUDF.jl:44-57): calculate_two_center_intgrals (TwoCenterIntegrals.jl:7), schwarz_screen_itegrals_df
 * (SchwarzScreening.jl:9), calculate_three_center_integrals (ThreeCenterIntegrals.jl:9 / ...Screened.jl) and the C++
 * libint engine behind them (deps/src/jeri-df-tei.hpp:22-96).  Conventions are the reference's: Cartesian functions in
 * libint/CCA order, libint Shell::renorm() on the raw basis-set-exchange coefficients (jeri-core.hpp:41-62), JuliaChem's
 * axial normalisation on every component (Globals.jl:20-44, EnergyHelpers.jl:320-411).
 *
 * jcdf_set_basis: which = 0 primary (shells must add up to N functions, l <= 3), 1 auxiliary (Q functions, l <= 4).
 *   am[nshell], nprim[nshell]; centers 3 x nshell column-major in bohr (after the centre-of-mass shift of
 *   JCBasis.jl:86-101); exponents / coefficients concatenated shell by shell, raw values as in the basis-set file
 *   (L shells already split into s + p, JCBasis.jl:245-290).                                                       */
int32_t jcdf_set_basis(jcdf_handle* h, int32_t which, int64_t nshell, const int32_t* am, const int32_t* nprim,
                       const double* centers, const double* exponents, const double* coefficients);
/* (P|Q) on the shard-0 device; stays there for jcdf_factor_and_form_B (no jcdf_upload_metric needed).
 * J_out: NULL or Q x Q column-major host buffer receiving the full symmetric matrix.                              */
int32_t jcdf_compute_two_center(jcdf_handle* h, double* J_out);
/* Schwarz screening with threshold sigma^2 / max_P (P|P) (SchwarzScreening.jl:22): shell pair kept iff |sum of the whole
 * (ab|ab) block| >= threshold (:44-50), function pair inside a kept shell pair kept iff |(mn|mn)| >= threshold (:60-66);
 * pair compaction outer p, inner q (:72-81).  Ends with jcdf_set_mask on the result.  sigma <= 0: unscreened (:97-111).
 * fn_mask_out: NULL or N x N bytes; pairs_kept_out: NULL or the count P.  Needs the metric (either route) first.    */
int32_t jcdf_compute_schwarz_mask(jcdf_handle* h, double sigma, uint8_t* fn_mask_out, int64_t* pairs_kept_out);
/* (P|mn) of every device's Q shard, written directly into the tile-packed layout (screened pairs and shell pairs the
 * Schwarz test dropped are never computed).  Follow with jcdf_factor_and_form_B.                                   */
int32_t jcdf_compute_three_center(jcdf_handle* h);

/* B = L^-1 (P|mn) for every device of this process with NO exchange between devices or ranks (replaces
 * the host-staged block exchange of calculate_B_GPU!, GPUDF.jl:918-997, and of the multi-rank CPU path,
 * ScreenedDF.jl:134-190): each device factors its own copy of the metric and recomputes the raw integrals of the
 * auxiliary rows h <= its shard in chunks, B_g = sum_h Linv[Q_g, Q_h] T_h.  Works for one process per GPU (torchrun / MPI
 * ranks), where jcdf_factor_and_form_B cannot reach the other ranks' integrals.  Needs both basis sets and the mask.  */
int32_t jcdf_compute_B(jcdf_handle* h);
/* One-electron matrices S, T, V = -sum_C Z_C (m| 1/|r - C| |n) over the primary basis and the nuclear repulsion — what
 * rhf_kernel hands the SCF loop (SCF.jl:100-135; libint one-body engines, deps/src/jeri-oei.hpp).  centers: 3 x natoms
 * column-major, bohr.  Outputs are N x N column-major host buffers (NULL = not wanted); H = T + V.                   */
int32_t jcdf_compute_one_electron(jcdf_handle* h, int32_t natoms, const double* charges, const double* centers,
                                  double* S_out, double* T_out, double* V_out, double* E_nuc_out);

/* ---- on-device SCF step: F, C, D and the DIIS history never leave the GPU between builds ------------
 * jcdf_scf_init replaces the set-up part of rhf_kernel (src/rhf/energy/SCF.jl:141-183): uploads S and H (N x N
 * column-major, host), builds the symmetric orthogonaliser X = U_good diag(s_good^-1/2) U_good^T on the device
 * (overlap eigenvalues below 1e-6 dropped, :141-163) and runs the hcore guess iteration(F = H) (:178-181).
 * ndiis = scf_flags["ndiis"] (0 disables DIIS); E_nuc is added to every reported energy (:524).
 * E_guess_out (NULL allowed) receives E_elec(hcore guess) + E_nuc.  Requires B to be formed (jcdf_fock_build ready). */
int32_t jcdf_scf_init(jcdf_handle* h, const double* S, const double* H, double E_nuc, int32_t ndiis, double* E_guess_out);
/* One pass of the `while !iter_converged` body of scf_cycles_kernel (SCF.jl:399-551): Fock build from the resident
 * C[:,1:occ] (:463), DIIS on e = FDS - (FDS)^T with the newest-first history (:471-501; DIIS(), EnergyHelpers.jl:234-258;
 * a singular system takes the "Faulty DIIS!" branch), dynamic damping (:504-508), iteration() (:510-512, :1072-1125),
 * dE and D_rms (:521-525).  result (NULL allowed) = { E, dE, D_rms, E_elec, t_fock, t_diis, t_iteration, t_total } (s). */
int32_t jcdf_scf_step(jcdf_handle* h, double* result);
/* Repeat jcdf_scf_step until |dE| <= dele and D_rms <= rmsd (SCF.jl:527-530, :551) or max_iterations.
 * history: NULL or double[3 * max_iterations] receiving (E, dE, D_rms) per iteration.                             */
int32_t jcdf_scf_run(jcdf_handle* h, int32_t max_iterations, double dele, double rmsd, double* E_out,
                     int32_t* iterations_out, int32_t* converged_out, double* history);
/* Copy a resident matrix back to the host: name = "F" (last extrapolated/damped Fock), "D", "C", "X", "S", "H"
 * (N x N column-major) or "evals" (N) — the entries of the Dict rhf_kernel returns (SCF.jl:251-259).              */
int32_t jcdf_scf_get(jcdf_handle* h, const char* name, double* out);

/* ---- one process per GPU (torchrun / MPI ranks > 1) --------------------------------------------------------------
 * The reference sums rank-partial Fock matrices with MPI.Allreduce! (DensityFitting.jl:68-71).  Here rank 0 calls
 * jcdf_comm_unique_id, ships the 128 bytes to the other ranks by any means, and every rank calls jcdf_comm_init_rank;
 * jcdf_fock_build then all-reduces across all devices of all ranks and H is added by global rank 0 only.          */
int32_t jcdf_comm_unique_id(uint8_t id_out[128]);
int32_t jcdf_comm_init_rank(jcdf_handle* h, int32_t n_ranks, int32_t rank, const uint8_t id[128]);

/* ---- introspection (tests, bench, DESIGN.md numbers) ------------------------------------------------------------- */
typedef struct jcdf_layout_info {
    int64_t N, Q, occ;
    int64_t n_pad;            /* N rounded up to the 32-wide tile edge                      */
    int64_t occ_pad;          /* occ rounded up to 8                                        */
    int64_t tiles_total;      /* lower-triangular 32x32 tiles incl. diagonal                */
    int64_t tiles_kept;       /* tiles holding at least one unscreened pair                 */
    int64_t pairs_kept;       /* P   = screened_indices_count (ordered pairs)               */
    int64_t pairs_tri;        /* P_tri = kept pairs with q >= p                             */
    int64_t packed_row_len;   /* doubles per auxiliary row in HBM = tiles_kept * 1024       */
    int64_t q_local;          /* auxiliary functions held by this process                   */
    int64_t bytes_B, bytes_W; /* device bytes, summed over this process's devices (bytes_B includes the
                               * partner-compacted second copy when the per-function family is active)   */
    int64_t kernel_launches_per_build; /* launches of this library's own kernels per build  */
    int64_t graph_replays;    /* builds (summed over devices) whose kernel sequence ran as a CUDA-graph replay */
    /* block-screened exchange (option "exchange_block_screen"; all 0 when it is off or the mask is NULL): */
    int64_t exchange_blocks_total;   /* JCTC.total_exchange_blocks: lower-triangular K blocks of the reference's grid      */
    int64_t exchange_blocks_kept;    /* JCTC.unscreened_exchange_blocks: length(blocks_to_calculate), ScreenedDF.jl:454      */
    int64_t exchange_block_width;    /* K_block_width (:392-397)                                                            */
    int64_t exchange_tiles_total;    /* this library's own 128 x 128 lower-triangular exchange tiles (always filled)       */
    int64_t exchange_tiles_skipped;  /* ... of which left out of the work lists: every element lies in a dropped block     */
    int64_t exchange_half_tiles_skipped; /* 64-row halves of the remaining off-diagonal tiles that lie wholly in dropped blocks
                                      * (their warps compute nothing)                                                      */
} jcdf_layout_info;
int32_t jcdf_get_layout(jcdf_handle* h, jcdf_layout_info* out);

/* Deterministic synthetic shard for benchmarks at sizes no host array can hold: fills the packed B of every device
 * directly in HBM with B[Q,(p,q)] = decay(p,q) * gauss(seed, Q, min(p,q), max(p,q)) / sqrt(Q_total) for kept pairs.
 * jcdf_synthetic_value returns the same value on the host so tests can rebuild the tensor for the oracle.        */
int32_t jcdf_fill_synthetic_B(jcdf_handle* h, uint64_t seed, double decay_length);
double jcdf_synthetic_value(uint64_t seed, int64_t Q_total, int64_t q_aux, int64_t p, int64_t q, double decay_length);

/* Tuning knobs (tests, bench).  Names: "coulomb_fused" (tile-family Coulomb: 2 = default, single-pass cooperative J
 * kernel on long column segments and two-pass sweeps on short ones; 1 = single pass whenever a row segment fits the
 * shared-memory ring; 0 = two-pass sweeps), "coulomb_overlap" (two-pass sweeps on a second
 * stream, concurrent with the tensor-core kernels; default 0), "exchange_slices" (0 = automatic split of the K
 * contraction index), "regroup" (1 = default: mask-driven internal function order), "ht_mode" (half-transform kernel:
 * 0 = default, chosen from the mask by executed DMMA count; 1 = 32 x 32 tiles with strip skipping; 2 = per-function
 * partner lists over a partner-compacted copy of B = calculate_W_screened's own blocking, ScreenedDF.jl:242-289),
 * "coulomb_pp" (1 = default: with the per-function kernel the Coulomb passes stream the compact copy too),
 * "coulomb_v_in_ht" (1 = default: V[Q] = sum_p sum_i C[p,i] W[Q,i,p] is taken from the half-transform's accumulators — by the
 * per-function kernel whenever the compact Coulomb is active, by the tile kernel when "coulomb_fused" is left at 2 — so the
 * Coulomb build needs no density pass and streams B once instead of twice; 0 = density + separate V pass),
 * "pp_layout" (consumer warp arrangement of the per-function kernel: -1 = default, chosen from occ; 0 = 2 x 4 warps,
 * orbitals padded to 32 per item; 1 = 8 x 1 warps, no padding beyond the 8-wide DMMA tile; 2 = two CTAs per SM of 2 x 2 warps
 * each, orbitals padded to 16 per item — measured slower than 0 / 1 on every benchmark shape, never chosen automatically),
 * "pp_max_bytes" (cap on the compact copy per shard; above it — or when cudaMalloc refuses — the automatic choice falls
 * back to the tile family, ht_mode = 2 reports JCDF_ERR_CUDA; 0 = no cap),
 * "coulomb_pp_fused" (0 = default: two streaming passes; 1 = one cooperative single-pass kernel, measured slower),
 * "exchange_balance" (1 = default: cost-balanced exchange work lists; 0 = round-robin),
 * "pp_order" (work lists of the per-function kernel: 1 = default, a CTA walks its items function by function, so CTAs that
 * run side by side gather the same rows of C; 0 = chunk by chunk — measured 3 % slower at w50),
 * "pp_bigreg" (1 = default: both half-transform kernels run 12 warps and moves registers from the producer's warpgroup to the
 * consumers with setmaxnreg, 40 / 232 per thread; 0 = 9 warps at the 168-register launch cap),
 * "exchange_producer" (1 = default: the exchange kernel runs 12 warps with a dedicated copy-producer warp and the same register
 * split; 0 = 8 warps, thread 0 of a DMMA warp also issues the copies),
 * "register_output" (0 = default; 1 = a host F_out that arrives twice in a row is page-locked in place with
 * cudaHostRegister until it changes or the handle is destroyed — scf_data.two_electron_fock is a pageable Julia Array, and
 * a copy into pageable memory runs at a fraction of the PCIe rate.  Opt-in: the caller must keep that buffer alive until
 * jcdf_destroy; julia/B200DF.jl holds a reference to it inside SCFGPUDataB200 for exactly that reason),
 * "graph" (1 = default: the kernel sequence of a build, prep .. epilogue, is captured into a CUDA graph on the first build
 * and replayed afterwards; 0 = every kernel launched eagerly),
 * "exchange_block_screen" (0 = default, off; n > 0 = the reference's keyword df_exchange_screen = true with
 * df_exchange_n_blocks = n, the caller having replaced 0 by the reference's default of 10: jcdf_set_mask builds
 * calculate_exchange_block_screen_matrix's K block grid — below 100 functions one block, else width N div n, the remainder
 * strip never screened — and its block_screen_matrix from the mask, ScreenedDF.jl:385-457, reported by jcdf_get_layout; the
 * exchange is then calculate_K_lower_diagonal_block, :459-545: K is computed for the kept blocks only.  The reference neither
 * computes NOR WRITES a dropped block, and its Coulomb copy skips it too because all its pairs are screened (:367-378), so
 * there scf_data.two_electron_fock keeps what it held — the previous iteration's matrix after the caller's DIIS and damping,
 * SCF.jl:471-508 — plus H on rank 0, and with n ranks the all-reduce sums n such copies (DensityFitting.jl:62-71).  libjcdf
 * reproduces exactly that: F_out is read before the build and each shard acts as one rank.  A NULL mask (dense statement)
 * has no K blocks and ignores the option),
 * "regroup", "ht_mode", "pp_layout", "pp_order", "pp_max_bytes", "exchange_slices", "exchange_balance" and
 * "exchange_block_screen" must be set before jcdf_set_mask.  A refused call (unknown name, value out of
 * range, too late) returns JCDF_ERR_INVALID and leaves the handle unchanged.                                          */
int32_t jcdf_set_option(jcdf_handle* h, const char* name, int64_t value);

/* Per-kernel device time of the last build of device `dev` in ms (CUDA events on the launching stream):
 * out[0..JCDF_N_STAGES) in the stage order above.                                                                */
int32_t jcdf_last_kernel_ms(jcdf_handle* h, int32_t dev, double* out);

/* Per-CTA busy time of the last exchange launch of device `dev`, in SM clock cycles (clock64 around the CTA's whole work
 * list): out[0..n) for the first n persistent CTAs; returns the number of CTAs launched.  Profiling aid for the
 * cost-balanced work lists (no reference counterpart).                                                            */
int32_t jcdf_exchange_cta_cycles(jcdf_handle* h, int32_t dev, int64_t* out, int32_t n);

/* Host-side view of the exchange kernel's work split (no device needed; used by the CPU test suite): the block of the
 * 128 x 128 output tile (ta, tb), ta >= tb, that warp `warp` (0..7) owns for a basis of N functions, in units of 8 rows /
 * columns: out5 = {first row unit, first column unit, row units, column units, staircase}; staircase = 1 when the
 * rectangle starts on the diagonal of a diagonal tile and only its column units <= row unit are computed; row and
 * column units are zero when the warp has no work.
 * The same function drives the kernel and the cost model of the work lists.                                        */
int32_t jcdf_debug_exchange_block(int32_t ta, int32_t tb, int32_t warp, int64_t N, int32_t* out5);

/* 1 when the library is the bounds-checked build (make CHECKED=1: every global store and bulk-copy source of the pipelined
 * kernels is compared with its buffer's extent, a violation prints its place and traps), else 0.  fire != 0 on a checked build
 * launches a kernel whose check must fail and returns the resulting CUDA error code (> 0): the proof that the checks are live;
 * the CUDA context is unusable afterwards.  compute-sanitizer is closed on the development pool; this build stands in for
 * its memcheck tool.                                                                                                   */
int32_t jcdf_checked_selftest(int32_t fire);
/* Checked build only (a no-op otherwise): schedule fuzzing.  With seed != 0 every mbarrier wait / arrive, bulk-copy issue and
 * grid-wide flag operation of the pipelined kernels is preceded, one time in four, by a pseudo-random sleep of up to 4 us;
 * results must not depend on the seed.  The stand-in for compute-sanitizer's racecheck, which is closed on the pool.        */
int32_t jcdf_checked_set_fuzz(jcdf_handle* h, uint32_t seed);

/* The CUDA stream (cudaStream_t, returned as void*) every kernel, copy and all-reduce of device `dev` is issued on —
 * so a caller can bracket builds with its own CUDA events or order its own work against them (the fork's CUDA.jl
 * path uses the task-local default stream and CUDA.synchronize() instead, GPUDF.jl:206-263).                     */
int32_t jcdf_get_stream(jcdf_handle* h, int32_t dev, void** stream_out);

/* No counterpart in the reference: added for callers that may not allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  One dense Fock build
 *     F = H + 2 J[D] - K[D],  D = C_occ C_occ^T,  over B = L^-1 (P|mn)
 * — the contraction of jcdf_fock_build on a single unscreened shard — that calls no cudaMalloc / cudaFree and no allocating
 * library routine: its scratch (the half-transformed W, the Coulomb vector, per-chunk partial Fock matrices) is a fixed
 * file-scope __device__ array of 1 722 368 bytes (csrc/no_alloc.cu) that every GPU loading the library holds until exit.
 * Sizes beyond JCDF_NO_ALLOC_MAX_* are refused with JCDF_ERR_INVALID (use jcdf_create + jcdf_fock_build there).
 *   B_dev      Q x N^2 column-major, leading dimension ldb >= Q: B[Q, p + N q], symmetric in (p, q) — the layout
 *              jcdf_set_B takes when no mask is set
 *   C_occ_dev  N x occ column-major, leading dimension ldc >= N;  H_dev  N x N column-major or NULL
 *   F_out_dev  N x N column-major, fully overwritten, exactly symmetric
 *   stream     cudaStream_t (NULL = the default stream) of the CURRENT device, to which all four pointers must belong; the
 *              four kernels are enqueued there and the call returns without synchronising.
 * All calls on one GPU share the array and it keeps the last call's contents (every word read is rewritten first, nothing
 * needs clearing): calls ordered on one stream are safe, calls that overlap on the GPU race.  The first call on each device
 * looks up the array's device address (cudaGetSymbolAddress); make that call outside a stream capture.                  */
#define JCDF_NO_ALLOC_MAX_N 48
#define JCDF_NO_ALLOC_MAX_Q 256
#define JCDF_NO_ALLOC_MAX_OCC 16
int32_t jcdf_fock_build_no_alloc(int64_t N, int64_t Q, int64_t occ, const double* B_dev, int64_t ldb,
                                 const double* C_occ_dev, int64_t ldc, const double* H_dev, double* F_out_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* JCDF_H */
