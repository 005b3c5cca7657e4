# B200DF.jl — the Julia side of the B200 drop-in for JuliaChem's density-fitted RHF Fock build.
#
# A maintainer adds this file next to GPUDF.jl (src/rhf/energy/DensityFitting/) and includes it from Energy.jl:20-29;
# INTEGRATION.md lists the other one-line edits (Constants.jl, SCF.jl, DensityFitting.jl).  Everything device-side lives
# in libjcdf.so (include/jcdf.h); this file only marshals pointers, in the style of the module's own raw BLAS call
# `call_gemm!` (ScreenedDF.jl:645-673).  It follows df_rhf_fock_build_GPU! (GPUDF.jl:11-304) stage by stage;
# juliachem.jl_b200/df.py + rhf.py are the same wrapper in Python (Julia is not present in the build image) and are what
# the test-suite drives.
#
# NOT EXECUTED in this repository's CI (no Julia toolchain in the image).  tests/test_julia_binding.py parses every
# `ccall` below and checks symbol, arity and C types against the table the Python wrapper binds (_lib.SIGNATURES), and
# every JCTC.<name> against src/shared/JCTiming.jl:50-107 when the reference checkout is present.

using Libdl

const libjcdf = get(ENV, "JULIACHEM_LIBJCDF", "libjcdf")   # deps/build.jl puts libjcdf.so next to libjeri.so

const JCDF_N_STAGES = 10
const JCDF_LAYOUT_INFO_FIELDS = 21   # jcdf_layout_info: 21 x int64_t (include/jcdf.h); 16-17 = exchange_blocks_total / _kept
# stage slot (0-based in jcdf.h) -> JCTiming key prefix (src/shared/JCTiming.jl:58-107).  Slot 6 (C_occ upload) is what
# the fork's GPU path books as GPU_non_zero_coeff_time of device 1 (GPUDF.jl:288); slot 9 (J scatter + H add + mirror)
# as gpu_copy_J_time (GPUDF.jl:289).
const JCDF_STAGE_KEYS = (JCTC.W_time, JCTC.K_time, JCTC.V_time, JCTC.J_time, JCTC.fock_MPI_time, JCTC.density_time,
    JCTiming_GPUkey(JCTC.GPU_non_zero_coeff_time, 1), JCTC.fock_gpu_cpu_copy_reduce_time, JCTC.fock_time,
    JCTC.gpu_copy_J_time)

# what scf_data.gpu_data holds for contraction_mode == "B200" (cf. SCFGPUData_cuda, GPUData_cuda.jl:4-38)
mutable struct SCFGPUDataB200 <: SCFGPUData
    handle::Ptr{Cvoid}
    number_of_devices_used::Int
    pinned_fock::Any   # scf_data.two_electron_fock, page-locked by the library ("register_output"): kept alive until the finalizer
    function SCFGPUDataB200()
        d = new(C_NULL, 0, nothing)
        finalizer(d) do x   # device memory lives exactly as long as scf_data
            x.handle == C_NULL || ccall((:jcdf_destroy, libjcdf), Cint, (Ptr{Cvoid},), x.handle)
            x.handle = C_NULL
        end
        return d
    end
end

@inline function jcdf_check(status::Cint)
    status == 0 && return
    error(unsafe_string(ccall((:jcdf_last_error, libjcdf), Cstring, ())))   # same convention as error(...) GPUDF.jl:40
end

# JCModules.Shell keeps the contraction coefficients AFTER calculate_coefficients(..., unnormalize = true) divided each of
# them by sqrt((pi/2a)^1.5 * {p: 1/(2*2a), d: 3/(4*(2a)^2), f: 15/(8*(2a)^3)}) (BasisStructs.jl:62-80); libint — and
# jcdf_set_basis — take the RAW basis-set-exchange values (JCBasis.jl:301-307, jeri-core.hpp:41-62).  Undo the division.
function jcdf_raw_coefficients(shell)
    raw = Vector{Float64}(undef, shell.nprim)
    for iprim in 1:shell.nprim
        two_exp = 2.0 * shell.exponents[iprim]
        f = (pi / two_exp)^1.5
        f *= shell.am == 2 ? 0.5 / two_exp : 1.0
        f *= shell.am == 3 ? 0.75 / (two_exp^2) : 1.0
        f *= shell.am == 4 ? 1.875 / (two_exp^3) : 1.0
        raw[iprim] = shell.coefficients[iprim] * sqrt(f)
    end
    return raw
end

# flat shell arrays of one Basis (BasisStructs.jl:170-178) -> jcdf_set_basis; which = 0 primary, 1 auxiliary
function jcdf_set_basis_from!(handle::Ptr{Cvoid}, which::Integer, basis)
    nshell = length(basis.shells)
    am = Vector{Cint}(undef, nshell)
    nprim = Vector{Cint}(undef, nshell)
    centers = Matrix{Float64}(undef, 3, nshell)
    exponents = Float64[]
    coefficients = Float64[]
    for (i, sh) in enumerate(basis.shells)
        am[i] = sh.am - 1                      # JuliaChem counts s = 1 (JCBasis.jl:301 passes am - 1 to libint as well)
        nprim[i] = sh.nprim
        centers[:, i] .= sh.atom_center        # bohr, after the centre-of-mass shift (JCBasis.jl:86-101)
        append!(exponents, sh.exponents[1:sh.nprim])
        append!(coefficients, jcdf_raw_coefficients(sh))
    end
    GC.@preserve am nprim centers exponents coefficients begin
        jcdf_check(ccall((:jcdf_set_basis, libjcdf), Cint,
            (Ptr{Cvoid}, Cint, Int64, Ptr{Cint}, Ptr{Cint}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
            handle, Cint(which), Int64(nshell), am, nprim, centers, exponents, coefficients))
    end
end

# Iteration-1 set-up with the reference's own host integral drivers (libint through JERI) and an H2D upload of their
# results — the route of GPUDF.jl:44-57.  One MPI rank only: jcdf_factor_and_form_B cannot reach another rank's (P|mn).
function setup_host_integrals_B200!(gd::SCFGPUDataB200, scf_data, jeri_engine_thread_df, jeri_engine_thread, basis_sets,
    scf_options, jc_timing, num_devices::Int)
    two_center_integrals = calculate_two_center_intgrals(jeri_engine_thread_df, basis_sets, scf_options)
    get_screening_metadata!(scf_data, scf_options.df_screening_sigma, jeri_engine_thread, two_center_integrals,
        basis_sets, jc_timing)
    sd = scf_data.screening_data
    # mask + pair compaction -> tile-compacted HBM layout (replaces setup_gpu_screening_data!, GPUDF.jl:347-420)
    GC.@preserve sd begin
        jcdf_check(ccall((:jcdf_set_mask, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Bool}, Ptr{Int64}), gd.handle,
            sd.basis_function_screen_matrix, sd.sparse_pq_index_map))
    end
    for device_id in 1:num_devices   # screened (P|mn) rows of each device's shard (GPUDF.jl:51-57)
        T3 = calculate_three_center_integrals(jeri_engine_thread_df, basis_sets, scf_options, scf_data,
            device_id - 1, num_devices, true)                                   # (Q_dev, P) column-major
        GC.@preserve T3 jcdf_check(ccall((:jcdf_upload_three_center, libjcdf), Cint,
            (Ptr{Cvoid}, Cint, Ptr{Float64}, Int64), gd.handle, Cint(device_id - 1), T3, Int64(size(T3, 1))))
    end
    # potrf + trtri + B = L^-1 T on the devices (replaces calculate_B_GPU!, GPUDF.jl:828-1008, incl. its CPU LAPACK step)
    jcdf_check(ccall((:jcdf_upload_metric, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}), gd.handle, two_center_integrals))
    jcdf_check(ccall((:jcdf_factor_and_form_B, libjcdf), Cint, (Ptr{Cvoid},), gd.handle))
end

# Iteration-1 set-up with the integral front-end of the library (default; any number of MPI ranks): (P|Q), the Schwarz
# mask with the reference's rule (SchwarzScreening.jl:9-83) and B = L^-1 (P|mn) are computed on the devices; every device
# recomputes the auxiliary rows it needs, so no rank ever waits for another rank's integrals (jcdf_compute_B).
function setup_device_integrals_B200!(gd::SCFGPUDataB200, scf_data, basis_sets, scf_options, n_ranks::Int, rank::Int, comm)
    jcdf_set_basis_from!(gd.handle, 0, basis_sets.primary)
    jcdf_set_basis_from!(gd.handle, 1, basis_sets.auxillary)
    jcdf_check(ccall((:jcdf_compute_two_center, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}), gd.handle, C_NULL))
    N = scf_data.μ
    # the reference's choice between its dense and its screened GPU algorithm (run_gpu_fock_build!, DensityFitting.jl:79-90):
    # dense = no Schwarz mask at all, which jcdf_compute_schwarz_mask does for sigma <= 0 (SchwarzScreening.jl:97-111)
    dense = scf_options.df_force_dense || (scf_options.df_use_adaptive && N < 800 && n_ranks == 1)
    sigma = dense ? 0.0 : Float64(scf_options.df_screening_sigma)
    fn_mask = Matrix{Bool}(undef, N, N)
    pairs_kept = Ref{Int64}(0)
    GC.@preserve fn_mask begin   # ends with jcdf_set_mask on the result
        jcdf_check(ccall((:jcdf_compute_schwarz_mask, libjcdf), Cint, (Ptr{Cvoid}, Float64, Ptr{Bool}, Ref{Int64}),
            gd.handle, sigma, fn_mask, pairs_kept))
    end
    scf_data.screening_data.basis_function_screen_matrix = fn_mask
    scf_data.screening_data.screened_indices_count = pairs_kept[]
    if n_ranks > 1   # NCCL communicator over all devices of all ranks; the 128-byte id travels by MPI
        id = zeros(UInt8, 128)
        rank == 0 && jcdf_check(ccall((:jcdf_comm_unique_id, libjcdf), Cint, (Ptr{UInt8},), id))
        MPI.Bcast!(id, 0, comm)
        jcdf_check(ccall((:jcdf_comm_init_rank, libjcdf), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{UInt8}), gd.handle,
            Cint(n_ranks), Cint(rank), id))
    end
    jcdf_check(ccall((:jcdf_compute_B, libjcdf), Cint, (Ptr{Cvoid},), gd.handle))
end

function df_rhf_fock_build_B200!(scf_data, jeri_engine_thread_df::Vector{T}, jeri_engine_thread::Vector{T2},
    basis_sets::CalculationBasisSets, occupied_orbital_coefficients, iteration, scf_options::SCFOptions,
    H::Array{Float64}, jc_timing::JCTiming) where {T<:DFRHFTEIEngine,T2<:RHFTEIEngine}

    comm = MPI.COMM_WORLD
    rank = MPI.Comm_rank(comm)
    n_ranks = MPI.Comm_size(comm)
    num_devices = scf_options.num_devices
    num_devices_global = num_devices * n_ranks
    gd = scf_data.gpu_data::SCFGPUDataB200

    if iteration == 1
        host_integrals = get(ENV, "JULIACHEM_B200_HOST_INTEGRALS", "0") == "1"
        # refuse BEFORE any collective or integral work (the reference's options the library does not take are refused
        # the same way: SCFOptions.jl:86-93)
        if host_integrals && n_ranks > 1
            error("B200 path: JULIACHEM_B200_HOST_INTEGRALS=1 needs one MPI rank (B formation from uploaded integrals does not cross ranks); unset it to compute the integrals on the devices")
        end
        if host_integrals && scf_options.df_force_dense
            error("B200 path: df_force_dense needs the device integral route (unset JULIACHEM_B200_HOST_INTEGRALS)")
        end
        navail = ccall((:jcdf_device_count, libjcdf), Cint, ())
        if num_devices > navail                                                     # GPUDF.jl:39-41
            error("Number of devices requested is greater than the number of devices available")
        end
        h = Ref{Ptr{Cvoid}}(C_NULL)
        jcdf_check(ccall((:jcdf_create, libjcdf), Cint, (Int64, Int64, Int64, Cint, Ptr{Cint}, Ref{Ptr{Cvoid}}),
            Int64(scf_data.μ), Int64(scf_data.A), Int64(scf_data.occ), Cint(num_devices), C_NULL, h))
        gd.handle = h[]
        gd.number_of_devices_used = num_devices
        # Q shards: the reference's contiguous-auxiliary-shell rule per global device (GPUDF.jl:50, DynamicLoad.jl:160-203)
        q_begin = Vector{Int64}(undef, num_devices)
        q_len = Vector{Int64}(undef, num_devices)
        for device_id in 1:num_devices
            global_device_id = device_id + rank * num_devices
            _, aux_indicies, _ = static_load_rank_indicies(global_device_id - 1, num_devices_global, basis_sets)
            q_begin[device_id] = first(aux_indicies) - 1
            q_len[device_id] = length(aux_indicies)
        end
        jcdf_check(ccall((:jcdf_set_shards, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Int64}, Ptr{Int64}), gd.handle, q_begin, q_len))
        if scf_options.df_screen_exchange
            # df_exchange_screen = true (calculate_exchange_block_screen_matrix + calculate_K_lower_diagonal_block,
            # ScreenedDF.jl:385-545): K only for the blocks of the df_exchange_n_blocks grid whose mask sub-block holds a pair.
            # The reference never writes a dropped block, so it keeps what scf_data.two_electron_fock held (the last F after
            # DIIS and damping) plus H; the library reproduces that by reading F_out before every build.  Set before the
            # mask; the dense statement of the contraction (df_force_dense / adaptive) has no K blocks and ignores it.
            n_blocks = scf_options.df_exchange_n_blocks == 0 ? SCF_Keywords.Screening.df_exchange_n_blocks_cpu_default :
                       scf_options.df_exchange_n_blocks
            jcdf_check(ccall((:jcdf_set_option, libjcdf), Cint, (Ptr{Cvoid}, Cstring, Int64), gd.handle,
                "exchange_block_screen", Int64(n_blocks)))
        end

        if host_integrals
            setup_host_integrals_B200!(gd, scf_data, jeri_engine_thread_df, jeri_engine_thread, basis_sets, scf_options,
                jc_timing, num_devices)
        else
            setup_device_integrals_B200!(gd, scf_data, basis_sets, scf_options, n_ranks, rank, comm)
        end
        jcdf_check(ccall((:jcdf_set_core_hamiltonian, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}), gd.handle, H))
        if scf_options.df_screen_exchange
            # what calculate_exchange_block_screen_matrix books (ScreenedDF.jl:409-410, 454), from the library's own block grid
            li = zeros(Int64, JCDF_LAYOUT_INFO_FIELDS)
            jcdf_check(ccall((:jcdf_get_layout, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Int64}), gd.handle, li))
            jc_timing.non_timing_data[JCTC.total_exchange_blocks] = string(li[16])
            jc_timing.non_timing_data[JCTC.unscreened_exchange_blocks] = string(li[17])
            if scf_options.df_exchange_n_blocks == 0                                 # the reference updates the option in place
                scf_options.df_exchange_n_blocks = SCF_Keywords.Screening.df_exchange_n_blocks_cpu_default   # :388-390
            end
            if scf_data.μ < 100
                scf_options.df_exchange_n_blocks = 1                                  # :392-394
            end
            jc_timing.non_timing_data[JCTC.df_exchange_n_blocks] = string(scf_options.df_exchange_n_blocks)
        end
        # two_electron_fock is allocated once per SCF (DensityFitting.jl:45) and is pageable: let the library page-lock it in
        # place after its second use; the reference held here keeps it alive until the handle is destroyed
        gd.pinned_fock = scf_data.two_electron_fock
        jcdf_check(ccall((:jcdf_set_option, libjcdf), Cint, (Ptr{Cvoid}, Cstring, Int64), gd.handle, "register_output", Int64(1)))
    end

    # every iteration: one call.  F_out is scf_data.two_electron_fock (caller-owned; DIIS mutates it afterwards).  The
    # library adds H on global shard 0 and all-reduces over every device of every rank, so the caller skips both
    # `F .+= H` (DensityFitting.jl:62-65) and MPI.Allreduce! (:68-71) for this mode.
    stage_seconds = zeros(Float64, JCDF_N_STAGES)
    C_occ = occupied_orbital_coefficients          # N x occ column-major, as sliced at DensityFitting.jl:49
    F = scf_data.two_electron_fock
    GC.@preserve C_occ F stage_seconds begin
        jcdf_check(ccall((:jcdf_fock_build, libjcdf), Cint,
            (Ptr{Cvoid}, Ptr{Float64}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
            gd.handle, C_occ, Int64(stride(C_occ, 2)), C_NULL, F, stage_seconds))
    end
    for (slot, key) in enumerate(JCDF_STAGE_KEYS)
        jc_timing.timings[JCTiming_key(key, iteration)] = stage_seconds[slot]
    end
end

# on-device replacement of iteration()'s linear algebra (SCF.jl:1072-1099): returns eigenvalues, fills C = X * U
function iteration_eigh_B200!(scf_data, F::Matrix{Float64}, ortho::Matrix{Float64}, F_eval::Vector{Float64}, C::Matrix{Float64})
    gd = scf_data.gpu_data::SCFGPUDataB200
    jcdf_check(ccall((:jcdf_eigh, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
        gd.handle, F, ortho, F_eval, C))
    return F_eval
end

# ---- on-device SCF step (SURVEY §8f-2) --------------------------------------------------------------------------------
# Replaces, for contraction_mode == "B200" with scf_flags["b200_resident_scf"] = true, the body of scf_cycles_kernel's
# `while !iter_converged` loop (SCF.jl:399-551): F, C, D and the DIIS history stay on the GPU; Julia only sees the
# scalars it prints (:555) and tests for convergence (:527-530).
function scf_init_B200!(scf_data, S::Matrix{Float64}, H::Matrix{Float64}, E_nuc::Float64, ndiis::Int)
    gd = scf_data.gpu_data::SCFGPUDataB200
    E_guess = Ref{Float64}(0.0)
    jcdf_check(ccall((:jcdf_scf_init, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Float64, Cint, Ref{Float64}),
        gd.handle, S, H, E_nuc, Cint(ndiis), E_guess))
    return E_guess[]
end

# returns (E, ΔE, D_rms, E_elec) and fills the fock/DIIS/iteration timings of this iteration
function scf_step_B200!(scf_data, iter::Int, jc_timing)
    gd = scf_data.gpu_data::SCFGPUDataB200
    r = zeros(Float64, 8)
    jcdf_check(ccall((:jcdf_scf_step, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}), gd.handle, r))
    jc_timing.timings[JCTiming_key(JCTC.fock_time, iter)] = r[5]
    jc_timing.timings[JCTiming_key(JCTC.iteration_time, iter)] = r[8]
    return r[1], r[2], r[3], r[4]
end

# the matrices rhf_kernel returns in its Dict (SCF.jl:251-259): name in ("F", "D", "C", "X", "S", "H")
function scf_get_B200(scf_data, name::String)
    gd = scf_data.gpu_data::SCFGPUDataB200
    N = scf_data.μ
    out = zeros(Float64, N, N)
    jcdf_check(ccall((:jcdf_scf_get, libjcdf), Cint, (Ptr{Cvoid}, Cstring, Ptr{Float64}), gd.handle, name, out))
    return out
end

# ---- allocation-free dense build (no counterpart in the reference) -------
# For callers that may not allocate (CUDA graph capture, serving loops): B (Q x N^2), C_occ (N x occ), H and F are device
# arrays of the CURRENT device — e.g. CUDA.jl CuArray{Float64}, passed as reinterpret(Ptr{Cvoid}, pointer(x)) — and
# `stream` a cudaStream_t (C_NULL = default stream).  Scratch is a fixed __device__ array inside libjcdf: N <= 48, Q <= 256,
# occ <= 16, refused beyond; calls that overlap on one GPU race on it.  The call enqueues and returns.
function fock_build_no_alloc_B200!(F_dev::Ptr{Cvoid}, B_dev::Ptr{Cvoid}, C_occ_dev::Ptr{Cvoid}, H_dev::Ptr{Cvoid},
    N::Int, Q::Int, occ::Int; ldb::Int = Q, ldc::Int = N, stream::Ptr{Cvoid} = C_NULL)
    jcdf_check(ccall((:jcdf_fock_build_no_alloc, libjcdf), Cint,
        (Int64, Int64, Int64, Ptr{Cvoid}, Int64, Ptr{Cvoid}, Int64, Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}),
        Int64(N), Int64(Q), Int64(occ), B_dev, Int64(ldb), C_occ_dev, Int64(ldc), H_dev, F_dev, stream))
end
