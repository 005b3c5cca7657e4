# B200DF.jl — the Julia side of the B200 drop-in for JuliaChem's density-fitted RHF Fock build.
#
# A maintainer adds this file next to GPUDF.jl (src/rhf/energy/DensityFitting/) and includes it from Energy.jl:20-29;
# INTEGRATION.md lists the three other one-line edits (Constants.jl, GPUData.jl, DensityFitting.jl).  Everything
# device-side lives in libjcdf.so (include/jcdf.h); this file only marshals pointers, in the style of the module's own
# raw BLAS call `call_gemm!` (ScreenedDF.jl:645-673).  It mirrors df_rhf_fock_build_GPU! (GPUDF.jl:11-304) stage by
# stage; juliachem.jl_b200/df.py is the same wrapper in Python (Julia is not present in the build image) and is what
# the test-suite drives.
#
# NOT EXECUTED in this repository's CI (no Julia toolchain in the image) — kept ccall-for-ccall in step with df.py.

using Libdl

const libjcdf = get(ENV, "JULIACHEM_LIBJCDF", "libjcdf")   # deps/build.jl puts libjcdf.so next to libjeri.so

const JCDF_N_STAGES = 10
# stage slot (0-based in jcdf.h) -> JCTiming constant (src/shared/JCTiming.jl:58-90)
const JCDF_STAGE_KEYS = (JCTC.W_time, JCTC.K_time, JCTC.V_time, JCTC.J_time, JCTC.fock_MPI_time, JCTC.density_time,
    JCTC.gpu_copy_coeff_time, JCTC.fock_gpu_cpu_copy_reduce_time, JCTC.fock_time, JCTC.gpu_copy_J_time)

# what scf_data.gpu_data holds for contraction_mode == "B200" (cf. SCFGPUData_cuda, GPUData_cuda.jl:4-38)
mutable struct SCFGPUDataB200 <: SCFGPUData
    handle::Ptr{Cvoid}
    number_of_devices_used::Int
    function SCFGPUDataB200()
        d = new(C_NULL, 0)
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
        navail = ccall((:jcdf_device_count, libjcdf), Cint, ())
        if num_devices > navail                                                     # GPUDF.jl:39-41
            error("Number of devices requested is greater than the number of devices available")
        end
        h = Ref{Ptr{Cvoid}}(C_NULL)
        jcdf_check(ccall((:jcdf_create, libjcdf), Cint, (Int64, Int64, Int64, Cint, Ptr{Cint}, Ref{Ptr{Cvoid}}),
            scf_data.μ, scf_data.A, scf_data.occ, num_devices, C_NULL, h))
        gd.handle = h[]
        gd.number_of_devices_used = num_devices

        # (P|Q) and the Schwarz mask stay in Julia / libint (GPUDF.jl:44-49)
        two_center_integrals = calculate_two_center_intgrals(jeri_engine_thread_df, basis_sets, scf_options)
        get_screening_metadata!(scf_data, scf_options.df_screening_sigma, jeri_engine_thread, two_center_integrals,
            basis_sets, jc_timing)
        sd = scf_data.screening_data

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

        # mask + pair compaction -> tile-compacted HBM layout (replaces setup_gpu_screening_data!, GPUDF.jl:347-420)
        GC.@preserve sd begin
            jcdf_check(ccall((:jcdf_set_mask, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Bool}, Ptr{Int64}), gd.handle,
                sd.basis_function_screen_matrix, sd.sparse_pq_index_map))
        end

        if n_ranks > 1   # NCCL communicator over all devices of all ranks; the id travels by MPI
            id = zeros(UInt8, 128)
            rank == 0 && jcdf_check(ccall((:jcdf_comm_unique_id, libjcdf), Cint, (Ptr{UInt8},), id))
            MPI.Bcast!(id, 0, comm)
            jcdf_check(ccall((:jcdf_comm_init_rank, libjcdf), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{UInt8}), gd.handle,
                n_ranks, rank, id))
        end

        # screened (P|mn) rows of each device's shard, computed on the host by libint exactly as today (GPUDF.jl:51-57)
        for device_id in 1:num_devices
            global_device_id = device_id + rank * num_devices
            T3 = calculate_three_center_integrals(jeri_engine_thread_df, basis_sets, scf_options, scf_data,
                global_device_id - 1, num_devices_global, true)                    # (Q_dev, P) column-major
            GC.@preserve T3 jcdf_check(ccall((:jcdf_upload_three_center, libjcdf), Cint,
                (Ptr{Cvoid}, Cint, Ptr{Float64}, Int64), gd.handle, device_id - 1, T3, size(T3, 1)))
        end
        # potrf + trtri + B = L^-1 T on the devices (replaces calculate_B_GPU!, GPUDF.jl:828-1008, incl. its CPU LAPACK step)
        jcdf_check(ccall((:jcdf_upload_metric, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}), gd.handle, two_center_integrals))
        n_ranks == 1 || error("B200 path: multi-rank B formation not available yet (SURVEY §8f-3); use one rank with num_devices GPUs")
        jcdf_check(ccall((:jcdf_factor_and_form_B, libjcdf), Cint, (Ptr{Cvoid},), gd.handle))
        jcdf_check(ccall((:jcdf_set_core_hamiltonian, libjcdf), Cint, (Ptr{Cvoid}, Ptr{Float64}), gd.handle, H))
    end

    # every iteration: one call.  F_out is scf_data.two_electron_fock (caller-owned; DIIS mutates it afterwards).
    stage_seconds = zeros(Float64, JCDF_N_STAGES)
    C_occ = occupied_orbital_coefficients          # N x occ column-major, as sliced at DensityFitting.jl:49
    F = scf_data.two_electron_fock
    GC.@preserve C_occ F stage_seconds begin
        jcdf_check(ccall((:jcdf_fock_build, libjcdf), Cint,
            (Ptr{Cvoid}, Ptr{Float64}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
            gd.handle, C_occ, stride(C_occ, 2), C_NULL, F, stage_seconds))
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
