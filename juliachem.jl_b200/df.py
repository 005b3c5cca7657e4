"""Host-side mirror of the reference's density-fitting seam, over the C-ABI of libjcdf.so.

Julia is not present in this image, so the thin layer the reference would keep in Julia
(src/rhf/energy/DensityFitting/DensityFitting.jl:23-76 `df_rhf_fock_build!` and the per-mode functions it dispatches
to) is mirrored here in Python with the same names, argument meaning and error behaviour; julia/B200DF.jl is the
Julia file a maintainer would `include` next to GPUDF.jl.  All arithmetic happens inside libjcdf.so
on the GPU — this module only marshals pointers, exactly like the `ccall` wrapper.

Reference types mirrored (fields used by this path only):
  SCFOptions      src/shared/SCFOptions.jl:2-22, keyword names src/shared/Constants.jl:3-78
  SCFData         src/shared/SCFData.jl:19-44
  ScreeningData   src/shared/SCFData.jl:1-17
  JCTiming        src/shared/JCTiming.jl:3-145
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import JCDFError, LayoutInfo, N_STAGES, STAGE_NAMES, check, load_library

CONTRACTION_MODE_B200 = "B200"  # new value beside "default"/"screened"/"dense"/"GPU"/"denseGPU" (Constants.jl:45-50)


# ------------------------------------------------------------------------------------------------------------------
# options / timing / state
# ------------------------------------------------------------------------------------------------------------------
@dataclass
class SCFOptions:
    """create_scf_options (SCFOptions.jl:47-139); defaults from Constants.jl."""
    density_fitting: bool = True
    contraction_mode: str = CONTRACTION_MODE_B200
    load: str = "static"
    guess: str = "hcore"
    df_exchange_n_blocks: int = 0
    df_screening_sigma: float = 1e-5      # "df_sigma"
    df_screen_exchange: bool = False      # "df_exchange_screen"
    df_force_dense: bool = False
    df_use_adaptive: bool = True
    num_devices: int = 1


class JCTiming:
    """timings dict keyed "<name>-<iteration>" (JCTiming.jl:111-145, JCTiming_key :139)."""

    def __init__(self):
        self.timings: dict[str, float] = {}
        self.options: dict[str, str] = {}
        self.non_timing_data: dict[str, str] = {}

    @staticmethod
    def key(name: str, iteration: int) -> str:
        return "%s-%d" % (name, iteration)


# libjcdf stage slot -> the reference's JCTiming constant (JCTiming.jl:58-65, 75-90)
_STAGE_TO_JCTC = {
    "W": "W_time", "K": "K_time", "V": "V_time", "J": "J_time", "allreduce": "fock_MPI_time",
    # h2d: what the fork books as JCTiming_GPUkey(JCTC.GPU_non_zero_coeff_time, 1) (GPUDF.jl:288, JCTiming.jl:88, 143-145)
    "density": "density_time", "h2d": "GPU_1_non_zero_coeff_time", "d2h": "fock_gpu_cpu_copy_reduce_time",
    "total": "fock_time", "epilogue": "gpu_copy_J_time",
}


def sparse_pq_index_map(fn_mask: np.ndarray) -> np.ndarray:
    """Pair compaction of schwarz_screen_itegrals_df (SchwarzScreening.jl:72-81): running 1-based index over kept
    pairs, outer loop p (column), inner loop q (row); 0 = screened.  Int64 (N, N) column-major."""
    m = np.asfortranarray(fn_mask).astype(bool)
    flat = m.ravel(order="F")
    out = np.zeros(flat.shape[0], dtype=np.int64)
    out[flat] = np.arange(1, int(flat.sum()) + 1, dtype=np.int64)
    return np.asfortranarray(out.reshape(m.shape, order="F"))


class ScreeningData:
    """ScreeningData (SCFData.jl:1-17) — the members the B200 path consumes."""

    def __init__(self, basis_function_screen_matrix: np.ndarray | None, N: int):
        if basis_function_screen_matrix is None:
            # setup_unscreened_screening_matricies (SchwarzScreening.jl:97-111)
            self.basis_function_screen_matrix = None
            self.sparse_pq_index_map = None
            self.screened_indices_count = N * N
        else:
            m = np.asfortranarray(np.asarray(basis_function_screen_matrix).astype(bool))
            if m.shape != (N, N):
                raise ValueError("basis_function_screen_matrix must be N x N")
            self.basis_function_screen_matrix = m
            self.sparse_pq_index_map = sparse_pq_index_map(m)
            self.screened_indices_count = int(m.sum())


def get_df_static_shell_indices(n_aux_shells: int, n_ranks: int, rank: int) -> range:
    """DynamicLoad.jl:160-171 (0-based half-open): n_shells // n_ranks contiguous shells, remainder to the last."""
    n = n_aux_shells // n_ranks
    begin = n * rank
    end = n_aux_shells if rank == n_ranks - 1 else begin + n
    return range(begin, end)


def static_load_rank_indicies(rank: int, n_ranks: int, aux_shell_nbas) -> tuple[range, range]:
    """DynamicLoad.jl:174-203: (shell range, auxiliary function range) of `rank`, 0-based half-open."""
    nb = np.asarray(aux_shell_nbas, dtype=np.int64)
    pos = np.concatenate([[0], np.cumsum(nb)])
    sh = get_df_static_shell_indices(len(nb), n_ranks, rank)
    if len(sh) == 0:
        return sh, range(0, 0)
    return sh, range(int(pos[sh.start]), int(pos[sh.stop]))


def balanced_function_shards(Q: int, n: int) -> list[range]:
    """Even split by function count (libjcdf's default when jcdf_set_shards is never called)."""
    return [range(Q * d // n, Q * (d + 1) // n) for d in range(n)]


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64_colmajor(a, shape=None) -> np.ndarray:
    a = np.asarray(a, dtype=np.float64)
    if not a.flags.f_contiguous:
        a = np.asfortranarray(a)
    if shape is not None and tuple(a.shape) != tuple(shape):
        raise ValueError("expected shape %s, got %s" % (shape, a.shape))
    return a


class B200Data:
    """What SCFGPUData holds for the fork's CUDA.jl path (src/shared/GPUData_cuda.jl:4-38), here an opaque libjcdf
    handle: all device memory lives behind it and is released with the object (lifetime = scf_data)."""

    def __init__(self, N: int, Q: int, occ: int, num_devices: int = 1, device_ids=None):
        self._lib = load_library()
        self._h = C.c_void_p()
        ids = None
        if device_ids is not None:
            ids = np.asarray(device_ids, dtype=np.int32)
            if len(ids) != num_devices:
                raise ValueError("device_ids must have num_devices entries")
        check(self._lib.jcdf_create(N, Q, occ, num_devices, _ptr(ids), C.byref(self._h)))
        self.N, self.Q, self.occ, self.num_devices = N, Q, occ, num_devices

    # -- setup -----------------------------------------------------------------------------------------------------
    def set_shards(self, ranges) -> None:
        qb = np.asarray([r.start for r in ranges], dtype=np.int64)
        ql = np.asarray([len(r) for r in ranges], dtype=np.int64)
        if len(qb) != self.num_devices:
            raise ValueError("one range per device")
        check(self._lib.jcdf_set_shards(self._h, _ptr(qb), _ptr(ql)))

    def set_mask(self, sd: ScreeningData) -> None:
        if sd.basis_function_screen_matrix is None:
            check(self._lib.jcdf_set_mask(self._h, None, None))
            return
        m = np.asfortranarray(sd.basis_function_screen_matrix.astype(np.uint8))
        imap = np.asfortranarray(sd.sparse_pq_index_map.astype(np.int64))
        check(self._lib.jcdf_set_mask(self._h, _ptr(m), _ptr(imap)))

    def set_option(self, name: str, value: int) -> None:
        check(self._lib.jcdf_set_option(self._h, name.encode(), int(value)))

    def upload_three_center(self, dev: int, T: np.ndarray) -> None:
        T = _f64_colmajor(T)
        check(self._lib.jcdf_upload_three_center(self._h, dev, _ptr(T), T.shape[0]))

    def upload_metric(self, J_PQ: np.ndarray) -> None:
        J = _f64_colmajor(J_PQ, (self.Q, self.Q))
        check(self._lib.jcdf_upload_metric(self._h, _ptr(J)))

    def factor_and_form_B(self) -> None:
        check(self._lib.jcdf_factor_and_form_B(self._h))

    def set_B(self, dev: int, B: np.ndarray) -> None:
        B = _f64_colmajor(B)
        check(self._lib.jcdf_set_B(self._h, dev, _ptr(B), B.shape[0]))

    def get_B(self, dev: int, q_len: int, P: int) -> np.ndarray:
        """Packed shard of device `dev` back in the reference (q_len x P) layout (jcdf_get_B)."""
        out = np.zeros((q_len, P), dtype=np.float64, order="F")
        check(self._lib.jcdf_get_B(self._h, dev, _ptr(out), max(q_len, 1)))
        return out

    def get_B_rows(self, dev: int, row_begin: int, row_count: int, P: int) -> np.ndarray:
        """Rows [row_begin, row_begin + row_count) of device `dev`'s shard in the reference layout (jcdf_get_B_rows)."""
        out = np.zeros((row_count, P), dtype=np.float64, order="F")
        check(self._lib.jcdf_get_B_rows(self._h, dev, row_begin, row_count, _ptr(out), max(row_count, 1)))
        return out

    def get_metric_inverse(self) -> np.ndarray:
        out = np.zeros((self.Q, self.Q), dtype=np.float64, order="F")
        check(self._lib.jcdf_get_metric_inverse(self._h, _ptr(out)))
        return out

    def fill_synthetic_B(self, seed: int, decay_length: float) -> None:
        check(self._lib.jcdf_fill_synthetic_B(self._h, seed, decay_length))

    def comm_init_rank(self, n_ranks: int, rank: int, unique_id: bytes) -> None:
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        check(self._lib.jcdf_comm_init_rank(self._h, n_ranks, rank, buf))

    # -- per iteration ---------------------------------------------------------------------------------------------
    def fock_build(self, C_occ: np.ndarray, H: np.ndarray | None, F_out: np.ndarray | None = None):
        """jcdf_fock_build with host arrays; returns (F, stage_seconds dict)."""
        N = self.N
        Cc = np.asarray(C_occ, dtype=np.float64)
        if Cc.ndim != 2 or Cc.shape[0] != N or Cc.shape[1] < self.occ:
            raise ValueError("C_occ must be N x occ")
        if Cc.strides[0] == 8 and Cc.strides[1] >= 8 * N and Cc.strides[1] % 8 == 0:
            ldc = Cc.strides[1] // 8  # a column-major view into a larger matrix: pass its leading dimension
        else:
            Cc = np.asfortranarray(Cc)
            ldc = N
        Hc = None if H is None else _f64_colmajor(H, (N, N))
        if F_out is None:
            F_out = np.empty((N, N), dtype=np.float64, order="F")
        elif not (F_out.flags.f_contiguous and F_out.dtype == np.float64 and F_out.shape == (N, N)):
            raise ValueError("F_out must be an N x N column-major float64 array")
        st = np.zeros(N_STAGES, dtype=np.float64)
        check(self._lib.jcdf_fock_build(self._h, _ptr(Cc), ldc, _ptr(Hc), _ptr(F_out), _ptr(st)))
        return F_out, dict(zip(STAGE_NAMES, st.tolist()))

    def fock_build_device(self, C_occ_ptr: int, ldc: int, H_ptr: int | None, F_ptr: int):
        """jcdf_fock_build_device: raw device addresses (e.g. torch.Tensor.data_ptr()) on this handle's first device."""
        st = np.zeros(N_STAGES, dtype=np.float64)
        check(self._lib.jcdf_fock_build_device(self._h, C.c_void_p(C_occ_ptr), ldc,
                                               C.c_void_p(H_ptr) if H_ptr else None, C.c_void_p(F_ptr), _ptr(st)))
        return dict(zip(STAGE_NAMES, st.tolist()))

    def eigh(self, F: np.ndarray, X: np.ndarray):
        """On-device X'FX, eigen, C = X U (iteration(), SCF.jl:1072-1099).  Returns (evals, C)."""
        N = self.N
        Fc, Xc = _f64_colmajor(F, (N, N)), _f64_colmajor(X, (N, N))
        ev = np.zeros(N, dtype=np.float64)
        Cm = np.zeros((N, N), dtype=np.float64, order="F")
        check(self._lib.jcdf_eigh(self._h, _ptr(Fc), _ptr(Xc), _ptr(ev), _ptr(Cm)))
        return ev, Cm

    # -- integral front-end on the GPU --------------------------------------------------------------
    def set_basis(self, which: int, am, nprim, centers, exponents, coefficients) -> None:
        """which = 0 primary / 1 auxiliary; centers (nshell, 3) in bohr; exponents/coefficients concatenated, raw
        basis-set-exchange values (the library applies libint's renorm + JuliaChem's axial factors)."""
        am = np.ascontiguousarray(am, dtype=np.int32)
        npr = np.ascontiguousarray(nprim, dtype=np.int32)
        cen = np.ascontiguousarray(np.asarray(centers, dtype=np.float64).reshape(len(am), 3))
        ex = np.ascontiguousarray(exponents, dtype=np.float64)
        co = np.ascontiguousarray(coefficients, dtype=np.float64)
        if len(ex) != int(npr.sum()) or len(co) != len(ex):
            raise ValueError("exponents/coefficients must hold sum(nprim) entries")
        check(self._lib.jcdf_set_basis(self._h, which, len(am), _ptr(am), _ptr(npr), _ptr(cen), _ptr(ex), _ptr(co)))

    def compute_two_center(self, want_host_copy: bool = True):
        """calculate_two_center_intgrals (TwoCenterIntegrals.jl:7) on the device; the metric stays there."""
        out = np.zeros((self.Q, self.Q), dtype=np.float64, order="F") if want_host_copy else None
        check(self._lib.jcdf_compute_two_center(self._h, _ptr(out)))
        return out

    def compute_schwarz_mask(self, sigma: float):
        """schwarz_screen_itegrals_df (SchwarzScreening.jl:9) on the device + jcdf_set_mask.  Returns (fn_mask, P)."""
        m = np.zeros((self.N, self.N), dtype=np.uint8, order="F")
        kept = C.c_int64()
        check(self._lib.jcdf_compute_schwarz_mask(self._h, float(sigma), _ptr(m), C.byref(kept)))
        return m.astype(bool), int(kept.value)

    def compute_three_center(self) -> None:
        """calculate_three_center_integrals (ThreeCenterIntegrals.jl:9) straight into the packed Q shards."""
        check(self._lib.jcdf_compute_three_center(self._h))

    def compute_B(self) -> None:
        """B = L^-1 (P|mn) per device by recomputing integrals instead of exchanging them (jcdf_compute_B)."""
        check(self._lib.jcdf_compute_B(self._h))

    def compute_one_electron(self, charges, centers):
        """S, T, V and E_nuc on the device (rhf_kernel's one-electron set-up, SCF.jl:100-135).  centers (natoms, 3) bohr."""
        Z = np.ascontiguousarray(charges, dtype=np.float64)
        cen = np.ascontiguousarray(np.asarray(centers, dtype=np.float64).reshape(len(Z), 3))
        N = self.N
        S, T, V = (np.zeros((N, N), dtype=np.float64, order="F") for _ in range(3))
        e = C.c_double()
        check(self._lib.jcdf_compute_one_electron(self._h, len(Z), _ptr(Z), _ptr(cen), _ptr(S), _ptr(T), _ptr(V), C.byref(e)))
        return S, T, V, e.value

    # -- on-device SCF step -------------------------------------------------------------------------
    def scf_init(self, S: np.ndarray, H: np.ndarray, E_nuc: float, ndiis: int = 10) -> float:
        """rhf_kernel's set-up (SCF.jl:141-183) on the device: X = S^-1/2, hcore guess.  Returns the guess energy."""
        N = self.N
        Sc, Hc = _f64_colmajor(S, (N, N)), _f64_colmajor(H, (N, N))
        e = C.c_double()
        check(self._lib.jcdf_scf_init(self._h, _ptr(Sc), _ptr(Hc), float(E_nuc), int(ndiis), C.byref(e)))
        return e.value

    def scf_step(self) -> dict:
        """One body of scf_cycles_kernel's loop (SCF.jl:399-551) with F, C, D resident on the GPU."""
        r = np.zeros(8, dtype=np.float64)
        check(self._lib.jcdf_scf_step(self._h, _ptr(r)))
        return dict(zip(("E", "dE", "D_rms", "E_elec", "t_fock", "t_diis", "t_iteration", "t_total"), r.tolist()))

    def scf_run(self, max_iterations: int = 50, dele: float = 1e-6, rmsd: float = 1e-6) -> dict:
        hist = np.zeros(3 * max_iterations, dtype=np.float64)
        e, it, cv = C.c_double(), C.c_int32(), C.c_int32()
        check(self._lib.jcdf_scf_run(self._h, max_iterations, dele, rmsd, C.byref(e), C.byref(it), C.byref(cv), _ptr(hist)))
        rows = hist.reshape(-1, 3)[:it.value]
        return {"E": e.value, "converged": bool(cv.value),
                "iterations": [(i + 1, float(r[0]), float(r[1]), float(r[2])) for i, r in enumerate(rows)]}

    def scf_get(self, name: str) -> np.ndarray:
        N = self.N
        out = np.zeros(N if name == "evals" else (N, N), dtype=np.float64, order="F")
        check(self._lib.jcdf_scf_get(self._h, name.encode(), _ptr(out)))
        return out

    # -- introspection ---------------------------------------------------------------------------------------------
    def layout(self) -> LayoutInfo:
        li = LayoutInfo()
        check(self._lib.jcdf_get_layout(self._h, C.byref(li)))
        return li

    def last_kernel_ms(self, dev: int = 0) -> dict:
        out = np.zeros(N_STAGES, dtype=np.float64)
        check(self._lib.jcdf_last_kernel_ms(self._h, dev, _ptr(out)))
        return dict(zip(STAGE_NAMES, out.tolist()))


    def exchange_cta_cycles(self, dev: int = 0) -> np.ndarray:
        """Busy cycles of every persistent CTA of the last exchange launch (profiling aid)."""
        out = np.zeros(1024, dtype=np.int64)
        n = self._lib.jcdf_exchange_cta_cycles(self._h, dev, _ptr(out), 1024)
        if n < 0:
            check(n)
        return out[:n]

    def checked_set_fuzz(self, seed: int) -> None:
        """Checked build only: seed of the schedule fuzzing (0 = off); results must not depend on it."""
        check(self._lib.jcdf_checked_set_fuzz(self._h, int(seed)))

    def stream(self, dev: int = 0) -> int:
        """cudaStream_t (as an integer) the builds of device `dev` run on — wrap with torch.cuda.ExternalStream."""
        out = C.c_void_p()
        check(self._lib.jcdf_get_stream(self._h, dev, C.byref(out)))
        return int(out.value or 0)

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.jcdf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def unique_id() -> bytes:
    buf = (C.c_uint8 * 128)()
    check(load_library().jcdf_comm_unique_id(buf))
    return bytes(buf)


def fock_build_no_alloc(N: int, Q: int, occ: int, B_ptr: int, ldb: int, C_occ_ptr: int, ldc: int, H_ptr: int | None,
                        F_ptr: int, stream: int = 0) -> None:
    """jcdf_fock_build_no_alloc (no counterpart in the reference; include/jcdf.h): dense F = H + 2J - K on raw device
    addresses of the current device with the scratch in the library's fixed __device__ array — N <= 48, Q <= 256,
    occ <= 16; enqueued on `stream`, not synchronised; no allocation on the call path."""
    check(load_library().jcdf_fock_build_no_alloc(N, Q, occ, C.c_void_p(B_ptr), ldb, C.c_void_p(C_occ_ptr), ldc,
                                                   C.c_void_p(H_ptr) if H_ptr else None, C.c_void_p(F_ptr),
                                                   C.c_void_p(stream) if stream else None))


def device_count() -> int:
    n = load_library().jcdf_device_count()
    if n < 0:
        check(n)
    return n


@dataclass
class SCFData:
    """SCFData (SCFData.jl:19-44): μ = N, A = Q, occ, the caller-owned two_electron_fock, and gpu_data."""
    mu: int = 0
    A: int = 0
    occ: int = 0
    two_electron_fock: np.ndarray | None = None
    screening_data: ScreeningData | None = None
    gpu_data: B200Data | None = None
    extra: dict = field(default_factory=dict)


class IntegralSource:
    """Stand-in for the JERI engines + basis_sets arguments of df_rhf_fock_build! (SCF.jl:387-388, 463): whatever
    produces the one-off inputs of iteration 1.  Methods mirror the reference's integral drivers."""

    N: int
    Q: int
    n_electrons: int

    def aux_shell_nbas(self):
        """functions per auxiliary shell (for static_load_rank_indicies); None = shard by function count"""
        return None

    def two_center_integrals(self) -> np.ndarray:  # calculate_two_center_intgrals (TwoCenterIntegrals.jl:7)
        raise NotImplementedError

    def screening_matrix(self, sigma: float):  # schwarz_screen_itegrals_df (SchwarzScreening.jl:9)
        return None

    def three_center_integrals(self, q_range: range, sd: ScreeningData) -> np.ndarray:
        """(len(q_range), P) column-major in the reference pair order (calculate_three_center_integrals,
        ThreeCenterIntegrals.jl:9 / ...Screened.jl:8)"""
        raise NotImplementedError

    def B_rows(self, q_range: range, sd: ScreeningData):
        """Optional shortcut: rows of B = L^-1 (P|mn) when the caller already holds them; None = not available."""
        return None

    def device_basis(self):
        """Optional: (primary, auxiliary) flat shell arrays, each the argument tuple of B200Data.set_basis — what
        basis_sets.primary / basis_sets.auxillary hold (BasisStructs.jl:170-185).  When given, (P|Q), the Schwarz mask and
        B are computed on the devices (julia/B200DF.jl: setup_device_integrals_B200!), for any number of ranks."""
        return None


def uses_dense_contraction(scf_options: SCFOptions, N: int, n_ranks: int) -> bool:
    """run_gpu_fock_build!'s choice between the dense and the screened statement of the contraction
    (DensityFitting.jl:79-90): df_force_dense, or df_use_adaptive on a small system (N < 800) with a single rank."""
    return bool(scf_options.df_force_dense or (scf_options.df_use_adaptive and N < 800 and n_ranks == 1))


DF_EXCHANGE_N_BLOCKS_CPU_DEFAULT = 10  # Constants.jl:13


def check_supported_options(scf_options: SCFOptions) -> None:
    """Options of the reference the B200 path cannot honour are refused, not ignored (SCFOptions.jl:86-93).
    df_exchange_n_blocks names the reference's K block grid (ScreenedDF.jl:385-417): without df_exchange_screen it
    changes no result, so any value is accepted; with it the grid decides which K blocks exist (exchange_block_screen)."""
    if scf_options.df_exchange_n_blocks < 0:
        raise ValueError("df_exchange_n_blocks must be >= 0")


def exchange_block_screen_blocks(scf_options: SCFOptions) -> int:
    """Value of libjcdf's option "exchange_block_screen" for these options: 0 unless df_exchange_screen is set, else the
    block count calculate_exchange_block_screen_matrix starts from (ScreenedDF.jl:388-390: 0 -> the CPU default of 10).
    The N < 100 -> one block rule (:392-394) is applied by the library, which knows N."""
    if not scf_options.df_screen_exchange:
        return 0
    return scf_options.df_exchange_n_blocks or DF_EXCHANGE_N_BLOCKS_CPU_DEFAULT


def df_rhf_fock_build_B200(scf_data: SCFData, integrals: IntegralSource, occupied_orbital_coefficients: np.ndarray,
                           iteration: int, scf_options: SCFOptions, H: np.ndarray, jc_timing: JCTiming,
                           rank: int = 0, n_ranks: int = 1, comm_id: bytes | None = None) -> np.ndarray:
    """The branch `contraction_mode == "B200"` of df_rhf_fock_build! (DensityFitting.jl:51-60) — same stage order as
    df_rhf_fock_build_GPU! (GPUDF.jl:11-304) with every device-side step behind libjcdf:

      iteration 1, device integrals (IntegralSource.device_basis given; any number of ranks):
                    basis sets -> jcdf_set_basis, jcdf_compute_two_center, jcdf_compute_schwarz_mask, jcdf_compute_B
      iteration 1, host integrals (one rank):
                    (P|Q)            -> jcdf_upload_metric            (GPUDF.jl:44, TwoCenterIntegrals.jl:7)
                    Schwarz mask     -> jcdf_set_mask                 (GPUDF.jl:45, ScreenedDF.jl:16)
                    shard ranges     -> jcdf_set_shards               (GPUDF.jl:50, :1026-1056)
                    (P|mn) per shard -> jcdf_upload_three_center      (GPUDF.jl:51-57)
                    B = L^-1 T       -> jcdf_factor_and_form_B        (GPUDF.jl:76, :828-1008; on-device potrf/trtri)
      every iter:   F = H + 2J - K   -> jcdf_fock_build               (GPUDF.jl:186-277 + DensityFitting.jl:62-71)

    `occupied_orbital_coefficients` is C[:, 1:occ] (DensityFitting.jl:49).  Returns scf_data.two_electron_fock, which
    the caller may mutate afterwards (DIIS, SCF.jl:472-508)."""
    if iteration == 1:
        check_supported_options(scf_options)
        N, Q = integrals.N, integrals.Q
        scf_data.mu, scf_data.A = N, Q
        scf_data.occ = integrals.n_electrons // 2  # DensityFitting.jl:37
        scf_data.two_electron_fock = np.zeros((N, N), dtype=np.float64, order="F")  # :45
        ndev = scf_options.num_devices
        if ndev > device_count():  # GPUDF.jl:39-41
            raise RuntimeError("Number of devices requested is greater than the number of devices available")
        if n_ranks > 1 and comm_id is None:
            raise ValueError("comm_id required when n_ranks > 1")
        dense = uses_dense_contraction(scf_options, N, n_ranks)
        sigma = 0.0 if dense else scf_options.df_screening_sigma
        gd = B200Data(N, Q, scf_data.occ, ndev)
        scf_data.gpu_data = gd
        # shard ranges: ranks x devices contiguous pieces (GPUDF.jl:1026-1056 splits a rank's range over its devices)
        nbas = integrals.aux_shell_nbas()
        total = n_ranks * ndev
        if nbas is not None:
            ranges = [static_load_rank_indicies(rank * ndev + d, total, nbas)[1] for d in range(ndev)]
        else:
            allr = balanced_function_shards(Q, total)
            ranges = allr[rank * ndev:(rank + 1) * ndev]
        gd.set_shards(ranges)
        # df_exchange_screen = true: K only for the blocks whose mask sub-block holds a pair; dropped blocks keep what
        # scf_data.two_electron_fock held (ScreenedDF.jl:459-545) — F_out is read before every build.  Before the mask;
        # the dense statement has no K blocks (DensityFitting.jl:185-224) and the library ignores the option there.
        gd.set_option("exchange_block_screen", exchange_block_screen_blocks(scf_options))
        basis = integrals.device_basis()
        if basis is not None:
            # integral front-end of the library: nothing but the basis sets crosses the boundary, any number of ranks
            gd.set_basis(0, *basis[0])
            gd.set_basis(1, *basis[1])
            gd.compute_two_center(False)
            mask, _ = gd.compute_schwarz_mask(sigma)  # ends with jcdf_set_mask
            scf_data.screening_data = ScreeningData(None if sigma <= 0 else mask, N)
            if n_ranks > 1:
                gd.comm_init_rank(n_ranks, rank, comm_id)
            gd.compute_B()
        else:
            mask = None if dense else integrals.screening_matrix(sigma)
            sd = ScreeningData(mask, N)
            scf_data.screening_data = sd
            gd.set_mask(sd)
            if n_ranks > 1:
                gd.comm_init_rank(n_ranks, rank, comm_id)
            have_B = True
            for d, r in enumerate(ranges):
                Brows = integrals.B_rows(r, sd)
                if Brows is None:
                    have_B = False
                    break
                gd.set_B(d, Brows)
            if not have_B:
                if n_ranks > 1:
                    raise RuntimeError("host-integral route with several ranks needs B_rows from the caller; give "
                                       "device_basis() to let every rank compute its own rows (jcdf_compute_B)")
                gd.upload_metric(integrals.two_center_integrals())
                for d, r in enumerate(ranges):
                    gd.upload_three_center(d, integrals.three_center_integrals(r, sd))
                gd.factor_and_form_B()
        if scf_options.df_screen_exchange:
            # what calculate_exchange_block_screen_matrix books (ScreenedDF.jl:409-410, 454)
            li = gd.layout()
            if li.exchange_blocks_total:  # 0: dense statement of the contraction, which has no K blocks
                n_blocks = 1 if N < 100 else exchange_block_screen_blocks(scf_options)
                jc_timing.non_timing_data["total_exchange_blocks"] = str(li.exchange_blocks_total)
                jc_timing.non_timing_data["unscreened_exchange_blocks"] = str(li.exchange_blocks_kept)
                jc_timing.non_timing_data["df_exchange_n_blocks"] = str(n_blocks)
    gd = scf_data.gpu_data
    if gd is None:
        raise RuntimeError("df_rhf_fock_build_B200 called with iteration > 1 before iteration 1")
    F, stages = gd.fock_build(occupied_orbital_coefficients, H, scf_data.two_electron_fock)
    for name, sec in stages.items():
        jc_timing.timings[JCTiming.key(_STAGE_TO_JCTC[name], iteration)] = sec
    return F
