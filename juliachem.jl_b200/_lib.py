"""ctypes binding of libjcdf.so (include/jcdf.h).  This is the same table of symbols the Julia wrapper
julia/B200DF.jl binds with `ccall`; nothing here computes — a missing library or symbol raises immediately
(there is no CPU fallback on the product path)."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# JCDF_LIBRARY selects another build of the same ABI (the bounds-checked libjcdf_checked.so, `make CHECKED=1`)
LIB_PATH = os.environ.get("JCDF_LIBRARY") or os.path.join(_HERE, "libjcdf.so")

N_STAGES = 10
STAGE_NAMES = ("W", "K", "V", "J", "allreduce", "density", "h2d", "d2h", "total", "epilogue")

JCDF_OK = 0
JCDF_ERR_INVALID = -1
JCDF_ERR_CUDA = -2
JCDF_ERR_NCCL = -3
JCDF_ERR_NOMEM = -4
JCDF_ERR_NUMERIC = -5


class LayoutInfo(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "N", "Q", "occ", "n_pad", "occ_pad", "tiles_total", "tiles_kept", "pairs_kept", "pairs_tri",
        "packed_row_len", "q_local", "bytes_B", "bytes_W", "kernel_launches_per_build", "graph_replays",
        "exchange_blocks_total", "exchange_blocks_kept", "exchange_block_width", "exchange_tiles_total",
        "exchange_tiles_skipped", "exchange_half_tiles_skipped")]


_p = C.c_void_p
_i32, _i64, _u64, _f64 = C.c_int32, C.c_int64, C.c_uint64, C.c_double

# name -> (restype, argtypes): every symbol include/jcdf.h declares
SIGNATURES = {
    "jcdf_last_error": (C.c_char_p, []),
    "jcdf_device_count": (_i32, []),
    "jcdf_create": (_i32, [_i64, _i64, _i64, _i32, _p, C.POINTER(_p)]),
    "jcdf_destroy": (_i32, [_p]),
    "jcdf_set_shards": (_i32, [_p, _p, _p]),
    "jcdf_set_mask": (_i32, [_p, _p, _p]),
    "jcdf_upload_three_center": (_i32, [_p, _i32, _p, _i64]),
    "jcdf_upload_three_center_device": (_i32, [_p, _i32, _p, _i64, _i64, _i64]),
    "jcdf_upload_metric": (_i32, [_p, _p]),
    "jcdf_factor_and_form_B": (_i32, [_p]),
    "jcdf_set_B": (_i32, [_p, _i32, _p, _i64]),
    "jcdf_set_B_device": (_i32, [_p, _i32, _p, _i64, _i64, _i64]),
    "jcdf_get_B": (_i32, [_p, _i32, _p, _i64]),
    "jcdf_get_B_rows": (_i32, [_p, _i32, _i64, _i64, _p, _i64]),
    "jcdf_get_metric_inverse": (_i32, [_p, _p]),
    "jcdf_set_core_hamiltonian": (_i32, [_p, _p]),
    "jcdf_fock_build": (_i32, [_p, _p, _i64, _p, _p, _p]),
    "jcdf_fock_build_device": (_i32, [_p, _p, _i64, _p, _p, _p]),
    "jcdf_eigh": (_i32, [_p, _p, _p, _p, _p]),
    "jcdf_comm_unique_id": (_i32, [_p]),
    "jcdf_comm_init_rank": (_i32, [_p, _i32, _i32, _p]),
    "jcdf_get_layout": (_i32, [_p, C.POINTER(LayoutInfo)]),
    "jcdf_fill_synthetic_B": (_i32, [_p, _u64, _f64]),
    "jcdf_synthetic_value": (_f64, [_u64, _i64, _i64, _i64, _i64, _f64]),
    "jcdf_last_kernel_ms": (_i32, [_p, _i32, _p]),
    "jcdf_exchange_cta_cycles": (_i32, [_p, _i32, _p, _i32]),
    "jcdf_debug_exchange_block": (_i32, [_i32, _i32, _i32, _i64, _p]),
    "jcdf_set_option": (_i32, [_p, C.c_char_p, _i64]),
    "jcdf_get_stream": (_i32, [_p, _i32, C.POINTER(_p)]),
    "jcdf_checked_selftest": (_i32, [_i32]),
    "jcdf_checked_set_fuzz": (_i32, [_p, C.c_uint32]),
    "jcdf_fock_build_no_alloc": (_i32, [_i64, _i64, _i64, _p, _i64, _p, _i64, _p, _p, _p]),
    "jcdf_set_basis": (_i32, [_p, _i32, _i64, _p, _p, _p, _p, _p]),
    "jcdf_compute_two_center": (_i32, [_p, _p]),
    "jcdf_compute_schwarz_mask": (_i32, [_p, _f64, _p, C.POINTER(_i64)]),
    "jcdf_compute_three_center": (_i32, [_p]),
    "jcdf_compute_B": (_i32, [_p]),
    "jcdf_compute_one_electron": (_i32, [_p, _i32, _p, _p, _p, _p, _p, C.POINTER(_f64)]),
    "jcdf_scf_init": (_i32, [_p, _p, _p, _f64, _i32, C.POINTER(_f64)]),
    "jcdf_scf_step": (_i32, [_p, _p]),
    "jcdf_scf_run": (_i32, [_p, _i32, _f64, _f64, C.POINTER(_f64), C.POINTER(_i32), C.POINTER(_i32), _p]),
    "jcdf_scf_get": (_i32, [_p, C.c_char_p, _p]),
}

_lib = None


class JCDFError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__("libjcdf status %d: %s" % (status, message))
        self.status = status


def _name_shared_nccl() -> None:
    """libjcdf binds NCCL with dlopen on its first multi-device build: an already-mapped copy if there is one, else the copy
    named in JCDF_NCCL_LIBRARY, else the loader's default.  Whatever is mapped first serves the whole process, and PyTorch
    cannot be imported over an NCCL older than the one it ships (undefined symbol at `import torch`), so when the pip package
    nvidia-nccl is installed its library is the one to share.  Only the environment variable is set here: nothing is loaded
    until a build actually needs a collective."""
    if os.environ.get("JCDF_NCCL_LIBRARY"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for loc in (spec.submodule_search_locations if spec else []):
            path = os.path.join(loc, "lib", "libnccl.so.2")
            if os.path.exists(path):
                os.environ["JCDF_NCCL_LIBRARY"] = path
                return
    except Exception:
        pass


def load_library() -> C.CDLL:
    """dlopen the in-tree libjcdf.so and bind every declared entry point.  Raises if the extension has not been
    built (run `python -c 'import __graft_entry__ as g; g.build()'` or `make -C juliachem.jl_b200/csrc`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libjcdf.so not built at %s — the B200 DF Fock build has no CPU fallback" % LIB_PATH)
    _name_shared_nccl()
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int) -> None:
    if status != 0:
        msg = load_library().jcdf_last_error()
        raise JCDFError(status, msg.decode("utf-8", "replace") if msg else "")
