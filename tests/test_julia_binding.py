"""The shipped Julia seam (juliachem.jl_b200/julia/B200DF.jl) cannot be executed here (no Julia toolchain), so it is checked
statically: every `ccall((:sym, libjcdf), ret, (argtypes...), args...)` is parsed and compared — symbol, return type,
arity, C type of every argument — with the table the Python wrapper binds (_lib.SIGNATURES, itself checked against
include/jcdf.h and the built library in test_abi_and_host.py); every JCTC.<name> must exist in the reference's
src/shared/JCTiming.jl (when the reference checkout is present) and in the list committed below; every reference
function the file calls must exist in the reference tree; and df.py must use the same timing keys and the same
stage order."""
import ctypes as C
import os
import re

import pytest

from helpers import ROOT

JL = os.path.join(ROOT, "juliachem.jl_b200", "julia", "B200DF.jl")
REF = "/root/reference"

# JCTC constants of src/shared/JCTiming.jl:50-107 this path may use (name -> key prefix), kept here so the check also runs
# on the GPU box, where the reference checkout does not exist
JCTC_KNOWN = {
    "W_time": "W_time-", "K_time": "K_time-", "V_time": "V_time-", "J_time": "J_time-", "fock_MPI_time": "fock_MPI_time-",
    "density_time": "density_time-", "GPU_non_zero_coeff_time": "GPU_-N-_non_zero_coeff_time-",
    "fock_gpu_cpu_copy_reduce_time": "fock_gpu_cpu_copy_reduce_time-", "fock_time": "fock_time-",
    "gpu_copy_J_time": "gpu_copy_J_time-", "iteration_time": "iteration_time-",
    # non_timing_data keys of calculate_exchange_block_screen_matrix (ScreenedDF.jl:409-410, 454; JCTiming.jl:42, 78-79)
    "total_exchange_blocks": "total_exchange_blocks", "unscreened_exchange_blocks": "unscreened_exchange_blocks",
    "df_exchange_n_blocks": "df_exchange_n_blocks",
}

JULIA_TO_C = {"Cint": C.c_int32, "Int64": C.c_int64, "Float64": C.c_double, "Cdouble": C.c_double, "UInt64": C.c_uint64,
              "Cstring": C.c_char_p}


def _split_top(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "({[":
            depth += 1
        elif ch in ")}]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def _ccalls(src):
    """[(symbol, return type, [argument types], number of actual arguments)]"""
    calls = []
    for m in re.finditer(r"ccall\(\(:(\w+), libjcdf\),", src):
        i = m.end()
        depth, j = 1, i
        while depth:
            depth += {"(": 1, ")": -1}.get(src[j], 0)
            j += 1
        parts = _split_top(src[i:j - 1])
        ret, types = parts[0], parts[1]
        assert types.startswith("(") and types.endswith(")"), (m.group(1), types)
        argtypes = _split_top(types[1:-1])
        calls.append((m.group(1), ret, argtypes, len(parts) - 2))
    return calls


def _matches(julia_type, ctype):
    if julia_type.startswith(("Ptr{", "Ref{")):
        return ctype is C.c_void_p or (isinstance(ctype, type) and issubclass(ctype, C._Pointer)) or ctype is C.c_char_p
    return JULIA_TO_C.get(julia_type) is ctype


def test_every_ccall_matches_the_bound_signature():
    from juliachem_jl_b200 import _lib
    src = open(JL).read()
    calls = _ccalls(src)
    assert len(calls) >= 20
    for sym, ret, argtypes, nargs in calls:
        assert sym in _lib.SIGNATURES, "B200DF.jl calls %s, which libjcdf does not export" % sym
        res, args = _lib.SIGNATURES[sym]
        assert _matches(ret, res) or (ret == "Cint" and res is C.c_int32), (sym, ret)
        assert len(argtypes) == len(args) == nargs, "%s: %d types, %d arguments, library takes %d" % (sym, len(argtypes), nargs, len(args))
        for k, (jt, ct) in enumerate(zip(argtypes, args)):
            assert _matches(jt, ct), "%s argument %d: Julia %s vs C %s" % (sym, k, jt, ct)
    used = {c[0] for c in calls}
    # the per-iteration call, both set-up routes and the multi-rank plumbing are all bound
    for sym in ("jcdf_create", "jcdf_destroy", "jcdf_set_shards", "jcdf_set_mask", "jcdf_upload_three_center",
                "jcdf_upload_metric", "jcdf_factor_and_form_B", "jcdf_set_basis", "jcdf_compute_two_center",
                "jcdf_compute_schwarz_mask", "jcdf_compute_B", "jcdf_comm_unique_id", "jcdf_comm_init_rank",
                "jcdf_set_core_hamiltonian", "jcdf_fock_build", "jcdf_eigh", "jcdf_scf_init", "jcdf_scf_step", "jcdf_scf_get"):
        assert sym in used, sym


def test_timing_keys_exist_in_the_reference_and_match_df_py():
    from juliachem_jl_b200 import _lib, df
    src = open(JL).read()
    names = set(re.findall(r"JCTC\.(\w+)", src))
    assert names and names <= set(JCTC_KNOWN), names - set(JCTC_KNOWN)
    if os.path.exists(os.path.join(REF, "src", "shared", "JCTiming.jl")):
        ref = open(os.path.join(REF, "src", "shared", "JCTiming.jl")).read()
        for n, prefix in JCTC_KNOWN.items():
            m = re.search(r"const %s\s*=\s*\"([^\"]*)\"" % n, ref)
            assert m, "JCTC.%s is not defined in src/shared/JCTiming.jl" % n
            assert m.group(1) == prefix
    # stage order of the tuple == stage order of jcdf.h == df.py's table
    tup = re.search(r"const JCDF_STAGE_KEYS = \((.*?)\)\n", src, flags=re.S).group(1)
    keys = []
    for item in _split_top(tup):
        m = re.match(r"JCTiming_GPUkey\(JCTC\.(\w+), (\d+)\)", item)
        keys.append(JCTC_KNOWN[m.group(1)].replace("-N-", m.group(2)) if m else JCTC_KNOWN[item.replace("JCTC.", "")])
    assert len(keys) == _lib.N_STAGES
    assert keys == [df._STAGE_TO_JCTC[s] + "-" for s in _lib.STAGE_NAMES]


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference checkout not present")
def test_reference_functions_and_fields_the_wrapper_uses_exist():
    src = open(JL).read()
    tree = ""
    for dp, _, files in os.walk(os.path.join(REF, "src")):
        for f in files:
            if f.endswith(".jl"):
                tree += open(os.path.join(dp, f), errors="replace").read()
    for fn in ("calculate_two_center_intgrals", "get_screening_metadata!", "calculate_three_center_integrals",
               "static_load_rank_indicies", "JCTiming_key", "JCTiming_GPUkey"):
        assert fn in src and re.search(r"function %s\(" % re.escape(fn), tree), fn
    for field in ("df_screening_sigma", "df_screen_exchange", "df_force_dense", "df_use_adaptive", "num_devices"):
        assert "scf_options.%s" % field in src and re.search(r"\b%s\s*::" % field, tree), field
    for field in ("basis_function_screen_matrix", "sparse_pq_index_map", "screened_indices_count", "two_electron_fock"):
        assert field in src and field in tree
    assert "abstract type SCFGPUData" in tree  # SCFGPUDataB200 <: SCFGPUData
    assert "n_ranks > 1" in src and "jcdf_compute_B" in src  # several ranks are served, not refused
    assert "gpu_copy_coeff_time" not in src
