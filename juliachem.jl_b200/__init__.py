"""juliachem.jl_b200 — B200-native (sm_100a) density-fitted RHF Fock build behind JuliaChem.jl's DF-RHF seam.

Holds only what the hot path needs: csrc/ (CUDA kernels + the C-ABI of include/jcdf.h -> libjcdf.so), the ctypes
binding (_lib.py), the host-side mirror of the reference's df_rhf_fock_build! seam (df.py) and the Julia wrapper a
maintainer would include (julia/B200DF.jl).  The directory name carries a dot, so import it through the root-level
shim `juliachem_jl_b200` (or importlib)."""
from ._lib import JCDFError, LayoutInfo, STAGE_NAMES, load_library, LIB_PATH  # noqa: F401
from .df import (  # noqa: F401
    B200Data, CONTRACTION_MODE_B200, IntegralSource, JCTiming, SCFData, SCFOptions, ScreeningData,
    balanced_function_shards, device_count, df_rhf_fock_build_B200, fock_build_no_alloc, get_df_static_shell_indices,
    sparse_pq_index_map, static_load_rank_indicies, unique_id,
)
