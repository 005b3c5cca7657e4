"""CPU-side checks: libjcdf.so loads and exports every symbol include/jcdf.h declares (no compute calls — there is no
GPU here), the host-side mirror of the reference interface (df.py) agrees with the oracle's restatement of the same
reference functions, and the product path has no CPU fallback."""
import ctypes
import os
import re
import sys

import numpy as np
import pytest

from helpers import ROOT, jc, orc, wl


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "jcdf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(jcdf_[A-Za-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = jc.load_library()
    from juliachem_jl_b200 import _lib
    names = _declared_symbols()
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), "libjcdf.so does not export %s" % n
        assert n in _lib.SIGNATURES, "ctypes table misses %s" % n
    assert sorted(_lib.SIGNATURES) == names


def test_status_enum_and_stage_slots_match_header():
    src = open(os.path.join(ROOT, "include", "jcdf.h")).read()
    from juliachem_jl_b200 import _lib
    for name, val in re.findall(r"(JCDF_ERR_[A-Z]+) = (-\d+)", src):
        assert getattr(_lib, name) == int(val)
    assert int(re.search(r"JCDF_N_STAGES = (\d+)", src).group(1)) == _lib.N_STAGES == len(_lib.STAGE_NAMES)


def test_no_gpu_means_error_not_fallback():
    """Without a CUDA device every entry point that needs one fails with a status + message; nothing computes on the CPU."""
    lib = jc.load_library()
    if lib.jcdf_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(jc.JCDFError):
        jc.B200Data(25, 96, 5, 1)
    assert lib.jcdf_last_error()


def test_product_package_never_imports_the_oracle():
    """No file of the product package may import, load or execute anything under oracle/: Python sources are parsed (import
    statements, importlib / __import__ / exec-style calls naming it, string literals pointing into oracle/), native and Julia
    sources are scanned for the path.  bench.py may touch it only inside its CPU legs."""
    import ast
    pkg = os.path.join(ROOT, "juliachem.jl_b200")

    def python_refs(path):
        tree = ast.parse(open(path).read(), path)
        bad = []
        for node in ast.walk(tree):
            if isinstance(node, ast.Import):
                bad += [a.name for a in node.names if a.name.split(".")[0] == "oracle"]
            elif isinstance(node, ast.ImportFrom):
                if (node.module or "").split(".")[0] == "oracle" and node.level == 0:
                    bad.append(node.module)
            elif isinstance(node, ast.Constant) and isinstance(node.value, str) and not isinstance(getattr(node, "_parent", None), ast.Expr):
                v = node.value
                if re.search(r"(^|[/\\\s'\"])oracle([/\\.]|$)", v) and len(v) < 200 and "\n" not in v:
                    bad.append("string %r" % v)
        return bad

    docstrings = set()
    for dp, _, files in os.walk(pkg):
        for f in files:
            path = os.path.join(dp, f)
            if f.endswith(".py"):
                tree = ast.parse(open(path).read())
                for node in ast.walk(tree):  # docstrings may mention the oracle; code and data strings may not
                    if isinstance(node, (ast.Module, ast.FunctionDef, ast.ClassDef, ast.AsyncFunctionDef)):
                        ds = ast.get_docstring(node, clean=False)
                        if ds:
                            docstrings.add(ds)
                refs = [r for r in python_refs(path) if not (r.startswith("string ") and any(r[8:-1] in d for d in docstrings))]
                assert not refs, (path, refs)
            elif f.endswith((".cu", ".cuh", ".cpp", ".h", ".jl")):
                code = re.sub(r"//.*|#.*", "", open(path).read())
                assert not re.search(r"oracle[/.]", code), path
    # bench.py: the only imports of the oracle sit inside the CPU-leg function
    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef):
            for sub in ast.walk(node):
                if isinstance(sub, (ast.Import, ast.ImportFrom)):
                    mod = sub.module if isinstance(sub, ast.ImportFrom) else sub.names[0].name
                    if (mod or "").split(".")[0] == "oracle":
                        assert node.name == "cpu_reference_builds_per_sec", node.name
    for node in tree.body:
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            mod = node.module if isinstance(node, ast.ImportFrom) else node.names[0].name
            assert (mod or "").split(".")[0] != "oracle"


def test_synthetic_value_host_twin_is_bit_identical():
    """jcdf_synthetic_value is a pure host function of libjcdf; workloads.synthetic_B_rows is its numpy twin."""
    lib = jc.load_library()
    w = wl.WORKLOADS["tiny"]
    rows = wl.synthetic_B_rows(w, range(3, 6), None)
    for (qa, p, q) in [(3, 0, 0), (4, 10, 3), (5, 69, 69), (3, 7, 60)]:
        v = lib.jcdf_synthetic_value(w.seed, w.Q, qa, p, q, w.decay_length)
        assert v == rows[qa - 3, p, q] == rows[qa - 3, q, p]


@pytest.mark.parametrize("n_shells,n_ranks", [(30, 1), (30, 4), (7, 8), (96, 3), (5, 5)])
def test_static_shard_rule_matches_oracle(n_shells, n_ranks):
    nbas = np.random.default_rng(n_shells).choice([1, 3, 6, 10, 15], size=n_shells)
    cover = []
    for r in range(n_ranks):
        assert jc.get_df_static_shell_indices(n_shells, n_ranks, r) == orc.get_df_static_shell_indices(n_shells, n_ranks, r)
        a, b = jc.static_load_rank_indicies(r, n_ranks, nbas), orc.static_load_rank_indicies(r, n_ranks, nbas)
        assert a == b
        cover += list(a[1])
    assert cover == list(range(int(nbas.sum())))  # contiguous, disjoint, complete


def test_pair_compaction_matches_oracle_and_reference_order():
    m = wl.screening_mask(wl.WORKLOADS["tiny"])
    a, b = jc.sparse_pq_index_map(m), orc.sparse_pq_index_map(m)
    assert np.array_equal(a, b) and a.dtype == np.int64
    # outer loop p (column), inner loop q (row), 1-based running index (SchwarzScreening.jl:72-81)
    k = 0
    for p in range(m.shape[0]):
        for q in range(m.shape[0]):
            if m[q, p]:
                k += 1
                assert a[q, p] == k
            else:
                assert a[q, p] == 0
    sd = jc.ScreeningData(m, m.shape[0])
    assert sd.screened_indices_count == k


def test_workload_shapes_and_algorithmic_work():
    for name, (N, Q, occ) in {"water": (25, 96, 5), "benzene_dimer": (600, 1632, 42), "w50": (1250, 4800, 250),
                              "c20h42": (1330, 3490, 81), "gly18": (3395, 9333, 275)}.items():
        w = wl.WORKLOADS[name]
        assert (w.N, w.Q, w.occ) == (N, Q, occ)
    work = wl.algorithmic_work(1250, 4800, 250, 1250 * 1250, 1250 * 1251 // 2)
    assert abs(work["w_flops"] + work["k_flops"] - 5.63e12) < 0.01e12   # SURVEY.md §8(d) worked value
    assert abs(work["j_bytes"] - 30.0e9) < 0.1e9
    for name in ("tiny", "small", "c20h42"):
        m = wl.screening_mask(wl.WORKLOADS[name])
        assert m.dtype == bool and np.array_equal(m, m.T) and m.diagonal().all()
    assert 0.2 < wl.screening_mask(wl.WORKLOADS["c20h42"]).mean() < 0.7


def test_balanced_shards_cover_Q():
    for Q, n in [(4800, 8), (131, 3), (7, 8)]:
        r = jc.balanced_function_shards(Q, n)
        assert r[0].start == 0 and r[-1].stop == Q and all(r[i].stop == r[i + 1].start for i in range(n - 1))


def test_run_basis_mirrors_jcbasis():
    """JCBasis.run (JCBasis.jl:47-170): Angstrom -> bohr, centre of mass at the origin, shells in atom order,
    n_electrons = sum Z - charge; shard ranges by auxiliary shells (DynamicLoad.jl:160-203)."""
    import json
    import os
    from juliachem_jl_b200 import rhf
    geo = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "w50_geometry.json")))
    mol = rhf.run_basis({"geometry": geo["geometry"], "symbols": geo["symbols"], "molecular_charge": 0}, {"basis": "cc-pVDZ"})
    assert mol["primary"].n_functions == 1250 and mol["auxiliary"].n_functions == 4800 and mol["n_electrons"] == 500
    m = np.array([rhf.ATOMIC_MASS[s] for s in mol["symbols"]])
    assert np.abs((m[:, None] * mol["centers"]).sum(axis=0)).max() < 1e-9
    d01 = np.linalg.norm(mol["centers"][0] - mol["centers"][1]) / rhf.ANGSTROM_TO_BOHR
    raw = np.array(geo["geometry"]).reshape(-1, 3)
    assert abs(d01 - np.linalg.norm(raw[0] - raw[1])) < 1e-12
    assert len(mol["primary"].am) == 50 * 12 and len(mol["auxiliary"].am) == 50 * 30
    with pytest.raises(KeyError):
        rhf.load_basis_library("cc-pVTZ")


@pytest.mark.parametrize("N", [129, 200, 256, 600, 1250, 1279, 1280, 3395])
def test_exchange_work_split_covers_the_lower_triangle_exactly_once(N):
    """Host view of the exchange kernel's warp blocks (jcdf_debug_exchange_block, the function the kernel and the work-list
    cost model share): over every lower-triangular 128 x 128 tile the eight warp rectangles cover every 8 x 8 DMMA tile
    that holds an element (row >= col, row < N, col < N) exactly once, never reach past N rounded up to 8, and on whole
    diagonal tiles the two warps of each SM sub-partition (w, w + 4) together own 34 of the 64 DMMA tiles a full tile
    costs per sub-partition (exactly the 136 tiles on or below the diagonal in total)."""
    import ctypes as C
    lib = jc.load_library()
    n128 = (N + 127) // 128
    out = (C.c_int32 * 5)()
    units = (N + 7) // 8
    for ta in range(n128):
        for tb in range(ta + 1):
            cover = np.zeros((16, 16), dtype=np.int32)
            per_warp = []
            for w in range(8):
                assert lib.jcdf_debug_exchange_block(ta, tb, w, N, out) == 0
                r0, c0, mt, nt, stair = out[0], out[1], out[2], out[3], out[4]
                assert 0 <= mt <= 8 and 0 <= nt <= 4
                if mt:
                    assert r0 + mt <= 16 and c0 + nt <= 16
                    block = np.ones((mt, nt), dtype=np.int32)
                    if stair:  # a rectangle that starts on the diagonal computes its lower staircase only
                        assert ta == tb and r0 == c0
                        block = np.tril(block)
                    cover[r0:r0 + mt, c0:c0 + nt] += block
                    per_warp.append(int(block.sum()))
                else:
                    per_warp.append(0)
            assert cover.max() <= 1  # no 8 x 8 tile computed twice
            rr, cc = np.meshgrid(16 * ta + np.arange(16), 16 * tb + np.arange(16), indexing="ij")
            needed = (rr < units) & (cc < units) & (rr >= cc)
            assert np.all(cover[needed] == 1)
            assert not np.any(cover[(rr >= units) | (cc >= units)])  # nothing beyond N rounded up to 8
            if ta == tb and 128 * (ta + 1) <= N:
                assert [per_warp[j] + per_warp[j + 4] for j in range(4)] == [34, 34, 34, 34]  # of the 64 of a full tile
                assert int(cover.sum()) == 136  # exactly the 8 x 8 tiles on or below the diagonal
            if ta != tb and 128 * (ta + 1) <= N:
                assert per_warp == [32] * 8


@pytest.mark.parametrize("N,n_blocks,hw", [(70, 10, 9), (120, 10, 18), (203, 10, 30), (150, 7, 25), (1250, 50, 200)])
def test_exchange_dropped_elements_equal_the_oracles_block_screen_matrix(N, n_blocks, hw):
    """The host helper bench.py and tools/block_screen_cases.py use to know which elements of F the block-screened exchange
    leaves uncomputed, against the oracle's restatement of calculate_exchange_block_screen_matrix (ScreenedDF.jl:385-457):
    one block below 100 functions, width N // n_blocks, never-screened remainder strip, counts of blocks."""
    from helpers import banded_mask
    mask = banded_mask(N, hw, seed=N)
    dropped, total, kept = wl.exchange_dropped_elements(mask, n_blocks)
    nb, width, batch = orc.exchange_block_layout(N, n_blocks, 10)
    todo, bsm = orc.exchange_block_screen_matrix(mask, nb, width, batch, True)
    ref = np.zeros((N, N), dtype=bool)
    ref[:nb * width, :nb * width] = ~np.kron(bsm | bsm.T, np.ones((width, width), dtype=bool))
    assert (total, kept) == (len(batch), len(todo))
    assert np.array_equal(dropped, ref) and np.array_equal(dropped, dropped.T)
    assert not (dropped & mask).any()          # a dropped block holds no surviving pair
    assert not wl.exchange_dropped_elements(mask, 0)[0].any()
    with pytest.raises(ValueError):
        wl.exchange_dropped_elements(None, n_blocks)


def test_shared_nccl_is_named_without_loading_anything(monkeypatch):
    """_lib._name_shared_nccl: JCDF_NCCL_LIBRARY is pointed at the pip package's libnccl.so.2 when there is one (so that a
    later `import torch` finds the NCCL it was built against already mapped, not an older system copy), a value the host
    set is kept, and nothing is dlopen'ed by naming it."""
    import importlib.util
    from juliachem_jl_b200 import _lib
    monkeypatch.setenv("JCDF_NCCL_LIBRARY", "/somewhere/libnccl.so.2")
    _lib._name_shared_nccl()
    assert os.environ["JCDF_NCCL_LIBRARY"] == "/somewhere/libnccl.so.2"
    monkeypatch.delenv("JCDF_NCCL_LIBRARY")
    _lib._name_shared_nccl()
    spec = importlib.util.find_spec("nvidia.nccl")
    if spec is None:
        assert "JCDF_NCCL_LIBRARY" not in os.environ
    else:
        path = os.environ["JCDF_NCCL_LIBRARY"]
        assert path.endswith(os.path.join("nvidia", "nccl", "lib", "libnccl.so.2")) and os.path.exists(path)
        assert "libnccl" not in open("/proc/self/maps").read() or "torch" in sys.modules


def test_layout_info_struct_matches_header_python_and_julia():
    """jcdf_layout_info is filled by the library and read by both wrappers: same fields in the same order in include/jcdf.h
    and _lib.LayoutInfo (all int64_t), and the Julia wrapper reads it as a vector of that many Int64 with the block counts at
    the 1-based positions it uses."""
    from juliachem_jl_b200 import _lib
    src = open(os.path.join(ROOT, "include", "jcdf.h")).read()
    body = re.search(r"typedef struct jcdf_layout_info \{(.*?)\} jcdf_layout_info;", src, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in re.findall(r"int64_t\s+([^;]+);", body):
        names += [n.strip() for n in decl.split(",")]
    assert names == [f[0] for f in _lib.LayoutInfo._fields_]
    assert ctypes.sizeof(_lib.LayoutInfo) == 8 * len(names)
    jl = open(os.path.join(ROOT, "juliachem.jl_b200", "julia", "B200DF.jl")).read()
    assert int(re.search(r"const JCDF_LAYOUT_INFO_FIELDS = (\d+)", jl).group(1)) == len(names)
    assert names.index("exchange_blocks_total") + 1 == 16 and "string(li[16])" in jl
    assert names.index("exchange_blocks_kept") + 1 == 17 and "string(li[17])" in jl
