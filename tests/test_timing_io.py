"""the HDF5 timing file of example_scripts/jc_timings_write.jl, written without an HDF5 library.  The writer
and the minimal reader are checked against each other, the table contents against the reference script's rules, and the
reader's object-header parsing against a file libhdf5 itself wrote (the reference's fock_GPU_1_ranks_reduced.h5, when the
reference checkout is present)."""
import os
import struct

import numpy as np
import pytest

from helpers import jc
from juliachem_jl_b200 import timing_io as tio


def _timing():
    t = jc.JCTiming()
    for it in (1, 2, 10):
        for name, v in (("W_time", 0.015), ("K_time", 0.055), ("fock_time", 0.0713), ("GPU_1_non_zero_coeff_time", 1e-4)):
            t.timings[jc.JCTiming.key(name, it)] = v * it
    t.timings["B_time"] = 0.9
    t.timings["screening_time"] = 0.12
    t.non_timing_data.update({"n_basis_functions": "1250", "n_auxiliary_basis_functions": "4800", "n_oo": "250",
                              "contraction_algorithm": "B200", "n_iterations": "29"})
    return t


def test_round_trip_and_reference_table_rules(tmp_path):
    t = _timing()
    path = str(tmp_path / "run.h5")
    opts = {"density_fitting": "true", "contraction_mode": "B200", "num_devices": 8, "df_screening_sigma": 1e-5}
    tio.save_jc_timings_to_hdf5(t, path, run_time=12.5, run_name="w50", converged=True, scf_energy=-3801.590582588,
                                options=opts, user_options={"contraction_mode": "B200"})
    got = tio.read_string_tables(path)
    assert sorted(got) == ["non_timing_data", "run_level_data", "scf_options", "scf_options-user", "timings"]
    # scf_options: 18 rows in the reference's fixed order (jc_timings_write.jl:84-140)
    assert [k for k, _ in got["scf_options"]] == list(tio.SCF_OPTION_KEYS) and len(got["scf_options"]) == 18
    assert dict(got["scf_options"])["contraction_mode"] == "B200" and dict(got["scf_options"])["num_devices"] == "8"
    # run_level_data: 12 rows (:165-205)
    rl = got["run_level_data"]
    assert [k for k, _ in rl][:4] == ["run_name", "run_time", "converged", "scf_energy"] and len(rl) == 12
    assert dict(rl)["converged"] == "true" and float(dict(rl)["scf_energy"]) == -3801.590582588
    assert dict(rl)["n_basis_functions"] == "1250" and dict(rl)["n_atoms"] == ""
    # timings: by iteration (numeric, so 10 follows 2), by key inside an iteration, non-iteration keys last (:40-82)
    keys = [k for k, _ in got["timings"]]
    assert keys[:4] == ["GPU_1_non_zero_coeff_time-1", "K_time-1", "W_time-1", "fock_time-1"]
    assert keys[4].endswith("-2") and keys[8].endswith("-10") and keys[-2:] == ["B_time", "screening_time"]
    assert len(keys) == len(t.timings)
    assert float(dict(got["timings"])["K_time-10"]) == t.timings["K_time-10"]
    assert got["non_timing_data"] == sorted(t.non_timing_data.items())


def test_container_structure(tmp_path):
    """Superblock version 0, 8-byte offsets, end-of-file address = file size, root symbol table cached in the superblock,
    symbol-node entries sorted by name, every structure 8-byte aligned."""
    path = str(tmp_path / "t.h5")
    tio.write_string_tables(path, {"b": [("k", "v")], "a": [], "c": [("x" * 37, "é")]})
    buf = open(path, "rb").read()
    assert buf[:8] == b"\x89HDF\r\n\x1a\n" and buf[8] == 0 and buf[13] == 8 and buf[14] == 8
    base, free, eof, drv = struct.unpack_from("<QQQQ", buf, 24)
    assert base == 0 and eof == len(buf) and free == drv == tio.UNDEF and len(buf) % 8 == 0
    got = tio.read_string_tables(path)
    assert list(got) == ["a", "b", "c"] and got["a"] == [] and got["c"] == [("x" * 37, "é")]
    with pytest.raises(ValueError):
        tio.write_string_tables(path, {str(i): [] for i in range(9)})


@pytest.mark.skipif(not os.path.exists("/root/reference/fock_GPU_1_ranks_reduced.h5"), reason="reference checkout not present")
def test_reader_parses_a_header_written_by_libhdf5():
    buf = open("/root/reference/fock_GPU_1_ranks_reduced.h5", "rb").read()
    dims, cls, esize, raw = tio.read_dataset_at(buf, 0x320)  # the "fock" dataset of the reference's dumped Fock matrix
    assert dims == (110, 110) and cls == 1 and esize == 8 and len(raw) == 110 * 110 * 8
    F = np.frombuffer(raw, dtype="<f8").reshape(110, 110)
    assert np.isfinite(F).all() and np.abs(F - F.T).max() < 1e-8
