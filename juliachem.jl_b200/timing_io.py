"""Timing file of a run: the HDF5 file the reference writes with example_scripts/jc_timings_write.jl:7-140
(`save_jc_timings_to_hdf5`) and its plotting scripts read back (test/create_plots-df-vs-rhf.jl) — five datasets, each an
(n x 2) table of strings [key, value]:

    scf_options, scf_options-user    18 rows in the reference's fixed order (:84-140)
    run_level_data                   run_name, run_time, converged, scf_energy + eight counts taken from non_timing_data (:165-205)
    timings                          every JCTiming.timings entry, sorted by iteration then key; non-iteration keys last (:40-82)
    non_timing_data                  sorted by key (:21-32)

In the Julia drop-in the file is written by the reference's own script from the JCTiming object B200DF.jl fills; this module
is the same writer for the Python mirror of the seam.  Neither h5py nor libhdf5 exists in the build image, so the container
format is written directly: HDF5 superblock version 0, one old-style root group (symbol table = B-tree node + local heap +
one symbol node), object headers version 1, contiguous fixed-length UTF-8 string datasets of rank 2 stored (2, n) row-major —
the layout HDF5.jl gives a Julia Array{String}(n, 2).  `read_string_tables` is the matching minimal reader (tests).
"""
from __future__ import annotations

import struct

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"
LEAF_K, INTERNAL_K = 4, 16

# JCTC option keys in the order save_scf_options writes them (jc_timings_write.jl:84-140; names JCTiming.jl:20-49)
SCF_OPTION_KEYS = ("density_fitting", "contraction_mode", "load", "guess", "energy_convergence", "density_convergence",
                   "df_energy_convergence", "df_density_convergence", "max_iterations", "df_max_iterations",
                   "df_exchange_n_blocks", "df_screening_sigma", "df_screen_exchange", "df_force_dense", "df_use_adaptive",
                   "num_devices", "df_use_K_sym", "df_K_sym_type")


# JCTC.n_basis_functions ... n_iterations (JCTiming.jl:20-29, 75): rows 5-12 of run_level_data
RUN_LEVEL_COUNTS = ("n_basis_functions", "n_auxiliary_basis_functions", "n_electrons", "n_oo", "n_atoms", "n_threads",
                    "n_ranks", "n_iterations")


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _message(mtype: int, data: bytes, flags: int = 0) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), flags) + data


def _object_header(messages: list[bytes]) -> bytes:
    body = b"".join(messages)
    # version 1, reserved, message count, object reference count, header size; messages start on an 8-byte boundary
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(body)) + body


def _string_table_dataset(rows: list[tuple[str, str]], data_address: int) -> tuple[bytes, bytes]:
    """(object header, raw data) of an (n x 2) string table stored as HDF5 dims (2, n), fixed-length null-padded UTF-8."""
    n = len(rows)
    enc = [k.encode("utf-8") for k, _ in rows] + [v.encode("utf-8") for _, v in rows]
    width = max([1] + [len(e) for e in enc])
    raw = b"".join(e.ljust(width, b"\0") for e in enc)
    dataspace = struct.pack("<BBB5x", 1, 2, 1) + struct.pack("<QQ", 2, n) + struct.pack("<QQ", 2, n)  # v1, rank 2, max dims
    datatype = struct.pack("<BBBBI", 0x13, 0x11, 0, 0, width)  # class 3 (string) v1; null-padded, UTF-8; element size
    fill = struct.pack("<BBBB", 2, 2, 2, 0)                    # v2: late allocation, never written, undefined
    layout = struct.pack("<BBQQ", 3, 1, data_address if n else UNDEF, len(raw))  # v3, contiguous
    header = _object_header([_message(0x0001, dataspace), _message(0x0003, datatype, 1), _message(0x0005, fill, 1),
                             _message(0x0008, layout)])
    return header, raw


def write_string_tables(path: str, tables: dict[str, list[tuple[str, str]]]) -> None:
    """One HDF5 file whose root group holds one (n x 2) string dataset per entry of `tables`."""
    names = sorted(tables)  # symbol-table entries are ordered by name
    if len(names) > 2 * LEAF_K:
        raise ValueError("at most %d datasets fit the single symbol node this writer emits" % (2 * LEAF_K))
    # local heap data segment: the empty name at offset 0, then the dataset names, each NUL-terminated and 8-byte aligned
    heap_data, name_off = _pad8(b"\0"), {}
    for nm in names:
        name_off[nm] = len(heap_data)
        heap_data += _pad8(nm.encode("utf-8") + b"\0")
    # fixed part of the file: superblock (96 bytes incl. the root symbol-table entry), root object header, B-tree node,
    # heap header + data, symbol node; then per dataset its object header and raw data
    sb_size = 8 + 8 + 4 + 4 + 4 * 8 + 40
    root_header_addr = sb_size
    root_header_size = 16 + 8 + 16
    btree_addr = root_header_addr + root_header_size
    btree_size = 24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8
    heap_addr = btree_addr + btree_size
    heap_data_addr = heap_addr + 32
    snod_addr = heap_data_addr + len(heap_data)
    snod_size = 8 + 2 * LEAF_K * 40
    cursor = snod_addr + snod_size
    blobs, entries = [], []
    for nm in names:
        probe, _ = _string_table_dataset(tables[nm], 0)
        hdr_addr = cursor
        data_addr = hdr_addr + len(probe)
        hdr, raw = _string_table_dataset(tables[nm], data_addr)
        assert len(hdr) == len(probe)
        raw = _pad8(raw)
        blobs.append(hdr + raw)
        entries.append((name_off[nm], hdr_addr))
        cursor = data_addr + len(raw)
    eof = cursor
    out = bytearray()
    out += SIGNATURE + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", LEAF_K, INTERNAL_K, 0)
    out += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    out += struct.pack("<QQII", 0, root_header_addr, 1, 0) + struct.pack("<QQ", btree_addr, heap_addr)  # root entry, cached stab
    assert len(out) == sb_size
    out += _object_header([_message(0x0011, struct.pack("<QQ", btree_addr, heap_addr))])
    assert len(out) == btree_addr
    # B-tree v1 group node, level 0: key 0 = the empty name, one child (the symbol node), key 1 = the largest name in it
    node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1 if names else 0, UNDEF, UNDEF)
    node += struct.pack("<QQQ", 0, snod_addr if names else UNDEF, name_off[names[-1]] if names else 0)
    out += node.ljust(btree_size, b"\0")
    out += b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 1, heap_data_addr) + heap_data  # free list head 1 = none
    assert len(out) == snod_addr
    snod = b"SNOD" + struct.pack("<BBH", 1, 0, len(names))
    for off, addr in entries:
        snod += struct.pack("<QQII16x", off, addr, 0, 0)
    out += snod.ljust(snod_size, b"\0")
    for b in blobs:
        out += b
    assert len(out) == eof
    with open(path, "wb") as f:
        f.write(bytes(out))


# ----------------------------------------------------------------------------------------------------------------------
# the reference's table contents
# ----------------------------------------------------------------------------------------------------------------------
def sort_timings(timings: dict[str, float]) -> list[tuple[str, str]]:
    """sort_timings (jc_timings_write.jl:40-82): keys "<name>-<iteration>" grouped by iteration (ascending), sorted by key
    inside an iteration; keys without a "-" part last, sorted."""
    by_iter: dict[int, list[tuple[str, float]]] = {}
    rest: dict[str, float] = {}
    for key, val in timings.items():
        parts = key.split("-")
        if len(parts) >= 2:
            by_iter.setdefault(int(parts[-1]), []).append((key, val))
        else:
            rest[key] = val
    rows = []
    for it in sorted(by_iter):
        rows += [(k, repr(float(v))) for k, v in sorted(by_iter[it])]
    rows += [(k, repr(float(rest[k]))) for k in sorted(rest)]
    return rows


def save_jc_timings_to_hdf5(jc_timing, file_path: str, run_time: float = 0.0, run_name: str = "", converged: bool = False,
                            scf_energy: float = 0.0, options: dict | None = None, user_options: dict | None = None) -> None:
    """save_jc_timings_to_hdf5 (jc_timings_write.jl:7-19) for a df.JCTiming object."""
    def option_rows(opts):
        opts = opts or {}
        return [(k, str(opts.get(k, ""))) for k in SCF_OPTION_KEYS]

    def jl(v):
        return "true" if v is True else "false" if v is False else str(v)

    ntd = getattr(jc_timing, "non_timing_data", {})
    tables = {
        "scf_options": option_rows(options if options is not None else getattr(jc_timing, "options", {})),
        "scf_options-user": option_rows(user_options if user_options is not None else getattr(jc_timing, "user_options", {})),
        # save_run_level_data (jc_timings_write.jl:165-205): 12 rows in this order; the counts come from non_timing_data
        "run_level_data": [("run_name", run_name), ("run_time", jl(float(run_time))), ("converged", jl(bool(converged))),
                           ("scf_energy", jl(float(scf_energy)))] +
                          [(k, str(ntd.get(k, ""))) for k in RUN_LEVEL_COUNTS],
        "timings": sort_timings(jc_timing.timings),
        "non_timing_data": sorted((k, str(v)) for k, v in getattr(jc_timing, "non_timing_data", {}).items()),
    }
    write_string_tables(file_path, tables)


# ----------------------------------------------------------------------------------------------------------------------
# minimal reader of the subset this module writes (tests; also parses version-1 object headers of files libhdf5 wrote)
# ----------------------------------------------------------------------------------------------------------------------
def _read_object_header_v1(buf: bytes, addr: int) -> list[tuple[int, bytes]]:
    version, _, nmsg, _, size = struct.unpack_from("<BBHII", buf, addr)
    if version != 1:
        raise ValueError("object header version %d not supported" % version)
    msgs, blocks = [], [(addr + 16, size)]
    while blocks and len(msgs) < nmsg:
        pos, left = blocks.pop(0)
        end = pos + left
        while pos + 8 <= end and len(msgs) < nmsg:
            mtype, msize, _ = struct.unpack_from("<HHB", buf, pos)
            data = buf[pos + 8:pos + 8 + msize]
            msgs.append((mtype, data))
            if mtype == 0x0010:  # continuation block
                blocks.append(struct.unpack_from("<QQ", data, 0))
            pos += 8 + msize
    return msgs


def read_dataset_at(buf: bytes, addr: int):
    """(dims, datatype class, element size, raw bytes) of the contiguous dataset whose version-1 object header is at addr."""
    dims, cls, esize, raw = None, None, None, None
    for mtype, data in _read_object_header_v1(buf, addr):
        if mtype == 0x0001:
            ver, rank, _ = struct.unpack_from("<BBB", data, 0)
            if ver != 1:
                raise ValueError("dataspace version %d not supported" % ver)
            dims = struct.unpack_from("<%dQ" % rank, data, 8)
        elif mtype == 0x0003:
            cls, esize = data[0] & 0x0F, struct.unpack_from("<I", data, 4)[0]
        elif mtype == 0x0008:
            ver, lclass = struct.unpack_from("<BB", data, 0)
            if ver != 3 or lclass != 1:
                raise ValueError("only contiguous version-3 layouts are supported")
            a, n = struct.unpack_from("<QQ", data, 2)
            raw = b"" if a == UNDEF else buf[a:a + n]
    return dims, cls, esize, raw


def read_string_tables(path: str) -> dict[str, list[tuple[str, str]]]:
    buf = open(path, "rb").read()
    if buf[:8] != SIGNATURE or buf[8] != 0:
        raise ValueError("not an HDF5 file with a version-0 superblock")
    btree_addr, heap_addr = struct.unpack_from("<QQ", buf, 24 + 32 + 24)
    assert buf[btree_addr:btree_addr + 4] == b"TREE" and buf[heap_addr:heap_addr + 4] == b"HEAP"
    heap_data = struct.unpack_from("<Q", buf, heap_addr + 24)[0]
    level, used = struct.unpack_from("<BH", buf, btree_addr + 5)
    assert level == 0
    out = {}
    for c in range(used):
        snod = struct.unpack_from("<Q", buf, btree_addr + 24 + 8 + 16 * c)[0]
        assert buf[snod:snod + 4] == b"SNOD"
        nsym = struct.unpack_from("<H", buf, snod + 6)[0]
        for e in range(nsym):
            off, addr = struct.unpack_from("<QQ", buf, snod + 8 + 40 * e)
            name = buf[heap_data + off:buf.index(b"\0", heap_data + off)].decode("utf-8")
            dims, cls, esize, raw = read_dataset_at(buf, addr)
            assert cls == 3 and len(dims) == 2 and dims[0] == 2
            cells = [raw[i * esize:(i + 1) * esize].rstrip(b"\0").decode("utf-8") for i in range(2 * dims[1])]
            out[name] = list(zip(cells[:dims[1]], cells[dims[1]:]))
    return out
