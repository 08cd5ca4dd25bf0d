"""The hand-written HDF5 files of host/h5min.h (H5::save / H5::load of the reference's h5.h) against an independent
reader: a pure-Python walk over the published HDF5 file format (superblock version 0, old-style root group:
B-tree + local heap + symbol-table nodes, version-1 object headers, contiguous layout).  h5py is not installed here;
this reader shares no code with the writer."""
import os
import struct

import numpy as np
import pytest

UNDEF = 0xFFFFFFFFFFFFFFFF


def parse_hdf5(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89HDF\r\n\x1a\n"
    assert b[8] == 0 and b[13] == 8 and b[14] == 8          # superblock version 0, 8-byte offsets and lengths
    leaf_k, internal_k = struct.unpack_from("<HH", b, 16)
    base, free, eof, driver = struct.unpack_from("<QQQQ", b, 24)
    assert base == 0 and free == UNDEF and driver == UNDEF and eof == len(b)
    name_off, header, cache, _, btree, heap = struct.unpack_from("<QQIIQQ", b, 56)
    assert cache == 1
    # the root object header repeats the two addresses in its symbol-table message
    assert b[header] == 1 and struct.unpack_from("<H", b, header + 2)[0] == 1
    mtype, msize = struct.unpack_from("<HH", b, header + 16)
    assert mtype == 0x11 and msize == 16 and struct.unpack_from("<QQ", b, header + 24) == (btree, heap)
    assert b[heap:heap + 4] == b"HEAP"
    heap_size, free_list, heap_data = struct.unpack_from("<QQQ", b, heap + 8)

    def name_at(off):
        end = b.index(b"\0", heap_data + off)
        return b[heap_data + off:end].decode()

    assert b[btree:btree + 4] == b"TREE" and b[btree + 4] == 0 and b[btree + 5] == 0
    n_children = struct.unpack_from("<H", b, btree + 6)[0]
    assert n_children <= 2 * internal_k
    out, names = {}, []
    keys = [struct.unpack_from("<Q", b, btree + 24 + 16 * k)[0] for k in range(n_children + 1)]
    for k in range(n_children):
        node = struct.unpack_from("<Q", b, btree + 32 + 16 * k)[0]
        assert b[node:node + 4] == b"SNOD" and b[node + 4] == 1
        n_sym = struct.unpack_from("<H", b, node + 6)[0]
        assert 0 < n_sym <= 2 * leaf_k or (n_children == 1 and n_sym == 0)
        for e in range(n_sym):
            off, obj = struct.unpack_from("<QQ", b, node + 8 + 40 * e)
            name = name_at(off)
            names.append(name)
            assert name_at(keys[k]) < name <= name_at(keys[k + 1]) or keys[k] == 0 and name <= name_at(keys[k + 1])
            out[name] = read_dataset(b, obj)
    assert names == sorted(names)                           # the library finds names by binary search
    return out


def read_dataset(b, obj):
    assert b[obj] == 1
    n_msgs = struct.unpack_from("<H", b, obj + 2)[0]
    size = struct.unpack_from("<I", b, obj + 8)[0]
    p, end = obj + 16, obj + 16 + size
    dims = address = nbytes = None
    is_double = False
    for _ in range(n_msgs):
        mtype, msize = struct.unpack_from("<HH", b, p)
        d = p + 8
        if mtype == 1:
            assert b[d] == 1
            rank = b[d + 1]
            dims = struct.unpack_from("<%dQ" % rank, b, d + 8)
        elif mtype == 3:
            assert b[d] == 0x11 and b[d + 1] & 1 == 0 and struct.unpack_from("<I", b, d + 4)[0] == 8
            bit_offset, precision, exp_loc, exp_size, man_loc, man_size, bias = struct.unpack_from("<HHBBBBI", b, d + 8)
            is_double = (precision, exp_loc, exp_size, man_loc, man_size, bias, b[d + 2]) == (64, 52, 11, 0, 52, 1023, 63)
        elif mtype == 8:
            assert b[d] == 3 and b[d + 1] == 1
            address, nbytes = struct.unpack_from("<QQ", b, d + 2)
        p += 8 + msize
        assert p <= end and msize % 8 == 0
    assert is_double and dims is not None and nbytes == 8 * int(np.prod(dims))
    return np.frombuffer(b, dtype="<f8", count=int(np.prod(dims)), offset=address).reshape(dims)


def test_saved_vectors_come_back_bit_for_bit_through_an_independent_reader(pkg, tmp_path):
    f = str(tmp_path / "checkpoint_save.h5")
    rng = np.random.default_rng(35)
    want = {}
    # the reference's checkpoint set for dim = 3 (step-35a.cc:889-899), in its order: the file grows by one dataset per call
    for name, n in [("u_n0", 1001), ("u_n_minus_10", 1001), ("u_n1", 1001), ("u_n_minus_11", 1001), ("u_n2", 1001),
                    ("u_n_minus_12", 1001), ("phi_n", 77), ("phi_n_minus_1", 77), ("pres_n", 77), ("pres_n_minus_1", 77)]:
        want[name] = rng.standard_normal(n)
        want[name][0] = -0.0
        pkg.h5_save(f, name, want[name])
        got = parse_hdf5(f)
        assert sorted(got) == sorted(want)
        for k in want:
            assert got[k].tobytes() == want[k].tobytes()
    for k in want:                                          # and through the library's own reader
        assert pkg.h5_load(f, k).tobytes() == want[k].tobytes()
    assert pkg.h5_load(f, "/phi_n").tobytes() == want["phi_n"].tobytes()   # the reference's block vectors use "/a0"


def test_many_datasets_empty_datasets_and_replacement(pkg, tmp_path):
    f = str(tmp_path / "many.h5")
    want = {}
    for i in range(21):                                     # more than one symbol-table node (8 entries each)
        want["a%d" % i] = np.arange(i, dtype=float) * 0.5   # a0 is empty
        pkg.h5_save(f, "a%d" % i, want["a%d" % i])
    want["a7"] = np.array([1.0, 2.0, 3.0])                  # the same name again: the new data replace the old
    pkg.h5_save(f, "a7", want["a7"])
    got = parse_hdf5(f)
    assert sorted(got) == sorted(want)
    for k in want:
        assert np.array_equal(got[k], want[k]) and np.array_equal(pkg.h5_load(f, k), want[k])


def test_missing_files_and_datasets_are_errors(pkg, tmp_path):
    f = str(tmp_path / "x.h5")
    with pytest.raises(pkg.S35Error):
        pkg.h5_load(f, "phi_n")
    pkg.h5_save(f, "phi_n", np.ones(3))
    with pytest.raises(pkg.S35Error) as e:
        pkg.h5_load(f, "pres_n")
    assert "not found" in str(e.value)
    open(f, "wb").write(b"not an hdf5 file at all")
    with pytest.raises(pkg.S35Error) as e:
        pkg.h5_load(f, "phi_n")
    assert "not an HDF5 file" in str(e.value)
    truncated = str(tmp_path / "t.h5")
    pkg.h5_save(truncated, "phi_n", np.ones(100))
    data = open(truncated, "rb").read()
    open(truncated, "wb").write(data[:len(data) - 200])
    with pytest.raises(pkg.S35Error):
        pkg.h5_load(truncated, "phi_n")
