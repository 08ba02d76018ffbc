"""utils/hdf5_lite: the mesh files' container format without h5py.  The reader is checked against a file written by
libhdf5 itself (scipy ships a MATLAB 7.3 = HDF5 sample), the writer against the reader, against the structure of that
genuine file (same superblock / heap / B-tree / symbol-node fields) and -- where h5py happens to be importable -- against
h5py in both directions."""
import os
import struct

import numpy as np
import pytest

import conftest  # noqa: F401
from utils import hdf5_lite as H


def _mesh(seed=0, n=25, f=32, e=112):
    r = np.random.default_rng(seed)
    return {"pos": r.standard_normal((n, 3)).astype(np.float32), "norm": r.standard_normal((n, 3)).astype(np.float32),
            "face": r.integers(0, n, (3, f)).astype(np.int64), "edge_index": r.integers(0, n, (2, e)).astype(np.int64)}


def test_round_trip_of_a_mesh_file(tmp_path):
    mesh = _mesh()
    path = H.write_hdf5(str(tmp_path / "mesh_000.hdf5"), mesh)
    back = H.read_hdf5(path)
    assert sorted(back) == sorted(mesh)
    for k, v in mesh.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape and np.array_equal(back[k], v), k
    assert np.array_equal(H.read_hdf5(path, ["face"])["face"], mesh["face"])
    with pytest.raises(KeyError):
        H.read_hdf5(path, ["missing"])


@pytest.mark.parametrize("dtype", ["f2", "f4", "f8", "i1", "i2", "i4", "i8", "u1", "u2", "u4", "u8"])
def test_every_numeric_dtype_scalars_and_empty_arrays(tmp_path, dtype):
    r = np.random.default_rng(3)
    data = {"a": (r.standard_normal((4, 5, 2)) * 50).astype(dtype), "scalar": np.array(7, dtype), "empty": np.zeros((0, 3), dtype),
            "vector": np.arange(17).astype(dtype)}
    back = H.read_hdf5(H.write_hdf5(str(tmp_path / "t.hdf5"), data))
    for k, v in data.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape and np.array_equal(back[k], v), k


def test_many_datasets_span_several_symbol_nodes(tmp_path):
    data = {f"d{i:03d}": np.full((i % 5 + 1,), i, np.int32) for i in range(37)}
    back = H.read_hdf5(H.write_hdf5(str(tmp_path / "many.hdf5"), data))
    assert sorted(back) == sorted(data) and all(np.array_equal(back[k], v) for k, v in data.items())


def test_unsupported_inputs_raise():
    with pytest.raises(H.Hdf5Unsupported):
        H.write_hdf5("/tmp/never.hdf5", {"b": np.zeros(3, bool)})
    with pytest.raises(H.Hdf5Unsupported):
        H.write_hdf5("/tmp/never.hdf5", {"s": np.array(["x"])})
    with pytest.raises(ValueError):
        H.write_hdf5("/tmp/never.hdf5", {"a/b": np.zeros(3)})


def _libhdf5_sample():
    try:
        import scipy.io.matlab.tests as t
    except Exception:
        return None
    p = os.path.join(os.path.dirname(t.__file__), "data", "testhdf5_7.4_GLNX86.mat")
    return p if os.path.exists(p) else None


@pytest.mark.skipif(_libhdf5_sample() is None, reason="scipy's bundled HDF5 sample is not installed")
def test_reader_on_a_file_written_by_libhdf5():
    """MATLAB 7.3 files are HDF5 behind a 512-byte user block: old-style group, v1 object headers, layout v2, a modification
    time and an attribute message beside the four the reader needs.  Content: linspace(0, 2 pi, 9) as a 9x1 double."""
    out = H.read_hdf5(_libhdf5_sample())
    assert list(out) == ["testdouble"]
    assert out["testdouble"].dtype == np.float64 and out["testdouble"].shape == (9, 1)
    assert np.allclose(out["testdouble"][:, 0], np.linspace(0, 2 * np.pi, 9), rtol=0, atol=1e-15)


@pytest.mark.skipif(_libhdf5_sample() is None, reason="scipy's bundled HDF5 sample is not installed")
def test_written_structures_match_the_genuine_files(tmp_path):
    """Field by field against the libhdf5-written sample: superblock, root symbol-table entry, heap header and free block,
    B-tree and symbol node headers -- the parts every HDF5 reader walks before it reaches a dataset."""
    ref = open(_libhdf5_sample(), "rb").read()[512:]
    own = open(H.write_hdf5(str(tmp_path / "x.hdf5"), {"testdouble": np.linspace(0, 2 * np.pi, 9)[:, None]}), "rb").read()
    assert own[:8] == ref[:8] == H.SIGNATURE
    assert own[8:20] == ref[8:20]                                  # versions 0, 8-byte offsets / lengths, K = 4 / 16
    assert struct.unpack_from("<Q", own, 32)[0] == struct.unpack_from("<Q", ref, 32)[0] == H.UNDEF          # no free-space info
    assert struct.unpack_from("<Q", own, 40)[0] == len(own)                                                # end-of-file address
    assert struct.unpack_from("<Q", own, 48)[0] == struct.unpack_from("<Q", ref, 48)[0] == H.UNDEF          # no driver block

    def walk(buf):
        name0, header, cache, _, btree, heap = struct.unpack_from("<QQIIQQ", buf, 56)
        assert (name0, cache) == (0, 1)
        version, nmsg, refs, size = struct.unpack_from("<BxHII", buf, header)
        kind, n, flags = struct.unpack_from("<HHB", buf, header + 16)
        assert (version, refs, kind, n) == (1, 1, 0x11, 16) and struct.unpack_from("<QQ", buf, header + 24) == (btree, heap)
        assert buf[heap:heap + 8] == b"HEAP\0\0\0\0"
        seg_size, free, seg = struct.unpack_from("<QQQ", buf, heap + 8)
        assert buf[seg:seg + 8] == b"\0" * 8 and buf[seg + 8:seg + 19] == b"testdouble\0"
        assert struct.unpack_from("<QQ", buf, seg + free) == (H.HEAP_FREE_NULL, seg_size - free)
        assert buf[btree:btree + 8] == b"TREE\0\0\x01\0" and struct.unpack_from("<QQ", buf, btree + 8) == (H.UNDEF, H.UNDEF)
        key0, snod, key1 = struct.unpack_from("<QQQ", buf, btree + 24)
        assert (key0, key1) == (0, 8) and buf[snod:snod + 8] == b"SNOD\x01\0\x01\0"
        name, obj, ctype = struct.unpack_from("<QQI", buf, snod + 8)
        assert (name, ctype) == (8, 0)
        version, nmsg, refs, size = struct.unpack_from("<BxHII", buf, obj)
        assert (version, refs) == (1, 1)
        msgs, q = {}, obj + 16
        for _ in range(nmsg):
            kind, n = struct.unpack_from("<HH", buf, q)
            msgs[kind] = buf[q + 8:q + 8 + n]
            q += 8 + n
        return msgs

    a, b = walk(own), walk(ref)
    assert a[0x0003] == b[0x0003]                                  # the IEEE double datatype message, byte for byte
    assert a[0x0001] == b[0x0001]                                  # dataspace v1: rank 2, dims 9 x 1
    assert a[0x0005][1:] == b[0x0005][1:]                          # fill value: late allocation, written if set, default value


def test_against_h5py_when_it_is_importable(tmp_path):
    h5py = pytest.importorskip("h5py")
    mesh = _mesh(1)
    with h5py.File(H.write_hdf5(str(tmp_path / "lite.hdf5"), mesh), "r") as f:
        assert sorted(f) == sorted(mesh) and all(np.array_equal(f[k][:], v) for k, v in mesh.items())
    with h5py.File(str(tmp_path / "h5py.hdf5"), "w") as f:
        for k, v in mesh.items():
            f.create_dataset(k, data=v)
    back = H.read_hdf5(str(tmp_path / "h5py.hdf5"))
    assert all(np.array_equal(back[k], v) for k, v in mesh.items())
