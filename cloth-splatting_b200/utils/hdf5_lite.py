"""A dependency-free reader / writer for the flat HDF5 files of the reference's mesh hand-off:

    with h5py.File(path, "w") as f:                      meshnet/generate_mesh_predictions.py:35-39,
        for key, value in mesh.items():                  scene_reconstruction/gaussian_mesh.py:463-465
            f.create_dataset(key, data=value.numpy())
    h5py.File(path, "r")["pos"][:] ...                   meshnet/data_utils.py:450-457 (load_mesh_from_h5py)

h5py / libhdf5 are not in this image, so the container format is written and parsed here from the HDF5 File Format
Specification (version 1.1 structures, what h5py's default `libver="earliest"` emits for such a file):

    superblock v0 -> root group object header (v1, Symbol Table message) -> local heap (names) + B-tree v1 (one leaf
    level) -> symbol table nodes ("SNOD") -> one v1 object header per dataset (Dataspace v1, Datatype v1 fixed / floating
    point, Fill Value v2, Data Layout v3 contiguous) -> raw little-endian data.

The reader additionally understands what other writers of the same era produce (user blocks, superblock v1, dataspace v2,
layout v1 / v2, compact layout, big-endian data, object header continuation blocks, multi-level B-trees); it is checked
against a file written by libhdf5 itself (scipy's bundled MATLAB 7.3 sample, tests/test_hdf5_lite.py).  Chunked or
filtered datasets, new-style groups (superblock v2+), compound / string / reference types raise `Hdf5Unsupported`."""
import struct

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF
LEAF_K, INTERNAL_K = 4, 16          # group B-tree ranks of the default file creation property list
HEAP_FREE_NULL = 1                   # H5HL_FREE_NULL: "no (further) free block" in a local heap's free list


class Hdf5Unsupported(ValueError):
    """The file uses a feature outside the flat contiguous-dataset subset this module covers."""


def _pad8(n):
    return (n + 7) & ~7


# ------------------------------------------------------------------------------------------------------------------
# writer
# ------------------------------------------------------------------------------------------------------------------
def _datatype_message(dt):
    dt = np.dtype(dt)
    if dt.byteorder == ">":
        raise Hdf5Unsupported("big-endian arrays are not written")
    if dt.kind in "iu":
        bits = 0x08 if dt.kind == "i" else 0x00                     # bit 3: two's complement signed
        head = struct.pack("<BBBBI", 0x10 | 0, bits, 0, 0, dt.itemsize)
        return head + struct.pack("<HH", 0, 8 * dt.itemsize)
    if dt.kind == "f" and dt.itemsize in (2, 4, 8):
        exp_bits, man_bits = {2: (5, 10), 4: (8, 23), 8: (11, 52)}[dt.itemsize]
        head = struct.pack("<BBBBI", 0x10 | 1, 0x20, 8 * dt.itemsize - 1, 0, dt.itemsize)  # implied-msb mantissa, sign bit position
        return head + struct.pack("<HHBBBBI", 0, 8 * dt.itemsize, man_bits, exp_bits, 0, man_bits, (1 << (exp_bits - 1)) - 1)
    raise Hdf5Unsupported(f"dtype {dt} (only integers and IEEE floats)")


def _message(kind, data, flags=0):
    data = data + b"\0" * (_pad8(len(data)) - len(data))
    return struct.pack("<HHB3x", kind, len(data), flags) + data


def _object_header(messages):
    body = b"".join(messages)
    return struct.pack("<BxHII4x", 1, len(messages), 1, len(body)) + body


def write_hdf5(path, datasets):
    """Writes {name: array} as the root-level contiguous datasets of a new HDF5 file (the `create_dataset(key, data=...)`
    loop of the reference).  Arrays are stored C-contiguous, little-endian, in their own dtype."""
    items = []
    for name in sorted(datasets, key=lambda s: s.encode()):       # symbol table nodes are ordered by strcmp
        arr = np.asarray(datasets[name])
        if arr.dtype == np.bool_:
            raise Hdf5Unsupported("bool arrays (h5py stores them as an enum type)")
        if not name or "/" in name or "\0" in name:
            raise ValueError(f"dataset name {name!r}")
        arr = arr.astype(arr.dtype.newbyteorder("<") if arr.dtype.byteorder == ">" else arr.dtype, order="C", copy=False)  # keeps 0-d
        items.append((name, arr))
    n_snod = max(1, -(-len(items) // (2 * LEAF_K)))
    if n_snod > 2 * INTERNAL_K:
        raise Hdf5Unsupported(f"{len(items)} datasets (one B-tree level holds {4 * LEAF_K * INTERNAL_K})")

    # local heap data segment: "" at offset 0, the names, one trailing free block
    heap, name_off = bytearray(8), {}
    for name, _ in items:
        name_off[name] = len(heap)
        raw = name.encode() + b"\0"
        heap += raw + b"\0" * (_pad8(len(raw)) - len(raw))
    free_at = len(heap)
    heap_size = _pad8(max(len(heap) + 16, 88))
    heap += struct.pack("<QQ", HEAP_FREE_NULL, heap_size - free_at) + b"\0" * (heap_size - free_at - 16)

    # layout of the file
    root_header_at = 96
    root_header_len = 16 + 24 + 8
    heap_at = root_header_at + root_header_len
    heap_data_at = heap_at + 32
    btree_at = heap_data_at + heap_size
    btree_len = 24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8
    snod_at = btree_at + btree_len
    snod_len = 8 + 2 * LEAF_K * 40
    cursor = snod_at + n_snod * snod_len
    headers, header_at = [], {}
    for name, arr in items:                                         # object headers first (their size does not depend on addresses)
        header_at[name] = cursor
        cursor += 16 + (8 + _pad8(8 + 8 * arr.ndim)) + (8 + _pad8(len(_datatype_message(arr.dtype)))) + (8 + 8) + (8 + 24)
    data_at = {}
    for name, arr in items:
        cursor = _pad8(cursor)
        data_at[name] = cursor if arr.nbytes else UNDEF
        cursor += arr.nbytes
    eof = cursor

    out = bytearray()
    out += SIGNATURE + struct.pack("<BBBxBBBxHHI", 0, 0, 0, 0, 8, 8, LEAF_K, INTERNAL_K, 0)
    out += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    out += struct.pack("<QQII", 0, root_header_at, 1, 0) + struct.pack("<QQ", btree_at, heap_at)
    assert len(out) == root_header_at
    out += _object_header([_message(0x0011, struct.pack("<QQ", btree_at, heap_at)), _message(0x0000, b"")])
    assert len(out) == heap_at
    out += b"HEAP" + struct.pack("<B3xQQQ", 0, heap_size, free_at, heap_data_at) + heap
    assert len(out) == btree_at
    groups = [items[i:i + 2 * LEAF_K] for i in range(0, len(items), 2 * LEAF_K)] or [[]]
    tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(groups), UNDEF, UNDEF) + struct.pack("<Q", 0)
    for i, grp in enumerate(groups):
        tree += struct.pack("<QQ", snod_at + i * snod_len, name_off[grp[-1][0]] if grp else 0)
    out += tree + b"\0" * (btree_len - len(tree))
    for grp in groups:
        node = b"SNOD" + struct.pack("<BxH", 1, len(grp))
        for name, _ in grp:
            node += struct.pack("<QQII16x", name_off[name], header_at[name], 0, 0)
        out += node + b"\0" * (snod_len - len(node))
    for name, arr in items:
        assert len(out) == header_at[name]
        out += _object_header([
            _message(0x0001, struct.pack("<BBB5x", 1, arr.ndim, 0) + b"".join(struct.pack("<Q", d) for d in arr.shape)),
            _message(0x0003, _datatype_message(arr.dtype), flags=1),
            _message(0x0005, struct.pack("<BBBBI", 2, 2, 2, 1, 0), flags=1),      # v2, late allocation, fill if set, default fill (size 0)
            _message(0x0008, struct.pack("<BBQQ", 3, 1, data_at[name], arr.nbytes))])
    for name, arr in items:
        out += b"\0" * (_pad8(len(out)) - len(out))
        if arr.nbytes:
            assert len(out) == data_at[name]
            out += arr.tobytes()
    assert len(out) == eof
    with open(path, "wb") as f:
        f.write(bytes(out))
    return path


# ------------------------------------------------------------------------------------------------------------------
# reader
# ------------------------------------------------------------------------------------------------------------------
class _File:
    def __init__(self, buf):
        self.buf = buf
        at = 0
        while buf[at:at + 8] != SIGNATURE:                         # a user block moves the superblock to 512, 1024, ...
            at = 512 if at == 0 else at * 2
            if at + 8 > len(buf):
                raise ValueError("not an HDF5 file (no superblock signature)")
        version = buf[at + 8]
        if version > 1:
            raise Hdf5Unsupported(f"superblock version {version} (new-style groups, written with libver='latest')")
        self.so, self.sl = buf[at + 13], buf[at + 14]
        if (self.so, self.sl) != (8, 8):
            raise Hdf5Unsupported(f"{self.so}-byte offsets / {self.sl}-byte lengths")
        p = at + 24 + (4 if version == 1 else 0)
        self.base, _, self.eof, _ = struct.unpack_from("<QQQQ", buf, p)
        if self.eof > len(buf) + self.base:                         # (MATLAB writes the absolute end behind its user block)
            raise ValueError("truncated HDF5 file")
        _, self.root_header, cache, _, btree, heap = struct.unpack_from("<QQIIQQ", buf, p + 32)
        self.root_cache = (btree, heap) if cache == 1 else None

    def at(self, address):
        return self.base + address

    # -- object headers ----------------------------------------------------------------------------------------------
    def messages(self, address):
        p = self.at(address)
        version, count, _, size = struct.unpack_from("<BxHII", self.buf, p)
        if version != 1:
            raise Hdf5Unsupported(f"object header version {version}" if self.buf[p:p + 4] != b"OHDR" else "object header version 2")
        blocks, out = [(p + 16, size)], []
        while blocks and len(out) < count:
            q, left = blocks.pop(0)
            while left >= 8 and len(out) < count:
                kind, n, flags = struct.unpack_from("<HHB", self.buf, q)
                data = self.buf[q + 8:q + 8 + n]
                if kind == 0x0010:                                  # continuation: (address, length) of the next block
                    a, ln = struct.unpack_from("<QQ", data)
                    blocks.append((self.at(a), ln))
                if flags & 0x02:
                    raise Hdf5Unsupported("shared header messages")
                out.append((kind, data))
                q += 8 + n
                left -= 8 + n
        return out

    # -- groups ------------------------------------------------------------------------------------------------------
    def heap_name(self, heap, offset):
        p = self.at(heap)
        if self.buf[p:p + 4] != b"HEAP":
            raise ValueError("bad local heap signature")
        data = self.at(struct.unpack_from("<Q", self.buf, p + 24)[0])
        end = self.buf.index(b"\0", data + offset)
        return self.buf[data + offset:end].decode()

    def group_entries(self, btree, heap):
        p = self.at(btree)
        sig = self.buf[p:p + 4]
        if sig == b"SNOD":
            n = struct.unpack_from("<H", self.buf, p + 6)[0]
            for i in range(n):
                name_off, header, _, _ = struct.unpack_from("<QQII", self.buf, p + 8 + 40 * i)
                yield self.heap_name(heap, name_off), header
            return
        if sig != b"TREE":
            raise ValueError("bad group B-tree signature")
        node_type, _, used = struct.unpack_from("<BBH", self.buf, p + 4)
        if node_type != 0:
            raise ValueError("not a group B-tree")
        for i in range(used):
            child = struct.unpack_from("<Q", self.buf, p + 24 + 8 + 16 * i)[0]
            yield from self.group_entries(child, heap)

    def root_entries(self):
        for kind, data in self.messages(self.root_header):
            if kind == 0x0011:
                return list(self.group_entries(*struct.unpack_from("<QQ", data)))
        if self.root_cache:
            return list(self.group_entries(*self.root_cache))
        raise Hdf5Unsupported("root group without a symbol table (new-style group)")

    # -- datasets ----------------------------------------------------------------------------------------------------
    def dataset(self, header):
        shape = dtype = data = None
        for kind, d in self.messages(header):
            if kind == 0x0001:
                version, rank, flags = d[0], d[1], d[2]
                start = 8 if version == 1 else 4
                if version == 2 and d[3] == 2:
                    raise Hdf5Unsupported("null dataspace")
                shape = struct.unpack_from(f"<{rank}Q", d, start)
            elif kind == 0x0003:
                dtype = _decode_datatype(d)
            elif kind == 0x0008:
                version = d[0]
                if version in (1, 2):
                    rank, cls = d[1], d[2]
                    if cls == 1:
                        data = ("contiguous", struct.unpack_from("<Q", d, 8)[0], None)
                    elif cls == 0:
                        n = struct.unpack_from("<I", d, 8 + 4 * rank)[0]
                        data = ("compact", d[12 + 4 * rank:12 + 4 * rank + n], None)
                    else:
                        raise Hdf5Unsupported("chunked dataset")
                elif version == 3:
                    cls = d[1]
                    if cls == 1:
                        a, n = struct.unpack_from("<QQ", d, 2)
                        data = ("contiguous", a, n)
                    elif cls == 0:
                        n = struct.unpack_from("<H", d, 2)[0]
                        data = ("compact", d[4:4 + n], None)
                    else:
                        raise Hdf5Unsupported("chunked dataset")
                else:
                    raise Hdf5Unsupported(f"data layout version {version}")
            elif kind == 0x000B:
                raise Hdf5Unsupported("filtered (compressed) dataset")
        if shape is None or dtype is None or data is None:
            return None                                             # a group or a named datatype, not a dataset
        count = int(np.prod(shape, dtype=np.int64)) if shape else 1
        nbytes = count * dtype.itemsize
        if data[0] == "compact":
            raw = bytes(data[1][:nbytes])
        elif data[1] == UNDEF or nbytes == 0:
            raw = b"\0" * nbytes                                    # never written: the (default, zero) fill value
        else:
            raw = bytes(self.buf[self.at(data[1]):self.at(data[1]) + nbytes])
            if len(raw) != nbytes:
                raise ValueError("dataset extends past the end of the file")
        return np.frombuffer(raw, dtype=dtype, count=count).reshape(shape).copy()


def _decode_datatype(d):
    cls, version = d[0] & 0x0F, d[0] >> 4
    bits0, size = d[1], struct.unpack_from("<I", d, 4)[0]
    order = ">" if bits0 & 1 else "<"
    if version not in (1, 2, 3):
        raise Hdf5Unsupported(f"datatype version {version}")
    if cls == 0:
        offset, precision = struct.unpack_from("<HH", d, 8)
        if offset or precision != 8 * size:
            raise Hdf5Unsupported("integer type with padding bits")
        return np.dtype(f"{order}{'i' if bits0 & 0x08 else 'u'}{size}")
    if cls == 1:
        offset, precision, exp_at, exp_bits, man_at, man_bits, bias = struct.unpack_from("<HHBBBBI", d, 8)
        if (size, exp_bits, man_bits, bias) not in ((2, 5, 10, 15), (4, 8, 23, 127), (8, 11, 52, 1023)) or offset or man_at:
            raise Hdf5Unsupported("non-IEEE floating-point type")
        return np.dtype(f"{order}f{size}")
    raise Hdf5Unsupported(f"datatype class {cls} (only fixed- and floating-point)")


def read_hdf5(path, names=None):
    """{name: array} of the root-level datasets (all, or `names`).  Raises KeyError for a requested name that is absent,
    like `h5py.File(path)[name]`."""
    with open(path, "rb") as f:
        hf = _File(f.read())
    entries = dict(hf.root_entries())
    out = {}
    for name in (entries if names is None else names):
        if name not in entries:
            raise KeyError(f"Unable to open object (object '{name}' doesn't exist)")
        arr = hf.dataset(entries[name])
        if arr is not None:
            out[name] = arr.astype(arr.dtype.newbyteorder("="), copy=False)
        elif names is not None:
            raise Hdf5Unsupported(f"'{name}' is not a dataset")
    return out
