"""Minimal HDF5 writer / reader for Keras weight files (`model.weights.h5` inside a `.keras` archive) — no h5py / libhdf5 here.

Scope: what `h5py.File(mode="w")` (libver "earliest", the default Keras uses) produces for a tree of groups holding small
n-dimensional little-endian float32 / float64 / int32 / int64 datasets: superblock version 0, version-1 object headers, "old
style" groups (symbol-table message -> version-1 B-tree of symbol-table nodes + local heap), contiguous (or compact) data
layout.  Written from the HDF5 File Format Specification (version 1.1 structures); UNPINNED: no HDF5 library and not a single
HDF5 file exists on the build machine, so the only checks are writer -> reader round trips and structural assertions against the
specification's field tables (tests/test_keras_io.py).  Chunked / filtered datasets, new-style (link-message) groups and
variable-length attributes are rejected with a clear error instead of being guessed at.
"""
from __future__ import annotations

import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"
LEAF_K, INTERNAL_K = 4, 16              # symbol-table node holds <= 2 * LEAF_K entries, a B-tree node <= 2 * INTERNAL_K children
_DTYPES = {np.dtype("<f4"): (1, 4), np.dtype("<f8"): (1, 8), np.dtype("<i4"): (0, 4), np.dtype("<i8"): (0, 8)}


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


# ------------------------------------------------------------------------------------------------ writer
def _datatype_message(dt: np.dtype) -> bytes:
    cls, size = _DTYPES[dt]
    if cls == 1:                        # floating point, IEEE little-endian: implied-msb mantissa normalisation (bits 4-5 = 2)
        sign = size * 8 - 1
        exp_size, man_size = (8, 23) if size == 4 else (11, 52)
        head = struct.pack("<BBBBI", 0x11, 0x20, sign, 0, size)
        prop = struct.pack("<HHBBBBI", 0, size * 8, man_size, exp_size, 0, man_size, (1 << (exp_size - 1)) - 1)
    else:                               # fixed point, little-endian, signed (bit 3)
        head = struct.pack("<BBBBI", 0x10, 0x08, 0, 0, size)
        prop = struct.pack("<HH", 0, size * 8)
    return head + prop


def _message(mtype: int, data: bytes) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), 0) + data


def _object_header(messages: list[bytes]) -> bytes:
    body = b"".join(messages)
    return struct.pack("<BxHII4x", 1, len(messages), 1, len(body)) + body


class _Node:
    def __init__(self):
        self.children: dict[str, "_Node"] = {}
        self.data: np.ndarray | None = None


def write_h5(path: str, datasets: dict[str, np.ndarray], groups: list[str] = ()) -> None:
    """datasets: {"a/b/vars/0": array}; groups: paths of (possibly empty) groups that must exist."""
    root = _Node()

    def node_at(p: str) -> _Node:
        n = root
        for part in [q for q in p.split("/") if q]:
            n = n.children.setdefault(part, _Node())
        return n
    for g in groups:
        node_at(g)
    for p, a in datasets.items():
        a = np.asarray(a)                            # (ascontiguousarray would turn a scalar into shape (1,))
        if not a.flags.c_contiguous:
            a = a.copy(order="C")
        if a.dtype.newbyteorder("<") not in _DTYPES and a.dtype not in _DTYPES:
            raise TypeError(f"{p}: dtype {a.dtype} not supported")
        node_at(p).data = a.astype(a.dtype.newbyteorder("<"), copy=False)

    out = bytearray(b"\0" * 96)                      # superblock, filled in at the end

    def alloc(b: bytes) -> int:
        while len(out) % 8:
            out.append(0)
        addr = len(out)
        out.extend(b)
        return addr

    def emit(n: _Node) -> int:
        """writes the object (children first) and returns the address of its object header"""
        if n.data is not None:
            a = n.data
            raw = a.tobytes()
            daddr = alloc(raw) if raw else UNDEF
            space = struct.pack("<BBB5x", 1, a.ndim, 0) + b"".join(struct.pack("<Q", d) for d in a.shape)
            layout = struct.pack("<BBQQ", 3, 1, daddr, len(raw))
            return alloc(_object_header([_message(0x0001, space), _message(0x0003, _datatype_message(a.dtype)), _message(0x0008, layout)]))
        names = sorted(n.children)                   # strcmp order == byte order for ASCII names
        addrs = {k: emit(n.children[k]) for k in names}
        heap = bytearray(b"\0" * 8)                  # offset 0: the empty string
        offs = {}
        for k in names:
            offs[k] = len(heap)
            heap.extend(_pad8(k.encode() + b"\0"))
        heap_data = alloc(bytes(heap))
        heap_addr = alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), 1, heap_data))      # free-list head 1 = H5HL_FREE_NULL: no free block
        snods, keys = [], [0]
        for i in range(0, len(names), 2 * LEAF_K):
            chunk = names[i:i + 2 * LEAF_K]
            ent = b"".join(struct.pack("<QQII16x", offs[k], addrs[k], 0, 0) for k in chunk)
            ent += b"\0" * (40 * (2 * LEAF_K - len(chunk)))
            snods.append(alloc(b"SNOD" + struct.pack("<BxH", 1, len(chunk)) + ent))
            keys.append(offs[chunk[-1]])
        if len(snods) > 2 * INTERNAL_K:
            raise ValueError("more than 256 links in one group")
        tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(snods), UNDEF, UNDEF)
        kc = b""
        for i in range(2 * INTERNAL_K):
            kc += struct.pack("<Q", keys[i] if i < len(keys) else 0)
            kc += struct.pack("<Q", snods[i] if i < len(snods) else 0)
        kc += struct.pack("<Q", keys[2 * INTERNAL_K] if len(keys) > 2 * INTERNAL_K else 0)
        n.btree = alloc(tree + kc)
        n.heap = heap_addr
        return alloc(_object_header([_message(0x0011, struct.pack("<QQ", n.btree, heap_addr))]))

    root_addr = emit(root)
    eof = len(out) + (-len(out) % 8)
    out.extend(b"\0" * (eof - len(out)))
    sb = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, LEAF_K, INTERNAL_K, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    sb += struct.pack("<QQII", 0, root_addr, 1, 0) + struct.pack("<QQ", root.btree, root.heap)     # root entry: cached B-tree / heap addresses
    assert len(sb) == 96
    out[:96] = sb
    with open(path, "wb") as f:
        f.write(bytes(out))


# ------------------------------------------------------------------------------------------------ reader
class H5Error(ValueError):
    pass


def read_h5(path: str) -> dict[str, np.ndarray]:
    """{"a/b/vars/0": array} of every dataset in the file (groups without datasets are listed with value None)."""
    buf = open(path, "rb").read()
    base = buf.find(SIGNATURE)
    if base != 0:
        raise H5Error("not an HDF5 file (or a user block precedes the superblock)")
    ver = buf[8]
    if ver not in (0, 1):
        raise H5Error(f"superblock version {ver}: only the classic (libver='earliest') layout is supported")
    if buf[13] != 8 or buf[14] != 8:
        raise H5Error("only 8-byte offsets / lengths are supported")
    p = 24 + (4 if ver == 1 else 0)
    _base_addr, _fs, _eof, _drv = struct.unpack_from("<QQQQ", buf, p)
    _name_off, root_hdr, _cache, _ = struct.unpack_from("<QQII", buf, p + 32)
    out: dict[str, np.ndarray] = {}

    def messages(addr: int):
        if buf[addr:addr + 4] == b"OHDR":
            raise H5Error("version-2 object header (libver='latest' file): not supported")
        version, nmsg, _ref, size = struct.unpack_from("<BxHII", buf, addr)
        if version != 1:
            raise H5Error(f"object header version {version}")
        blocks = [(addr + 16, size)]
        got = []
        while blocks and len(got) < nmsg:
            q, left = blocks.pop(0)
            end = q + left
            while q + 8 <= end and len(got) < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", buf, q)
                data = buf[q + 8:q + 8 + msize]
                if mtype == 0x0010:                                    # continuation: more messages elsewhere
                    blocks.append(struct.unpack_from("<QQ", data, 0))
                got.append((mtype, data))
                q += 8 + msize
        return got

    def heap_name(heap_addr: int, off: int) -> str:
        if buf[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5Error("bad local heap")
        data_addr = struct.unpack_from("<Q", buf, heap_addr + 24)[0]
        e = buf.index(b"\0", data_addr + off)
        return buf[data_addr + off:e].decode()

    def walk_btree(addr: int, heap: int, path: str):
        if buf[addr:addr + 4] == b"SNOD":
            nsym = struct.unpack_from("<H", buf, addr + 6)[0]
            for i in range(nsym):
                noff, ohdr = struct.unpack_from("<QQ", buf, addr + 8 + 40 * i)
                visit(ohdr, f"{path}/{heap_name(heap, noff)}" if path else heap_name(heap, noff))
            return
        if buf[addr:addr + 4] != b"TREE":
            raise H5Error("bad group B-tree node")
        ntype, _level, used = struct.unpack_from("<BBH", buf, addr + 4)
        if ntype != 0:
            raise H5Error("not a group B-tree")
        for i in range(used):
            child = struct.unpack_from("<Q", buf, addr + 24 + 8 + 16 * i)[0]
            walk_btree(child, heap, path)

    def visit(addr: int, path: str):
        msgs = messages(addr)
        kinds = {t for t, _ in msgs}
        if 0x0011 in kinds:                                            # group
            btree, heap = struct.unpack_from("<QQ", dict(msgs)[0x0011], 0)
            if path:
                out.setdefault(path, None)
            walk_btree(btree, heap, path)
            return
        if 0x0002 in kinds or 0x0006 in kinds:
            raise H5Error("new-style group (link messages): not supported")
        if 0x0008 not in kinds:
            return                                                     # e.g. a committed datatype: ignore
        m = dict(msgs)
        sp = m[0x0001]
        sver, rank, sflags = sp[0], sp[1], sp[2]
        doff = 8 if sver == 1 else 4
        shape = struct.unpack_from(f"<{rank}Q", sp, doff) if rank else ()
        dtm = m[0x0003]
        cls, size = dtm[0] & 0x0F, struct.unpack_from("<I", dtm, 4)[0]
        if dtm[1] & 1:
            raise H5Error(f"{path}: big-endian data")
        if cls == 1:
            dt = {4: "<f4", 8: "<f8", 2: "<f2"}.get(size)
        elif cls == 0:
            dt = ("<i" if dtm[1] & 0x08 else "<u") + str(size)
        else:
            raise H5Error(f"{path}: datatype class {cls} not supported")
        lay = m[0x0008]
        if lay[0] != 3:
            raise H5Error(f"{path}: data layout message version {lay[0]}")
        count = int(np.prod(shape)) if rank else 1
        if lay[1] == 1:                                                # contiguous
            daddr, dsize = struct.unpack_from("<QQ", lay, 2)
            raw = b"" if daddr == UNDEF else buf[daddr:daddr + dsize]
        elif lay[1] == 0:                                              # compact
            dsize = struct.unpack_from("<H", lay, 2)[0]
            raw = lay[4:4 + dsize]
        else:
            raise H5Error(f"{path}: chunked datasets are not supported")
        if 0x000B in kinds:
            raise H5Error(f"{path}: filtered dataset")
        out[path] = np.frombuffer(raw, dt, count).reshape(shape).copy()

    visit(root_hdr, "")
    return out
