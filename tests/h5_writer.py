"""TEST INFRASTRUCTURE: a minimal HDF5 writer, independent of the package's reader, following the HDF5 File Format
Specification (version 3.0): enough to lay out a Keras ``model.weights.h5`` -- nested groups and float datasets.

Two dialects:
  * ``write_v0``: superblock version 0, version-1 object headers, old-style groups (local heap + v1 B-tree + symbol table
    nodes, 8 symbols per node) -- what h5py writes by default; datasets contiguous or chunked (v1 chunk B-tree).
  * ``write_v2``: superblock version 2, version-2 object headers ("OHDR", Jenkins lookup3 checksums), compact groups
    (link messages), contiguous datasets -- what ``libver='latest'`` produces for small groups.

A tree is a dict: {"name": subtree-dict | ndarray}.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


class _Buf:
    def __init__(self, reserve):
        self.b = bytearray(reserve)

    def put(self, data: bytes, align=8) -> int:
        self.b += b"\0" * (-len(self.b) % align)
        addr = len(self.b)
        self.b += data
        return addr


def _dtype_msg(dt: np.dtype) -> bytes:
    size = dt.itemsize
    # class 1 (floating point), version 1; bit field: little-endian, mantissa normalisation 2 (implied msb), sign position
    if size == 4:
        props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
        bits = bytes([0x20, 31, 0])
    else:
        props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
        bits = bytes([0x20, 63, 0])
    return bytes([0x11]) + bits + struct.pack("<I", size) + props


# ======================================================================================================================
# dialect 0: superblock v0, object header v1, symbol-table groups
# ======================================================================================================================
def _msg_v1(mtype, body, flags=0):
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _ohdr_v1(msgs):
    data = b"".join(msgs)
    return struct.pack("<BxHII4x", 1, len(msgs), 1, len(data)) + data


def _dataset_v1(buf, arr, chunks=None):
    arr = np.ascontiguousarray(arr)
    dt = arr.dtype.newbyteorder("<")
    rank = arr.ndim
    space = struct.pack("<BBBx4x", 1, rank, 0) + b"".join(struct.pack("<Q", s) for s in arr.shape)
    fill = struct.pack("<BBBB", 2, 2, 2, 0)                                # fill value v2: alloc late, never written, undefined
    if chunks is None:
        raw = buf.put(arr.astype(dt).tobytes())
        layout = struct.pack("<BBQQ", 3, 1, raw, arr.nbytes)
    else:
        keys = []
        grid = [range(0, s, c) for s, c in zip(arr.shape, chunks)]
        for offs in np.ndindex(*[len(g) for g in grid]):
            o = [g[i] for g, i in zip(grid, offs)]
            block = np.zeros(chunks, dtype=dt)
            sl = tuple(slice(a, min(a + c, s)) for a, c, s in zip(o, chunks, arr.shape))
            block[tuple(slice(0, s.stop - s.start) for s in sl)] = arr[sl]
            keys.append((o, buf.put(block.tobytes()), block.nbytes))
        node = b"TREE" + struct.pack("<BBHQQ", 1, 0, len(keys), UNDEF, UNDEF)
        for o, addr, nbytes in keys:
            node += struct.pack("<II", nbytes, 0) + b"".join(struct.pack("<Q", v) for v in o) + struct.pack("<Q", 0)
            node += struct.pack("<Q", addr)
        node += struct.pack("<II", 0, 0) + b"".join(struct.pack("<Q", s) for s in arr.shape) + struct.pack("<Q", 0)   # final key
        bt = buf.put(node)
        layout = struct.pack("<BBBQ", 3, 2, rank + 1, bt) + b"".join(struct.pack("<I", c) for c in chunks) + struct.pack("<I", dt.itemsize)
    return buf.put(_ohdr_v1([_msg_v1(1, space), _msg_v1(3, _dtype_msg(dt), 1), _msg_v1(5, fill), _msg_v1(8, layout)]))


def _group_v1(buf, tree, chunks):
    entries = []
    for name in sorted(tree, key=lambda s: s.encode()):
        child = tree[name]
        addr = _group_v1(buf, child, chunks) if isinstance(child, dict) else _dataset_v1(buf, child, chunks)
        entries.append((name, addr))
    # local heap: offset 0 holds the empty string, names follow, 8-aligned
    heap_data, offs = bytearray(b"\0" * 8), {}
    for name, _ in entries:
        offs[name] = len(heap_data)
        heap_data += _pad8(name.encode() + b"\0")
    free_off = len(heap_data)
    heap_data += struct.pack("<QQ", 1, 16)                                  # one free block: next = 1 (none), size 16
    seg = buf.put(bytes(heap_data))
    heap = buf.put(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free_off, seg))
    # symbol table nodes of at most 8 symbols, one level-0 B-tree node over them
    snods, keys = [], [0]
    for i in range(0, max(len(entries), 1), 8):
        part = entries[i:i + 8]
        node = b"SNOD" + struct.pack("<BxH", 1, len(part))
        for name, addr in part:
            node += struct.pack("<QQII16x", offs[name], addr, 0, 0)
        node += b"\0" * (40 * (8 - len(part)))
        snods.append(buf.put(node))
        keys.append(offs[part[-1][0]] if part else 0)
    bt = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(snods), UNDEF, UNDEF)
    for k, s in zip(keys, snods):
        bt += struct.pack("<QQ", k, s)
    bt += struct.pack("<Q", keys[-1])
    bt += b"\0" * (16 * (32 - len(snods)))                                   # room for 2K = 32 children
    btree = buf.put(bt)
    return buf.put(_ohdr_v1([_msg_v1(0x11, struct.pack("<QQ", btree, heap))]))


def write_v0(tree, chunks=None) -> bytes:
    buf = _Buf(96)                                                          # superblock: 24 + 32 + 40 bytes
    root = _group_v1(buf, tree, chunks)
    sb = SIG + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", 4, 16, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, len(buf.b), UNDEF)
    sb += struct.pack("<QQII16x", 0, root, 0, 0)                             # root group symbol table entry
    buf.b[:len(sb)] = sb
    return bytes(buf.b)


# ======================================================================================================================
# dialect 2: superblock v2, object header v2, compact (link message) groups
# ======================================================================================================================
def _rot(x, k):
    return ((x << k) | (x >> (32 - k))) & 0xFFFFFFFF


def lookup3(data: bytes, init=0) -> int:
    """Bob Jenkins' lookup3 ``hashlittle`` -- the metadata checksum of HDF5 (H5_checksum_lookup3)."""
    M = 0xFFFFFFFF
    a = b = c = (0xDEADBEEF + len(data) + init) & M
    p, n = 0, len(data)
    while n > 12:
        a = (a + int.from_bytes(data[p:p + 4], "little")) & M
        b = (b + int.from_bytes(data[p + 4:p + 8], "little")) & M
        c = (c + int.from_bytes(data[p + 8:p + 12], "little")) & M
        a = (a - c) & M; a ^= _rot(c, 4); c = (c + b) & M
        b = (b - a) & M; b ^= _rot(a, 6); a = (a + c) & M
        c = (c - b) & M; c ^= _rot(b, 8); b = (b + a) & M
        a = (a - c) & M; a ^= _rot(c, 16); c = (c + b) & M
        b = (b - a) & M; b ^= _rot(a, 19); a = (a + c) & M
        c = (c - b) & M; c ^= _rot(b, 4); b = (b + a) & M
        p += 12; n -= 12
    tail = data[p:] + b"\0" * 12
    if n == 0:
        return c
    a = (a + int.from_bytes(tail[0:4], "little")) & M
    b = (b + int.from_bytes(tail[4:8], "little")) & M
    c = (c + int.from_bytes(tail[8:12], "little")) & M
    c ^= b; c = (c - _rot(b, 14)) & M
    a ^= c; a = (a - _rot(c, 11)) & M
    b ^= a; b = (b - _rot(a, 25)) & M
    c ^= b; c = (c - _rot(b, 16)) & M
    a ^= c; a = (a - _rot(c, 4)) & M
    b ^= a; b = (b - _rot(a, 14)) & M
    c ^= b; c = (c - _rot(b, 24)) & M
    return c


def _msg_v2(mtype, body, flags=0):
    return struct.pack("<BHB", mtype, len(body), flags) + body


def _ohdr_v2(msgs):
    data = b"".join(msgs)
    head = b"OHDR" + bytes([2, 0x02]) + struct.pack("<I", len(data)) + data      # flags: 4-byte chunk-0 size field
    return head + struct.pack("<I", lookup3(head))


def _dataset_v2(buf, arr):
    arr = np.ascontiguousarray(arr)
    dt = arr.dtype.newbyteorder("<")
    raw = buf.put(arr.astype(dt).tobytes())
    space = struct.pack("<BBBB", 2, arr.ndim, 0, 1) + b"".join(struct.pack("<Q", s) for s in arr.shape)
    layout = struct.pack("<BBQQ", 3, 1, raw, arr.nbytes)
    fill = struct.pack("<BB", 3, 0x09)                                       # fill value v3: alloc late, never written
    return buf.put(_ohdr_v2([_msg_v2(1, space), _msg_v2(3, _dtype_msg(dt), 1), _msg_v2(5, fill), _msg_v2(8, layout)]))


def _group_v2(buf, tree):
    msgs = [_msg_v2(2, struct.pack("<BBQQ", 0, 0, UNDEF, UNDEF)),           # link info: no dense storage
            _msg_v2(0x0A, struct.pack("<BBHH", 0, 0x01, max(8, len(tree)), 6))]   # group info: all links stay compact
    for name in tree:
        child = tree[name]
        addr = _group_v2(buf, child) if isinstance(child, dict) else _dataset_v2(buf, child)
        nb = name.encode()
        msgs.append(_msg_v2(6, struct.pack("<BBB", 1, 0x00, len(nb)) + nb + struct.pack("<Q", addr)))
    return buf.put(_ohdr_v2(msgs))


def write_v2(tree) -> bytes:
    buf = _Buf(48)
    root = _group_v2(buf, tree)
    sb = SIG + bytes([2, 8, 8, 0]) + struct.pack("<QQQQ", 0, UNDEF, len(buf.b), root)
    sb += struct.pack("<I", lookup3(sb))
    buf.b[:len(sb)] = sb
    return bytes(buf.b)
