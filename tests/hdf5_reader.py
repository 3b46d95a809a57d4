"""Independent reader for the subset of HDF5 that `iterative_learning_nmpc_b200.hdf5_min` writes -- test infrastructure.

It walks the file the way libhdf5 does (superblock -> root symbol table entry -> object header -> Symbol Table message
-> B-tree -> symbol table nodes -> local heap names -> dataset object headers -> raw data) and validates every signature,
version and size on the way, so a layout mistake in the writer fails here instead of producing a plausible array.  It
shares no code with the writer.  It is not libhdf5: a file this accepts has been checked against the published format
specification as restated below, nothing more.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


class Hdf5FormatError(Exception):
    pass


def _need(cond, msg):
    if not cond:
        raise Hdf5FormatError(msg)


def _u(buf, off, n):
    return int.from_bytes(buf[off:off + n], "little")


def _cstr(buf, off):
    end = buf.index(b"\0", off)
    return buf[off:end].decode("ascii")


def _object_messages(buf, addr):
    """Version-1 object header at `addr` -> list of (type, flags, payload bytes)."""
    _need(buf[addr] == 1, f"object header version {buf[addr]} at {addr}")
    nmsg = _u(buf, addr + 2, 2)
    _need(_u(buf, addr + 4, 4) >= 1, "object reference count")
    size = _u(buf, addr + 8, 4)
    pos, end, msgs = addr + 16, addr + 16 + size, []
    _need(end <= len(buf), "object header runs past the end of the file")
    while pos < end and len(msgs) < nmsg:
        mtype, msize, flags = _u(buf, pos, 2), _u(buf, pos + 2, 2), buf[pos + 4]
        _need(msize % 8 == 0, "version-1 header messages are padded to 8 bytes")
        _need(mtype != 0x0010, "continuation messages are not expected here")
        msgs.append((mtype, flags, bytes(buf[pos + 8:pos + 8 + msize])))
        pos += 8 + msize
    _need(len(msgs) == nmsg and pos == end, "message count / header size mismatch")
    return msgs


def _heap(buf, addr):
    _need(buf[addr:addr + 4] == b"HEAP" and buf[addr + 4] == 0, "local heap signature / version")
    size, free_head, data_addr = _u(buf, addr + 8, 8), _u(buf, addr + 16, 8), _u(buf, addr + 24, 8)
    _need(size % 8 == 0, "heap data segment is 8-byte aligned")
    _need(free_head == 1 or free_head + 16 <= size, "free-list head: 1 (libhdf5's end marker) or a block inside the segment")
    _need(data_addr + size <= len(buf), "local heap data segment")
    return bytes(buf[data_addr:data_addr + size])


def _group_entries(buf, btree_addr, heap):
    """Group B-tree (node type 0) -> [(name, object header address)] in tree order."""
    _need(buf[btree_addr:btree_addr + 4] == b"TREE", "B-tree signature")
    _need(buf[btree_addr + 4] == 0, "B-tree node type (group = 0)")
    level, used = buf[btree_addr + 5], _u(buf, btree_addr + 6, 2)
    _need(_u(buf, btree_addr + 8, 8) == UNDEF and _u(buf, btree_addr + 16, 8) == UNDEF, "single node: no siblings")
    out, pos = [], btree_addr + 24
    for i in range(used):
        child = _u(buf, pos + 8 + 16 * i, 8)
        last_key = _u(buf, pos + 16 + 16 * i, 8)
        if level > 0:
            sub = _group_entries(buf, child, heap)
        else:
            _need(buf[child:child + 4] == b"SNOD" and buf[child + 4] == 1, "symbol table node signature / version")
            n = _u(buf, child + 6, 2)
            sub = []
            for e in range(n):
                eo = child + 8 + 40 * e
                sub.append((_cstr(heap, _u(buf, eo, 8)), _u(buf, eo + 8, 8)))
                _need(_u(buf, eo + 16, 4) == 0, "dataset entries carry no cached data")
        _need(sub and _cstr(heap, last_key) == sub[-1][0], "right key of a child = its largest name")
        out += sub
    names = [n for n, _ in out]
    _need(names == sorted(names) and len(set(names)) == len(names), "names sorted and unique")
    return out


def _dataset(buf, addr):
    msgs = {t: (f, p) for t, f, p in _object_messages(buf, addr)}
    for t in (0x0001, 0x0003, 0x0008):
        _need(t in msgs, f"dataset header lacks message {t:#06x}")
    sp = msgs[0x0001][1]
    _need(sp[0] == 1 and sp[2] == 0, "dataspace version 1 without maximum dimensions")
    rank = sp[1]
    shape = tuple(_u(sp, 8 + 8 * i, 8) for i in range(rank))
    dt = msgs[0x0003][1]
    version, cls, size = dt[0] >> 4, dt[0] & 15, _u(dt, 4, 4)
    _need(version == 1 and size == 8, "datatype version 1, 8 bytes")
    if cls == 0:
        _need(dt[1] & 1 == 0, "little-endian")
        _need(dt[1] & 8, "signed")
        _need(_u(dt, 8, 2) == 0 and _u(dt, 10, 2) == 64, "bit offset / precision")
        dtype = np.dtype("<i8")
    elif cls == 1:
        _need(dt[1] & 1 == 0 and dt[1] & 0x40 == 0, "little-endian")
        _need((dt[1] >> 4) & 3 == 2, "mantissa normalisation: msb implied")
        _need(dt[2] == 63, "sign bit position")
        _need(_u(dt, 8, 2) == 0 and _u(dt, 10, 2) == 64, "bit offset / precision")
        _need((dt[12], dt[13], dt[14], dt[15]) == (52, 11, 0, 52), "IEEE binary64 field layout")
        _need(_u(dt, 16, 4) == 1023, "exponent bias")
        dtype = np.dtype("<f8")
    else:
        raise Hdf5FormatError(f"datatype class {cls}")
    if 0x0005 in msgs:
        fv = msgs[0x0005][1]
        _need(fv[0] == 2 and fv[3] == 0, "fill value version 2, undefined")
    lay = msgs[0x0008][1]
    _need(lay[0] == 3 and lay[1] == 1, "data layout version 3, contiguous")
    data_addr, nbytes = _u(lay, 2, 8), _u(lay, 10, 8)
    count = int(np.prod(shape)) if shape else 1
    _need(nbytes == 8 * count, "layout size = element count x 8")
    _need(data_addr != UNDEF and data_addr + nbytes <= len(buf), "raw data inside the file")
    return np.frombuffer(buf, dtype=dtype, count=count, offset=data_addr).reshape(shape).copy()


def read_hdf5(path):
    with open(path, "rb") as f:
        buf = f.read()
    _need(buf[:8] == b"\x89HDF\r\n\x1a\n", "format signature")
    _need(buf[8] == 0, "superblock version 0")
    _need(buf[9] == 0 and buf[10] == 0 and buf[12] == 0, "free-space / root group / shared header versions")
    _need(buf[13] == 8 and buf[14] == 8, "8-byte offsets and lengths")
    leaf_k, internal_k = _u(buf, 16, 2), _u(buf, 18, 2)
    _need(leaf_k > 0 and internal_k > 0, "group K values")
    _need(_u(buf, 24, 8) == 0, "base address")
    _need(_u(buf, 40, 8) == len(buf), "end-of-file address = file size")
    # root symbol table entry at 56: link name offset, object header address, cache type, reserved, scratch
    root_hdr, cache = _u(buf, 64, 8), _u(buf, 72, 4)
    msgs = _object_messages(buf, root_hdr)
    st = [p for t, _, p in msgs if t == 0x0011]
    _need(len(st) == 1, "root group has one Symbol Table message")
    btree, heap_addr = _u(st[0], 0, 8), _u(st[0], 8, 8)
    if cache == 1:
        _need((_u(buf, 80, 8), _u(buf, 88, 8)) == (btree, heap_addr), "cached B-tree / heap addresses agree with the header")
    heap = _heap(buf, heap_addr)
    _need(heap[0] == 0, "heap offset 0 is the empty string")
    entries = _group_entries(buf, btree, heap)
    _need(len(entries) <= 2 * leaf_k, "one symbol table node holds at most 2K entries")
    return {name: _dataset(buf, addr) for name, addr in entries}
