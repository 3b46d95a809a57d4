"""Minimal HDF5 writer for the reference's `Database` file (Behavior_Cloning/utils/database.py:156-203): a flat set of
root-level, contiguous, little-endian float64 / int64 datasets.

`h5py` is not part of this image, so `dataset.save_hdf5` falls back to this writer when it cannot import it.  The file is
laid out by hand after the HDF5 File Format Specification, version 1.1 structures only (the ones every libhdf5 since
1.0 reads):

    superblock version 0                  signature, sizes of offsets / lengths = 8, group K values, root symbol-table entry
    root group                            version-1 object header with one Symbol Table message -> B-tree + local heap
    local heap ("HEAP")                   the dataset names, null-terminated, 8-byte aligned
    group B-tree node ("TREE", level 0)   one child: the symbol table node
    symbol table node ("SNOD")            one 40-byte entry per dataset, sorted by name
    per dataset                           version-1 object header: Dataspace v1, Datatype v1 (IEEE F64LE / I64LE),
                                          Fill Value v2 (undefined), Data Layout v3 (contiguous), then the raw data

At most 8 datasets (one symbol table node with the default leaf K = 4); no groups, attributes, chunking or compression.
In this repository the output is checked by `tests/hdf5_reader.py`, an independent reader written from the same
specification; it has NOT been read back by libhdf5 / h5py here (neither is installed).
"""
from __future__ import annotations

import struct
from typing import Dict

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"
GROUP_LEAF_K, GROUP_INTERNAL_K = 4, 16
HEAP_FREE_NULL = 1


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _message(mtype: int, data: bytes, flags: int = 0) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), flags) + data


def _object_header(messages) -> bytes:
    body = b"".join(messages)
    # version 1, reserved, number of messages, reference count 1, header size; the messages start 8-byte aligned
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(body)) + body


def _dataspace(shape) -> bytes:
    return struct.pack("<BBBB4x", 1, len(shape), 0, 0) + b"".join(struct.pack("<Q", int(d)) for d in shape)


def _datatype(dtype: np.dtype) -> bytes:
    if dtype == np.float64:
        # class 1 (floating point), version 1; little-endian, mantissa normalisation "msb implied", sign at bit 63
        return struct.pack("<BBBBI", 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
    if dtype == np.int64:
        # class 0 (fixed point), version 1; little-endian, two's complement
        return struct.pack("<BBBBI", 0x10, 0x08, 0x00, 0x00, 8) + struct.pack("<HH", 0, 64)
    raise TypeError(f"dtype {dtype} is not part of the Database layout")


def _dataset_header(arr: np.ndarray, data_addr: int) -> bytes:
    fill = struct.pack("<BBBB", 2, 1, 0, 0)             # version 2, early allocation, written at allocation, undefined
    layout = struct.pack("<BBQQ", 3, 1, data_addr, arr.nbytes)      # version 3, contiguous
    return _object_header([_message(0x0001, _dataspace(arr.shape)), _message(0x0003, _datatype(arr.dtype), flags=1),
                           _message(0x0005, fill), _message(0x0008, layout)])


def write_hdf5(path: str, datasets: Dict[str, np.ndarray]) -> str:
    """Write `datasets` (name -> float64 / int64 array) as root-level contiguous datasets of one HDF5 file."""
    names = sorted(datasets)                            # the symbol table node is ordered by name (strcmp)
    if not 0 < len(names) <= 2 * GROUP_LEAF_K:
        raise ValueError(f"1..{2 * GROUP_LEAF_K} datasets")
    arrs = {}
    for n in names:
        a = np.asarray(datasets[n])
        a = a.astype(np.int64) if a.dtype.kind in "iub" else a.astype(np.float64)
        arrs[n] = np.ascontiguousarray(a if a.ndim else a.reshape(1)).astype(a.dtype.newbyteorder("<"))
    # local heap data segment: the empty name of the root at offset 0, then the dataset names
    heap_data, name_off = bytearray(_pad8(b"\0")), {}
    for n in names:
        name_off[n] = len(heap_data)
        heap_data += _pad8(n.encode("ascii") + b"\0")
    heap_data = bytes(heap_data)

    # ---- addresses: superblock | root header | B-tree node | heap header | heap data | SNOD | dataset headers | data ----
    root_hdr_len = len(_object_header([_message(0x0011, struct.pack("<QQ", 0, 0))]))
    btree_len = 24 + (2 * GROUP_INTERNAL_K + 1) * 8 + 2 * GROUP_INTERNAL_K * 8
    heap_hdr_len = 32
    snod_len = 8 + 2 * GROUP_LEAF_K * 40
    a_root = 96
    a_btree = a_root + root_hdr_len
    a_heap = a_btree + btree_len
    a_heap_data = a_heap + heap_hdr_len
    a_snod = a_heap_data + len(heap_data)
    addr = a_snod + snod_len
    a_hdr, a_data = {}, {}
    for n in names:
        a_hdr[n] = addr
        addr += len(_dataset_header(arrs[n], 0))
    for n in names:
        addr += -addr % 8
        a_data[n] = addr
        addr += arrs[n].nbytes
    eof = addr

    out = bytearray()
    # superblock, version 0
    out += SIGNATURE + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0)
    out += struct.pack("<HHI", GROUP_LEAF_K, GROUP_INTERNAL_K, 0)
    out += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    out += struct.pack("<QQII", 0, a_root, 1, 0) + struct.pack("<QQ", a_btree, a_heap)     # root symbol table entry
    assert len(out) == a_root
    out += _object_header([_message(0x0011, struct.pack("<QQ", a_btree, a_heap))])
    assert len(out) == a_btree
    # B-tree node of the root group: one child (the symbol table node); key 0 = the empty name, key 1 = the last name
    node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF) + struct.pack("<QQQ", 0, a_snod, name_off[names[-1]])
    out += node + b"\0" * (btree_len - len(node))
    assert len(out) == a_heap
    # free-list head: libhdf5 encodes "no free block" as 1 (H5HL_FREE_NULL) and rejects any other value >= the segment
    # size, the undefined address of the specification's text included
    out += b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), HEAP_FREE_NULL, a_heap_data)
    out += heap_data
    assert len(out) == a_snod
    snod = b"SNOD" + struct.pack("<BBH", 1, 0, len(names))
    for n in names:
        snod += struct.pack("<QQII16x", name_off[n], a_hdr[n], 0, 0)
    out += snod + b"\0" * (snod_len - len(snod))
    for n in names:
        assert len(out) == a_hdr[n]
        out += _dataset_header(arrs[n], a_data[n])
    for n in names:
        out += b"\0" * (a_data[n] - len(out))
        out += arrs[n].tobytes()
    assert len(out) == eof
    with open(path, "wb") as f:
        f.write(out)
    return path
