"""TEST INFRASTRUCTURE ONLY: writes small HDF5 files in the classic layout h5py produces with default settings
(superblock 0, symbol-table root group, version-1 object headers, data layout 3 -- contiguous or chunked with a B-tree v1
chunk index -- and optional shuffle + deflate filters), following the HDF5 File Format Specification.  h5py is not in
this image, so the reader in transmil-deepgraft_b200/h5bag.py is exercised on files made here (and on one file
written by the real HDF5 library, see tests/test_h5bag.py).  The chunked form is what the reference's converters write:
`create_dataset(key, shape, maxshape=(None,)+shape[1:], chunks=(1,)+shape[1:])` (code/datasets/utils/move_to_hdf5.py:19).
"""
import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


class _File:
    def __init__(self, userblock=0):
        self.base = userblock
        self.data = bytearray(userblock)

    def alloc(self, blob, align=8):
        """Appends `blob`, returns its address relative to the base address."""
        while (len(self.data) - self.base) % align:
            self.data.append(0)
        addr = len(self.data) - self.base
        self.data += blob
        return addr

    def patch(self, addr, blob):
        a = self.base + addr
        self.data[a:a + len(blob)] = blob


def _msg(mtype, body, flags=0):
    body = body + b"\0" * (-len(body) % 8)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _dataspace(shape, maxshape):
    body = struct.pack("<BBB5x", 1, len(shape), 1 if maxshape is not None else 0)
    body += b"".join(struct.pack("<Q", s) for s in shape)
    if maxshape is not None:
        body += b"".join(struct.pack("<Q", UNDEF if m is None else m) for m in maxshape)
    return _msg(0x1, body)


def _datatype(dt):
    dt = np.dtype(dt)
    big = 1 if dt.byteorder == ">" else 0
    if dt.kind == "f":
        # class 1 version 1; bit field: byte order, padding, mantissa normalisation 2 (implied msb) in bits 4-5, sign location
        exp, man = {2: (5, 10), 4: (8, 23), 8: (11, 52)}[dt.itemsize]
        bits = dt.itemsize * 8
        body = struct.pack("<BBBBI", 0x11, 0x20 | big, bits - 1, 0, dt.itemsize)
        body += struct.pack("<HHBBBBI", 0, bits, man, exp, 0, man, (1 << (exp - 1)) - 1)
    elif dt.kind in "iu":
        body = struct.pack("<BBBBI", 0x10, (8 if dt.kind == "i" else 0) | big, 0, 0, dt.itemsize)
        body += struct.pack("<HH", 0, dt.itemsize * 8)
    else:
        raise TypeError(dt)
    return _msg(0x3, body, flags=1)


def _filters(shuffle, deflate, itemsize):
    items = []
    if shuffle:
        items.append((2, [itemsize]))
    if deflate is not None:
        items.append((1, [deflate]))
    body = struct.pack("<BB6x", 1, len(items))
    for fid, cd in items:
        body += struct.pack("<HHHH", fid, 0, 1, len(cd)) + b"".join(struct.pack("<I", v) for v in cd)
        if len(cd) % 2:
            body += b"\0" * 4
    return _msg(0xB, body), items


def _object_header(f, messages, split=False):
    """Version-1 object header; `split` moves the last message into a continuation block (as the library does when a
    header outgrows its first allocation)."""
    if split and len(messages) > 1:
        tail = f.alloc(messages[-1])
        cont = _msg(0x10, struct.pack("<QQ", tail, len(messages[-1])))
        first, n = messages[:-1] + [cont], len(messages) + 1
    else:
        first, n = messages, len(messages)
    blob = b"".join(first)
    return f.alloc(struct.pack("<BxHII4x", 1, n, 1, len(blob)) + blob)


def _chunk_btree(f, entries, rank, cap):
    """entries: [(offsets tuple, address, stored bytes)] in row-major order -> address of the root node."""
    def key(nbytes, offs):
        return struct.pack("<II", nbytes, 0) + b"".join(struct.pack("<Q", o) for o in offs) + struct.pack("<Q", 0)

    level, nodes = 0, [(e[0], e[1], e[2]) for e in entries]          # (first offsets, child address, bytes)
    while True:
        groups = [nodes[i:i + cap] for i in range(0, len(nodes), cap)] or [[]]
        addrs, size = [], 8 + 16 + cap * (8 + 8 * (rank + 1) + 8) + 8 + 8 * (rank + 1)
        for g in groups:
            addrs.append(f.alloc(b"\0" * size))
        for gi, g in enumerate(groups):
            blob = b"TREE" + struct.pack("<BBH", 1, level, len(g))
            blob += struct.pack("<QQ", addrs[gi - 1] if gi else UNDEF, addrs[gi + 1] if gi + 1 < len(groups) else UNDEF)
            for offs, child, nbytes in g:
                blob += key(nbytes, offs) + struct.pack("<Q", child)
            last = g[-1][0] if g else (0,) * rank
            blob += key(0, tuple(o + 1 for o in last))
            f.patch(addrs[gi], blob)
        if len(groups) == 1:
            return addrs[0]
        nodes = [(g[0][0], addrs[gi], g[0][2]) for gi, g in enumerate(groups)]
        level += 1


def _write_dataset(f, arr, chunks=None, maxshape=None, shuffle=False, deflate=None, btree_cap=64, interleave=False,
                   split_header=False, skip_chunks=(), fillvalue=None):
    arr = np.ascontiguousarray(arr)
    msgs = [_dataspace(arr.shape, maxshape), _datatype(arr.dtype)]
    if fillvalue is not None:                       # fill value message version 2: alloc time 2, write time 0, defined, size, value
        raw = np.asarray(fillvalue, arr.dtype).tobytes()
        msgs.append(_msg(0x5, struct.pack("<BBBBI", 2, 2, 0, 1, len(raw)) + raw))
    if chunks is None:
        addr = f.alloc(arr.tobytes()) if arr.size else UNDEF
        msgs.append(_msg(0x8, struct.pack("<BBQQ", 3, 1, addr, arr.nbytes)))
        return _object_header(f, msgs, split_header)
    fmsg, pipeline = _filters(shuffle, deflate, arr.itemsize) if (shuffle or deflate is not None) else (None, [])
    rank = arr.ndim
    grid = [-(-s // c) for s, c in zip(arr.shape, chunks)]
    entries = []
    for idx in np.ndindex(*grid):
        if idx in skip_chunks:
            continue
        offs = tuple(i * c for i, c in zip(idx, chunks))
        block = np.zeros(chunks, arr.dtype)
        sel = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, chunks, arr.shape))
        block[tuple(slice(0, s.stop - s.start) for s in sel)] = arr[sel]
        raw = block.tobytes()
        for fid, cd in pipeline:
            if fid == 2:
                raw = np.frombuffer(raw, np.uint8).reshape(-1, cd[0]).T.tobytes()
            elif fid == 1:
                raw = zlib.compress(raw, cd[0])
        if interleave and len(entries) % 3 == 2:
            f.alloc(b"\xAA" * 24)                                    # something else between two chunks (a B-tree node, say)
        entries.append((offs, f.alloc(raw, align=1 if not interleave else 8), len(raw)))
    root = _chunk_btree(f, entries, rank, btree_cap) if entries else UNDEF
    layout = struct.pack("<BBBQ", 3, 2, rank + 1, root) + b"".join(struct.pack("<I", c) for c in chunks)
    layout += struct.pack("<I", arr.itemsize)
    msgs.append(_msg(0x8, layout))
    if fmsg is not None:
        msgs.append(fmsg)
    return _object_header(f, msgs, split_header)


def write_h5(path, datasets, userblock=0, **kw):
    """datasets: name -> array, or name -> (array, dict of per-dataset options: chunks, maxshape, shuffle, deflate,
    btree_cap, interleave, split_header, skip_chunks, fillvalue)."""
    f = _File(userblock)
    f.alloc(b"\0" * 96)                                              # superblock, patched at the end
    links = {}
    for name, spec in datasets.items():
        arr, opts = spec if isinstance(spec, tuple) else (spec, {})
        links[name] = _write_dataset(f, arr, **{**kw, **opts})
    # local heap: offset 0 = "" (the B-tree's left-most key), then the names; 8-byte aligned entries
    heap_data, offsets = bytearray(b"\0" * 8), {}
    for name in sorted(links):
        offsets[name] = len(heap_data)
        raw = name.encode() + b"\0"
        heap_data += raw + b"\0" * (-len(raw) % 8)
    seg = f.alloc(bytes(heap_data))
    heap = f.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), UNDEF, seg))
    snod = b"SNOD" + struct.pack("<BxH", 1, len(links))
    for name in sorted(links):
        snod += struct.pack("<QQII16x", offsets[name], links[name], 0, 0)
    snod_addr = f.alloc(snod)
    last = offsets[sorted(links)[-1]] if links else 0
    tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF) + struct.pack("<QQQ", 0, snod_addr, last)
    tree_addr = f.alloc(tree)
    root = _object_header(f, [_msg(0x11, struct.pack("<QQ", tree_addr, heap))])
    eof = len(f.data) - f.base
    sb = SIG + struct.pack("<BBBxBBBxHHI", 0, 0, 0, 0, 8, 8, 4, 16, 0)
    sb += struct.pack("<QQQQ", f.base, UNDEF, eof, UNDEF)
    sb += struct.pack("<QQII", 0, root, 1, 0) + struct.pack("<QQ", tree_addr, heap)
    assert len(sb) == 96
    f.patch(0, sb)
    with open(path, "wb") as out:
        out.write(f.data)
    return path
