"""Feature-bag files of the reference, read without h5py (the data format on the input side of the path).

The reference stores one bag per slide as an HDF5 file with two datasets, `features [n, D]` and `coords [n, 2]`, and
reads them whole: `hdf5_file['features'][:]`, `hdf5_file['coords'][:]` (code/datasets/feature_dataloader.py:252-254),
then `torch.from_numpy` (:273).  The files are written by h5py with default settings as chunked, resizable datasets
(`create_dataset(key, shape, maxshape=(None,)+shape[1:], chunks=(1,)+shape[1:])`, code/datasets/utils/move_to_hdf5.py:19),
i.e. the classic on-disk layout of the HDF5 library: superblock version 0/1, symbol-table groups (B-tree v1 "TREE" +
"SNOD" nodes + local "HEAP"), version-1 object headers, data layout message version 3 (compact / contiguous / chunked
with a B-tree v1 chunk index), optional deflate / shuffle / fletcher32 filters.  This module reads exactly that subset,
straight into a caller-owned (pinned) host buffer so that the bag goes file -> pinned memory -> device with no
intermediate copy, and raises H5FormatError on anything else (superblock v2/v3, version-2 object headers, dense
groups, layout v4 chunk indexes, non-numeric datatypes).  h5py / libhdf5 are not in this image; the format is restated
from the published HDF5 File Format Specification v1/v2.  Pinning: tests/test_h5bag.py parses a file written by the
real HDF5 library that ships with scipy's test data (superblock, group walk, object headers, datatypes, contiguous
layout) and files written by tests/h5_writer.py (chunked / filtered layouts), which follows the same specification.
"""
import mmap
import os
import zlib

import numpy as np

__all__ = ["H5FormatError", "FeatureFile", "Dataset", "read_bag", "read_bag_rows", "iter_bags"]

SIGNATURE = b"\x89HDF\r\n\x1a\n"
MSG_DATASPACE, MSG_DATATYPE, MSG_FILL_OLD, MSG_FILL, MSG_LAYOUT, MSG_FILTERS = 0x1, 0x3, 0x4, 0x5, 0x8, 0xB
MSG_CONTINUATION, MSG_SYMBOL_TABLE = 0x10, 0x11
FILTER_DEFLATE, FILTER_SHUFFLE, FILTER_FLETCHER32 = 1, 2, 3


class H5FormatError(ValueError):
    """The file is not HDF5, is damaged, or uses a part of the format this reader does not implement."""


class _Buf:
    """Little-endian cursor over the mapped file; every address is relative to the superblock's base address."""

    def __init__(self, mm, base=0, osz=8, lsz=8):
        self.mm, self.base, self.osz, self.lsz = mm, base, osz, lsz
        self.undef = (1 << (8 * osz)) - 1

    def bytes(self, addr, n):
        a = self.base + addr
        if addr < 0 or a + n > len(self.mm):
            raise H5FormatError(f"address {addr:#x}+{n} is outside the file ({len(self.mm)} bytes)")
        return self.mm[a:a + n]

    def array(self, addr, dtype, count):
        """Zero-copy numpy view of file bytes (bulk data); the caller drops it before the file is closed."""
        dtype = np.dtype(dtype)
        a = self.base + addr
        if addr < 0 or a + count * dtype.itemsize > len(self.mm):
            raise H5FormatError(f"address {addr:#x}+{count * dtype.itemsize} is outside the file ({len(self.mm)} bytes)")
        return np.frombuffer(self.mm, dtype, count, a)

    def uint(self, addr, n):
        return int.from_bytes(self.bytes(addr, n), "little")

    def offset(self, addr):
        return self.uint(addr, self.osz)

    def length(self, addr):
        return self.uint(addr, self.lsz)

    def expect(self, addr, tag):
        if self.bytes(addr, 4) != tag:
            raise H5FormatError(f"expected {tag!r} at {addr:#x}, found {bytes(self.bytes(addr, 4))!r}")


def _find_superblock(mm):
    """The signature sits at byte 0 or at 512, 1024, 2048, ... (a user block precedes it)."""
    pos = 0
    while pos + 8 <= len(mm):
        if mm[pos:pos + 8] == SIGNATURE:
            return pos
        pos = 512 if pos == 0 else pos * 2
    raise H5FormatError("not an HDF5 file (no signature)")


def _read_superblock(mm):
    """-> (_Buf, root object header address, root B-tree address or None, root heap address or None)."""
    sb = _find_superblock(mm)
    ver = mm[sb + 8]
    if ver not in (0, 1):
        raise H5FormatError(f"superblock version {ver} is not supported (files written with libver='latest'); "
                            "the reference writes h5py defaults = version 0")
    osz, lsz = mm[sb + 13], mm[sb + 14]
    if osz not in (2, 4, 8) or lsz not in (2, 4, 8):
        raise H5FormatError("bad size-of-offsets / size-of-lengths")
    p = sb + 24 + (4 if ver == 1 else 0)          # v1 adds indexed-storage internal node K (2) + reserved (2)
    rd = lambda q: int.from_bytes(mm[q:q + osz], "little")
    base = rd(p)
    p += 4 * osz                                   # base, free-space info, end of file, driver info
    # root group symbol table entry: link name offset, object header address, cache type, reserved, scratch[16]
    hdr = rd(p + osz)
    cache = int.from_bytes(mm[p + 2 * osz:p + 2 * osz + 4], "little")
    btree = heap = None
    if cache == 1:
        btree, heap = rd(p + 2 * osz + 8), rd(p + 2 * osz + 8 + osz)
    # every other address in the file is relative to `base` (= the size of the user block, 0 without one)
    return _Buf(mm, base, osz, lsz), hdr, btree, heap


def _messages(buf, addr):
    """All header messages of the version-1 object header at `addr`: list of (type, flags, payload address, size)."""
    if buf.bytes(addr, 4) == b"OHDR":
        raise H5FormatError("version-2 object headers are not supported (file written with libver='latest')")
    ver = buf.uint(addr, 1)
    if ver != 1:
        raise H5FormatError(f"object header version {ver} at {addr:#x}")
    nmsg = buf.uint(addr + 2, 2)
    blocks = [(addr + 16, buf.uint(addr + 8, 4))]   # 12-byte prefix padded to 8-byte alignment
    out = []
    while blocks and len(out) < nmsg:
        p, size = blocks.pop(0)
        end = p + size
        while p + 8 <= end and len(out) < nmsg:
            mtype, msize, flags = buf.uint(p, 2), buf.uint(p + 2, 2), buf.uint(p + 4, 1)
            body = p + 8
            if mtype == MSG_CONTINUATION:
                blocks.append((buf.offset(body), buf.length(body + buf.osz)))
            out.append((mtype, flags, body, msize))
            p = body + msize
    return out


def _heap_name(buf, heap_addr, off):
    buf.expect(heap_addr, b"HEAP")
    data = buf.offset(heap_addr + 8 + 2 * buf.lsz)
    seg = buf.length(heap_addr + 8)
    raw = buf.bytes(data + off, min(seg - off, 1024))
    end = bytes(raw).find(b"\0")
    return bytes(raw[:end if end >= 0 else len(raw)]).decode("utf-8")


def _group_entries(buf, btree, heap):
    """name -> object header address of every link of a symbol-table group."""
    out = {}

    def node(addr):
        buf.expect(addr, b"TREE")
        ntype, level, used = buf.uint(addr + 4, 1), buf.uint(addr + 5, 1), buf.uint(addr + 6, 2)
        if ntype != 0:
            raise H5FormatError("group B-tree has the wrong node type")
        p = addr + 8 + 2 * buf.osz
        for i in range(used):                       # key_i (L) child_i (O) ... key_used
            child = buf.offset(p + buf.lsz + i * (buf.lsz + buf.osz))
            if level > 0:
                node(child)
            else:
                snod(child)

    def snod(addr):
        buf.expect(addr, b"SNOD")
        n = buf.uint(addr + 6, 2)
        esz = 2 * buf.osz + 24
        for i in range(n):
            e = addr + 8 + i * esz
            out[_heap_name(buf, heap, buf.offset(e))] = buf.offset(e + buf.osz)

    node(btree)
    return out


def _parse_datatype(buf, p):
    cv, b0, size = buf.uint(p, 1), buf.uint(p + 1, 1), buf.uint(p + 4, 4)
    cls = cv & 0xF
    order = ">" if (b0 & 1) else "<"
    if cls == 0 and size in (1, 2, 4, 8):            # fixed point: bit 3 = signed
        return np.dtype(f"{order}{'i' if b0 & 8 else 'u'}{size}")
    if cls == 1 and size in (2, 4, 8):               # IEEE floating point
        return np.dtype(f"{order}f{size}")
    raise H5FormatError(f"datatype class {cls} of size {size} is not a numeric array type")


def _parse_dataspace(buf, p):
    ver, rank, flags = buf.uint(p, 1), buf.uint(p + 1, 1), buf.uint(p + 2, 1)
    if ver == 1:
        q = p + 8
    elif ver == 2:
        q = p + 4
    else:
        raise H5FormatError(f"dataspace message version {ver}")
    shape = tuple(buf.length(q + i * buf.lsz) for i in range(rank))
    maxshape = None
    if flags & 1:
        q += rank * buf.lsz
        m = [buf.length(q + i * buf.lsz) for i in range(rank)]
        maxshape = tuple(None if v == (1 << (8 * buf.lsz)) - 1 else v for v in m)
    return shape, maxshape


def _parse_filters(buf, p):
    """-> list of (filter id, client data) in pipeline (write) order."""
    ver, n = buf.uint(p, 1), buf.uint(p + 1, 1)
    q = p + (8 if ver == 1 else 2)
    out = []
    for _ in range(n):
        fid = buf.uint(q, 2)
        if ver == 1 or fid >= 256:
            namelen = buf.uint(q + 2, 2)
            q += 2
        else:
            namelen = 0
        ncd = buf.uint(q + 4, 2)
        q += 6
        q += (namelen + 7) // 8 * 8 if ver == 1 else namelen
        cd = [buf.uint(q + 4 * i, 4) for i in range(ncd)]
        q += 4 * ncd
        if ver == 1 and ncd % 2:
            q += 4
        out.append((fid, cd))
    return out


def _unfilter(raw, filters, mask, itemsize):
    """Undo the pipeline on one chunk: last applied filter first; bit i of `mask` = filter i was skipped."""
    data = bytes(raw)
    for i in reversed(range(len(filters))):
        if mask >> i & 1:
            continue
        fid, cd = filters[i]
        if fid == FILTER_DEFLATE:
            data = zlib.decompress(data)
        elif fid == FILTER_SHUFFLE:
            es = cd[0] if cd else itemsize
            n = len(data) // es
            body = np.frombuffer(data, np.uint8, n * es).reshape(es, n).T.tobytes()
            data = body + data[n * es:]
        elif fid == FILTER_FLETCHER32:
            data = data[:-4]
        else:
            raise H5FormatError(f"filter {fid} is not supported (deflate, shuffle and fletcher32 are)")
    return data


class Dataset:
    """One numeric array of the file.  `ds[:]` / `ds[...]` / `ds[a:b]` return numpy arrays (what the reference reads,
    feature_dataloader.py:253-254); `read_direct(out)` fills a caller-owned C-contiguous array (a pinned tensor's
    `.numpy()`), converting the element type on the way if `out.dtype` differs."""

    def __init__(self, buf, name, addr):
        self._buf, self.name = buf, name
        self.shape = self.dtype = self.maxshape = None
        self.chunks, self._filters, self._layout = None, [], None
        self.fillvalue, fills = 0, []
        for mtype, _flags, p, size in _messages(buf, addr):
            if mtype == MSG_DATASPACE:
                self.shape, self.maxshape = _parse_dataspace(buf, p)
            elif mtype == MSG_DATATYPE:
                self.dtype = _parse_datatype(buf, p)
            elif mtype == MSG_FILTERS:
                self._filters = _parse_filters(buf, p)
            elif mtype == MSG_LAYOUT:
                self._parse_layout(p, size)
            elif mtype in (MSG_FILL, MSG_FILL_OLD):
                fills.append((mtype, p, size))
        if self.shape is None or self.dtype is None or self._layout is None:
            raise H5FormatError(f"'{name}' is not a dataset (dataspace / datatype / layout message missing)")
        if self._layout[0] == "chunked" and self._layout[3] != self.dtype.itemsize:
            raise H5FormatError("chunk element size disagrees with the datatype")
        for mtype, p, size in fills:                # what unallocated chunks / never-written datasets read as (default 0)
            raw = self._fill_bytes(mtype, p, size)
            if raw is not None and len(raw) == self.dtype.itemsize:
                self.fillvalue = np.frombuffer(raw, self.dtype, 1)[0]

    def _fill_bytes(self, mtype, p, size):
        buf = self._buf
        if mtype == MSG_FILL_OLD:                   # size (4), value
            n = buf.uint(p, 4)
            return bytes(buf.bytes(p + 4, n)) if 0 < n <= size - 4 else None
        ver = buf.uint(p, 1)
        if ver in (1, 2):                           # version, alloc time, write time, defined, [size (4), value]
            if ver == 2 and buf.uint(p + 3, 1) == 0:
                return None
            n = buf.uint(p + 4, 4) if size >= 8 else 0
            return bytes(buf.bytes(p + 8, n)) if 0 < n <= size - 8 else None
        if ver == 3:                                # version, flags (bit 5: value defined), [size (4), value]
            if not buf.uint(p + 1, 1) & 0x20:
                return None
            n = buf.uint(p + 2, 4)
            return bytes(buf.bytes(p + 6, n)) if 0 < n <= size - 6 else None
        return None

    def _parse_layout(self, p, size):
        buf = self._buf
        ver, cls = buf.uint(p, 1), buf.uint(p + 1, 1)
        if ver in (1, 2):
            # HDF5 < 1.6.3: version, dimensionality, class, reserved[5], address (not for compact), 4-byte dims
            # (dataset dims for contiguous; chunk dims + element size for chunked, as in version 3)
            ndim, cls = buf.uint(p + 1, 1), buf.uint(p + 2, 1)
            q = p + 8
            if cls == 1:
                dims = [buf.uint(q + buf.osz + 4 * i, 4) for i in range(ndim)]
                self._layout = ("contiguous", buf.offset(q), int(np.prod(dims, dtype=np.int64)))
            elif cls == 2:
                dims = [buf.uint(q + buf.osz + 4 * i, 4) for i in range(ndim)]
                self.chunks = tuple(dims[:-1])
                self._layout = ("chunked", buf.offset(q), ndim, dims[-1])
            else:
                n = buf.uint(q + 4 * ndim, 4)
                self._layout = ("compact", q + 4 * ndim + 4, n)
            return
        if ver not in (3, 4):
            raise H5FormatError(f"data layout message version {ver} is not supported")
        if cls == 0:
            n = buf.uint(p + 2, 2)
            self._layout = ("compact", p + 4, n)
        elif cls == 1:
            self._layout = ("contiguous", buf.offset(p + 2), buf.length(p + 2 + buf.osz))
        elif cls == 2 and ver == 3:
            ndim = buf.uint(p + 2, 1)               # rank + 1: the last entry is the element size
            btree = buf.offset(p + 3)
            dims = [buf.uint(p + 3 + buf.osz + 4 * i, 4) for i in range(ndim)]
            self.chunks = tuple(dims[:-1])
            self._layout = ("chunked", btree, ndim, dims[-1])
        else:
            raise H5FormatError(f"data layout class {cls} of message version {ver} is not supported "
                                "(version-4 chunk indexes are written only with libver='latest')")

    # ---- chunk index ------------------------------------------------------------------------------------------------
    def _chunk_list(self):
        """Every allocated chunk, in B-tree (= row-major offset) order:
        (offsets int64 [m, rank], file address int64 [m], stored bytes [m], filter mask [m])."""
        buf = self._buf
        _, root, ndim, _ = self._layout
        ksz = 8 + 8 * ndim
        leaf = np.dtype([("nbytes", "<u4"), ("mask", "<u4"), ("offs", "<u8", (ndim,)), ("child", f"<u{buf.osz}")])
        assert leaf.itemsize == ksz + buf.osz
        parts = []

        def node(addr):
            buf.expect(addr, b"TREE")
            ntype, level, used = buf.uint(addr + 4, 1), buf.uint(addr + 5, 1), buf.uint(addr + 6, 2)
            if ntype != 1:
                raise H5FormatError("chunk B-tree has the wrong node type")
            rec = buf.array(addr + 8 + 2 * buf.osz, leaf, used)        # key_i child_i ... (the closing key is not needed)
            if level > 0:
                for child in rec["child"].tolist():
                    node(child)
            else:
                parts.append(rec.copy())

        if root != buf.undef:
            node(root)
        rec = np.concatenate(parts) if parts else np.zeros(0, leaf)
        return (rec["offs"][:, :-1].astype(np.int64), rec["child"].astype(np.int64), rec["nbytes"].astype(np.int64),
                rec["mask"].astype(np.int64))

    # ---- reads ------------------------------------------------------------------------------------------------------
    def read_direct(self, out):
        """The whole dataset into a caller-owned array."""
        return self.read_rows(0, self.shape[0] if self.shape else 0, out)

    def read_rows(self, r0, r1, out=None):
        """Rows [r0, r1) of the leading axis into `out` (shape (r1-r0,) + shape[1:]; allocated when None).  Only the chunks
        that hold those rows are touched: a rank of a token-sharded bag reads its own rows of a 1.6 GB file, not the file."""
        if not self.shape:
            raise H5FormatError("scalar datasets are not supported")
        n = self.shape[0]
        if not (0 <= r0 <= r1 <= n):
            raise IndexError(f"rows [{r0}, {r1}) outside a dataset of {n} rows")
        want = (r1 - r0,) + tuple(self.shape[1:])
        if out is None:
            out = np.empty(want, self.dtype.newbyteorder("="))
        if not isinstance(out, np.ndarray) or not out.flags.c_contiguous or not out.flags.writeable:
            raise ValueError("read_direct / read_rows need a writeable C-contiguous numpy array")
        if tuple(out.shape) != want:
            raise ValueError(f"destination has shape {tuple(out.shape)}, rows [{r0}, {r1}) of dataset '{self.name}' have {want}")
        buf, kind = self._buf, self._layout[0]
        per_row = int(np.prod(self.shape[1:], dtype=np.int64))
        if out.size == 0:
            return out
        if kind in ("compact", "contiguous"):
            addr = self._layout[1]
            if kind == "contiguous" and addr == buf.undef:      # never written: fill value (0)
                out[...] = self.fillvalue
                return out
            src = buf.array(addr + r0 * per_row * self.dtype.itemsize, self.dtype, (r1 - r0) * per_row)
            np.copyto(out, src.reshape(want), casting="unsafe")
            return out
        chunks, isz = self.chunks, self.dtype.itemsize
        celems = int(np.prod(chunks, dtype=np.int64))
        cbytes = celems * isz
        offs, addrs, sizes, masks = self._chunk_list()
        keep = (offs < np.asarray(self.shape, np.int64)).all(axis=1)      # drop chunks left behind by a shrink
        keep &= (offs[:, 0] < r1) & (offs[:, 0] + chunks[0] > r0)         # chunks that hold rows of the range
        offs, addrs, sizes, masks = offs[keep], addrs[keep], sizes[keep], masks[keep]
        grid = [-(-s // c) for s, c in zip(self.shape, chunks)]
        grid[0] = (r1 - 1) // chunks[0] - r0 // chunks[0] + 1
        if len(addrs) < int(np.prod(grid, dtype=np.int64)):
            out[...] = self.fillvalue                            # unallocated chunks read as the fill value (default 0)
        if len(addrs) == 0:
            return out
        row_chunks = all(c == s for c, s in zip(chunks[1:], self.shape[1:]))
        aligned = r0 % chunks[0] == 0 and (r1 % chunks[0] == 0 or r1 == n)
        if row_chunks and aligned and not self._filters:
            # chunks span whole rows (the reference's (1, D) chunks): consecutive chunks that are also adjacent in the file
            # (h5py appends them in order) are one contiguous run of the output -> one copy per run, not per chunk
            flat = out.reshape(r1 - r0, -1)
            rows = offs[:, 0]
            brk = np.flatnonzero((np.diff(addrs) != cbytes) | (np.diff(rows) != chunks[0])) + 1
            starts = np.concatenate(([0], brk)).tolist()
            ends = np.concatenate((brk, [len(addrs)])).tolist()
            for a, b in zip(starts, ends):
                c0 = int(rows[a])
                c1 = min(c0 + (b - a) * chunks[0], r1)
                src = buf.array(int(addrs[a]), self.dtype, (b - a) * celems)[:(c1 - c0) * per_row]
                np.copyto(flat[c0 - r0:c1 - r0], src.reshape(c1 - c0, per_row), casting="unsafe")
            return out
        lo = (r0,) + (0,) * (len(self.shape) - 1)
        hi = (r1,) + tuple(self.shape[1:])
        for i in range(len(addrs)):
            o = offs[i].tolist()
            if self._filters:
                data = _unfilter(buf.bytes(int(addrs[i]), int(sizes[i])), self._filters, int(masks[i]), isz)
                block = np.frombuffer(data, self.dtype, celems).reshape(chunks)
            else:
                block = buf.array(int(addrs[i]), self.dtype, celems).reshape(chunks)
            beg = [max(a, l) for a, l in zip(o, lo)]
            end = [min(a + c, h) for a, c, h in zip(o, chunks, hi)]
            dst = tuple(slice(b - l, e - l) for b, e, l in zip(beg, end, lo))
            cut = tuple(slice(b - a, e - a) for b, e, a in zip(beg, end, o))
            np.copyto(out[dst], block[cut], casting="unsafe")
        return out

    def __getitem__(self, idx):
        if isinstance(idx, slice) and idx.step in (None, 1) and self.shape:
            r0, r1, _ = idx.indices(self.shape[0])
            return self.read_rows(r0, max(r0, r1))
        full = np.empty(self.shape, self.dtype.newbyteorder("="))
        self.read_direct(full)
        if idx is Ellipsis or idx == ():
            return full
        return full[idx]

    def __len__(self):
        return self.shape[0]

    def __repr__(self):
        return f'<h5bag dataset "{self.name}": shape {self.shape}, type "{self.dtype.str}", chunks {self.chunks}>'


class FeatureFile:
    """`with FeatureFile(path) as f: f['features'][:]` -- the subset of `h5py.File(path, 'r')` the reference's loaders use
    (feature_dataloader.py:252-254): `keys()`, `in`, `f[name]` (also 'group/name'), context manager."""

    def __init__(self, path, mode="r"):
        if mode != "r":
            raise ValueError("FeatureFile is read-only")
        self.filename = os.fspath(path)
        self._fh = open(self.filename, "rb")
        try:
            if os.fstat(self._fh.fileno()).st_size < 8:
                raise H5FormatError("not an HDF5 file (too short)")
            self._mm = mmap.mmap(self._fh.fileno(), 0, access=mmap.ACCESS_READ)
            self._buf, self._root_hdr, bt, hp = _read_superblock(self._mm)
            self._root = self._links(self._root_hdr, bt, hp)
        except Exception:
            self.close()
            raise

    def _links(self, hdr, btree=None, heap=None):
        if btree is None:
            for mtype, _f, p, _s in _messages(self._buf, hdr):
                if mtype == MSG_SYMBOL_TABLE:
                    btree, heap = self._buf.offset(p), self._buf.offset(p + self._buf.osz)
                    break
            else:
                raise H5FormatError("object is not a symbol-table group (link-message groups need libver='latest' support)")
        return _group_entries(self._buf, btree, heap)

    def keys(self):
        return list(self._root)

    def __contains__(self, name):
        return name in self._root

    def __iter__(self):
        return iter(self._root)

    def __getitem__(self, name):
        links, parts = self._root, [p for p in name.split("/") if p]
        for i, part in enumerate(parts):
            if part not in links:
                raise KeyError(f"Unable to open object (object '{part}' doesn't exist)")     # h5py's KeyError wording
            addr = links[part]
            if i + 1 < len(parts):
                links = self._links(addr)
        return Dataset(self._buf, name, addr)

    def close(self):
        mm, self._mm = getattr(self, "_mm", None), None
        self._buf = None
        if mm is not None:
            mm.close()
        if self._fh is not None:
            self._fh.close()
            self._fh = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def read_bag(path, pin=None, dtype=None, features="features", coords="coords", out=None):
    """One slide: (wsi_bag [n, D], batch_coords [n, 2] or None) as torch tensors, the pair FeatureBagLoader.get_data builds
    (feature_dataloader.py:252-254,273).  The feature rows are read from the file directly into the returned host tensor --
    pinned when `pin` (default: when CUDA is available), so `StreamedInference` / `.cuda(non_blocking=True)` copies it
    asynchronously.  `dtype` None keeps the stored element type (float32 feature files stay float32, float16 files stay
    float16 and are consumed by fc1 as halves); a torch dtype converts while reading.  `out`: a caller-owned contiguous
    host tensor with at least n*D elements (a reused pinned staging buffer: a first-touch allocation costs more than the
    read itself -- 164 MB read in 28 ms into a warm buffer, 120-160 ms into a fresh one); the bag is returned as a view of it."""
    import torch
    if pin is None:
        pin = torch.cuda.is_available()
    with FeatureFile(path) as f:
        ds = f[features]
        want = dtype if dtype is not None else torch.from_numpy(np.empty(0, ds.dtype.newbyteorder("="))).dtype
        if out is not None:
            count = int(np.prod(ds.shape, dtype=np.int64))
            if out.is_cuda or not out.is_contiguous() or out.numel() < count:
                raise ValueError(f"`out` must be a contiguous host tensor with at least {count} elements")
            bag = out.view(-1)[:count].view(ds.shape)
        else:
            bag = torch.empty(ds.shape, dtype=want, pin_memory=bool(pin))
        ds.read_direct(bag.numpy())
        xy = torch.from_numpy(f[coords][:]) if coords in f else None
    return bag, xy


def read_bag_rows(path, ranges, pin=None, dtype=None, features="features"):
    """Concatenation of the row ranges [(a, b), ...] of one bag's `features`, read range by range into one (pinned) host
    tensor -- what a rank of the token-sharded path holds of a giant bag (`ShardPlan.take`: its own rows plus the duplicated
    head rows it recomputes), without reading the rest of the file."""
    import torch
    if pin is None:
        pin = torch.cuda.is_available()
    ranges = [(int(a), int(b)) for a, b in ranges]
    with FeatureFile(path) as f:
        ds = f[features]
        want = dtype if dtype is not None else torch.from_numpy(np.empty(0, ds.dtype.newbyteorder("="))).dtype
        out = torch.empty((sum(b - a for a, b in ranges),) + tuple(ds.shape[1:]), dtype=want, pin_memory=bool(pin))
        pos = 0
        for a, b in ranges:
            ds.read_rows(a, b, out[pos:pos + b - a].numpy())
            pos += b - a
    return out


def iter_bags(paths, pin=None, dtype=None, with_coords=False, workers=0, ahead=2):
    """Bags of a list of feature files, in order -- the iterable `StreamedInference(model)(...)` consumes.

    workers > 0: files are read by that many threads up to `workers + ahead` bags ahead of the consumer (the role of the
    reference's DataLoader workers, code/datasets/data_interface.py); a 164 MB bag takes ~30 ms to read from the page cache
    against ~2 ms of forward, so the slide loop is reader-bound without it.  The big copies release the GIL.  Order is kept."""
    paths = list(paths)
    pick = (lambda r: r) if with_coords else (lambda r: r[0])
    if workers <= 0:
        for path in paths:
            yield pick(read_bag(path, pin=pin, dtype=dtype))
        return
    from collections import deque
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=workers, thread_name_prefix="h5bag") as pool:
        pending, nxt = deque(), 0
        while pending or nxt < len(paths):
            while nxt < len(paths) and len(pending) < workers + max(0, ahead):
                pending.append(pool.submit(read_bag, paths[nxt], pin, dtype))
                nxt += 1
            yield pick(pending.popleft().result())
