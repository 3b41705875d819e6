"""Minimal HDF5 reader / writer — just enough for BERT pre-training shards, written against the HDF5 file
format specification because neither h5py nor libhdf5 exist on this image.

Reader: superblock v0/v1 (what ``h5py.File(..., 'w')`` writes by default), old-style groups (symbol-table message ->
v1 B-tree -> SNOD -> local heap), v1 object headers with continuation blocks, fixed-point and IEEE float datatypes,
simple dataspaces, data layout v3 in its three classes: compact, contiguous and chunked (v1 chunk B-tree) with the
deflate (gzip) and shuffle filters — the combination NVIDIA's ``create_pretraining_data.py`` produces with
``compression='gzip'``.  Only the root group's datasets are read.

Writer: the same subset (contiguous, or chunked + deflate when ``compress=True``), so shards written here can be read
back by this module and follow the layout libhdf5 expects.

``read_hdf5(path) -> {name: np.ndarray}``; ``write_hdf5(path, {name: np.ndarray}, compress=False)``.
"""
from __future__ import annotations

import struct
import zlib
from typing import Dict, List, Tuple

import numpy as np

SIG = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


# =====================================================================================================
# reader
# =====================================================================================================
class _Reader:
    def __init__(self, buf: bytes):
        self.b = buf
        if buf[:8] != SIG:
            raise ValueError("not an HDF5 file (bad signature)")
        ver = buf[8]
        if ver not in (0, 1):
            raise ValueError("HDF5 superblock version %d is not supported by hdf5_lite (re-save with libver='earliest' "
                             "or convert the shard to .b2s)" % ver)
        self.so, self.sl = buf[13], buf[14]
        if (self.so, self.sl) != (8, 8):
            raise ValueError("only 8-byte offsets/lengths are supported")
        off = 24 if ver == 0 else 28
        self.base = self.u64(off)
        root_entry = off + 32
        # symbol table entry: name offset, header address, cache type, reserved, scratch
        self.root_header = self.u64(root_entry + 8)
        cache = self.u32(root_entry + 16)
        if cache == 1:
            self.root_btree, self.root_heap = self.u64(root_entry + 24), self.u64(root_entry + 32)
        else:
            self.root_btree = self.root_heap = None

    def u8(self, o): return self.b[o]
    def u16(self, o): return struct.unpack_from("<H", self.b, o)[0]
    def u32(self, o): return struct.unpack_from("<I", self.b, o)[0]
    def u64(self, o): return struct.unpack_from("<Q", self.b, o)[0]

    # ---- object header v1 ---------------------------------------------------------------------------
    def messages(self, addr) -> List[Tuple[int, int, int]]:
        """[(type, data_offset, size)] of an object header, following continuation messages."""
        addr += self.base
        if self.b[addr:addr + 4] == b"OHDR":
            raise ValueError("version-2 object headers are not supported by hdf5_lite")
        if self.u8(addr) != 1:
            raise ValueError("unsupported object header version %d" % self.u8(addr))
        nmsg = self.u16(addr + 2)
        hsize = self.u32(addr + 8)
        blocks = [(addr + 16, hsize)]
        out = []
        while blocks and len(out) < nmsg + 64:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end:
                mtype, msize = self.u16(pos), self.u16(pos + 2)
                data = pos + 8
                if mtype == 0x0010:   # continuation
                    blocks.append((self.base + self.u64(data), self.u64(data + 8)))
                else:
                    out.append((mtype, data, msize))
                pos = data + msize
        return out

    # ---- old-style group ----------------------------------------------------------------------------
    def group_entries(self) -> Dict[str, int]:
        btree, heap = self.root_btree, self.root_heap
        if btree is None:
            for mtype, data, _ in self.messages(self.root_header):
                if mtype == 0x0011:
                    btree, heap = self.u64(data), self.u64(data + 8)
        if btree is None:
            raise ValueError("root group has no symbol table (new-style groups are not supported by hdf5_lite)")
        heap += self.base
        assert self.b[heap:heap + 4] == b"HEAP"
        heap_data = self.base + self.u64(heap + 24)
        entries: Dict[str, int] = {}
        self._walk_group(btree, heap_data, entries)
        return entries

    def _name(self, heap_data, off) -> str:
        p = heap_data + off
        e = self.b.index(b"\0", p)
        return self.b[p:e].decode("utf-8")

    def _walk_group(self, node, heap_data, out):
        node += self.base
        assert self.b[node:node + 4] == b"TREE", "bad B-tree node"
        level, used = self.u8(node + 5), self.u16(node + 6)
        pos = node + 24
        for i in range(used):
            child = self.u64(pos + 8)
            pos += 16
            if level > 0:
                self._walk_group(child, heap_data, out)
            else:
                s = child + self.base
                assert self.b[s:s + 4] == b"SNOD", "bad symbol node"
                n = self.u16(s + 6)
                for j in range(n):
                    e = s + 8 + 40 * j
                    out[self._name(heap_data, self.u64(e))] = self.u64(e + 8)

    # ---- dataset ------------------------------------------------------------------------------------
    def dataset(self, header_addr) -> np.ndarray:
        shape = dtype = layout = None
        filters: List[int] = []
        for mtype, data, msize in self.messages(header_addr):
            if mtype == 0x0001:
                ver, rank, flags = self.u8(data), self.u8(data + 1), self.u8(data + 2)
                o = data + (8 if ver == 1 else 4)
                shape = tuple(self.u64(o + 8 * i) for i in range(rank))
            elif mtype == 0x0003:
                cv = self.u8(data)
                cls, bits0 = cv & 0x0F, self.u8(data + 1)
                size = self.u32(data + 4)
                endian = ">" if bits0 & 1 else "<"
                if cls == 0:
                    dtype = np.dtype("%s%s%d" % (endian, "i" if bits0 & 0x08 else "u", size))
                elif cls == 1:
                    dtype = np.dtype("%sf%d" % (endian, size))
                else:
                    raise ValueError("unsupported HDF5 datatype class %d" % cls)
            elif mtype == 0x0008:
                layout = (data, msize)
            elif mtype == 0x000B:
                ver, nf = self.u8(data), self.u8(data + 1)
                p = data + (8 if ver == 1 else 2)
                for _ in range(nf):
                    fid = self.u16(p)
                    if ver == 1 or fid >= 256:
                        nlen = self.u16(p + 2); ncv = self.u16(p + 6); p += 8 + ((nlen + 7) // 8 * 8 if ver == 1 else nlen)
                    else:
                        ncv = self.u16(p + 4); p += 6
                    p += 4 * ncv
                    if ver == 1 and ncv % 2:
                        p += 4
                    filters.append(fid)
        if shape is None or dtype is None or layout is None:
            raise ValueError("dataset header is missing dataspace / datatype / layout")
        data, _ = layout
        ver = self.u8(data)
        if ver != 3:
            raise ValueError("data layout message version %d is not supported" % ver)
        cls = self.u8(data + 1)
        count = int(np.prod(shape)) if shape else 1
        if cls == 0:
            size = self.u16(data + 2)
            return np.frombuffer(self.b, dtype=dtype, count=count, offset=data + 4).reshape(shape).copy()
        if cls == 1:
            addr = self.u64(data + 2)
            if addr == UNDEF:
                return np.zeros(shape, dtype=dtype)
            return np.frombuffer(self.b, dtype=dtype, count=count, offset=self.base + addr).reshape(shape).copy()
        if cls == 2:
            ndims = self.u8(data + 2)            # rank + 1
            btree = self.u64(data + 3)
            cdims = tuple(self.u32(data + 11 + 4 * i) for i in range(ndims - 1))
            out = np.zeros(shape, dtype=dtype)
            if btree != UNDEF:
                self._walk_chunks(btree, ndims, cdims, dtype, filters, out)
            return out
        raise ValueError("unknown layout class %d" % cls)

    def _walk_chunks(self, node, ndims, cdims, dtype, filters, out):
        node += self.base
        assert self.b[node:node + 4] == b"TREE", "bad chunk B-tree node"
        level, used = self.u8(node + 5), self.u16(node + 6)
        key = 8 + 8 * ndims
        pos = node + 24
        for _ in range(used):
            csize, fmask = self.u32(pos), self.u32(pos + 4)
            offs = tuple(self.u64(pos + 8 + 8 * i) for i in range(ndims - 1))
            child = self.u64(pos + key)
            pos += key + 8
            if level > 0:
                self._walk_chunks(child, ndims, cdims, dtype, filters, out)
                continue
            raw = self.b[self.base + child:self.base + child + csize]
            for i, fid in reversed(list(enumerate(filters))):
                if fmask & (1 << i):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    es = dtype.itemsize
                    n = len(raw) // es
                    raw = np.frombuffer(raw, dtype=np.uint8)[: n * es].reshape(es, n).T.tobytes()
                elif fid == 3:
                    raw = raw[:-4]               # fletcher32 checksum: strip
                else:
                    raise ValueError("unsupported HDF5 filter id %d" % fid)
            chunk = np.frombuffer(raw, dtype=dtype, count=int(np.prod(cdims))).reshape(cdims)
            sl = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, out.shape))
            out[sl] = chunk[tuple(slice(0, s.stop - s.start) for s in sl)]


def read_hdf5(path: str) -> Dict[str, np.ndarray]:
    with open(path, "rb") as f:
        r = _Reader(f.read())
    return {name: r.dataset(addr) for name, addr in r.group_entries().items()}


# =====================================================================================================
# writer
# =====================================================================================================
def _pad8(b: bytes) -> bytes:
    return b + b"\0" * ((-len(b)) % 8)


def _msg(mtype: int, data: bytes) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), 0) + data


def _dtype_msg(dt: np.dtype) -> bytes:
    if dt.kind in "iu":
        bits0 = (0x08 if dt.kind == "i" else 0) | (1 if dt.byteorder == ">" else 0)
        return struct.pack("<B3BI", 0x10 | 0, bits0, 0, 0, dt.itemsize) + struct.pack("<HH", 0, dt.itemsize * 8)
    if dt.kind == "f":
        if dt.itemsize == 4:
            props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
            bits = (0x20, 0x1F, 0)
        else:
            props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
            bits = (0x20, 0x3F, 0)
        return struct.pack("<B3BI", 0x10 | 1, *bits, dt.itemsize) + props
    raise ValueError("hdf5_lite writes integer and float datasets only")


def write_hdf5(path: str, arrays: Dict[str, np.ndarray], compress: bool = False, chunk_rows: int = 1024) -> None:
    names = sorted(arrays.keys())
    LEAF_K, INT_K = 16, 16
    if len(names) > 2 * LEAF_K:
        raise ValueError("hdf5_lite writes at most %d datasets per file" % (2 * LEAF_K))
    out = bytearray(b"\0" * 96)                     # superblock placeholder

    def alloc(data: bytes) -> int:
        while len(out) % 8:
            out.append(0)
        addr = len(out)
        out.extend(data)
        return addr

    # ---- datasets -------------------------------------------------------------------------------------
    headers = {}
    for name in names:
        arr = np.ascontiguousarray(arrays[name])
        dt = arr.dtype.newbyteorder("<") if arr.dtype.byteorder == ">" else arr.dtype
        arr = arr.astype(dt, copy=False)
        rank = arr.ndim
        space = struct.pack("<BBB5x", 1, rank, 0) + b"".join(struct.pack("<Q", d) for d in arr.shape)
        msgs = _msg(0x0001, space) + _msg(0x0003, _dtype_msg(dt))
        if not compress or rank == 0 or arr.size == 0:
            daddr = alloc(arr.tobytes()) if arr.size else UNDEF
            msgs += _msg(0x0008, struct.pack("<BBQQ", 3, 1, daddr, arr.nbytes))
        else:
            cdims = (min(chunk_rows, arr.shape[0]),) + tuple(arr.shape[1:])
            keys = []
            for r0 in range(0, arr.shape[0], cdims[0]):
                chunk = np.zeros(cdims, dtype=dt)
                piece = arr[r0:r0 + cdims[0]]
                chunk[: piece.shape[0]] = piece
                raw = zlib.compress(chunk.tobytes(), 4)
                caddr = alloc(raw)
                keys.append((len(raw), (r0,) + (0,) * (rank - 1), caddr))
            if len(keys) > 2 * INT_K:
                raise ValueError("hdf5_lite writer: too many chunks; raise chunk_rows")
            node = bytearray(b"TREE" + struct.pack("<BBHQQ", 1, 0, len(keys), UNDEF, UNDEF))
            for size, offs, caddr in keys:
                node += struct.pack("<II", size, 0) + b"".join(struct.pack("<Q", o) for o in offs) + struct.pack("<Q", 0)
                node += struct.pack("<Q", caddr)
            node += struct.pack("<II", 0, 0) + struct.pack("<Q", arr.shape[0]) + b"".join(struct.pack("<Q", 0) for _ in range(rank))
            full = 24 + (2 * INT_K + 1) * (8 + 8 * (rank + 1)) + 2 * INT_K * 8
            node += b"\0" * (full - len(node))
            baddr = alloc(bytes(node))
            lay = struct.pack("<BBB", 3, 2, rank + 1) + struct.pack("<Q", baddr)
            lay += b"".join(struct.pack("<I", c) for c in cdims) + struct.pack("<I", dt.itemsize)
            msgs += _msg(0x0008, lay)
            filt = struct.pack("<BB6x", 1, 1) + struct.pack("<HHHH", 1, 8, 1, 1) + b"deflate\0" + struct.pack("<II", 4, 0)
            msgs += _msg(0x000B, filt)
        nmsg = 3 + (1 if compress and rank > 0 and arr.size else 0)
        hdr = struct.pack("<BBHII4x", 1, 0, nmsg, 1, len(msgs)) + msgs
        headers[name] = alloc(hdr)

    # ---- root group: local heap, symbol node, B-tree, object header ------------------------------------------
    heap_data = bytearray(b"\0" * 8)
    name_off = {}
    for name in names:
        name_off[name] = len(heap_data)
        heap_data += _pad8(name.encode("utf-8") + b"\0")
    free_off = len(heap_data)
    heap_data += struct.pack("<QQ", 1, 16)        # one free block: next = H5HL_FREE_NULL, size = 16
    data_addr = alloc(bytes(heap_data))
    heap_addr = alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free_off, data_addr))
    snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(names)))
    for name in names:
        snod += struct.pack("<QQII16x", name_off[name], headers[name], 0, 0)
    snod += b"\0" * (8 + 40 * 2 * LEAF_K - len(snod))
    snod_addr = alloc(bytes(snod))
    tree = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF))
    tree += struct.pack("<QQQ", 0, snod_addr, name_off[names[-1]] if names else 0)
    tree += b"\0" * (24 + (2 * INT_K + 1) * 8 + 2 * INT_K * 8 - len(tree))
    btree_addr = alloc(bytes(tree))
    root_msgs = _msg(0x0011, struct.pack("<QQ", btree_addr, heap_addr))
    root_hdr = alloc(struct.pack("<BBHII4x", 1, 0, 1, 1, len(root_msgs)) + root_msgs)

    eof = len(out)
    sb = SIG + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", LEAF_K, INT_K, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    sb += struct.pack("<QQII", 0, root_hdr, 1, 0) + struct.pack("<QQ", btree_addr, heap_addr)
    assert len(sb) == 96
    out[:96] = sb
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(bytes(out))
    import os
    os.replace(tmp, path)
