"""Minimal read-only HDF5 reader for the reference's fixture files (TEST INFRASTRUCTURE ONLY).

The reference stores its de-facto golden vectors as HDF5 written by Julia's HDF5.jl
(`MakeSynthData.jl:17-27`, `kin40kExperiment.jl:93-107`).  h5py is not installed in this
image, so this module parses just the subset those files use:

  * superblock version 0, 8-byte offsets/lengths
  * old-style groups (symbol table: v1 B-tree + local heap + SNOD nodes)
  * version-1 object headers (with continuation blocks)
  * simple dataspaces, little-endian IEEE float / fixed-point integer datatypes
  * contiguous (layout class 1) and compact (class 0) storage, layout message v3

HDF5 stores dimensions slowest-first (C order).  Julia writes its column-major arrays by
reversing the dimension list, so a Julia array of size (a,b,c) appears here with HDF5 dims
(c,b,a).  `read()` returns the array in *Julia orientation*: shape (a,b,c), Fortran order.

This file is part of the oracle: it may be imported by tests/ and by the fixture generator
(tests/golden/make_golden.py) only; nothing in the product path reads HDF5.
"""
from __future__ import annotations

import struct
import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5File:
    def __init__(self, path: str):
        with open(path, "rb") as fh:
            self.buf = fh.read()
        if self.buf[:8] != _SIG:
            raise ValueError(f"{path}: not an HDF5 file")
        ver = self.buf[8]
        if ver != 0:
            raise ValueError(f"{path}: superblock version {ver} unsupported")
        size_off, size_len = self.buf[13], self.buf[14]
        if (size_off, size_len) != (8, 8):
            raise ValueError("only 8-byte offsets/lengths supported")
        # superblock v0: 8 sig, 8 version bytes, 2+2 group K, 4 flags, 4x8 addresses, root entry
        root_entry = 24 + 32
        self.datasets = {}
        self._walk_entry(root_entry, "")

    # -- low-level helpers -------------------------------------------------
    def _u(self, off: int, n: int) -> int:
        return int.from_bytes(self.buf[off:off + n], "little")

    def _walk_entry(self, off: int, prefix: str):
        """Symbol-table entry: name offset(8) header addr(8) cache type(4) reserved(4) scratch(16)."""
        hdr = self._u(off + 8, 8)
        cache = self._u(off + 16, 4)
        if cache == 1:
            btree, heap = self._u(off + 24, 8), self._u(off + 32, 8)
            self._walk_group(btree, heap, prefix)
        else:
            msgs = self._object_header(hdr)
            stab = [m for m in msgs if m[0] == 0x11]
            if stab:
                body = stab[0][1]
                self._walk_group(self._u(body, 8), self._u(body + 8, 8), prefix)

    def _heap_data(self, heap: int) -> int:
        assert self.buf[heap:heap + 4] == b"HEAP"
        return self._u(heap + 8 + 16, 8)

    def _walk_group(self, btree: int, heap: int, prefix: str):
        data = self._heap_data(heap)
        for snod in self._btree_leaves(btree):
            assert self.buf[snod:snod + 4] == b"SNOD"
            nsym = self._u(snod + 6, 2)
            for i in range(nsym):
                e = snod + 8 + 40 * i
                name_off = self._u(e, 8)
                end = self.buf.index(b"\x00", data + name_off)
                name = self.buf[data + name_off:end].decode()
                hdr = self._u(e + 8, 8)
                cache = self._u(e + 16, 4)
                full = f"{prefix}/{name}" if prefix else name
                if cache == 1:
                    self._walk_group(self._u(e + 24, 8), self._u(e + 32, 8), full)
                    continue
                msgs = self._object_header(hdr)
                types = {m[0] for m in msgs}
                if 0x11 in types:
                    body = [m for m in msgs if m[0] == 0x11][0][1]
                    self._walk_group(self._u(body, 8), self._u(body + 8, 8), full)
                elif 0x08 in types:
                    self.datasets[full] = msgs

    def _btree_leaves(self, node: int):
        assert self.buf[node:node + 4] == b"TREE", "bad B-tree node"
        level = self.buf[node + 5]
        nent = self._u(node + 6, 2)
        # header: sig4 type1 level1 entries2 left8 right8, then key0 child0 key1 ...
        p = node + 24
        for i in range(nent):
            child = self._u(p + 8, 8)
            p += 16
            if level == 0:
                yield child
            else:
                yield from self._btree_leaves(child)

    def _object_header(self, addr: int):
        ver = self.buf[addr]
        if ver != 1:
            raise ValueError(f"object header version {ver} unsupported")
        nmsg = self._u(addr + 2, 2)
        size = self._u(addr + 8, 4)
        blocks = [(addr + 16, size)]
        msgs = []
        while blocks and len(msgs) < nmsg:
            p, remaining = blocks.pop(0)
            end = p + remaining
            while p + 8 <= end and len(msgs) < nmsg:
                mtype = self._u(p, 2)
                msize = self._u(p + 2, 2)
                body = p + 8
                if mtype == 0x10:  # continuation
                    blocks.append((self._u(body, 8), self._u(body + 8, 8)))
                msgs.append((mtype, body, msize))
                p = body + msize
        return msgs

    # -- dataset decoding ---------------------------------------------------
    def keys(self):
        return sorted(self.datasets)

    def read(self, name: str) -> np.ndarray:
        msgs = self.datasets[name]
        dims = dtype = None
        raw = None
        for mtype, body, msize in msgs:
            if mtype == 0x01:
                ver, rank, flags = self.buf[body], self.buf[body + 1], self.buf[body + 2]
                p = body + (8 if ver == 1 else 4)
                dims = [self._u(p + 8 * i, 8) for i in range(rank)]
            elif mtype == 0x03:
                cls = self.buf[body] & 0x0F
                bits0 = self.buf[body + 1]
                size = self._u(body + 4, 4)
                if bits0 & 1:
                    raise ValueError("big-endian data unsupported")
                if cls == 1:
                    dtype = np.dtype(f"<f{size}")
                elif cls == 0:
                    signed = bool(bits0 & 0x08)
                    dtype = np.dtype(f"<{'i' if signed else 'u'}{size}")
                else:
                    raise ValueError(f"datatype class {cls} unsupported")
            elif mtype == 0x08:
                ver = self.buf[body]
                if ver != 3:
                    raise ValueError(f"layout message version {ver} unsupported")
                lclass = self.buf[body + 1]
                if lclass == 1:
                    a, s = self._u(body + 2, 8), self._u(body + 10, 8)
                    raw = b"" if a == _UNDEF else self.buf[a:a + s]
                elif lclass == 0:
                    s = self._u(body + 2, 2)
                    raw = self.buf[body + 4:body + 4 + s]
                else:
                    raise ValueError("chunked storage unsupported")
        if dims is None or dtype is None or raw is None:
            raise ValueError(f"{name}: incomplete dataset header")
        arr = np.frombuffer(raw, dtype=dtype, count=int(np.prod(dims)) if dims else 1)
        # HDF5 dims are C-order; Julia orientation is the reversed shape in Fortran order.
        return arr.reshape(dims[::-1], order="F").copy(order="F") if dims else arr.reshape(())


def read_all(path: str) -> dict:
    f = H5File(path)
    return {k: f.read(k) for k in f.keys()}
