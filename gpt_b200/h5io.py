"""HDF5 files in the layout the reference's scripts write and read with HDF5.jl
(`kin40kExperiment.jl:93-107`: `write(file,"testRMSE",testRMSE)`, `write(file,"w_store",...)`, `h5read(file,"U_store")`;
`MakeSynthData.jl:17-27`): root-level datasets of Float64 / Int32 / Int64 arrays.

h5py is not part of this image, so the subset those files use is written (and read) directly:
superblock version 0 with 8-byte offsets, an old-style root group (symbol table: v1 B-tree, local heap, SNOD leaves
of 2*4 entries), version-1 object headers carrying dataspace (v1, with maximum dimensions), datatype (v1 little-endian
IEEE float / two's-complement integer), fill value (v2) and contiguous layout (v3) messages — byte for byte the
messages found in the reference's own `testRMSE_PP.h5`.

HDF5 lists dimensions slowest-first; HDF5.jl writes a column-major Julia array by reversing its dimension list.
`write_h5` takes arrays in JULIA orientation (shape (a,b,c), element [i,j,k] as Julia indexes it) and `read_h5`
returns them that way, so `U_store` written here is `h5read(file,"U_store")` there.
"""
from __future__ import annotations

import struct

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF
_LEAF_K, _INT_K = 4, 16


def _pad8(b: bytes) -> bytes:
    return b + b"\x00" * (-len(b) % 8)


def _msg(mtype: int, body: bytes, flags: int = 0) -> bytes:
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _datatype(dt: np.dtype) -> bytes:
    if dt == np.float64:   # class 1 v1, IEEE little-endian, implied mantissa bit, sign at 63 (bytes of testRMSE_PP.h5)
        return bytes.fromhex("11203f00") + struct.pack("<I", 8) + bytes.fromhex("00004000340b0034ff030000")
    if dt == np.float32:
        return bytes.fromhex("11201f00") + struct.pack("<I", 4) + bytes.fromhex("00002000170800177f000000")
    if dt.kind in "iu" and dt.itemsize in (1, 2, 4, 8):
        bits0 = 0x08 if dt.kind == "i" else 0x00
        return bytes([0x10, bits0, 0, 0]) + struct.pack("<I", dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    raise TypeError(f"unsupported dtype {dt}")


def _object_header(messages: bytes, nmsg: int) -> bytes:
    return struct.pack("<BxHII4x", 1, nmsg, 1, len(messages)) + messages


def write_h5(path: str, datasets: dict) -> None:
    """Writes {name: array} as root-level contiguous datasets.  Arrays are taken in Julia orientation (see module
    docstring); float64 / float32 / integer dtypes.  Names are plain (no '/'); at most 256 datasets."""
    names = sorted(datasets, key=lambda s: s.encode())
    if any("/" in nm or not nm for nm in names):
        raise ValueError("dataset names must be non-empty and must not contain '/'")
    nsnod = max(1, -(-len(names) // (2 * _LEAF_K)))
    if nsnod > 2 * _INT_K:
        raise ValueError("too many datasets for one B-tree node")
    arrays = {nm: np.asarray(datasets[nm]) for nm in names}
    for nm, a in arrays.items():
        if a.dtype == np.bool_:
            arrays[nm] = a.astype(np.int8)
        elif a.dtype.kind == "f" and a.dtype not in (np.float64, np.float32):
            arrays[nm] = a.astype(np.float64)
        elif a.dtype.byteorder == ">":
            arrays[nm] = a.astype(a.dtype.newbyteorder("<"))

    # ---- local heap data: "" at offset 0, then the names (null-terminated, 8-byte aligned), then one free block
    heap_data = b"\x00" * 8
    name_off = {}
    for nm in names:
        name_off[nm] = len(heap_data)
        heap_data += _pad8(nm.encode() + b"\x00")
    free_off = len(heap_data)
    heap_data += struct.pack("<QQ", 1, 32) + b"\x00" * 16            # free block: next = 1 (none), size 32

    # ---- addresses
    pos = 96                                                           # superblock (56) + root symbol-table entry (40)
    root_hdr = pos; pos += 16 + 24
    btree = pos; pos += 24 + (2 * _INT_K + 1) * 8 + 2 * _INT_K * 8
    heap = pos; pos += 32
    heap_seg = pos; pos += len(heap_data)
    snods = []
    for _ in range(nsnod):
        snods.append(pos); pos += 8 + 2 * _LEAF_K * 40
    hdr_addr, data_addr, hdr_bytes = {}, {}, {}
    for nm in names:                                                    # object headers first, raw data after them
        a = arrays[nm]
        dims = list(a.shape)[::-1]
        space = struct.pack("<BBB5x", 1, len(dims), 1) + b"".join(struct.pack("<Q", d) for d in dims) * 2
        hdr_bytes[nm] = (space, _datatype(a.dtype))
        hdr_addr[nm] = pos
        pos += 16 + len(_msg(1, space)) + len(_msg(3, hdr_bytes[nm][1])) + len(_msg(5, b"\x02\x02\x02\x01\x00\x00\x00\x00")) + \
            len(_msg(8, b"\x00" * 18))
    for nm in names:
        pos += -pos % 8
        data_addr[nm] = pos
        pos += arrays[nm].size * arrays[nm].dtype.itemsize
    eof = pos

    out = bytearray(eof)
    # ---- superblock v0 and the root group's symbol-table entry (cache type 1: B-tree and heap addresses)
    out[0:8] = _SIG
    out[8:16] = bytes([0, 0, 0, 0, 0, 8, 8, 0])
    out[16:24] = struct.pack("<HHI", _LEAF_K, _INT_K, 0)
    out[24:56] = struct.pack("<QQQQ", 0, _UNDEF, eof, _UNDEF)
    out[56:96] = struct.pack("<QQII", 0, root_hdr, 1, 0) + struct.pack("<QQ", btree, heap)
    out[root_hdr:root_hdr + 40] = _object_header(_msg(0x11, struct.pack("<QQ", btree, heap)), 1)
    # ---- B-tree node (group node, level 0): key[0] = "", child i = SNOD i, key[i+1] = largest name in child i
    node = bytearray(b"TREE" + struct.pack("<BBH", 0, 0, nsnod) + struct.pack("<QQ", _UNDEF, _UNDEF))
    node += struct.pack("<Q", 0)
    for i in range(nsnod):
        chunk = names[i * 2 * _LEAF_K:(i + 1) * 2 * _LEAF_K]
        node += struct.pack("<QQ", snods[i], name_off[chunk[-1]] if chunk else 0)
    out[btree:btree + len(node)] = node
    # ---- local heap
    out[heap:heap + 32] = b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free_off, heap_seg)
    out[heap_seg:heap_seg + len(heap_data)] = heap_data
    # ---- symbol-table nodes
    for i in range(nsnod):
        chunk = names[i * 2 * _LEAF_K:(i + 1) * 2 * _LEAF_K]
        sn = bytearray(b"SNOD" + struct.pack("<BxH", 1, len(chunk)))
        for nm in chunk:
            sn += struct.pack("<QQII16x", name_off[nm], hdr_addr[nm], 0, 0)
        out[snods[i]:snods[i] + len(sn)] = sn
    # ---- datasets
    for nm in names:
        a = arrays[nm]
        raw = a.ravel(order="F").tobytes()                              # first Julia index fastest
        space, dtype = hdr_bytes[nm]
        msgs = (_msg(1, space) + _msg(3, dtype, flags=1) + _msg(5, b"\x02\x02\x02\x01\x00\x00\x00\x00", flags=1) +
                _msg(8, struct.pack("<BBQQ", 3, 1, data_addr[nm] if raw else _UNDEF, len(raw))))
        h = _object_header(msgs, 4)
        out[hdr_addr[nm]:hdr_addr[nm] + len(h)] = h
        out[data_addr[nm]:data_addr[nm] + len(raw)] = raw
    with open(path, "wb") as fh:
        fh.write(out)


def read_h5(path: str) -> dict:
    """Reads every dataset of a file in the subset described above (the reference's data and result files: PPdata.h5,
    10000SynthData.h5, TensorSynthData*.h5, testRMSE_*.h5, and anything write_h5 wrote), groups flattened to
    "group/name".  Arrays come back in Julia orientation, Fortran-ordered."""
    with open(path, "rb") as fh:
        buf = fh.read()
    if buf[:8] != _SIG or buf[8] != 0 or (buf[13], buf[14]) != (8, 8):
        raise ValueError(f"{path}: not an HDF5 file with a version-0 superblock and 8-byte offsets")

    def u(off, n):
        return int.from_bytes(buf[off:off + n], "little")

    def messages(addr):
        if buf[addr] != 1:
            raise ValueError("object header version %d unsupported" % buf[addr])
        nmsg, blocks, out = u(addr + 2, 2), [(addr + 16, u(addr + 8, 4))], []
        while blocks and len(out) < nmsg:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end and len(out) < nmsg:
                mtype, msize = u(p, 2), u(p + 2, 2)
                if mtype == 0x10:
                    blocks.append((u(p + 8, 8), u(p + 16, 8)))
                out.append((mtype, p + 8, msize))
                p += 8 + msize
        return out

    def leaves(node):
        if buf[node:node + 4] != b"TREE":
            raise ValueError("bad B-tree node")
        level, nent, p = buf[node + 5], u(node + 6, 2), node + 24
        for _ in range(nent):
            child = u(p + 8, 8)
            p += 16
            if level == 0:
                yield child
            else:
                yield from leaves(child)

    found = {}

    def group(btree, heap, prefix):
        seg = u(heap + 24, 8)
        for snod in leaves(btree):
            for i in range(u(snod + 6, 2)):
                e = snod + 8 + 40 * i
                a = seg + u(e, 8)
                name = buf[a:buf.index(b"\x00", a)].decode()
                full = f"{prefix}/{name}" if prefix else name
                if u(e + 16, 4) == 1:
                    group(u(e + 24, 8), u(e + 32, 8), full)
                    continue
                msgs = messages(u(e + 8, 8))
                stab = [m for m in msgs if m[0] == 0x11]
                if stab:
                    group(u(stab[0][1], 8), u(stab[0][1] + 8, 8), full)
                elif any(m[0] == 0x08 for m in msgs):
                    found[full] = msgs

    group(u(56 + 24, 8), u(56 + 32, 8), "")
    out = {}
    for name, msgs in found.items():
        dims = dtype = raw = None
        for mtype, body, _ in msgs:
            if mtype == 0x01:
                p = body + (8 if buf[body] == 1 else 4)
                dims = [u(p + 8 * i, 8) for i in range(buf[body + 1])]
            elif mtype == 0x03:
                cls, size = buf[body] & 0x0F, u(body + 4, 4)
                if buf[body + 1] & 1 or cls not in (0, 1):
                    raise ValueError(f"{name}: unsupported datatype")
                dtype = np.dtype(f"<f{size}") if cls == 1 else np.dtype(f"<{'i' if buf[body + 1] & 8 else 'u'}{size}")
            elif mtype == 0x08:
                if buf[body] != 3 or buf[body + 1] not in (0, 1):
                    raise ValueError(f"{name}: unsupported storage layout")
                if buf[body + 1] == 1:
                    a, s = u(body + 2, 8), u(body + 10, 8)
                    raw = b"" if a == _UNDEF else buf[a:a + s]
                else:
                    raw = buf[body + 4:body + 4 + u(body + 2, 2)]
        if dims is None or dtype is None or raw is None:
            raise ValueError(f"{name}: incomplete dataset header")
        arr = np.frombuffer(raw, dtype=dtype, count=int(np.prod(dims)) if dims else 1)
        out[name] = arr.reshape(dims[::-1], order="F").copy(order="F") if dims else arr.reshape(())
    return out
