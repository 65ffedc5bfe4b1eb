"""A small read-only HDF5 reader — just enough to open the Keras 2.7 model files the reference saves with
`generator.save('*.h5')` (GAN/MTSS_WGAN_GP.py:286; read back at autoencoder_v4.ipynb:1277 with `load_model`).
h5py is not available where this package runs, and all a generator needs from the file is its layer list (the root
attribute `model_config`, a JSON string) and its float32 weight arrays.

Supported, because that is what h5py's default ("earliest") format writes for such files: superblock version 0,
version-1 object headers, groups as symbol tables (B-tree v1 + local heap), contiguous or compact datasets of
little-endian fixed-point / floating-point types, attributes (message versions 1-3) holding numbers, fixed-length
strings or variable-length strings (global heap).  Anything else raises `H5Error` naming what was met.
"""
from __future__ import annotations

import struct

import numpy as np


class H5Error(ValueError):
    pass


_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class _Type:
    def __init__(self, cls, size, dtype=None, base=None, is_vlen_str=False):
        self.cls, self.size, self.dtype, self.base, self.is_vlen_str = cls, size, dtype, base, is_vlen_str


class File:
    def __init__(self, path):
        with open(path, "rb") as f:
            self.buf = f.read()
        b = self.buf
        if b[:8] != _SIG:
            raise H5Error("not an HDF5 file")
        if b[8] != 0:
            raise H5Error(f"superblock version {b[8]} (only version 0 is read)")
        if b[13] != 8 or b[14] != 8:
            raise H5Error("only 8-byte offsets and lengths are read")
        self.base = struct.unpack_from("<Q", b, 24)[0]
        # root group symbol table entry: link name offset, object header address, cache type, reserved, scratch
        self.root = Group(self, struct.unpack_from("<Q", b, 56 + 8)[0], "/")

    # ---- low level
    def u(self, fmt, off):
        return struct.unpack_from("<" + fmt, self.buf, off)

    def messages(self, addr):
        """(type, flags, offset of the data, size) of every message of a version-1 object header."""
        b = self.buf
        if b[addr] != 1:
            raise H5Error(f"object header version {b[addr]} at {addr} (only version 1 is read)")
        n_msg, = self.u("H", addr + 2)
        hdr_size, = self.u("I", addr + 8)
        blocks = [(addr + 16, hdr_size)]
        out = []
        while blocks and len(out) < n_msg:
            off, size = blocks.pop(0)
            end = off + size
            while off + 8 <= end and len(out) < n_msg:
                mtype, msize, flags = self.u("HHB", off)
                data = off + 8
                if mtype == 0x0010:                                  # continuation: (offset, length) of more messages
                    c_off, c_len = self.u("QQ", data)
                    blocks.append((self.base + c_off, c_len))
                out.append((mtype, flags, data, msize))
                off = data + msize
        return out

    def heap_string(self, heap_addr, offset):
        if self.buf[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5Error("local heap signature missing")
        data_addr, = self.u("Q", heap_addr + 24)
        start = self.base + data_addr + offset
        return self.buf[start:self.buf.index(b"\0", start)].decode()

    def global_heap_object(self, coll_addr, index):
        b = self.buf
        if b[coll_addr:coll_addr + 4] != b"GCOL":
            raise H5Error("global heap signature missing")
        size, = self.u("Q", coll_addr + 8)
        off, end = coll_addr + 16, coll_addr + size
        while off + 16 <= end:
            idx, _, _, osize = self.u("HHIQ", off)
            if idx == 0:
                break
            if idx == index:
                return b[off + 16:off + 16 + osize]
            off += 16 + ((osize + 7) & ~7)
        raise H5Error(f"global heap object {index} not found")

    # ---- datatypes, dataspaces
    def datatype(self, off):
        cv, b0, b1, b2, size = self.u("BBBBI", off)
        cls, ver = cv & 15, cv >> 4
        if cls == 0:                                                  # fixed point
            if b0 & 1:
                raise H5Error("big-endian integers are not read")
            return _Type(cls, size, np.dtype(("<i" if b0 & 8 else "<u") + str(size)))
        if cls == 1:                                                  # floating point
            if b0 & 1:
                raise H5Error("big-endian floats are not read")
            return _Type(cls, size, np.dtype("<f" + str(size)))
        if cls == 3:                                                  # fixed-length string
            return _Type(cls, size, np.dtype("S" + str(size)))
        if cls == 9:                                                  # variable length; type 1 = string
            return _Type(cls, size, None, is_vlen_str=(b0 & 15) == 1)
        raise H5Error(f"datatype class {cls} (version {ver}) is not read")

    def dataspace(self, off):
        ver, rank, flags = self.u("BBB", off)
        if ver == 1:
            dims_off = off + 8
        elif ver == 2:
            dims_off = off + 4
        else:
            raise H5Error(f"dataspace version {ver}")
        return tuple(self.u("Q", dims_off + 8 * i)[0] for i in range(rank))

    def decode(self, typ, shape, raw_off):
        n = int(np.prod(shape)) if shape else 1
        if typ.is_vlen_str:
            vals = []
            for i in range(n):
                _, coll, idx = self.u("IQI", raw_off + 16 * i)
                vals.append(self.global_heap_object(self.base + coll, idx).decode())
            return vals[0] if not shape else np.array(vals, dtype=object).reshape(shape)
        if typ.dtype is None:
            raise H5Error("variable-length data other than strings is not read")
        a = np.frombuffer(self.buf, dtype=typ.dtype, count=n, offset=raw_off)
        if typ.cls == 3:
            a = np.array([s.split(b"\0")[0].decode() for s in a], dtype=object)
        return (a[0] if not shape else a.reshape(shape).copy())


class _Node:
    def __init__(self, f: File, addr: int, name: str):
        self.file, self.addr, self.name = f, addr, name
        self._msgs = f.messages(addr)

    @property
    def attrs(self) -> dict:
        f, out = self.file, {}
        for mtype, _, off, size in self._msgs:
            if mtype != 0x000C:
                continue
            ver = f.buf[off]
            if ver == 1:
                nsz, tsz, ssz = f.u("HHH", off + 2)
                p = off + 8
                pad = lambda v: (v + 7) & ~7
                name = f.buf[p:p + nsz].split(b"\0")[0].decode()
                t_off = p + pad(nsz)
                s_off = t_off + pad(tsz)
                d_off = s_off + pad(ssz)
            elif ver in (2, 3):
                nsz, tsz, ssz = f.u("HHH", off + 2)
                p = off + 8 + (1 if ver == 3 else 0)
                name = f.buf[p:p + nsz].split(b"\0")[0].decode()
                t_off, s_off, d_off = p + nsz, p + nsz + tsz, p + nsz + tsz + ssz
            else:
                raise H5Error(f"attribute message version {ver}")
            out[name] = f.decode(f.datatype(t_off), f.dataspace(s_off), d_off)
        return out


class Dataset(_Node):
    def read(self) -> np.ndarray:
        f = self.file
        typ = shape = None
        layout = None
        for mtype, _, off, size in self._msgs:
            if mtype == 0x0003:
                typ = f.datatype(off)
            elif mtype == 0x0001:
                shape = f.dataspace(off)
            elif mtype == 0x0008:
                layout = off
            elif mtype == 0x000B:
                raise H5Error(f"dataset {self.name}: filtered (compressed) data is not read")
        if typ is None or shape is None or layout is None:
            raise H5Error(f"dataset {self.name}: incomplete header")
        ver, cls = f.u("BB", layout)
        if ver != 3:
            raise H5Error(f"dataset {self.name}: data layout version {ver}")
        if cls == 1:                                                  # contiguous
            addr, _ = f.u("QQ", layout + 2)
            if addr == _UNDEF:
                return np.zeros(shape, typ.dtype)
            return f.decode(typ, shape, f.base + addr)
        if cls == 0:                                                  # compact: the data sits in the message
            return f.decode(typ, shape, layout + 4)
        raise H5Error(f"dataset {self.name}: chunked layout is not read")

    @property
    def shape(self):
        for mtype, _, off, _ in self._msgs:
            if mtype == 0x0001:
                return self.file.dataspace(off)
        return None


class Group(_Node):
    def _entries(self):
        f = self.file
        stab = [(off) for mtype, _, off, _ in self._msgs if mtype == 0x0011]
        if not stab:
            return {}
        btree, heap = f.u("QQ", stab[0])
        out = {}

        def walk(addr):
            b = f.buf
            if b[addr:addr + 4] == b"TREE":
                ntype, level, used = f.u("BBH", addr + 4)
                if ntype != 0:
                    raise H5Error("group B-tree node of the wrong type")
                p = addr + 24
                for i in range(used):
                    child, = f.u("Q", p + 8 + 16 * i)             # key_i, child_i, key_i+1, ...
                    walk(f.base + child)
            elif b[addr:addr + 4] == b"SNOD":
                n, = f.u("H", addr + 6)
                for i in range(n):
                    e = addr + 8 + 40 * i
                    name_off, obj = f.u("QQ", e)
                    out[f.heap_string(f.base + heap, name_off)] = f.base + obj
            else:
                raise H5Error("unexpected node in a group's B-tree")

        walk(f.base + btree)
        return out

    def keys(self):
        return list(self._entries().keys())

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            ent = node._entries()
            if part not in ent:
                raise KeyError(f"{part!r} not in group {node.name!r}")
            addr = ent[part]
            child_name = node.name.rstrip("/") + "/" + part
            is_group = any(m[0] == 0x0011 for m in node.file.messages(addr))
            node = (Group if is_group else Dataset)(node.file, addr, child_name)
        return node

    def datasets(self, prefix=""):
        """{path relative to this group: ndarray} of every dataset below it."""
        out = {}
        for k in self.keys():
            child = self[k]
            if isinstance(child, Group):
                out.update(child.datasets(prefix + k + "/"))
            else:
                out[prefix + k] = child.read()
        return out
