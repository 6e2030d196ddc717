"""CONCEPT snapshot reader without h5py.

Replaces the reference's h5py readers
  read_positions(filename)   FraP_kTilDelta.py:55-60  -> f['components/matter/pos'][:]
  read_boxsize(filename)     FraP_kTilDelta.py:63-68  -> f.attrs['boxsize']
(inline copies: Power_compare.py:143-144, Plotter.py:150-151).

CONCEPT writes HDF5 superblock-v0 files whose datasets are contiguous and
uncompressed, so a small walker over the v1 object headers / group B-trees is
enough: superblock -> root group symbol table -> (B-tree, local heap) -> object
headers -> {dataspace, datatype, layout, attribute} messages.  Only what those
files use is implemented; anything else raises ValueError (the reference raises
from h5py in the same situations).
"""
from __future__ import annotations

import os
import struct

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class _H5:
    def __init__(self, path):
        if not os.path.isfile(path):
            raise FileNotFoundError(path)
        self.path = path
        self.buf = np.memmap(path, dtype=np.uint8, mode="r")
        b = bytes(self.buf[:96])
        if b[:8] != _SIG:
            raise ValueError(f"{path}: not an HDF5 file")
        if b[8] != 0 or b[13] != 8 or b[14] != 8:
            raise ValueError(f"{path}: only superblock v0 with 8-byte offsets is supported")
        self.base = struct.unpack_from("<Q", b, 24)[0]
        # root group symbol-table entry starts at byte 56
        self.root_header = struct.unpack_from("<Q", b, 56 + 8)[0]

    # -- raw access -------------------------------------------------------
    def rd(self, off, n):
        off += self.base
        return bytes(self.buf[off:off + n])

    # -- object headers (version 1) ---------------------------------------
    def messages(self, addr):
        """Yield (type, flags, body) for every message of a v1 object header."""
        hdr = self.rd(addr, 16)
        if hdr[0] != 1:
            raise ValueError(f"{self.path}: object header v{hdr[0]} unsupported")
        nmsg = struct.unpack_from("<H", hdr, 2)[0]
        size = struct.unpack_from("<I", hdr, 8)[0]
        blocks = [(addr + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            off, left = blocks.pop(0)
            while left >= 8 and len(out) < nmsg:
                mtype, msize, flags = struct.unpack_from("<HHB", self.rd(off, 8), 0)
                body = self.rd(off + 8, msize)
                off += 8 + msize
                left -= 8 + msize
                if mtype == 0x10:  # continuation
                    caddr, clen = struct.unpack_from("<QQ", body, 0)
                    blocks.append((caddr, clen))
                out.append((mtype, flags, body))
        return out

    # -- groups -------------------------------------------------------------
    def children(self, addr):
        """name -> object header address for an old-style group."""
        for mtype, _f, body in self.messages(addr):
            if mtype == 0x11:
                btree, heap = struct.unpack_from("<QQ", body, 0)
                return self._walk_btree(btree, self._heap_data(heap))
        raise ValueError(f"{self.path}: object at {addr} is not a group")

    def _heap_data(self, heap):
        h = self.rd(heap, 32)
        if h[:4] != b"HEAP":
            raise ValueError("bad local heap")
        return struct.unpack_from("<Q", h, 24)[0]

    def _name(self, heap_data, off):
        raw = self.rd(heap_data + off, 256)
        return raw[:raw.index(b"\0")].decode("utf-8")

    def _walk_btree(self, node, heap_data):
        out = {}
        h = self.rd(node, 24)
        if h[:4] == b"TREE":
            level = h[5]
            used = struct.unpack_from("<H", h, 6)[0]
            body = self.rd(node + 24, (2 * used + 1) * 8)
            for i in range(used):
                child = struct.unpack_from("<Q", body, 8 + 16 * i)[0]
                if level > 0:
                    out.update(self._walk_btree(child, heap_data))
                else:
                    out.update(self._symbol_node(child, heap_data))
            return out
        raise ValueError("bad group B-tree node")

    def _symbol_node(self, addr, heap_data):
        h = self.rd(addr, 8)
        if h[:4] != b"SNOD":
            raise ValueError("bad symbol node")
        n = struct.unpack_from("<H", h, 6)[0]
        out = {}
        for i in range(n):
            e = self.rd(addr + 8 + 40 * i, 40)
            name_off, ohdr = struct.unpack_from("<QQ", e, 0)
            out[self._name(heap_data, name_off)] = ohdr
        return out

    def resolve(self, path):
        addr = self.root_header
        for part in [p for p in path.split("/") if p]:
            kids = self.children(addr)
            if part not in kids:
                raise KeyError(f"{self.path}: no object {path!r}")
            addr = kids[part]
        return addr

    # -- datatypes / dataspaces ----------------------------------------------
    @staticmethod
    def _dtype(body):
        cls = body[0] & 0x0F
        bits0 = body[1]
        size = struct.unpack_from("<I", body, 4)[0]
        order = ">" if bits0 & 1 else "<"
        if cls == 1:
            return np.dtype(f"{order}f{size}")
        if cls == 0:
            signed = (bits0 >> 3) & 1
            return np.dtype(f"{order}{'i' if signed else 'u'}{size}")
        if cls == 3:
            return np.dtype(f"S{size}")
        raise ValueError(f"HDF5 datatype class {cls} unsupported")

    @staticmethod
    def _shape(body):
        ver, rank, flags = body[0], body[1], body[2]
        off = 8 if ver == 1 else 4
        return tuple(struct.unpack_from("<Q", body, off + 8 * i)[0] for i in range(rank))

    def dataset(self, path):
        addr = self.resolve(path)
        shape = dtype = None
        data_addr = None
        for mtype, _f, body in self.messages(addr):
            if mtype == 0x01:
                shape = self._shape(body)
            elif mtype == 0x03:
                dtype = self._dtype(body)
            elif mtype == 0x08:
                ver = body[0]
                if ver == 3:
                    if body[1] != 1:
                        raise ValueError("only contiguous datasets are supported")
                    data_addr = struct.unpack_from("<Q", body, 2)[0]
                elif ver in (1, 2):
                    if body[2] != 1:
                        raise ValueError("only contiguous datasets are supported")
                    data_addr = struct.unpack_from("<Q", body, 8)[0]
                else:
                    raise ValueError(f"layout message v{ver} unsupported")
        if shape is None or dtype is None or data_addr is None or data_addr == _UNDEF:
            raise ValueError(f"{self.path}: {path!r} is not a contiguous dataset")
        n = int(np.prod(shape)) if shape else 1
        off = self.base + data_addr
        arr = np.frombuffer(self.buf, dtype=dtype, count=n, offset=off).reshape(shape)
        return np.array(arr, dtype=dtype.newbyteorder("="))

    def attrs(self, path="/"):
        addr = self.resolve(path)
        out = {}
        for mtype, _f, body in self.messages(addr):
            if mtype != 0x0C:
                continue
            ver = body[0]
            nlen, tlen, slen = struct.unpack_from("<HHH", body, 2)
            pos = 8 if ver == 1 else 9
            pad = (lambda x: (x + 7) & ~7) if ver == 1 else (lambda x: x)
            name = body[pos:pos + nlen].split(b"\0")[0].decode("utf-8")
            pos += pad(nlen)
            tbody = body[pos:pos + tlen]
            pos += pad(tlen)
            sbody = body[pos:pos + slen]
            pos += pad(slen)
            try:
                dt = self._dtype(tbody)
            except ValueError:
                continue
            shape = self._shape(sbody) if sbody[1] else ()
            n = int(np.prod(shape)) if shape else 1
            val = np.frombuffer(body, dtype=dt, count=n, offset=pos)
            if dt.kind == "S":
                out[name] = val[0].split(b"\0")[0].decode("utf-8")
            else:
                out[name] = val.reshape(shape)[()] if shape else val[0]
        return out


def read_positions(filename):
    """(Np,3) float64 positions, `f['components/matter/pos'][:]` (FraP_kTilDelta.py:55-60)."""
    return _H5(filename).dataset("components/matter/pos")


def read_momenta(filename):
    """(Np,3) float64 comoving momenta `components/matter/mom`."""
    return _H5(filename).dataset("components/matter/mom")


def read_boxsize(filename):
    """`f.attrs['boxsize']` (FraP_kTilDelta.py:63-68)."""
    a = _H5(filename).attrs("/")
    if "boxsize" not in a:
        raise KeyError(f"{filename}: no 'boxsize' attribute")
    return float(a["boxsize"])


def read_attrs(filename, path="/"):
    """All scalar attributes of an object (root: a, H0, boxsize, Ωb, Ωcdm, units)."""
    return _H5(filename).attrs(path)


def find_hdf5_files(root):
    """Sorted recursive list of *.hdf5 under root (FraP_kTilDelta.py:30-41)."""
    files = []
    for d, _dirs, names in os.walk(root):
        for fname in names:
            if fname.endswith(".hdf5"):
                files.append(os.path.join(d, fname))
    return sorted(files)
