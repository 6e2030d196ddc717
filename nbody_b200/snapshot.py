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


# ------------------------------------------------------------------------------------------
# writer: the same subset of HDF5 the reader walks, laid out like the files
# CONCEPT/h5py wrote for the reference (superblock v0, old-style groups = B-tree + local heap +
# symbol node, v1 object headers, contiguous little-endian datasets).  Differences from those
# files: string attributes are fixed-length (class 3) instead of variable-length, and no padding
# NIL messages are written.
_F64_TYPE = bytes.fromhex("11203f000800000000004000340b0034ff030000")      # IEEE f64 LE, 20 bytes
_I64_TYPE = bytes.fromhex("100800000800000000004000")                      # signed 64-bit LE, 12 bytes
_SCALAR_SPACE = bytes.fromhex("0100000000000000")
_LEAF_K, _INTERNAL_K = 4, 16


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _attr_message(name, value):
    if isinstance(value, str):
        raw = value.encode("utf-8") + b"\0"
        dt = struct.pack("<BBBBI", 0x13, 0x10, 0, 0, len(raw))                 # fixed string, UTF-8
        data = raw
    elif isinstance(value, (int, np.integer)) and not isinstance(value, bool):
        dt, data = _I64_TYPE, struct.pack("<q", int(value))
    else:
        dt, data = _F64_TYPE, struct.pack("<d", float(value))
    nm = name.encode("utf-8") + b"\0"
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(_SCALAR_SPACE))
    body += _pad8(nm) + _pad8(dt) + _pad8(_SCALAR_SPACE) + data
    return 0x0C, 0, _pad8(body)


def _object_header(messages):
    body = b"".join(struct.pack("<HHB3x", t, len(b), f) + b for t, f, b in messages)
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(body)) + body


class _Writer:
    def __init__(self):
        self.blocks = []          # (offset, bytes)
        self.end = 96             # superblock

    def alloc(self, size):
        off = self.end
        self.end += (size + 7) & ~7
        return off

    def put(self, off, data):
        self.blocks.append((off, data))

    def group(self, children, attrs=()):
        """children: {name: object header address}.  Returns (header address, btree, heap)."""
        names = sorted(children, key=lambda s: s.encode("utf-8"))
        if len(names) > 2 * _LEAF_K:
            raise ValueError("too many links for one symbol node")
        heap_data, offs = bytearray(8), {}
        for nm in names:
            offs[nm] = len(heap_data)
            heap_data += _pad8(nm.encode("utf-8") + b"\0")
        if len(heap_data) < 24:                       # room for one free-list entry like libhdf5 leaves
            heap_data += bytes(24 - len(heap_data))
        a_hdr_size = 16 + (8 + 16) + sum(8 + len(_attr_message(k, v)[2]) for k, v in attrs)
        a_hdr = self.alloc(a_hdr_size)
        a_tree = self.alloc(24 + (2 * _INTERNAL_K + 1) * 8 + 2 * _INTERNAL_K * 8)
        a_heap = self.alloc(32)
        a_hdat = self.alloc(len(heap_data))
        a_snod = self.alloc(8 + 2 * _LEAF_K * 40)
        msgs = [(0x11, 0, struct.pack("<QQ", a_tree, a_heap))] + [_attr_message(k, v) for k, v in attrs]
        self.put(a_hdr, _object_header(msgs))
        tree = bytearray(24 + (2 * _INTERNAL_K + 1) * 8 + 2 * _INTERNAL_K * 8)
        struct.pack_into("<4sBBHQQ", tree, 0, b"TREE", 0, 0, 1 if names else 0, _UNDEF, _UNDEF)
        if names:
            struct.pack_into("<QQQ", tree, 24, 0, a_snod, offs[names[-1]])
        self.put(a_tree, bytes(tree))
        self.put(a_heap, struct.pack("<4sB3xQQQ", b"HEAP", 0, len(heap_data), 1, a_hdat))   # 1 = no free block
        self.put(a_hdat, bytes(heap_data))
        snod = bytearray(8 + 2 * _LEAF_K * 40)
        struct.pack_into("<4sBBH", snod, 0, b"SNOD", 1, 0, len(names))
        for i, nm in enumerate(names):
            struct.pack_into("<QQII16x", snod, 8 + 40 * i, offs[nm], children[nm], 0, 0)
        self.put(a_snod, bytes(snod))
        return a_hdr, a_tree, a_heap

    def dataset(self, arr):
        arr = np.ascontiguousarray(arr, dtype="<f8")
        space = struct.pack("<BBB5x", 1, arr.ndim, 1) + b"".join(struct.pack("<Q", d) for d in arr.shape) * 2
        a_data = self.alloc(arr.nbytes)
        msgs = [(0x01, 0, _pad8(space)), (0x03, 1, _pad8(_F64_TYPE)), (0x05, 1, bytes.fromhex("0201020100000000")),
                (0x08, 1, _pad8(struct.pack("<BBQQ", 3, 1, a_data, arr.nbytes)))]
        hdr = _object_header(msgs)
        a_hdr = self.alloc(len(hdr))
        self.put(a_hdr, hdr)
        self.put(a_data, arr)
        return a_hdr

    def write(self, path, root):
        a_root, a_tree, a_heap = root
        sb = bytearray(96)
        sb[:8] = _SIG
        struct.pack_into("<BBBBBBBBHHI", sb, 8, 0, 0, 0, 0, 0, 8, 8, 0, _LEAF_K, _INTERNAL_K, 0)
        struct.pack_into("<QQQQ", sb, 24, 0, _UNDEF, self.end, _UNDEF)
        struct.pack_into("<QQII", sb, 56, 0, a_root, 1, 0)
        struct.pack_into("<QQ", sb, 80, a_tree, a_heap)
        tmp = path + ".tmp"
        with open(tmp, "wb") as f:
            f.truncate(self.end)
            f.write(bytes(sb))
            for off, data in self.blocks:
                f.seek(off)
                f.write(data if isinstance(data, (bytes, bytearray)) else memoryview(data).cast("B"))
        os.replace(tmp, path)


def write_snapshot(filename, pos, mom, boxsize, a=1.0, mass=1.0, H0=None, Omega_b=None, Omega_cdm=None,
                   units=("Mpc", "Gyr", "10¹⁰ m☉"), species="matter"):
    """Write a snapshot with the reference files' layout: root attributes a, boxsize, H0, Ωb, Ωcdm,
    `unit length/time/mass`; group components/matter with attributes N, mass, species and the
    (Np,3) float64 datasets pos and mom.  read_positions / read_momenta / read_boxsize / read_attrs
    read it back (h5py is not in this image; the layout follows the HDF5 1.x file-format
    specification for superblock v0, see the reader).
    UNVALIDATED AGAINST libhdf5: no HDF5 library has opened one of these files yet.  Known departures from
    the files CONCEPT writes: fixed-length instead of variable-length strings, non-ASCII attribute names
    ('Ωb', 'Ωcdm') in version-1 attribute messages without a character-set field, no modification-time / NIL
    messages.  A consumer that reads through h5py should be checked once against a file written here."""
    pos = np.ascontiguousarray(pos, dtype=np.float64)
    mom = np.ascontiguousarray(mom, dtype=np.float64)
    if pos.ndim != 2 or pos.shape[1] != 3 or mom.shape != pos.shape:
        raise ValueError("pos and mom must both be (Np,3)")
    w = _Writer()
    d_pos, d_mom = w.dataset(pos), w.dataset(mom)
    matter, _, _ = w.group({"pos": d_pos, "mom": d_mom},
                           attrs=[("species", species), ("mass", float(mass)), ("N", int(pos.shape[0]))])
    comps, _, _ = w.group({species: matter})
    root_attrs = [("unit time", units[1]), ("unit length", units[0]), ("unit mass", units[2])]
    for key, val in (("H0", H0), ("a", a), ("boxsize", boxsize), ("Ωb", Omega_b), ("Ωcdm", Omega_cdm)):
        if val is not None:
            root_attrs.append((key, float(val)))
    root = w.group({"components": comps}, attrs=root_attrs)
    w.write(filename, root)
    return filename
