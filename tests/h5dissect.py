"""A small, independent HDF5 file parser (test infrastructure; numpy + struct only).

It follows the published HDF5 File Format Specification (version 3.0) for the structures
libhdf5 writes with default settings and that the reference's files use: superblock v0-v3,
version 1 and 2 object headers, symbol-table groups (TREE / SNOD / HEAP), compact link
messages, contiguous / compact / chunked (version 1 B-tree) layouts, filter pipelines,
fixed / float / string / variable-length-string / enum datatypes, attributes, global heaps.

It exists so that the C++ reader/writer of the product (`csrc/cli/h5lite.hpp`) is checked by
a second implementation: both are pinned against the two libhdf5-written files the reference
keeps for its own tests (`dx2/tests/data/empty.h5`, `cut_strong.refl`, copied to
`tests/golden/`), and the files the product writes are walked by this parser.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


class H5Error(Exception):
    pass


def lookup3(data: bytes, initval: int = 0) -> int:
    """The metadata checksum of the version 2 structures (Jenkins lookup3 hashlittle, bytes taken
    little-endian).  Pinned in tests/test_h5lite.py against the checksums libhdf5 stored in the
    version 2 object headers of cut_strong.refl."""
    M = 0xFFFFFFFF
    rot = lambda x, k: ((x << k) | (x >> (32 - k))) & M
    n = len(data)
    a = b = c = (0xDEADBEEF + n + initval) & M
    k = 0
    while n > 12:
        a = (a + int.from_bytes(data[k:k + 4], "little")) & M
        b = (b + int.from_bytes(data[k + 4:k + 8], "little")) & M
        c = (c + int.from_bytes(data[k + 8:k + 12], "little")) & M
        a = (a - c) & M; a ^= rot(c, 4); c = (c + b) & M
        b = (b - a) & M; b ^= rot(a, 6); a = (a + c) & M
        c = (c - b) & M; c ^= rot(b, 8); b = (b + a) & M
        a = (a - c) & M; a ^= rot(c, 16); c = (c + b) & M
        b = (b - a) & M; b ^= rot(a, 19); a = (a + c) & M
        c = (c - b) & M; c ^= rot(b, 4); b = (b + a) & M
        n -= 12
        k += 12
    if n == 0:
        return c
    tail = bytes(data[k:k + n]) + bytes(12 - n)
    a = (a + int.from_bytes(tail[0:4], "little")) & M
    b = (b + int.from_bytes(tail[4:8], "little")) & M
    c = (c + int.from_bytes(tail[8:12], "little")) & M
    c ^= b; c = (c - rot(b, 14)) & M
    a ^= c; a = (a - rot(c, 11)) & M
    b ^= a; b = (b - rot(a, 25)) & M
    c ^= b; c = (c - rot(b, 16)) & M
    a ^= c; a = (a - rot(c, 4)) & M
    b ^= a; b = (b - rot(a, 14)) & M
    c ^= b; c = (c - rot(b, 24)) & M
    return c


def with_checksum(block: bytes) -> bytes:
    return bytes(block) + struct.pack("<I", lookup3(bytes(block)))


class Datatype:
    def __init__(self, cls, size, np_dtype=None, vlen_string=False, base=None, enum=None, raw=b""):
        self.cls, self.size, self.np_dtype = cls, size, np_dtype
        self.vlen_string, self.base, self.enum, self.raw = vlen_string, base, enum, raw

    def __repr__(self):
        if self.vlen_string:
            return "vlen-str"
        return f"{self.np_dtype}" if self.np_dtype is not None else f"class{self.cls}/{self.size}"


def parse_datatype(b, off=0):
    """-> (Datatype, bytes consumed)"""
    cv = b[off]
    cls, ver = cv & 15, cv >> 4
    bits = b[off + 1] | (b[off + 2] << 8) | (b[off + 3] << 16)
    size = struct.unpack_from("<I", b, off + 4)[0]
    p = off + 8
    if cls == 0:  # fixed point
        order = ">" if bits & 1 else "<"
        signed = bool(bits & 8)
        p += 4
        return Datatype(0, size, np.dtype(f"{order}{'i' if signed else 'u'}{size}"), raw=bytes(b[off:p])), p - off
    if cls == 1:  # floating point
        order = ">" if bits & 1 else "<"
        p += 12
        return Datatype(1, size, np.dtype(f"{order}f{size}"), raw=bytes(b[off:p])), p - off
    if cls == 3:  # fixed-length string
        return Datatype(3, size, np.dtype(f"S{size}"), raw=bytes(b[off:p])), p - off
    if cls == 9:  # variable length
        base, n = parse_datatype(b, p)
        p += n
        return Datatype(9, size, None, vlen_string=(bits & 15) == 1, base=base, raw=bytes(b[off:p])), p - off
    if cls == 8:  # enumeration
        base, n = parse_datatype(b, p)
        p += n
        nmemb = bits & 0xFFFF
        names = []
        for _ in range(nmemb):
            e = b.index(0, p)
            names.append(bytes(b[p:e]).decode())
            ln = e - p + 1
            p += (ln + 7) // 8 * 8 if ver < 3 else ln
        vals = np.frombuffer(bytes(b[p:p + nmemb * base.size]), base.np_dtype)
        p += nmemb * base.size
        return Datatype(8, size, base.np_dtype, enum=dict(zip(names, vals.tolist())), raw=bytes(b[off:p])), p - off
    if cls in (2, 4, 5, 6, 7, 10):  # time, bitfield, opaque, compound, reference, array: opaque bytes here
        return Datatype(cls, size, np.dtype(f"V{size}"), raw=bytes(b[off:off + 8])), 8
    raise H5Error(f"datatype class {cls} not handled")


def parse_dataspace(b, off=0):
    ver = b[off]
    if ver == 1:
        rank, flags = b[off + 1], b[off + 2]
        p = off + 8
    elif ver == 2:
        rank, flags, typ = b[off + 1], b[off + 2], b[off + 3]
        p = off + 4
        if typ == 2:
            return None, None  # null dataspace
    else:
        raise H5Error(f"dataspace version {ver}")
    dims = list(struct.unpack_from(f"<{rank}Q", b, p))
    p += 8 * rank
    maxdims = None
    if flags & 1:
        maxdims = list(struct.unpack_from(f"<{rank}Q", b, p))
    return dims, maxdims


class Obj:
    """One object header: its messages, decoded lazily."""

    def __init__(self, f, addr):
        self.f, self.addr = f, addr
        self.msgs = f._read_header(addr)  # [(type, flags, bytes)]

    def first(self, t):
        for typ, _, data in self.msgs:
            if typ == t:
                return data
        return None

    @property
    def is_group(self):
        return self.first(0x11) is not None or self.first(0x02) is not None

    @property
    def is_dataset(self):
        return self.first(0x08) is not None

    # -- group ---------------------------------------------------------------------------
    def links(self):
        """name -> ('hard', addr) | ('soft', path) | ('external', file, path)"""
        out = {}
        st = self.first(0x11)
        if st is not None:
            btree, heap = struct.unpack_from("<QQ", st, 0)
            hdata = self.f._local_heap(heap)
            for name_off, ohdr, cache, scratch in self.f._group_btree(btree):
                name = hdata[name_off:hdata.index(b"\0", name_off)].decode()
                if cache == 2:
                    so = struct.unpack_from("<I", scratch, 0)[0]
                    out[name] = ("soft", hdata[so:hdata.index(b"\0", so)].decode())
                else:
                    out[name] = ("hard", ohdr)
        for typ, _, d in self.msgs:
            if typ == 0x06:
                name, val = self.f._parse_link(d)
                out[name] = val
        li = self.first(0x02)
        if li is not None:
            flags = li[1]
            p = 2 + (8 if flags & 1 else 0)
            fheap = struct.unpack_from("<Q", li, p)[0]
            if fheap != UNDEF:
                raise H5Error("dense link storage (fractal heap) is not handled by this parser")
        return out

    # -- attributes ------------------------------------------------------------------------
    def attrs(self):
        out = {}
        for typ, _, d in self.msgs:
            if typ != 0x0C:
                continue
            ver = d[0]
            nlen, tlen, slen = struct.unpack_from("<HHH", d, 2)
            p = 8
            if ver == 3:
                p = 9
            pad = (lambda n: (n + 7) // 8 * 8) if ver == 1 else (lambda n: n)
            name = bytes(d[p:p + nlen]).split(b"\0")[0].decode()
            p += pad(nlen)
            dt, _ = parse_datatype(d, p)
            p += pad(tlen)
            dims, _ = parse_dataspace(d, p)
            p += pad(slen)
            out[name] = self.f._decode(dt, dims, d[p:])
        return out

    # -- dataset ---------------------------------------------------------------------------
    def dtype(self):
        return parse_datatype(self.first(0x03))[0]

    def shape(self):
        return parse_dataspace(self.first(0x01))[0]

    def filters(self):
        d = self.first(0x0B)
        if d is None:
            return []
        ver, n = d[0], d[1]
        p = 8 if ver == 1 else 2
        out = []
        for _ in range(n):
            fid = struct.unpack_from("<H", d, p)[0]
            p += 2
            nlen = 0
            if ver == 1 or fid >= 256:
                nlen = struct.unpack_from("<H", d, p)[0]
                p += 2
            flags, ncd = struct.unpack_from("<HH", d, p)
            p += 4
            name = bytes(d[p:p + nlen]).split(b"\0")[0].decode()
            p += (nlen + 7) // 8 * 8 if ver == 1 else nlen
            cd = list(struct.unpack_from(f"<{ncd}I", d, p))
            p += 4 * ncd
            if ver == 1 and ncd % 2:
                p += 4
            out.append({"id": fid, "name": name, "flags": flags, "client_data": cd})
        return out

    def layout(self):
        d = self.first(0x08)
        ver = d[0]
        if ver != 3:
            raise H5Error(f"layout version {ver} not handled by this parser")
        cls = d[1]
        if cls == 0:
            size = struct.unpack_from("<H", d, 2)[0]
            return {"class": "compact", "data": bytes(d[4:4 + size])}
        if cls == 1:
            addr, size = struct.unpack_from("<QQ", d, 2)
            return {"class": "contiguous", "addr": addr, "size": size}
        if cls == 2:
            nd = d[2]
            addr = struct.unpack_from("<Q", d, 3)[0]
            cdims = list(struct.unpack_from(f"<{nd}I", d, 11))
            return {"class": "chunked", "btree": addr, "chunk": cdims[:-1], "elem": cdims[-1]}
        raise H5Error(f"layout class {cls}")

    def chunks(self):
        """[(offset tuple, filter mask, size, addr)] of a chunked dataset (version 1 B-tree)."""
        lay = self.layout()
        if lay["btree"] == UNDEF:
            return []
        return self.f._chunk_btree(lay["btree"], len(lay["chunk"]) + 1)

    def read_chunk_raw(self, offset):
        for off, mask, size, addr in self.chunks():
            if tuple(off) == tuple(offset):
                return mask, self.f.buf[addr:addr + size]
        raise KeyError(offset)

    def read(self):
        dt, dims, lay = self.dtype(), self.shape(), self.layout()
        n = int(np.prod(dims)) if dims else 1
        if lay["class"] == "compact":
            raw = lay["data"]
        elif lay["class"] == "contiguous":
            raw = b"" if lay["addr"] == UNDEF else self.f.buf[lay["addr"]:lay["addr"] + n * dt.size]
            if lay["addr"] == UNDEF:
                raw = bytes(n * dt.size)
        else:
            if self.filters():
                raise H5Error("filtered chunks: use read_chunk_raw")
            out = np.zeros(dims, dt.np_dtype)
            for off, _, size, addr in self.chunks():
                c = np.frombuffer(self.f.buf[addr:addr + size], dt.np_dtype).reshape(lay["chunk"])
                sl = tuple(slice(o, min(o + cd, d)) for o, cd, d in zip(off, lay["chunk"], dims))
                out[sl] = c[tuple(slice(0, s.stop - s.start) for s in sl)]
            return out
        return self.f._decode(dt, dims, raw)


class File:
    def __init__(self, path):
        with open(path, "rb") as fh:
            self.buf = fh.read()
        if self.buf[:8] != SIG:
            raise H5Error("not an HDF5 file (no signature at offset 0)")
        self.version = self.buf[8]
        b = self.buf
        if self.version in (0, 1):
            self.size_off, self.size_len = b[13], b[14]
            self.leaf_k, self.int_k = struct.unpack_from("<HH", b, 16)
            p = 24 if self.version == 0 else 28
            self.base, self.free, self.eof, self.driver = struct.unpack_from("<QQQQ", b, p)
            p += 32
            self.root_entry = struct.unpack_from("<QQII16s", b, p)
            self.root_addr = self.root_entry[1]
        elif self.version in (2, 3):
            self.size_off, self.size_len = b[9], b[10]
            self.base, self.ext, self.eof, self.root_addr = struct.unpack_from("<QQQQ", b, 12)
        else:
            raise H5Error(f"superblock version {self.version}")
        if self.size_off != 8 or self.size_len != 8:
            raise H5Error("only 8-byte offsets and lengths are handled")
        self.root = Obj(self, self.root_addr)

    # -- object headers ----------------------------------------------------------------------
    def _read_header(self, addr):
        b = self.buf
        msgs = []
        if b[addr:addr + 4] == b"OHDR":
            flags = b[addr + 5]
            p = addr + 6
            if flags & 0x20:
                p += 16
            if flags & 0x10:
                p += 4
            szlen = 1 << (flags & 3)
            chunk0 = int.from_bytes(b[p:p + szlen], "little")
            p += szlen
            blocks = [(p, chunk0)]
            track = bool(flags & 4)
            while blocks:
                p, n = blocks.pop(0)
                end = p + n
                while p + 4 + (2 if track else 0) <= end:
                    typ = b[p]
                    size = struct.unpack_from("<H", b, p + 1)[0]
                    mflags = b[p + 3]
                    p += 4 + (2 if track else 0)
                    data = b[p:p + size]
                    p += size
                    if typ == 0x10:
                        o, ln = struct.unpack_from("<QQ", data, 0)
                        blocks.append((o + 4, ln - 8))  # skip "OCHK", drop the checksum
                    elif typ != 0:
                        msgs.append((typ, mflags, data))
            return msgs
        ver = b[addr]
        if ver != 1:
            raise H5Error(f"object header version {ver} at {addr:#x}")
        nmsg = struct.unpack_from("<H", b, addr + 2)[0]
        hsize = struct.unpack_from("<I", b, addr + 8)[0]
        blocks = [(addr + 16, hsize)]
        while blocks and len(msgs) < nmsg + 64:
            p, n = blocks.pop(0)
            end = p + n
            while p + 8 <= end:
                typ, size = struct.unpack_from("<HH", b, p)
                mflags = b[p + 4]
                data = b[p + 8:p + 8 + size]
                p += 8 + size
                if typ == 0x10:
                    o, ln = struct.unpack_from("<QQ", data, 0)
                    blocks.append((o, ln))
                elif typ != 0:
                    msgs.append((typ, mflags, data))
        return msgs

    # -- groups ------------------------------------------------------------------------------
    def _local_heap(self, addr):
        b = self.buf
        if b[addr:addr + 4] != b"HEAP":
            raise H5Error("local heap signature")
        size, _free, data = struct.unpack_from("<QQQ", b, addr + 8)
        return b[data:data + size]

    def _group_btree(self, addr):
        b = self.buf
        if b[addr:addr + 4] == b"SNOD":
            n = struct.unpack_from("<H", b, addr + 6)[0]
            for i in range(n):
                name_off, ohdr, cache, _r, scratch = struct.unpack_from("<QQII16s", b, addr + 8 + 40 * i)
                yield name_off, ohdr, cache, scratch
            return
        if b[addr:addr + 4] != b"TREE":
            raise H5Error(f"group B-tree signature at {addr:#x}")
        used = struct.unpack_from("<H", b, addr + 6)[0]
        p = addr + 24
        for i in range(used):
            child = struct.unpack_from("<Q", b, p + 8 + 16 * i)[0]
            yield from self._group_btree(child)

    def _chunk_btree(self, addr, ndims):
        b = self.buf
        if b[addr:addr + 4] != b"TREE" or b[addr + 4] != 1:
            raise H5Error("chunk B-tree signature")
        level = b[addr + 5]
        used = struct.unpack_from("<H", b, addr + 6)[0]
        keysz = 8 + 8 * ndims
        p = addr + 24
        out = []
        for i in range(used):
            size, mask = struct.unpack_from("<II", b, p)
            off = struct.unpack_from(f"<{ndims}Q", b, p + 8)[:-1]
            child = struct.unpack_from("<Q", b, p + keysz)[0]
            if level == 0:
                out.append((off, mask, size, child))
            else:
                out.extend(self._chunk_btree(child, ndims))
            p += keysz + 8
        return out

    def _parse_link(self, d):
        flags = d[1]
        p = 2
        ltype = 0
        if flags & 8:
            ltype = d[p]
            p += 1
        if flags & 4:
            p += 8
        if flags & 16:
            p += 1
        ln = 1 << (flags & 3)
        nlen = int.from_bytes(d[p:p + ln], "little")
        p += ln
        name = bytes(d[p:p + nlen]).decode()
        p += nlen
        if ltype == 0:
            return name, ("hard", struct.unpack_from("<Q", d, p)[0])
        vlen = struct.unpack_from("<H", d, p)[0]
        v = bytes(d[p + 2:p + 2 + vlen])
        if ltype == 1:
            return name, ("soft", v.decode())
        if ltype == 64:
            parts = v[1:].split(b"\0")
            return name, ("external", parts[0].decode(), parts[1].decode())
        raise H5Error(f"link type {ltype}")

    # -- data --------------------------------------------------------------------------------
    def _global_heap_object(self, addr, index):
        b = self.buf
        if b[addr:addr + 4] != b"GCOL":
            raise H5Error("global heap signature")
        total = struct.unpack_from("<Q", b, addr + 8)[0]
        p = addr + 16
        while p < addr + total:
            idx, _ref, _r, size = struct.unpack_from("<HHIQ", b, p)
            if idx == index:
                return b[p + 16:p + 16 + size]
            if idx == 0:
                break
            p += 16 + (size + 7) // 8 * 8
        raise H5Error(f"global heap object {index} not found at {addr:#x}")

    def _decode(self, dt, dims, raw):
        n = int(np.prod(dims)) if dims else 1
        if dt.vlen_string:
            out = []
            for i in range(n):
                ln, addr, idx = struct.unpack_from("<IQI", raw, 16 * i)
                out.append(bytes(self._global_heap_object(addr, idx)[:ln]).decode() if addr else "")
            return out if dims else out[0]
        a = np.frombuffer(bytes(raw[:n * dt.size]), dt.np_dtype)
        return a.reshape(dims) if dims else a[0]

    # -- navigation ----------------------------------------------------------------------------
    def get(self, path):
        o = self.root
        for part in [p for p in path.split("/") if p]:
            links = o.links()
            if part not in links:
                raise KeyError(path)
            kind = links[part]
            if kind[0] == "hard":
                o = Obj(self, kind[1])
            elif kind[0] == "soft":
                o = self.get(kind[1])
            else:
                raise H5Error(f"{path}: external link to {kind[1]}:{kind[2]}")
        return o

    def walk(self, path="/", obj=None):
        obj = obj or self.root
        yield path, obj
        if obj.is_group:
            for name, kind in sorted(obj.links().items()):
                if kind[0] == "hard":
                    yield from self.walk(path.rstrip("/") + "/" + name, Obj(self, kind[1]))
