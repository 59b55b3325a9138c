"""h5lite — a small read-only HDF5 reader, enough for the netCDF-4 files of rrtmgp-data (coefficient and LUT
files, RFMIP/all-sky inputs and reference fluxes).  The reference reads them through libnetcdf
(include_test/Netcdf_interface.h:256-300); neither libnetcdf/libhdf5 nor h5py/netCDF4 exist in this image, so
the on-disk format is read directly:

  superblock v0/v1 (symbol-table groups: B-tree v1 + local heap, object header v1) and v2/v3 (object header
  v2, link messages compact or "dense" in a fractal heap); datasets contiguous, compact or chunked (B-tree v1
  chunk index) with the deflate / shuffle / fletcher32 filters; fixed-point, IEEE float, fixed-length and
  variable-length string element types; attributes compact or dense.

Only what netCDF-4 writes is handled; anything else raises H5Error naming the feature (never a silent wrong
read).  Variables are returned in the file's (C, row-major) dimension order, like Netcdf_interface.h does
before its callers reverse the dimensions.
"""
from __future__ import annotations

import struct
import zlib
from typing import Dict, List, Optional, Tuple

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
NON_COORD = "_nc4_non_coord_"


class H5Error(RuntimeError):
    pass


def _guarded(fn):
    """A corrupt file surfaces as H5Error, whatever low-level exception its bytes provoke."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *a, **k):
        try:
            return fn(self, *a, **k)
        except (H5Error, KeyError):
            raise
        except (struct.error, IndexError, ValueError, TypeError, OverflowError, UnicodeDecodeError, zlib.error,
                MemoryError, RecursionError) as e:
            raise H5Error(f"{getattr(self, 'path', '?')}: corrupt or unsupported HDF5 structure ({type(e).__name__}: {e})") from e
    return wrapper


class _Dataset:
    def __init__(self):
        self.shape: Tuple[int, ...] = ()
        self.dtype: Optional[np.dtype] = None
        self.layout = None          # ("contiguous", addr, size) | ("compact", bytes) | ("chunked", btree, chunk_dims)
        self.filters: List[int] = []
        self.fill: Optional[bytes] = None
        self.is_dataset = False
        self.links: Dict[str, int] = {}     # groups: name -> object header address
        self.symtab = None                  # old-style group: (btree, heap)
        self.link_heap = None               # dense links: (fractal heap, name index B-tree) addresses
        self.unsupported = None             # datatype class this reader does not decode
        self.vlen_string = False            # variable-length strings: elements are global-heap references
        self.attrs: Dict[str, object] = {}  # compact attributes of decodable type (numbers, fixed-length strings)


class File:
    @_guarded
    def __init__(self, path: str):
        with open(path, "rb") as f:
            self.b = f.read()
        self.path = path
        if self.b[:8] != b"\x89HDF\r\n\x1a\n":
            raise H5Error(f"{path}: not an HDF5 file (netCDF classic files are read by ncfiles.read_classic)")
        ver = self.b[8]
        if ver in (0, 1):
            if self.b[13] != 8 or self.b[14] != 8:
                raise H5Error("only 8-byte offsets/lengths are supported")
            p = 24 + (4 if ver == 1 else 0)
            p += 4 * 8                                  # base, free-space, end-of-file, driver info addresses
            # root group symbol table entry: link name offset, object header address, cache type, reserved, scratch
            _, root, cache = struct.unpack_from("<QQI", self.b, p)
            self.root = root
        elif ver in (2, 3):
            if self.b[9] != 8 or self.b[10] != 8:
                raise H5Error("only 8-byte offsets/lengths are supported")
            self.root = struct.unpack_from("<Q", self.b, 12 + 3 * 8)[0]
        else:
            raise H5Error(f"superblock version {ver} is not supported")
        self._root = self._object(self.root)
        # libnetcdf stores a variable whose name clashes with a dimension it does not span as "_nc4_non_coord_<name>"
        self.vars: Dict[str, int] = {(n[len(NON_COORD):] if n.startswith(NON_COORD) else n): a
                                     for n, a in self._group_links(self._root)}

    # ------------------------------------------------------------------ public
    def names(self) -> List[str]:
        return sorted(self.vars)

    def __contains__(self, name: str) -> bool:
        return name in self.vars

    @_guarded
    def shape(self, name: str) -> Tuple[int, ...]:
        return self._dataset(name).shape

    @_guarded
    def dims(self, name: str) -> Tuple[str, ...]:
        """netCDF dimension names of a variable, slowest first: through DIMENSION_LIST (object references to the
        dimension scales), else _Netcdf4Coordinates / _Netcdf4Dimid; a dimension scale is its own dimension."""
        a = self.attrs(name)
        if a.get("CLASS") == "DIMENSION_SCALE" and len(self.shape(name)) == 1 and "DIMENSION_LIST" not in a:
            return (name,)
        by_addr = {addr: n for n, addr in self.vars.items()}
        if "DIMENSION_LIST" in a:
            return tuple(by_addr[r] for r in a["DIMENSION_LIST"])
        if "_Netcdf4Coordinates" in a:
            ids = {}
            for n in self.vars:
                an = self.attrs(n)
                if an.get("CLASS") == "DIMENSION_SCALE" and "_Netcdf4Dimid" in an:
                    ids[int(an["_Netcdf4Dimid"])] = n
            return tuple(ids[int(i)] for i in np.atleast_1d(a["_Netcdf4Coordinates"]))
        if not self.shape(name):
            return ()
        raise H5Error(f"'{name}': dimension names are not recorded in a form this reader decodes")

    @_guarded
    def is_pure_dimension(self, name: str) -> bool:
        """A netCDF dimension without a coordinate variable (libnetcdf stores a placeholder dataset)."""
        return str(self.attrs(name).get("NAME", "")).startswith("This is a netCDF dimension but not a netCDF variable")

    @_guarded
    def attrs(self, name: str) -> Dict[str, object]:
        """Attributes of a variable stored in its object header with a numeric or fixed-length string type."""
        if name not in self.vars:
            raise KeyError(f"{self.path}: no variable '{name}'")
        return self._object(self.vars[name]).attrs

    @_guarded
    def get(self, name: str) -> np.ndarray:
        """The variable in file (C) dimension order.  Fixed-length strings come back as bytes arrays ('S<n>')."""
        d = self._dataset(name)
        n = int(np.prod(d.shape)) if d.shape else 1
        nbytes = n * d.dtype.itemsize
        kind = d.layout[0]
        if kind == "contiguous":
            _, addr, size = d.layout
            raw = self._fill(d, nbytes) if addr == UNDEF else self.b[addr:addr + nbytes]
        elif kind == "compact":
            raw = d.layout[1][:nbytes]
        else:
            return self._read_chunked(d)
        if len(raw) < nbytes:
            raise H5Error(f"{name}: short read")
        if d.vlen_string:
            return self._vlen_strings(raw, n).reshape(d.shape)
        return np.frombuffer(raw, dtype=d.dtype, count=n).reshape(d.shape).copy()

    # ------------------------------------------------------------------ objects
    def _dataset(self, name: str) -> _Dataset:
        if name not in self.vars:
            raise KeyError(f"{self.path}: no variable '{name}'")
        d = self._object(self.vars[name])
        if not d.is_dataset:
            raise H5Error(f"'{name}' is a group, not a variable (netCDF-4 groups are not supported)")
        if d.dtype is None or d.layout is None:
            raise H5Error(f"'{name}': datatype class {d.unsupported} (variable-length / user-defined) is not supported"
                          if d.unsupported is not None else f"'{name}': incomplete object header")
        return d

    def _object(self, addr: int) -> _Dataset:
        d = _Dataset()
        if self.b[addr:addr + 4] == b"OHDR":
            self._ohdr_v2(addr, d)
        elif self.b[addr] == 1:
            self._ohdr_v1(addr, d)
        else:
            raise H5Error(f"object header at {addr}: unknown version")
        return d

    def _ohdr_v1(self, addr: int, d: _Dataset) -> None:
        _, _, nmsg, _, hsize = struct.unpack_from("<BBHII", self.b, addr)
        blocks = [(addr + 16, hsize)]
        seen = 0
        while blocks and seen < nmsg:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end and seen < nmsg:
                mtype, msize, mflags = struct.unpack_from("<HHB", self.b, p)
                body = p + 8
                self._message(mtype, body, msize, d, blocks)
                p = body + msize
                seen += 1

    def _ohdr_v2(self, addr: int, d: _Dataset) -> None:
        flags = self.b[addr + 5]
        p = addr + 6
        if flags & 0x20:
            p += 16
        if flags & 0x10:
            p += 4
        w = 1 << (flags & 3)
        size = int.from_bytes(self.b[p:p + w], "little")
        p += w
        track_order = bool(flags & 0x04)
        blocks = [(p, size)]
        while blocks:
            p, size = blocks.pop(0)
            end = p + size                       # messages end where the gap/checksum starts (checksum after `size`)
            while p + 4 <= end:
                mtype, msize, mflags = struct.unpack_from("<BHB", self.b, p)
                p += 4 + (2 if track_order else 0)
                if p + msize > end:
                    break
                self._message(mtype, p, msize, d, blocks, v2=True)
                p += msize

    def _message(self, mtype: int, p: int, size: int, d: _Dataset, blocks: list, v2: bool = False) -> None:
        b = self.b
        if mtype == 0x01:                        # dataspace
            ver, rank, flags = b[p], b[p + 1], b[p + 2]
            q = p + (8 if ver == 1 else 4)
            d.shape = tuple(struct.unpack_from("<Q", b, q + 8 * i)[0] for i in range(rank))
            d.is_dataset = True
        elif mtype == 0x03:                      # datatype
            d.dtype = self._datatype(p)
            if d.dtype is None:
                if (b[p] & 0x0F) == 9 and (b[p + 1] & 0x0F) == 1:      # variable-length string (netCDF NC_STRING)
                    d.vlen_string, d.dtype = True, np.dtype("V16")
                else:
                    d.unsupported = b[p] & 0x0F
        elif mtype == 0x05:                      # fill value
            ver = b[p]
            if ver in (1, 2):
                defined = b[p + 3]
                if ver == 1 or defined:
                    n = struct.unpack_from("<I", b, p + 4)[0]
                    d.fill = b[p + 8:p + 8 + n]
            elif ver == 3:
                if b[p + 1] & 0x20:
                    n = struct.unpack_from("<I", b, p + 2)[0]
                    d.fill = b[p + 6:p + 6 + n]
        elif mtype == 0x08:                      # data layout
            ver = b[p]
            if ver == 3:
                cls = b[p + 1]
                if cls == 0:
                    n = struct.unpack_from("<H", b, p + 2)[0]
                    d.layout = ("compact", b[p + 4:p + 4 + n])
                elif cls == 1:
                    addr, sz = struct.unpack_from("<QQ", b, p + 2)
                    d.layout = ("contiguous", addr, sz)
                elif cls == 2:
                    nd = b[p + 2]
                    addr = struct.unpack_from("<Q", b, p + 3)[0]
                    dims = struct.unpack_from("<" + "I" * nd, b, p + 11)
                    d.layout = ("chunked", addr, tuple(dims[:-1]))
                else:
                    raise H5Error(f"data layout class {cls} is not supported")
            elif ver in (1, 2):
                nd, cls = b[p + 1], b[p + 2]
                q = p + 8
                addr = UNDEF
                if cls != 0:
                    addr = struct.unpack_from("<Q", b, q)[0]
                    q += 8
                dims = struct.unpack_from("<" + "I" * nd, b, q)
                q += 4 * nd
                if cls == 1:
                    d.layout = ("contiguous", addr, 0)
                elif cls == 2:
                    d.layout = ("chunked", addr, tuple(dims[:-1]))
                else:
                    n = struct.unpack_from("<I", b, q)[0]
                    d.layout = ("compact", b[q + 4:q + 4 + n])
            else:
                raise H5Error(f"data layout message version {ver} (HDF5 >= 1.10 chunk indexes) is not supported")
        elif mtype == 0x0B:                      # filter pipeline
            ver, nf = b[p], b[p + 1]
            q = p + (8 if ver == 1 else 2)
            for _ in range(nf):
                fid = struct.unpack_from("<H", b, q)[0]
                if ver == 1 or fid >= 256:
                    nlen = struct.unpack_from("<H", b, q + 2)[0]
                    q += 2
                else:
                    nlen = 0
                flags, ncd = struct.unpack_from("<HH", b, q + 2)
                q += 6
                if ver == 1:
                    nlen = (nlen + 7) // 8 * 8
                q += nlen + 4 * ncd
                if ver == 1 and ncd % 2:
                    q += 4
                d.filters.append(fid)
        elif mtype == 0x0C:                      # attribute (compact storage)
            self._attribute(p, d)
        elif mtype == 0x15:                      # attribute info: attributes in dense storage
            flags = b[p + 1]
            q = p + 2 + (2 if flags & 1 else 0)
            heap, index = struct.unpack_from("<QQ", b, q)
            if heap != UNDEF:
                for at, _ in self._heap_objects(heap, index, 8, 0):
                    self._attribute(at, d)
        elif mtype == 0x06:                      # link
            name, addr = self._link(p)
            if addr is not None:
                d.links[name] = addr
        elif mtype == 0x02:                      # link info (dense storage)
            flags = b[p + 1]
            q = p + 2 + (8 if flags & 1 else 0)
            heap, index = struct.unpack_from("<QQ", b, q)
            if heap != UNDEF:
                d.link_heap = (heap, index)
        elif mtype == 0x11:                      # symbol table (old-style group)
            d.symtab = struct.unpack_from("<QQ", b, p)
        elif mtype == 0x10:                      # continuation
            off, ln = struct.unpack_from("<QQ", b, p)
            if v2:
                if b[off:off + 4] != b"OCHK":
                    raise H5Error("bad object header continuation block")
                blocks.append((off + 4, ln - 8))
            else:
                blocks.append((off, ln))

    def _attribute(self, p: int, d: _Dataset) -> None:
        b = self.b
        ver = b[p]
        nsz, tsz, ssz = struct.unpack_from("<HHH", b, p + 2)
        q = p + 8 + (1 if ver == 3 else 0)
        pad = (lambda n: (n + 7) // 8 * 8) if ver == 1 else (lambda n: n)
        name = b[q:q + nsz].split(b"\0")[0].decode("utf-8", "replace"); q += pad(nsz)
        dt = self._datatype(q); tq = q; q += pad(tsz)
        sver, rank = b[q], b[q + 1]
        sq = q + (8 if sver == 1 else 4)
        shape = tuple(struct.unpack_from("<Q", b, sq + 8 * i)[0] for i in range(rank))
        q += pad(ssz)
        if dt is None:
            if name == "DIMENSION_LIST" and (b[tq] & 0x0F) == 9:
                # one variable-length sequence of object references per dimension: the dimension scales' headers
                n = int(np.prod(shape)) if shape else 1
                refs = []
                for i in range(n):
                    ln, gaddr, gidx = struct.unpack_from("<IQI", b, q + 16 * i)
                    refs.append(struct.unpack_from("<Q", self._global_heap_object(gaddr, gidx), 0)[0] if ln else None)
                d.attrs["DIMENSION_LIST"] = refs
            return                                   # other variable-length strings / references: skipped
        n = int(np.prod(shape)) if shape else 1
        a = np.frombuffer(b[q:q + n * dt.itemsize], dtype=dt, count=n)
        if dt.kind == "S":
            v = [x.split(b"\0")[0].decode("utf-8", "replace") for x in a.tolist()]
            d.attrs[name] = v[0] if n == 1 else v
        else:
            d.attrs[name] = a.reshape(shape).copy() if shape else a[0]

    def _vlen_strings(self, raw: bytes, n: int) -> np.ndarray:
        out = np.empty(n, dtype=object)
        for i in range(n):
            ln, gaddr, gidx = struct.unpack_from("<IQI", raw, 16 * i)
            out[i] = self._global_heap_object(gaddr, gidx)[:ln].decode("utf-8", "replace") if ln else ""
        return out

    def _global_heap_object(self, addr: int, index: int) -> bytes:
        b = self.b
        if b[addr:addr + 4] != b"GCOL":
            raise H5Error("bad global heap collection")
        size = struct.unpack_from("<Q", b, addr + 8)[0]
        q, end = addr + 16, addr + size
        while q + 16 <= end:
            idx, _, _, osz = struct.unpack_from("<HHIQ", b, q)
            if idx == 0:
                break
            if idx == index:
                return b[q + 16:q + 16 + osz]
            q += 16 + (osz + 7) // 8 * 8
        raise H5Error("global heap object not found")

    def _datatype(self, p: int) -> Optional[np.dtype]:
        b = self.b
        cls = b[p] & 0x0F
        bits0 = b[p + 1]
        size = struct.unpack_from("<I", b, p + 4)[0]
        order = ">" if bits0 & 1 else "<"
        if cls == 0:
            return np.dtype(f"{order}{'i' if bits0 & 0x08 else 'u'}{size}")
        if cls == 1:
            return np.dtype(f"{order}f{size}")
        if cls == 3:
            return np.dtype(f"S{size}")
        return None                             # reported when (and only when) the variable is read

    # ------------------------------------------------------------------ groups
    def _link(self, p: int) -> Tuple[str, Optional[int]]:
        b = self.b
        flags = b[p + 1]
        q = p + 2
        ltype = 0
        if flags & 0x08:
            ltype = b[q]; q += 1
        if flags & 0x04:
            q += 8
        if flags & 0x10:
            q += 1
        w = 1 << (flags & 3)
        n = int.from_bytes(b[q:q + w], "little"); q += w
        name = b[q:q + n].decode(); q += n
        if ltype != 0:
            return name, None                   # soft / external links: not followed
        return name, struct.unpack_from("<Q", b, q)[0]

    def _group_links(self, g: _Dataset):
        yield from g.links.items()
        if g.link_heap is not None:
            yield from self._dense_links(*g.link_heap)
        if g.symtab is not None:
            yield from self._symtab_links(*g.symtab)

    def _dense_links(self, heap: int, name_index: int):
        """Link messages of a group in "dense" storage: the version-2 B-tree `name_index` (record type 5: name hash +
        heap id) lists the managed objects of the fractal heap `heap`; each is one link message."""
        for at, ln in self._heap_objects(heap, name_index, 5, 4):
            name, addr = self._link(at)
            if addr is not None:
                yield name, addr

    def _heap_objects(self, heap: int, name_index: int, want_type: int, hid_offset: int):
        """(file address, length) of the managed fractal-heap objects a version-2 B-tree indexes; the heap id sits
        `hid_offset` bytes into each record (type 5, links: after the hash; type 8, attributes: first)."""
        b = self.b
        if b[heap:heap + 4] != b"FRHP":
            raise H5Error("bad fractal heap header")
        p = heap + 5
        idlen, filt_len, flags = struct.unpack_from("<HHB", b, p)
        if filt_len:
            raise H5Error("filtered fractal heaps are not supported")
        p += 5 + 4                               # + maximum size of managed objects
        p += 8 * 12                              # next huge id, huge btree, free space, fs manager, 8 statistics fields
        width, start_size, max_direct, max_heap_bits, start_rows = struct.unpack_from("<HQQHH", b, p)
        p += 2 + 8 + 8 + 2 + 2
        root, cur_rows = struct.unpack_from("<QH", b, p)
        off_bytes = (max_heap_bits + 7) // 8
        len_bytes = min(idlen - 1 - off_bytes, 8)
        if root == UNDEF:
            return

        # (heap offset of the block, size, file address) of every direct block
        blocks = []

        def row_size(r):
            return start_size if r < 2 else start_size << (r - 1)

        def direct(addr, size):
            if b[addr:addr + 4] != b"FHDB":
                raise H5Error("bad fractal heap direct block")
            boff = int.from_bytes(b[addr + 13:addr + 13 + off_bytes], "little")
            blocks.append((boff, size, addr))

        if cur_rows == 0:
            direct(root, start_size)
        else:
            if b[root:root + 4] != b"FHIB":
                raise H5Error("bad fractal heap indirect block")
            q = root + 5 + 8 + off_bytes
            max_direct_rows = 2 + (int(max_direct).bit_length() - int(start_size).bit_length())
            for r in range(cur_rows):
                for _ in range(width):
                    child = struct.unpack_from("<Q", b, q)[0]; q += 8
                    if child == UNDEF:
                        continue
                    if r >= max_direct_rows:
                        raise H5Error("nested indirect fractal heap blocks are not supported")
                    direct(child, row_size(r))

        def managed(hid: bytes):
            if (hid[0] >> 4) & 3 != 0:
                raise H5Error("huge / tiny fractal heap objects are not supported")
            off = int.from_bytes(hid[1:1 + off_bytes], "little")
            ln = int.from_bytes(hid[1 + off_bytes:1 + off_bytes + len_bytes], "little")
            for boff, size, addr in blocks:
                if boff <= off < boff + size:
                    return addr + (off - boff), ln
            raise H5Error("fractal heap object outside every direct block")

        # version-2 B-tree, record type 5
        if b[name_index:name_index + 4] != b"BTHD":
            raise H5Error("bad version-2 B-tree header")
        btype, node_size, rec_size, depth = struct.unpack_from("<BIHH", b, name_index + 5)
        root_node, nroot = struct.unpack_from("<QH", b, name_index + 16)
        if btype != want_type or rec_size < hid_offset + idlen:
            raise H5Error("unexpected version-2 B-tree record type")
        leaf_max = (node_size - 10) // rec_size
        nrec_bytes = (int(leaf_max).bit_length() + 7) // 8

        def records(addr, n):
            return [b[addr + 6 + i * rec_size + hid_offset: addr + 6 + i * rec_size + hid_offset + idlen] for i in range(n)]

        def walk(addr, n, level):
            if level == 0:
                if b[addr:addr + 4] != b"BTLF":
                    raise H5Error("bad version-2 B-tree leaf")
                yield from records(addr, n)
                return
            if level > 1:
                raise H5Error("version-2 B-trees deeper than 1 are not supported")
            if b[addr:addr + 4] != b"BTIN":
                raise H5Error("bad version-2 B-tree internal node")
            yield from records(addr, n)
            q = addr + 6 + n * rec_size
            for _ in range(n + 1):
                child = struct.unpack_from("<Q", b, q)[0]
                cn = int.from_bytes(b[q + 8:q + 8 + nrec_bytes], "little")
                q += 8 + nrec_bytes
                yield from walk(child, cn, level - 1)

        if root_node == UNDEF:
            return
        for hid in walk(root_node, nroot, depth):
            yield managed(hid)

    def _symtab_links(self, btree: int, heap: int):
        b = self.b
        if b[heap:heap + 4] != b"HEAP":
            raise H5Error("bad local heap")
        data = struct.unpack_from("<Q", b, heap + 24)[0]

        def name_at(off):
            e = b.index(b"\0", data + off)
            return b[data + off:e].decode()

        def node(addr):
            if b[addr:addr + 4] == b"TREE":
                ntype, level, used = struct.unpack_from("<BBH", b, addr + 4)
                if ntype != 0:
                    raise H5Error("group B-tree expected")
                q = addr + 8 + 16
                for i in range(used):
                    child = struct.unpack_from("<Q", b, q + 8)[0]     # key(L) child(O)
                    q += 16
                    yield from node(child)
            elif b[addr:addr + 4] == b"SNOD":
                nsym = struct.unpack_from("<H", b, addr + 6)[0]
                q = addr + 8
                for _ in range(nsym):
                    noff, ohdr = struct.unpack_from("<QQ", b, q)
                    q += 40
                    yield name_at(noff), ohdr
            else:
                raise H5Error("bad group B-tree node")
        yield from node(btree)

    # ------------------------------------------------------------------ data
    def _fill(self, d: _Dataset, nbytes: int) -> bytes:
        if d.fill and len(d.fill) == d.dtype.itemsize:
            return d.fill * (nbytes // d.dtype.itemsize)
        return bytes(nbytes)

    def _read_chunked(self, d: _Dataset) -> np.ndarray:
        _, btree, cdims = d.layout
        for f in d.filters:
            if f not in (1, 2, 3):
                raise H5Error(f"filter {f} is not supported (deflate, shuffle, fletcher32 are)")
        shape = d.shape
        nd = len(shape)
        isz = d.dtype.itemsize
        out = np.frombuffer(self._fill(d, int(np.prod(shape)) * isz), dtype=d.dtype).reshape(shape).copy()
        if btree == UNDEF:
            return out
        b = self.b
        cbytes = int(np.prod(cdims)) * isz

        def node(addr):
            if b[addr:addr + 4] != b"TREE":
                raise H5Error("bad chunk B-tree node")
            ntype, level, used = struct.unpack_from("<BBH", b, addr + 4)
            if ntype != 1:
                raise H5Error("chunk B-tree expected")
            q = addr + 8 + 16
            ksz = 8 + 8 * (nd + 1)
            for _ in range(used):
                csize, mask = struct.unpack_from("<II", b, q)
                offs = struct.unpack_from("<" + "Q" * nd, b, q + 8)
                child = struct.unpack_from("<Q", b, q + ksz)[0]
                q += ksz + 8
                if level > 0:
                    node(child)
                    continue
                raw = b[child:child + csize]
                for k, f in reversed(list(enumerate(d.filters))):
                    if mask & (1 << k):
                        continue
                    if f == 3:
                        raw = raw[:-4]
                    elif f == 1:
                        raw = zlib.decompress(raw)
                    elif f == 2:
                        n = len(raw) // isz
                        raw = np.frombuffer(raw, np.uint8).reshape(isz, n).T.tobytes()
                if len(raw) < cbytes:
                    raise H5Error("short chunk")
                chunk = np.frombuffer(raw, dtype=d.dtype, count=cbytes // isz).reshape(cdims)
                sl_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, shape))
                sl_in = tuple(slice(0, s.stop - s.start) for s in sl_out)
                out[sl_out] = chunk[sl_in]
        node(btree)
        return out


def read_all(path: str) -> Dict[str, np.ndarray]:
    f = File(path)
    out = {}
    for n in f.names():
        try:
            out[n] = f.get(n)
        except H5Error:
            raise
    return out
