"""Keras `.h5` weight files for the verification head, without h5py.

The reference checkpoints its heads with `siamese_net.save_weights(modelName + ".h5")` and reloads them with
`load_weights` (/root/reference/code/siamese.py:114-125, code/siamese3.py:88-99). Keras 2.1.2 writes, through
h5py/libhdf5 [external: keras/engine/topology.py `save_weights_to_hdf5_group`]:

    /            attrs  layer_names = [b"input_1", ..., b"dense_3"]  (fixed-length byte strings),
                        backend, keras_version
    /<layer>     attrs  weight_names = [b"dense_1/kernel:0", b"dense_1/bias:0"]  (empty for weightless layers)
    /<layer>/<weight name>   float32 dataset (the "/" in the weight name nests one more group)

and a full `model.save()` file holds the same tree under `/model_weights`. h5py and libhdf5 are absent from
this image, so this module is a small reader (and writer) of the subset of the HDF5 file format [external:
"HDF5 File Format Specification Version 2.0/3.0", The HDF Group] that such files use with libhdf5's default
`libver="earliest"` settings:

    superblock v0/v1 (v2/v3 accepted), symbol-table groups (B-tree v1 + local heap + SNOD), version-1
    object headers with continuation blocks, dataspace v1/v2, datatypes: fixed-point, IEEE float, fixed-
    length strings, variable-length strings (global heap), data layout v3 compact / contiguous / chunked
    without filters (v1 chunk B-tree), attribute messages v1/v2/v3.

Filtered (compressed) datasets, v2 object headers and link-message groups (libver="latest") raise
`H5FormatError` naming the feature (data layout messages v1/v2 of libhdf5 <= 1.6 are read too). PARITY NOTE: the
reader is pinned against ONE file written by a real libhdf5 (tests/golden/libhdf5_written_testdouble.mat, SciPy's
MATLAB-7.3 fixture: user block, symbol-table group, v1 object header, float dataset, string attribute). No
Keras/h5py-written file exists in this image: the Keras tree layout is restated from Keras 2.1.2 and round-trips
only against `write_keras_weights` below (same specification, many structural variants), so cross-library
compatibility of the WRITER is unverified.
"""
from __future__ import annotations

import struct

import numpy as np

SIG = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class H5FormatError(ValueError):
    pass


def _pad8(n):
    return (n + 7) & ~7


# ------------------------------------------------------------------------------------------ reader
class _Type:
    __slots__ = ("cls", "size", "dtype", "vlen_str", "strpad")

    def __init__(self, cls, size, dtype=None, vlen_str=False, strpad=0):
        self.cls, self.size, self.dtype, self.vlen_str, self.strpad = cls, size, dtype, vlen_str, strpad


class H5Object:
    """A group or dataset: `.attrs` (dict), `.children` (dict, groups) or `.value` (ndarray, datasets)."""

    def __init__(self):
        self.attrs = {}
        self.children = None
        self.value = None

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            if node.children is None or part not in node.children:
                raise KeyError(path)
            node = node.children[part]
        return node

    def __contains__(self, path):
        try:
            self[path]
            return True
        except KeyError:
            return False


class H5File:
    """Read-only view of the HDF5 subset described in the module docstring. `root` is an H5Object."""

    def __init__(self, path_or_bytes):
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            self.buf = bytes(path_or_bytes)
        else:
            with open(path_or_bytes, "rb") as f:
                self.buf = f.read()
        self._gcol = {}
        self._seen = {}
        self.root = self._superblock()

    # -- primitives
    def _u(self, off, n):
        if off < 0 or off + n > len(self.buf):
            raise H5FormatError(f"read of {n} bytes at {off} past the end of the file ({len(self.buf)} bytes)")
        return int.from_bytes(self.buf[off:off + n], "little")

    def _bytes(self, off, n):
        if off < 0 or off + n > len(self.buf):
            raise H5FormatError(f"read of {n} bytes at {off} past the end of the file ({len(self.buf)} bytes)")
        return self.buf[off:off + n]

    def _superblock(self):
        base = 0
        while base < len(self.buf) and self.buf[base:base + 8] != SIG:
            base = 512 if base == 0 else base * 2  # the signature may sit at 0, 512, 1024, ...
        if base >= len(self.buf):
            raise H5FormatError("not an HDF5 file (signature not found)")
        ver = self._u(base + 8, 1)
        if ver in (0, 1):
            self.O, self.L = self._u(base + 13, 1), self._u(base + 14, 1)
            p = base + 24 + (4 if ver == 1 else 0)
            self.base = self._u(p, self.O)
            ent = p + 4 * self.O  # base, free-space, end-of-file, driver-info addresses
            ohdr = self._u(ent + self.O, self.O)
        elif ver in (2, 3):
            self.O, self.L = self._u(base + 9, 1), self._u(base + 10, 1)
            self.base = self._u(base + 12, self.O)
            ohdr = self._u(base + 12 + 3 * self.O, self.O)
        else:
            raise H5FormatError(f"superblock version {ver} not supported")
        if self.O != 8 or self.L != 8:
            raise H5FormatError(f"offset/length sizes {self.O}/{self.L} not supported (8/8 expected)")
        return self._object(ohdr)

    # -- object headers
    def _messages(self, addr):
        a = addr + self.base
        if self._bytes(a, 4) == b"OHDR":
            raise H5FormatError("version-2 object headers (libver='latest') not supported")
        ver = self._u(a, 1)
        if ver != 1:
            raise H5FormatError(f"object header version {ver} at {addr} not supported")
        nmsg, size = self._u(a + 2, 2), self._u(a + 8, 4)
        blocks = [(a + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            p, n = blocks.pop(0)
            end = p + n
            while p + 8 <= end and len(out) < nmsg:
                mtype, msize, mflags = self._u(p, 2), self._u(p + 2, 2), self._u(p + 4, 1)
                body = p + 8
                if mflags & 2:
                    raise H5FormatError("shared header messages not supported")
                if mtype == 0x10:
                    blocks.append((self._u(body, 8) + self.base, self._u(body + 8, 8)))
                out.append((mtype, body, msize))
                p = body + msize
        return out

    def _object(self, addr):
        if addr in self._seen:
            return self._seen[addr]
        obj = H5Object()
        self._seen[addr] = obj
        space = dtype = layout = None
        for mtype, body, msize in self._messages(addr):
            if mtype == 0x11:  # symbol table: old-style group
                obj.children = {}
                self._btree_group(self._u(body, 8), self._u(body + 8, 8), obj.children)
            elif mtype in (0x02, 0x06):
                raise H5FormatError("link-message groups (libver='latest') not supported")
            elif mtype == 0x01:
                space = self._dataspace(body)
            elif mtype == 0x03:
                dtype = self._datatype(body)[0]
            elif mtype == 0x08:
                layout = (body, msize)
            elif mtype == 0x0B:
                raise H5FormatError("filtered (compressed) datasets not supported")
            elif mtype == 0x0C:
                name, val = self._attribute(body)
                obj.attrs[name] = val
            elif mtype == 0x15:
                raise H5FormatError("dense attribute storage not supported")
        if layout is not None:
            if space is None or dtype is None:
                raise H5FormatError(f"dataset at {addr} lacks a dataspace or datatype message")
            obj.value = self._dataset(layout[0], space, dtype)
        return obj

    def _btree_group(self, btree, heap, out):
        h = heap + self.base
        if self._bytes(h, 4) != b"HEAP":
            raise H5FormatError(f"local heap signature missing at {heap}")
        seg = self._u(h + 8 + 2 * self.L, self.O) + self.base

        def name_at(off):
            end = self.buf.index(b"\0", seg + off)
            return self.buf[seg + off:end].decode("utf8")

        def node(addr):
            a = addr + self.base
            sig = self._bytes(a, 4)
            if sig == b"TREE":
                ntype, used = self._u(a + 4, 1), self._u(a + 6, 2)
                if ntype != 0:
                    raise H5FormatError("B-tree node of the wrong type under a group")
                p = a + 8 + 2 * self.O
                for i in range(used):
                    node(self._u(p + self.L + i * (self.L + self.O), self.O))
            elif sig == b"SNOD":
                n = self._u(a + 6, 2)
                for i in range(n):
                    e = a + 8 + i * (2 * self.O + 24)
                    out[name_at(self._u(e, self.O))] = self._object(self._u(e + self.O, self.O))
            else:
                raise H5FormatError(f"expected TREE or SNOD at {addr}")

        node(btree)

    # -- messages
    def _dataspace(self, p):
        ver, rank, flags = self._u(p, 1), self._u(p + 1, 1), self._u(p + 2, 1)
        if ver == 1:
            q = p + 8
        elif ver == 2:
            if self._u(p + 3, 1) == 2:
                return None  # null dataspace
            q = p + 4
        else:
            raise H5FormatError(f"dataspace version {ver} not supported")
        return tuple(self._u(q + i * self.L, self.L) for i in range(rank))

    def _datatype(self, p):
        """-> (_Type, bytes consumed)."""
        cv, b0, size = self._u(p, 1), self._u(p + 1, 1), self._u(p + 4, 4)
        cls = cv & 15
        order = ">" if b0 & 1 else "<"
        if cls == 0:
            signed = "i" if b0 & 8 else "u"
            return _Type(0, size, np.dtype(f"{order}{signed}{size}")), 8 + 4
        if cls == 1:
            if size not in (2, 4, 8):
                raise H5FormatError(f"{size}-byte floats not supported")
            return _Type(1, size, np.dtype(f"{order}f{size}")), 8 + 12
        if cls == 3:
            return _Type(3, size, np.dtype(f"S{size}"), strpad=b0 & 15), 8
        if cls == 9:
            base, n = self._datatype(p + 8)
            if (b0 & 15) != 1:
                raise H5FormatError("variable-length sequences (non-string) not supported")
            return _Type(9, size, None, vlen_str=True), 8 + n
        raise H5FormatError(f"datatype class {cls} not supported")

    def _global_heap_object(self, addr, index):
        if addr not in self._gcol:
            a = addr + self.base
            if self._bytes(a, 4) != b"GCOL":
                raise H5FormatError(f"global heap signature missing at {addr}")
            total = self._u(a + 8, self.L)
            objs, p = {}, a + 8 + self.L
            while p + 8 + self.L <= a + total:
                idx, size = self._u(p, 2), self._u(p + 8, self.L)
                if idx == 0:
                    break
                objs[idx] = self._bytes(p + 8 + self.L, size)
                p += 8 + self.L + _pad8(size)
            self._gcol[addr] = objs
        try:
            return self._gcol[addr][index]
        except KeyError:
            raise H5FormatError(f"global heap object {index} missing in the collection at {addr}") from None

    def _decode(self, raw, shape, t):
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if shape is None:
            return np.zeros((0,), np.float64)
        if t.vlen_str:
            vals = []
            for i in range(n):
                q = i * t.size
                ln = int.from_bytes(raw[q:q + 4], "little")
                addr = int.from_bytes(raw[q + 4:q + 12], "little")
                idx = int.from_bytes(raw[q + 12:q + 16], "little")
                vals.append(self._global_heap_object(addr, idx)[:ln] if ln else b"")
            arr = np.array(vals, dtype=object).reshape(shape)
        else:
            arr = np.frombuffer(raw, dtype=t.dtype, count=n).reshape(shape)
            if t.cls == 3:
                arr = np.char.rstrip(arr, b" ") if t.strpad == 2 else arr  # numpy strips the NUL padding itself
            elif t.dtype.byteorder == ">":
                arr = arr.astype(t.dtype.newbyteorder("<"))
            arr = arr.copy()
        return arr if shape else arr.reshape(())[()]

    def _attribute(self, p):
        ver = self._u(p, 1)
        nsz, tsz, ssz = self._u(p + 2, 2), self._u(p + 4, 2), self._u(p + 6, 2)
        if ver == 1:
            q = p + 8
            name = self._bytes(q, nsz)
            q += _pad8(nsz)
            t = self._datatype(q)[0]
            q += _pad8(tsz)
            shape = self._dataspace(q)
            q += _pad8(ssz)
        elif ver in (2, 3):
            if self._u(p + 1, 1) & 3:
                raise H5FormatError("shared attribute datatypes/dataspaces not supported")
            q = p + 8 + (1 if ver == 3 else 0)
            name = self._bytes(q, nsz)
            q += nsz
            t = self._datatype(q)[0]
            q += tsz
            shape = self._dataspace(q)
            q += ssz
        else:
            raise H5FormatError(f"attribute message version {ver} not supported")
        name = name.split(b"\0", 1)[0].decode("utf8")
        if shape is None:
            return name, np.zeros((0,), np.float64)
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        return name, self._decode(self._bytes(q, n * t.size), shape, t)

    def _dataset(self, p, shape, t):
        ver, cls = self._u(p, 1), self._u(p + 1, 1)
        if shape is None:
            return np.zeros((0,), t.dtype or object)
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if ver in (1, 2):  # libhdf5 <= 1.6 (e.g. MATLAB 7.3 files): version, rank, class, 5 reserved bytes
            ndim, cls = self._u(p + 1, 1), self._u(p + 2, 1)
            if cls == 1:
                addr = self._u(p + 8, self.O)
                raw = bytes(n * t.size) if addr == UNDEF else self._bytes(addr + self.base, n * t.size)
            elif cls == 0:
                q = p + 8 + (self.O if ver == 1 else 0) + 4 * ndim
                if self._u(q, 4) < n * t.size:
                    raise H5FormatError("compact dataset shorter than its dataspace")
                raw = self._bytes(q + 4, n * t.size)
            else:
                raise H5FormatError("chunked datasets not supported")
            return self._decode(raw, shape, t)
        if ver != 3:
            raise H5FormatError(f"data layout message version {ver} not supported")
        if cls == 0:
            size = self._u(p + 2, 2)
            if size < n * t.size:
                raise H5FormatError("compact dataset shorter than its dataspace")
            raw = self._bytes(p + 4, n * t.size)
        elif cls == 1:
            addr, size = self._u(p + 2, self.O), self._u(p + 2 + self.O, self.L)
            if addr == UNDEF:
                raw = bytes(n * t.size)  # never written: fill value (zeros)
            else:
                if size < n * t.size:
                    raise H5FormatError("contiguous dataset shorter than its dataspace")
                raw = self._bytes(addr + self.base, n * t.size)
        elif cls == 2:
            ndim = self._u(p + 2, 1)  # rank + 1: the last "dimension" is the element size
            btree = self._u(p + 3, self.O)
            cdims = tuple(self._u(p + 3 + self.O + 4 * i, 4) for i in range(ndim - 1))
            if len(cdims) != len(shape):
                raise H5FormatError("chunk rank differs from the dataspace rank")
            raw = self._chunked(btree, shape, cdims, t)
        else:
            raise H5FormatError(f"data layout class {cls} not supported")
        return self._decode(raw, shape, t)

    def _chunked(self, btree, shape, cdims, t):
        """Unfiltered chunked storage: walk the version-1 B-tree (node type 1) and paste every chunk (stored at
        its full chunk size, cropped at the dataset's edges) into a zero-filled array."""
        out = np.zeros(shape, dtype=np.dtype(f"V{t.size}"))
        csize = int(np.prod(cdims, dtype=np.int64)) * t.size
        rank = len(shape)

        def node(addr):
            a = addr + self.base
            if self._bytes(a, 4) != b"TREE" or self._u(a + 4, 1) != 1:
                raise H5FormatError(f"expected a chunk B-tree node at {addr}")
            level, used = self._u(a + 5, 1), self._u(a + 6, 2)
            q = a + 8 + 2 * self.O
            klen = 8 + 8 * (rank + 1)
            for i in range(used):
                k = q + i * (klen + self.O)
                nbytes, fmask = self._u(k, 4), self._u(k + 4, 4)
                off = tuple(self._u(k + 8 + 8 * d, 8) for d in range(rank))
                child = self._u(k + klen, self.O)
                if level > 0:
                    node(child)
                    continue
                if fmask or nbytes != csize:
                    raise H5FormatError("filtered (compressed) chunks not supported")
                chunk = np.frombuffer(self._bytes(child + self.base, csize), dtype=out.dtype).reshape(cdims)
                sl_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(off, cdims, shape))
                sl_in = tuple(slice(0, s.stop - s.start) for s in sl_out)
                out[sl_out] = chunk[sl_in]

        if btree != UNDEF and out.size:
            node(btree)
        return out.tobytes()


def _names(attr):
    if isinstance(attr, (bytes, str)):
        attr = [attr]
    return [a.decode("utf8") if isinstance(a, bytes) else str(a) for a in np.asarray(attr).ravel().tolist()]


def read_keras_weights(path):
    """-> list of (layer name, [(weight name, ndarray), ...]) in `layer_names` order, weightless layers
    included with an empty list — the traversal of Keras 2.1.2 `load_weights_from_hdf5_group` [external]."""
    f = H5File(path)
    root = f.root
    if "layer_names" not in root.attrs and "model_weights" in root:  # a full `model.save()` file
        root = root["model_weights"]
    if "layer_names" not in root.attrs:
        raise H5FormatError("no 'layer_names' attribute: not a Keras weight file")
    out = []
    for lname in _names(root.attrs["layer_names"]):
        g = root[lname]
        wn = g.attrs.get("weight_names", [])
        ws = []
        for w in (_names(wn) if np.size(wn) else []):
            v = g[w].value
            if v is None:
                raise H5FormatError(f"{lname}/{w} is not a dataset")
            ws.append((w, v))
        out.append((lname, ws))
    return out


def load_head_weights(path, D_in, H1, H2, C):
    """The six arrays [W1, b1, W2, b2, W3, b3] of the Dense(H1)-Dense(H2)-Dense(C) head
    (code/siamese.py:29-32, code/siamese3.py:23-25) from a Keras `.h5` weight file. The head is the last
    three (kernel, bias) pairs of the file (encoder layers, when the file has them, come first); their
    shapes must chain D_in -> H1 -> H2 -> C, as Keras' own shape check in `load_weights` would demand."""
    layers = read_keras_weights(path)
    flat = [v for _, ws in layers for _, v in ws]
    if len(flat) < 6:
        raise H5FormatError(f"{len(flat)} weight arrays in the file, the head needs 6")
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in flat[-6:]]
    # prefer the LAST three layers whose weights are a (kernel, bias) pair chaining D_in -> H1 -> H2 -> C, wherever
    # they sit in layer_names (a positional "last six arrays" would be wrong if a weighted layer followed the head)
    pairs = [[np.ascontiguousarray(v, dtype=np.float32) for _, v in ws] for _, ws in layers
             if len(ws) == 2 and np.ndim(ws[0][1]) == 2 and np.ndim(ws[1][1]) == 1]
    for i in range(len(pairs) - 3, -1, -1):
        cand = pairs[i] + pairs[i + 1] + pairs[i + 2]
        if [a.shape for a in cand] == [(D_in, H1), (H1,), (H1, H2), (H2,), (H2, C), (C,)]:
            arrs = cand
            break
    want = [(D_in, H1), (H1,), (H1, H2), (H2,), (H2, C), (C,)]
    got = [a.shape for a in arrs]
    if got != want:
        raise H5FormatError(f"head shapes {got} do not match {want}")
    return arrs


# ------------------------------------------------------------------------------------------ writer
class _Writer:
    """Lays out a file the way libhdf5 (libver='earliest') does: superblock v0, symbol-table groups,
    version-1 object headers. `variant` switches structural alternatives a reader must also accept:
    attribute message version (1|2|3), dataspace version (1|2), object-header continuation blocks,
    compact datasets, variable-length string attributes, and the superblock version (0|1|2)."""

    LEAF_K, INTERNAL_K = 4, 16

    def __init__(self, attr_version=1, space_version=1, continuation=False, compact=False, vlen_names=False,
                 superblock=0, chunks=None):
        self.buf = bytearray()
        self.av, self.sv, self.cont, self.compact, self.vlen, self.sb = (
            attr_version, space_version, continuation, compact, vlen_names, superblock)
        self.chunks = chunks  # e.g. 3: every dataset of rank >= 1 is stored in chunks of (up to) 3 per dimension
        self.gheap = []  # variable-length string payloads, written as one global heap collection at the end
        self.gheap_fix = []

    def alloc(self, data):
        while len(self.buf) % 8:
            self.buf.append(0)
        off = len(self.buf)
        self.buf += data
        return off

    # -- message bodies
    def m_space(self, shape):
        if self.sv == 1:
            return struct.pack("<BBBB4x", 1, len(shape), 0, 0) + b"".join(struct.pack("<Q", d) for d in shape)
        return struct.pack("<BBBB", 2, len(shape), 0, 1 if shape else 0) + b"".join(struct.pack("<Q", d) for d in shape)

    @staticmethod
    def m_type(dt):
        dt = np.dtype(dt)
        if dt.kind == "f":
            exp, man, bias = {2: (5, 10, 15), 4: (8, 23, 127), 8: (11, 52, 1023)}[dt.itemsize]
            bits = dt.itemsize * 8
            return struct.pack("<BBBBI", 0x11, 0x20, bits - 1, 0, dt.itemsize) + struct.pack(
                "<HHBBBBI", 0, bits, man, exp, 0, man, bias)
        if dt.kind in "iu":
            return struct.pack("<BBBBI", 0x10, 8 if dt.kind == "i" else 0, 0, 0, dt.itemsize) + struct.pack(
                "<HH", 0, dt.itemsize * 8)
        if dt.kind == "S":
            return struct.pack("<BBBBI", 0x13, 1, 0, 0, max(dt.itemsize, 1))  # null-padded, ASCII
        raise TypeError(dt)

    @staticmethod
    def m_vlen_str():
        base = struct.pack("<BBBBI", 0x13, 0, 0, 0, 1)
        return struct.pack("<BBBBI", 0x19, 1, 1, 0, 16) + base  # VL string, null-terminated, UTF-8

    def m_attr(self, name, value):
        nm = name.encode("utf8") + b"\0"
        if isinstance(value, (list, tuple)) and len(value) == 0:
            value = np.zeros((0,), np.float64)  # what h5py stores for `attrs[...] = []`
        arr = np.asarray(value)
        if self.vlen and arr.dtype.kind == "S":
            t = self.m_vlen_str()
            data = bytearray()
            for s in arr.ravel().tolist():
                self.gheap.append(s)
                data += struct.pack("<I", len(s))
                self.gheap_fix.append((None, len(data)))  # patched after the message is placed
                data += struct.pack("<QI", 0, len(self.gheap))
            data = bytes(data)
        else:
            t = self.m_type(arr.dtype)
            data = arr.astype(arr.dtype.newbyteorder("<") if arr.dtype.kind in "fiu" else arr.dtype).tobytes()
        s = self.m_space(arr.shape)
        if self.av == 1:
            body = struct.pack("<BBHHH", 1, 0, len(nm), len(t), len(s))
            head = body + nm.ljust(_pad8(len(nm)), b"\0") + t.ljust(_pad8(len(t)), b"\0") + s.ljust(_pad8(len(s)), b"\0")
        else:
            body = struct.pack("<BBHHH", self.av, 0, len(nm), len(t), len(s)) + (b"\0" if self.av == 3 else b"")
            head = body + nm + t + s
        return head, data

    # -- object headers
    def object_header(self, msgs):
        """msgs: list of (type, body bytes[, vlen fix-up list]) -> address."""
        enc = []
        for m in msgs:
            body = m[1].ljust(_pad8(len(m[1])), b"\0")
            enc.append(struct.pack("<HHB3x", m[0], len(body), 0) + body)
        if self.cont and len(enc) > 1:
            first, rest = enc[:1], enc[1:]
            rest_bytes = b"".join(rest)
            hdr_msgs = b"".join(first) + struct.pack("<HHB3x", 0x10, 16, 0)
            off = self.alloc(struct.pack("<BBHII4x", 1, 0, len(enc) + 1, 1, len(hdr_msgs) + 16) + hdr_msgs + bytes(16))
            coff = self.alloc(rest_bytes)
            struct.pack_into("<QQ", self.buf, off + 16 + len(hdr_msgs), coff, len(rest_bytes))
            starts, p = [off + 16], coff
            for e in rest:
                starts.append(p)
                p += len(e)
        else:
            allm = b"".join(enc)
            off = self.alloc(struct.pack("<BBHII4x", 1, 0, len(enc), 1, len(allm)) + allm)
            starts, p = [], off + 16
            for e in enc:
                starts.append(p)
                p += len(e)
        return off, starts

    def attr_msgs(self, attrs):
        out = []
        for k, v in attrs.items():
            n0 = len(self.gheap_fix)
            head, data = self.m_attr(k, v)
            out.append((0x0C, head + data, len(head), n0, len(self.gheap_fix)))
        return out

    def place(self, msgs):
        off, starts = self.object_header([(m[0], m[1]) for m in msgs])
        for m, st in zip(msgs, starts):
            if m[0] == 0x0C and len(m) > 2:
                for i in range(m[3], m[4]):
                    self.gheap_fix[i] = (st + 8 + m[2] + self.gheap_fix[i][1], None)
        return off

    def dataset(self, arr):
        arr = np.asarray(arr, order="C")  # (ascontiguousarray would turn a 0-d array into shape (1,))
        raw = arr.astype(arr.dtype.newbyteorder("<")).tobytes()
        if self.chunks and arr.ndim >= 1 and arr.size:
            layout = self.chunked_layout(arr)
        elif self.compact:
            layout = struct.pack("<BBH", 3, 0, len(raw)) + raw
        else:
            layout = struct.pack("<BBQQ", 3, 1, self.alloc(raw) if raw else UNDEF, len(raw))
        fill = struct.pack("<BBBB", 2, 2, 2, 0)
        return self.place([(0x01, self.m_space(arr.shape)), (0x03, self.m_type(arr.dtype)), (0x05, fill),
                           (0x08, layout)])

    def chunked_layout(self, arr):
        import itertools
        cd = tuple(min(self.chunks, d) for d in arr.shape)
        rank, es = arr.ndim, arr.dtype.itemsize
        csize = int(np.prod(cd)) * es
        entries = []
        for off in itertools.product(*[range(0, d, c) for d, c in zip(arr.shape, cd)]):
            chunk = np.zeros(cd, arr.dtype)
            sl = tuple(slice(o, min(o + c, d)) for o, c, d in zip(off, cd, arr.shape))
            chunk[tuple(slice(0, s.stop - s.start) for s in sl)] = arr[sl]
            entries.append((off, self.alloc(chunk.astype(arr.dtype.newbyteorder("<")).tobytes())))
        if len(entries) > 64:
            raise ValueError("too many chunks for a one-level chunk B-tree")

        def key(off, nbytes):
            return struct.pack("<II", nbytes, 0) + b"".join(struct.pack("<Q", o) for o in off) + struct.pack("<Q", 0)

        node = b"TREE" + struct.pack("<BBHQQ", 1, 0, len(entries), UNDEF, UNDEF)
        for off, addr in entries:
            node += key(off, csize) + struct.pack("<Q", addr)
        node += key(arr.shape, 0)  # the closing key: one past the last chunk
        btree = self.alloc(node)
        return struct.pack("<BBB", 3, 2, rank + 1) + struct.pack("<Q", btree) + b"".join(
            struct.pack("<I", c) for c in cd) + struct.pack("<I", es)

    def group(self, children, attrs):
        """children: {name: object-header address} -> (object header address, btree, heap)."""
        names = sorted(children, key=lambda s: s.encode("utf8"))
        heap_data = bytearray(8)  # offset 0: the empty string
        offs = {}
        for n in names:
            offs[n] = len(heap_data)
            b = n.encode("utf8") + b"\0"
            heap_data += b.ljust(_pad8(len(b)), b"\0")
        seg = self.alloc(bytes(heap_data))
        heap = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 1, seg))
        cap = 2 * self.LEAF_K
        leaves = [names[i:i + cap] for i in range(0, len(names), cap)] or [[]]
        if len(leaves) > 2 * self.INTERNAL_K:
            raise ValueError("too many entries for a one-level group B-tree")
        keys, kids = [0], []
        for leaf in leaves:
            ents = b"".join(struct.pack("<QQII16x", offs[n], children[n], 0, 0) for n in leaf)
            kids.append(self.alloc((b"SNOD" + struct.pack("<BBH", 1, 0, len(leaf)) + ents).ljust(8 + cap * 40, b"\0")))
            keys.append(offs[leaf[-1]] if leaf else 0)
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(kids), UNDEF, UNDEF) + struct.pack("<Q", keys[0])
        for k, c in zip(keys[1:], kids):
            node += struct.pack("<QQ", c, k)
        btree = self.alloc(node.ljust(24 + 2 * self.INTERNAL_K * 8 + (2 * self.INTERNAL_K + 1) * 8, b"\0"))
        ohdr = self.place([(0x11, struct.pack("<QQ", btree, heap))] + self.attr_msgs(attrs))
        return ohdr, btree, heap

    def finish(self, root):
        ohdr, btree, heap = root
        if self.gheap:
            objs = bytearray()
            for i, s in enumerate(self.gheap, 1):
                objs += struct.pack("<HH4xQ", i, 1, len(s)) + s.ljust(_pad8(len(s)), b"\0")
            free = 16 + 8  # object 0: the free space that ends the collection
            total = 16 + len(objs) + free
            col = self.alloc(b"GCOL" + struct.pack("<B3xQ", 1, total) + bytes(objs) + struct.pack("<HH4xQ", 0, 0, free) + bytes(8))
            for pos, _ in self.gheap_fix:
                struct.pack_into("<Q", self.buf, pos, col)
        eof = len(self.buf)
        if self.sb in (0, 1):
            sb = SIG + struct.pack("<BBBBBBBBHHI", self.sb, 0, 0, 0, 0, 8, 8, 0, self.LEAF_K, self.INTERNAL_K, 0)
            if self.sb == 1:
                sb += struct.pack("<HH", 32, 0)
            sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF) + struct.pack("<QQII", 0, ohdr, 1, 0) + struct.pack("<QQ", btree, heap)
        else:
            sb = SIG + struct.pack("<BBBB", 2, 8, 8, 0) + struct.pack("<QQQQ", 0, UNDEF, eof, ohdr) + bytes(4)
        assert len(sb) <= self.reserved
        self.buf[:len(sb)] = sb
        return bytes(self.buf)


def write_keras_weights(path, layers, backend="tensorflow", keras_version="2.1.2", model_weights_group=False,
                        **variant):
    """Write `layers` = [(layer name, [(weight name, ndarray), ...]), ...] in the tree Keras 2.1.2
    `save_weights` produces (module docstring). Returns the bytes written. `variant`: see `_Writer`."""
    w = _Writer(**variant)
    w.reserved = 96 + 8
    w.buf += bytes(w.reserved)  # the superblock is filled in last (it records the end-of-file address)

    def tree(entries, attrs):
        """entries: {name: ndarray | (dict, attrs)}"""
        kids = {}
        for name, e in entries.items():
            kids[name] = tree(*e)[0] if isinstance(e, tuple) else w.dataset(e)
        return w.group(kids, attrs)

    top = {}
    for lname, ws in layers:
        sub = {}
        for wname, arr in ws:
            parts = wname.split("/")
            d = sub
            for p in parts[:-1]:
                d = d.setdefault(p, ({}, {}))[0]
            d[parts[-1]] = np.asarray(arr)
        top[lname] = (sub, {"weight_names": np.array([n.encode("utf8") for n, _ in ws], dtype="S") if ws else []})
    attrs = {"layer_names": np.array([n.encode("utf8") for n, _ in layers], dtype="S"),
             "backend": np.bytes_(backend.encode("utf8")), "keras_version": np.bytes_(keras_version.encode("utf8"))}
    if model_weights_group:
        root = tree({"model_weights": (top, attrs)}, {"keras_version": attrs["keras_version"], "backend": attrs["backend"]})
    else:
        root = tree(top, attrs)
    data = w.finish(root)
    with open(path, "wb") as f:
        f.write(data)
    return data


def save_head_weights(path, arrs, out_act="softmax"):
    """Write the head's six arrays under the layer names the reference's graph gets in a fresh Keras
    session (code/siamese.py:24-33: two Inputs, the |l-r| Lambda, three Dense layers)."""
    W1, b1, W2, b2, W3, b3 = [np.asarray(a, np.float32) for a in arrs]
    layers = [("input_1", []), ("input_2", []), ("lambda_1", [])]
    for i, (W, b) in enumerate(((W1, b1), (W2, b2), (W3, b3)), 1):
        layers.append((f"dense_{i}", [(f"dense_{i}/kernel:0", W), (f"dense_{i}/bias:0", b)]))
    if out_act == "softmax":
        layers.append(("activation_1", []))  # code/siamese.py:32: Activation('softmax') is a (weightless) layer of its own
    return write_keras_weights(path, layers)
