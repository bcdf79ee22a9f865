"""A small reader for the HDF5 files RocAlphaGo produces -- no libhdf5, no h5py.

Why it exists: the reference stores trained weights (Keras 1.x `save_weights`, read back by
`NeuralNetBase.load_model`, nn_util.py:60-82) and converted games (`GameConverter.sgfs_to_hdf5`,
game_converter.py:42-141) as HDF5.  This image has no HDF5 library, so the subset of the format
those writers use (h5py defaults, HDF5 1.8 "earliest" layout) is parsed here directly:

  * superblock version 0/1, version-1 object headers with continuation blocks
  * old-style groups: symbol-table message -> v1 B-tree -> symbol nodes -> local heap names
  * datasets: fixed-point / IEEE float / fixed-length string types, simple and scalar dataspaces,
    compact, contiguous and chunked layouts (v1 chunk B-tree), filters deflate (1), shuffle (2),
    fletcher32 (3, checksum dropped unverified) and LZF (32000, h5py's `compression="lzf"`)
  * attributes (message versions 1-3) of the same types, e.g. Keras' `layer_names`/`weight_names`

Anything else (new-style groups, variable-length types, compound types, external storage) raises
`Hdf5Error` naming the feature.  The layout follows the public "HDF5 File Format Specification
Version 2.0"; nothing here is derived from the reference's code.

    f = hdf5_lite.File(path)
    f.keys(); f.attrs["layer_names"]; f["convolution2d_1"]["convolution2d_1_W"][...]
"""
import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"


class Hdf5Error(ValueError):
    pass


def lzf_decompress(src, out_len):
    """liblzf's format: control byte < 32 -> literal run of ctrl+1 bytes; otherwise a back
    reference of length (ctrl >> 5) + 2 (7 = extended by the next byte) at distance
    ((ctrl & 31) << 8 | next byte) + 1."""
    src = memoryview(src)
    out = bytearray(out_len)
    ip, op, n = 0, 0, len(src)
    while ip < n:
        ctrl = src[ip]
        ip += 1
        if ctrl < 32:
            run = ctrl + 1
            if op + run > out_len or ip + run > n:
                raise Hdf5Error("corrupt LZF stream (literal run past the end)")
            out[op:op + run] = src[ip:ip + run]
            ip += run
            op += run
        else:
            length = ctrl >> 5
            if length == 7:
                length += src[ip]
                ip += 1
            dist = ((ctrl & 31) << 8 | src[ip]) + 1
            ip += 1
            length += 2
            ref = op - dist
            if ref < 0 or op + length > out_len:
                raise Hdf5Error("corrupt LZF stream (reference outside the output)")
            if dist >= length:
                out[op:op + length] = out[ref:ref + length]
            else:                                   # overlapping: the pattern of `dist` bytes repeats
                pattern = bytes(out[ref:op])
                reps = -(-length // dist)
                out[op:op + length] = (pattern * reps)[:length]
            op += length
    if op != out_len:
        raise Hdf5Error("LZF stream ended after %d of %d bytes" % (op, out_len))
    return bytes(out)


def _unshuffle(data, itemsize):
    if itemsize <= 1:
        return data
    a = np.frombuffer(data, np.uint8)
    n = len(a) // itemsize
    body = a[:n * itemsize].reshape(itemsize, n).T.reshape(-1)
    return body.tobytes() + a[n * itemsize:].tobytes()


class _Reader(object):
    def __init__(self, buf):
        self.buf = buf
        self.O = self.L = 8

    def u(self, pos, size):
        return int.from_bytes(self.buf[pos:pos + size], "little")

    def off(self, pos):
        return self.u(pos, self.O)

    def length(self, pos):
        return self.u(pos, self.L)


def _pad8(n):
    return (n + 7) & ~7


class _Datatype(object):
    def __init__(self, r, pos):
        b0 = r.buf[pos]
        self.version, self.cls = b0 >> 4, b0 & 15
        bits = r.u(pos + 1, 3)
        self.size = r.u(pos + 4, 4)
        if self.cls == 0:                                    # fixed point
            order = ">" if bits & 1 else "<"
            signed = bool(bits & 8)
            if self.size not in (1, 2, 4, 8):
                raise Hdf5Error("integer type of %d bytes" % self.size)
            self.dtype = np.dtype("%s%s%d" % (order, "i" if signed else "u", self.size))
        elif self.cls == 1:                                  # IEEE float
            order = ">" if bits & 1 else "<"
            if self.size not in (2, 4, 8):
                raise Hdf5Error("floating-point type of %d bytes" % self.size)
            self.dtype = np.dtype("%sf%d" % (order, self.size))
        elif self.cls == 3:                                  # fixed-length string
            self.dtype = np.dtype("S%d" % self.size)
        elif self.cls == 9:
            raise Hdf5Error("variable-length types are not supported")
        else:
            raise Hdf5Error("datatype class %d is not supported" % self.cls)


def _dataspace(r, pos):
    version, rank, flags = r.buf[pos], r.buf[pos + 1], r.buf[pos + 2]
    if version == 1:
        p = pos + 8
    elif version == 2:
        if r.buf[pos + 3] == 2:
            raise Hdf5Error("null dataspace")
        p = pos + 4
    else:
        raise Hdf5Error("dataspace message version %d" % version)
    return tuple(r.length(p + i * r.L) for i in range(rank))


class _Object(object):
    """An object header: the messages that matter, collected across continuation blocks."""

    def __init__(self, r, addr):
        self.r, self.addr = r, addr
        self.messages = []                                   # (type, position of the data, size)
        if r.buf[addr:addr + 4] == b"OHDR":
            raise Hdf5Error("version-2 object headers (libver='latest') are not supported")
        if r.buf[addr] != 1:
            raise Hdf5Error("object header version %d at %d" % (r.buf[addr], addr))
        total = r.u(addr + 2, 2)
        blocks = [(addr + 16, r.u(addr + 8, 4))]
        while blocks and len(self.messages) < total:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and len(self.messages) < total:
                mtype, msize = r.u(pos, 2), r.u(pos + 2, 2)
                data = pos + 8
                self.messages.append((mtype, data, msize))
                if mtype == 0x10:
                    blocks.append((r.off(data), r.length(data + r.O)))
                pos = data + msize
        self._attrs = None

    def find(self, mtype):
        for t, pos, size in self.messages:
            if t == mtype:
                return pos, size
        return None

    @property
    def attrs(self):
        if self._attrs is None:
            self._attrs = {}
            for t, pos, size in self.messages:
                if t == 0x0C:
                    name, value = self._attribute(pos)
                    self._attrs[name] = value
        return self._attrs

    def _attribute(self, pos):
        r = self.r
        version = r.buf[pos]
        name_size, dt_size, ds_size = r.u(pos + 2, 2), r.u(pos + 4, 2), r.u(pos + 6, 2)
        if version == 1:
            p = pos + 8
            step = _pad8
        elif version in (2, 3):
            if r.buf[pos + 1] & 3:
                raise Hdf5Error("shared attribute datatypes/dataspaces are not supported")
            p = pos + 8 + (1 if version == 3 else 0)
            step = lambda n: n
        else:
            raise Hdf5Error("attribute message version %d" % version)
        name = bytes(r.buf[p:p + name_size]).split(b"\0")[0].decode("utf-8")
        p += step(name_size)
        dt = _Datatype(r, p)
        p += step(dt_size)
        shape = _dataspace(r, p)
        p += step(ds_size)
        count = int(np.prod(shape)) if shape else 1
        value = np.frombuffer(r.buf, dt.dtype, count, p).reshape(shape).copy()
        return name, (value if shape else value.reshape(())[()])


class Dataset(object):
    def __init__(self, obj, name):
        self._obj, self.name = obj, name
        r = obj.r
        dt = obj.find(0x03)
        ds = obj.find(0x01)
        lay = obj.find(0x08)
        if dt is None or ds is None or lay is None:
            raise Hdf5Error("%s is not a dataset" % name)
        self._dt = _Datatype(r, dt[0])
        self.dtype = self._dt.dtype
        self.shape = _dataspace(r, ds[0])
        self._layout = lay[0]
        self.attrs = obj.attrs

    def __len__(self):
        return self.shape[0]

    def _filters(self):
        hit = self._obj.find(0x0B)
        if hit is None:
            return []
        r, pos = self._obj.r, hit[0]
        version, count = r.buf[pos], r.buf[pos + 1]
        p = pos + (8 if version == 1 else 2)
        out = []
        for _ in range(count):
            fid = r.u(p, 2)
            if version == 1 or fid >= 256:
                name_len = r.u(p + 2, 2)
                p += 4
            else:
                name_len = 0
                p += 2
            nvals = r.u(p + 2, 2)
            p += 4
            p += _pad8(name_len) if version == 1 else name_len
            vals = [r.u(p + 4 * i, 4) for i in range(nvals)]
            p += 4 * nvals
            if version == 1 and nvals % 2:
                p += 4
            out.append((fid, vals))
        return out

    def _decode_chunk(self, raw, filters, mask, nbytes):
        for i in range(len(filters) - 1, -1, -1):
            if mask >> i & 1:
                continue                                     # the writer skipped this (optional) filter for this chunk
            fid, vals = filters[i]
            if fid == 1:
                raw = zlib.decompress(raw)
            elif fid == 2:
                raw = _unshuffle(raw, vals[0] if vals else self.dtype.itemsize)
            elif fid == 3:
                raw = raw[:-4]
            elif fid == 32000:
                raw = lzf_decompress(raw, vals[2] if len(vals) > 2 else nbytes)
            else:
                raise Hdf5Error("filter %d is not supported" % fid)
        return raw

    def read(self):
        r, pos = self._obj.r, self._layout
        version, cls = r.buf[pos], r.buf[pos + 1]
        if version != 3:
            raise Hdf5Error("data layout message version %d" % version)
        count = int(np.prod(self.shape)) if self.shape else 1
        if cls == 0:                                         # compact
            size = r.u(pos + 2, 2)
            flat = np.frombuffer(r.buf, self.dtype, count, pos + 4) if size else np.zeros(count, self.dtype)
        elif cls == 1:                                       # contiguous
            addr = r.off(pos + 2)
            flat = np.zeros(count, self.dtype) if addr == UNDEF else np.frombuffer(r.buf, self.dtype, count, addr)
        elif cls == 2:
            return self._read_chunked(pos)
        else:
            raise Hdf5Error("data layout class %d" % cls)
        return flat.reshape(self.shape).copy()

    def _read_chunked(self, pos):
        r = self._obj.r
        rank = r.buf[pos + 2] - 1
        btree = r.off(pos + 3)
        p = pos + 3 + r.O
        chunk = tuple(r.u(p + 4 * i, 4) for i in range(rank))
        if rank != len(self.shape):
            raise Hdf5Error("chunk rank %d for a rank-%d dataset" % (rank, len(self.shape)))
        out = np.zeros(self.shape, self.dtype)
        if btree == UNDEF or out.size == 0:
            return out
        filters = self._filters()
        nbytes = int(np.prod(chunk)) * self.dtype.itemsize

        def walk(addr):
            if r.buf[addr:addr + 4] != b"TREE" or r.buf[addr + 4] != 1:
                raise Hdf5Error("bad chunk B-tree node at %d" % addr)
            level, used = r.buf[addr + 5], r.u(addr + 6, 2)
            q = addr + 8 + 2 * r.O
            key = 8 + 8 * (rank + 1)
            for _ in range(used):
                size, mask = r.u(q, 4), r.u(q + 4, 4)
                start = tuple(r.u(q + 8 + 8 * i, 8) for i in range(rank))
                child = r.off(q + key)
                q += key + r.O
                if level > 0:
                    walk(child)
                    continue
                raw = self._decode_chunk(bytes(r.buf[child:child + size]), filters, mask, nbytes)
                block = np.frombuffer(raw, self.dtype, int(np.prod(chunk))).reshape(chunk)
                sel = tuple(slice(s, min(s + c, d)) for s, c, d in zip(start, chunk, self.shape))
                if any(s.start >= s.stop for s in sel):
                    continue
                out[sel] = block[tuple(slice(0, s.stop - s.start) for s in sel)]

        walk(btree)
        return out

    def __getitem__(self, key):
        data = self.read()
        if key is Ellipsis or key == ():
            return data if self.shape else data.reshape(())[()]
        return data[key]

    def __array__(self, dtype=None, copy=None):
        a = self.read()
        return a.astype(dtype) if dtype is not None else a


class Group(object):
    def __init__(self, r, obj, name):
        self._r, self._obj, self.name = r, obj, name
        self.attrs = obj.attrs
        self._links = None

    def _load(self):
        if self._links is not None:
            return
        r = self._r
        self._links = {}
        hit = self._obj.find(0x11)
        if hit is None:
            if self._obj.find(0x02) is not None or self._obj.find(0x06) is not None:
                raise Hdf5Error("new-style groups (link messages) are not supported")
            return
        btree, heap = r.off(hit[0]), r.off(hit[0] + r.O)
        if r.buf[heap:heap + 4] != b"HEAP":
            raise Hdf5Error("bad local heap at %d" % heap)
        names = r.off(heap + 8 + 2 * r.L)

        def walk(addr):
            sig = r.buf[addr:addr + 4]
            if sig == b"TREE":
                if r.buf[addr + 4] != 0:
                    raise Hdf5Error("group B-tree node of type %d" % r.buf[addr + 4])
                used = r.u(addr + 6, 2)
                q = addr + 8 + 2 * r.O + r.L                 # first child pointer (after key 0)
                for _ in range(used):
                    walk(r.off(q))
                    q += r.O + r.L
            elif sig == b"SNOD":
                count = r.u(addr + 6, 2)
                q = addr + 8
                for _ in range(count):
                    n0 = names + r.off(q)
                    n1 = r.buf.index(b"\0", n0)
                    cache_type = r.u(q + 2 * r.O, 4)
                    if cache_type == 2:
                        raise Hdf5Error("symbolic links are not supported")
                    self._links[bytes(r.buf[n0:n1]).decode("utf-8")] = r.off(q + r.O)
                    q += 2 * r.O + 24
            else:
                raise Hdf5Error("bad group node at %d" % addr)

        if btree != UNDEF:
            walk(btree)

    def keys(self):
        self._load()
        return sorted(self._links)

    def __contains__(self, name):
        self._load()
        return name in self._links

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        self._load()
        return len(self._links)

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            if not isinstance(node, Group):
                raise KeyError(path)
            node._load()
            if part not in node._links:
                raise KeyError("%s (no such object in %s)" % (path, node.name))
            obj = _Object(node._r, node._links[part])
            child = (node.name.rstrip("/") + "/" + part)
            node = Dataset(obj, child) if obj.find(0x08) is not None else Group(node._r, obj, child)
        return node


class File(Group):
    def __init__(self, path_or_bytes):
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            buf = bytes(path_or_bytes)
        else:
            with open(path_or_bytes, "rb") as fh:
                buf = fh.read()
        if buf[:8] != SIGNATURE:
            raise Hdf5Error("not an HDF5 file (or a user block precedes the superblock)")
        r = _Reader(buf)
        version = buf[8]
        if version not in (0, 1):
            raise Hdf5Error("superblock version %d (written with libver='latest'?) is not supported" % version)
        r.O, r.L = buf[13], buf[14]
        if r.O != 8 or r.L != 8:
            raise Hdf5Error("offset/length sizes %d/%d are not supported" % (r.O, r.L))
        p = 24 + (4 if version == 1 else 0)
        if r.off(p) != 0:
            raise Hdf5Error("non-zero base address")
        root_entry = p + 4 * r.O
        root = _Object(r, r.off(root_entry + r.O))
        Group.__init__(self, r, root, "/")

    def close(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


# ---- writer ---------------------------------------------------------------------------------------
#
# The same subset, written the way libhdf5 1.8 lays a new file out with default properties:
# superblock version 0, old-style groups (one B-tree node + one symbol node + a local heap per group),
# version-1 object headers, contiguous datasets, version-1 attribute messages.  Verified by reading
# the files back with the reader above (which is itself checked against files h5py wrote); no libhdf5
# is available on this image to cross-check with.

def _datatype_message(dtype):
    dtype = np.dtype(dtype)
    if dtype.kind == "S":
        return struct.pack("<B3BI", 0x13, 0x01, 0, 0, max(dtype.itemsize, 1))          # null-padded ASCII
    if dtype.byteorder == ">":
        raise Hdf5Error("big-endian arrays are not written")
    if dtype.kind in "iu":
        bits0 = 0x08 if dtype.kind == "i" else 0
        return struct.pack("<B3BIHH", 0x10, bits0, 0, 0, dtype.itemsize, 0, 8 * dtype.itemsize)
    if dtype.kind == "f":
        exp_loc, exp_size, man_size, bias = {2: (10, 5, 10, 15), 4: (23, 8, 23, 127), 8: (52, 11, 52, 1023)}[dtype.itemsize]
        return struct.pack("<B3BIHHBBBBI", 0x11, 0x20, 8 * dtype.itemsize - 1, 0, dtype.itemsize,
                           0, 8 * dtype.itemsize, exp_loc, exp_size, 0, man_size, bias)
    raise Hdf5Error("dtype %s is not written" % dtype)


def _dataspace_message(shape):
    return struct.pack("<BBB5x", 1, len(shape), 0) + b"".join(struct.pack("<Q", d) for d in shape)


def _pad(b, to=8):
    return b + b"\0" * (-len(b) % to)


class _WNode(object):
    def __init__(self):
        self.attrs = {}


class WGroup(_WNode):
    def __init__(self):
        _WNode.__init__(self)
        self.children = {}

    def create_group(self, name):
        if "/" in name or not name:
            raise Hdf5Error("bad object name %r ('/' separates path components)" % name)
        g = self.children[name] = WGroup()
        return g

    def require_group(self, name):
        return self.children[name] if name in self.children else self.create_group(name)

    def create_dataset(self, name, data):
        if "/" in name or not name:
            raise Hdf5Error("bad object name %r" % name)
        d = self.children[name] = WDataset(np.asarray(data))
        return d

    def __setitem__(self, name, data):
        self.create_dataset(name, data)

    def __getitem__(self, name):
        return self.children[name]


class WDataset(_WNode):
    def __init__(self, data):
        _WNode.__init__(self)
        if data.dtype.kind == "U":
            data = np.char.encode(data, "utf-8")
        self.data = data if isinstance(data, np.memmap) else np.asarray(data, order="C")


class Writer(WGroup):
    """w = Writer(); w.attrs[...] = ...; g = w.create_group("x"); g["y"] = array; w.save(path)

    Dataset arrays may be `np.memmap`s: `save` streams them to the file in slices, so a dataset
    larger than memory (converted game records) can be assembled from a scratch file."""

    def save(self, path):
        with open(path, "w+b") as fh:
            self._emit(fh)

    def tobytes(self):
        import io
        fh = io.BytesIO()
        self._emit(fh)
        return fh.getvalue()

    def _emit(self, fh):
        biggest = [4]

        def scan(g):
            biggest[0] = max(biggest[0], len(g.children))
            for c in g.children.values():
                if isinstance(c, WGroup):
                    scan(c)
        scan(self)
        # one symbol-table node holds 2 * leaf_k entries (leaf_k is a 16-bit field of the superblock); a group
        # with more children gets several nodes under its B-tree node, which has room for 2 * internal_k of them
        leaf_k = min(32767, max(4, (biggest[0] + 1) // 2))
        internal_k = 16
        if biggest[0] > 2 * leaf_k * 2 * internal_k:
            raise Hdf5Error("too many objects in one group (%d; at most %d)" % (biggest[0], 2 * leaf_k * 2 * internal_k))
        fh.write(b"\0" * 96)                                  # superblock, filled in at the end

        def alloc(data):
            pos = fh.seek(0, 2)
            if pos % 8:
                fh.write(b"\0" * (8 - pos % 8))
                pos += 8 - pos % 8
            fh.write(data)
            return pos

        def alloc_array(a):
            if a.nbytes == 0:
                return UNDEF
            flat = a.reshape(-1)
            step = max(1, (64 << 20) // max(a.dtype.itemsize, 1))
            addr = alloc(flat[:step].tobytes())
            for i in range(step, flat.size, step):
                fh.write(flat[i:i + step].tobytes())
            return addr

        def header(messages):
            body = b""
            for mtype, data in messages:
                data = _pad(data)
                if len(data) > 65528:
                    raise Hdf5Error("header message of %d bytes (attributes must stay below 64 KiB)" % len(data))
                body += struct.pack("<HHB3x", mtype, len(data), 0) + data
            return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(body)) + body

        def attribute_messages(node):
            msgs = []
            for name, value in node.attrs.items():
                value = np.asarray(value)
                if value.dtype.kind == "U":
                    value = np.char.encode(value, "utf-8")
                if value.dtype.kind == "S" and value.dtype.itemsize == 0:
                    value = value.astype("S1")
                value = np.asarray(value, order="C")
                nm = name.encode("utf-8") + b"\0"
                dt, ds = _datatype_message(value.dtype), _dataspace_message(value.shape)
                msgs.append((0x0C, struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + _pad(nm) + _pad(dt) + _pad(ds) + value.tobytes()))
            return msgs

        def write_dataset(d):
            addr = alloc_array(d.data)
            msgs = [(0x01, _dataspace_message(d.data.shape)), (0x03, _datatype_message(d.data.dtype)),
                    (0x05, struct.pack("<BBBB", 2, 2, 2, 0)),                            # fill value: late allocation, none defined
                    (0x08, struct.pack("<BBQQ", 3, 1, addr, d.data.nbytes))] + attribute_messages(d)
            return alloc(header(msgs))

        def write_group(g):
            names = sorted(g.children, key=lambda s: s.encode("utf-8"))
            entries = []
            for nm in names:
                c = g.children[nm]
                entries.append((nm, write_group(c)[0] if isinstance(c, WGroup) else write_dataset(c)))
            heap_data, offsets = bytearray(8), []             # offset 0: the empty name
            for nm, _ in entries:
                offsets.append(len(heap_data))
                heap_data += _pad(nm.encode("utf-8") + b"\0")
            free = len(heap_data)
            heap_data += struct.pack("<QQ", 1, 16)             # one free block; 1 = end of the free list
            data_addr = alloc(bytes(heap_data))
            heap_addr = alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free, data_addr))
            node_size = 24 + (2 * internal_k + 1) * 8 + 2 * internal_k * 8
            per = 2 * leaf_k
            tree_body = struct.pack("<Q", 0)                   # key 0: the empty name, below every child
            n_nodes = 0
            for i in range(0, len(entries), per):
                chunk, offs = entries[i:i + per], offsets[i:i + per]
                snod = [b"SNOD" + struct.pack("<BBH", 1, 0, len(chunk))]
                for (nm, addr), off in zip(chunk, offs):
                    snod.append(struct.pack("<QQII16x", off, addr, 0, 0))
                snod = b"".join(snod)
                snod_addr = alloc(snod + b"\0" * (8 + per * 40 - len(snod)))
                tree_body += struct.pack("<QQ", snod_addr, offs[-1])       # child, then the key of its largest name
                n_nodes += 1
            tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, n_nodes, UNDEF, UNDEF) + tree_body
            tree_addr = alloc(tree + b"\0" * (node_size - len(tree)))
            hdr = alloc(header([(0x11, struct.pack("<QQ", tree_addr, heap_addr))] + attribute_messages(g)))
            return hdr, tree_addr, heap_addr

        root, tree_addr, heap_addr = write_group(self)
        end = alloc(b"")
        sb = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, leaf_k, internal_k, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, end, UNDEF)
        sb += struct.pack("<QQII", 0, root, 1, 0) + struct.pack("<QQ", tree_addr, heap_addr)
        assert len(sb) == 96
        fh.seek(0)
        fh.write(sb)
