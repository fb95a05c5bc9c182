"""A small, dependency-free HDF5 writer and reader for Keras model files (README.md:237-246).

``save_model_hdf5(model, "trained-<partition>.hdf5")`` in the reference produces a Keras-2 HDF5
file: root attributes ``model_config`` / ``keras_version`` / ``backend``, a ``model_weights`` group
with a ``layer_names`` attribute, one group per layer with a ``weight_names`` attribute, and one
float32 data set per weight.  h5py / libhdf5 are not available on the target image, so this module
writes that layout directly in the classic HDF5 on-disk format (the one libhdf5 writes by default and
every HDF5 reader understands):

* superblock version 0, 8-byte offsets and lengths;
* groups as symbol tables: version-1 object header -> symbol-table message -> one B-tree node
  (``TREE``) with one leaf (``SNOD``) -> names in a local heap (``HEAP``); the leaf capacity
  (superblock field "group leaf node K") is chosen so that every group fits one leaf;
* data sets: version-1 object header with dataspace (v1), datatype (v1), fill value (v2) and
  contiguous data layout (v3) messages; raw data little-endian, C order;
* attributes: version-1 attribute messages holding fixed-length, null-padded strings (scalars or 1-D
  arrays — what h5py writes for ``bytes`` / ``numpy.bytes_`` values) or numeric arrays.

The reader understands exactly this subset plus what h5py adds when it writes the same content
(object-header continuation blocks, multi-leaf group B-trees, variable-length strings through the
global heap), which is enough to load Keras-2 weight files written by h5py.
"""
from __future__ import annotations

import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"

MSG_DATASPACE, MSG_DATATYPE, MSG_FILL, MSG_LAYOUT, MSG_ATTRIBUTE = 0x1, 0x3, 0x5, 0x8, 0xC
MSG_CONTINUATION, MSG_SYMBOL_TABLE = 0x10, 0x11


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


# ---------------------------------------------------------------------------------------------
# in-memory tree
# ---------------------------------------------------------------------------------------------
class Group:
    def __init__(self):
        self.attrs: dict[str, object] = {}
        self.children: dict[str, "Group | np.ndarray"] = {}

    def create_group(self, name: str) -> "Group":
        g = Group()
        self.children[name] = g
        return g

    def create_dataset(self, name: str, data) -> None:
        self.children[name] = np.require(np.asarray(data), requirements="C")    # keeps 0-d arrays 0-d

    def __getitem__(self, name):
        node = self
        for part in name.strip("/").split("/"):
            node = node.children[part]
        return node

    def __contains__(self, name):
        return name in self.children


# ---------------------------------------------------------------------------------------------
# writer
# ---------------------------------------------------------------------------------------------
def _datatype_of(arr: np.ndarray) -> bytes:
    """Datatype message body (version 1)."""
    dt = arr.dtype
    if dt.kind == "S":
        # class 3 (string), null-padded, ASCII
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(dt.itemsize, 1))
    if dt.kind == "f" and dt.itemsize in (4, 8):
        if dt.itemsize == 4:
            props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
            return struct.pack("<BBBBI", 0x11, 0x20, 31, 0, 4) + props
        props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
        return struct.pack("<BBBBI", 0x11, 0x20, 63, 0, 8) + props
    if dt.kind in "iu" and dt.itemsize in (1, 2, 4, 8):
        flags = 0x08 if dt.kind == "i" else 0x00          # bit 3: two's complement signed
        return struct.pack("<BBBBI", 0x10, flags, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    raise TypeError(f"hdf5 writer: unsupported dtype {dt}")


def _dataspace_of(shape: tuple[int, ...]) -> bytes:
    """Dataspace message body (version 1, no maximum dimensions)."""
    return struct.pack("<BBBB4x", 1, len(shape), 0, 0) + b"".join(struct.pack("<Q", int(d)) for d in shape)


def _as_attr_array(value) -> np.ndarray:
    if isinstance(value, str):
        value = value.encode("utf8")
    if isinstance(value, (bytes, bytearray)):
        return np.array(bytes(value), dtype=f"S{max(len(value), 1)}")
    if isinstance(value, (list, tuple)) and value and all(isinstance(v, (str, bytes)) for v in value):
        enc = [v.encode("utf8") if isinstance(v, str) else bytes(v) for v in value]
        return np.array(enc, dtype=f"S{max(max(map(len, enc)), 1)}")
    if isinstance(value, (list, tuple)) and not value:
        return np.zeros((0,), dtype="S1")
    arr = np.asarray(value)
    if arr.dtype.kind == "U":
        arr = np.char.encode(arr, "utf8")
    return arr


def _message(mtype: int, body: bytes, flags: int = 0) -> bytes:
    body = _pad8(body)
    if len(body) > 0xFFFF:
        raise ValueError("hdf5 writer: an object-header message (an attribute, most likely) exceeds 64 KiB")
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _attribute(name: str, value) -> bytes:
    arr = _as_attr_array(value)
    nm = name.encode("utf8") + b"\0"
    dt, ds = _datatype_of(arr), _dataspace_of(arr.shape)
    raw = np.require(arr, requirements="C").tobytes()
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) + _pad8(ds) + raw
    return _message(MSG_ATTRIBUTE, body)


def _object_header(messages: list[bytes]) -> bytes:
    data = b"".join(messages)
    # version, reserved, number of messages, reference count, size of the message block; 4 bytes of padding
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(data)) + data


class _Writer:
    def __init__(self, leaf_k: int):
        self.buf = bytearray()
        self.leaf_k = leaf_k
        self.internal_k = 16

    def alloc(self, n: int) -> int:
        self.buf += b"\0" * (-len(self.buf) % 8)
        addr = len(self.buf)
        self.buf += b"\0" * n
        return addr

    def put(self, addr: int, data: bytes):
        self.buf[addr:addr + len(data)] = data

    def write_dataset(self, arr: np.ndarray, attrs: dict) -> int:
        arr = np.require(np.asarray(arr), requirements="C")
        if arr.dtype.byteorder == ">":
            arr = arr.astype(arr.dtype.newbyteorder("<"))
        raw = arr.tobytes()
        data_addr = self.alloc(len(raw)) if raw else UNDEF
        if raw:
            self.put(data_addr, raw)
        msgs = [_message(MSG_DATASPACE, _dataspace_of(arr.shape)),
                _message(MSG_DATATYPE, _datatype_of(arr), flags=1),          # constant message
                _message(MSG_FILL, struct.pack("<BBBB", 2, 2, 2, 0)),       # late allocation, fill if set, undefined
                _message(MSG_LAYOUT, struct.pack("<BBQQ", 3, 1, data_addr, len(raw)))]
        msgs += [_attribute(k, v) for k, v in attrs.items()]
        hdr = _object_header(msgs)
        addr = self.alloc(len(hdr))
        self.put(addr, hdr)
        return addr

    def write_group(self, g: Group) -> tuple[int, int, int]:
        """Returns (object header address, B-tree address, local heap address)."""
        names = sorted(g.children, key=lambda s: s.encode("utf8"))
        if len(names) > 2 * self.leaf_k:
            raise ValueError("hdf5 writer: group larger than one symbol-table leaf")
        entries = []
        for nm in names:
            child = g.children[nm]
            if isinstance(child, Group):
                oh, bt, hp = self.write_group(child)
                entries.append((nm, oh, 1, struct.pack("<QQ", bt, hp)))
            else:
                entries.append((nm, self.write_dataset(child, {}), 0, b"\0" * 16))
        # local heap: offset 0 holds the empty string, names follow, each padded to 8 bytes
        heap_data = bytearray(b"\0" * 8)
        offsets = []
        for nm, *_ in entries:
            offsets.append(len(heap_data))
            heap_data += _pad8(nm.encode("utf8") + b"\0")
        heap_data_addr = self.alloc(len(heap_data))
        self.put(heap_data_addr, bytes(heap_data))
        heap_addr = self.alloc(32)
        # free-list head 1 = "no free block" (H5HL_FREE_NULL)
        self.put(heap_addr, b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 1, heap_data_addr))
        # symbol-table node (leaf)
        snod_size = 8 + 2 * self.leaf_k * 40
        snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(entries)))
        for (nm, oh, cache, scratch), off in zip(entries, offsets):
            snod += struct.pack("<QQII", off, oh, cache, 0) + scratch
        snod += b"\0" * (snod_size - len(snod))
        btree_addr = UNDEF
        tree_size = 24 + (2 * self.internal_k + 1) * 8 + 2 * self.internal_k * 8
        btree_addr = self.alloc(tree_size)
        tree = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, 1 if entries else 0, UNDEF, UNDEF))
        if entries:
            snod_addr = self.alloc(snod_size)
            self.put(snod_addr, bytes(snod))
            tree += struct.pack("<QQQ", 0, snod_addr, offsets[-1])   # key 0 = "", child, key 1 = largest name
        tree += b"\0" * (tree_size - len(tree))
        self.put(btree_addr, bytes(tree))
        msgs = [_message(MSG_SYMBOL_TABLE, struct.pack("<QQ", btree_addr, heap_addr))]
        msgs += [_attribute(k, v) for k, v in g.attrs.items()]
        hdr = _object_header(msgs)
        oh_addr = self.alloc(len(hdr))
        self.put(oh_addr, hdr)
        return oh_addr, btree_addr, heap_addr


def _max_group_size(g: Group) -> int:
    return max([len(g.children)] + [_max_group_size(c) for c in g.children.values() if isinstance(c, Group)])


def write_file(path, root: Group) -> None:
    """Serialise ``root`` (groups, float/int data sets, attributes) as an HDF5 file."""
    leaf_k = max(4, (_max_group_size(root) + 1) // 2)
    w = _Writer(leaf_k)
    w.alloc(96)                                   # superblock, filled in last
    oh, bt, hp = w.write_group(root)
    w.buf += b"\0" * (-len(w.buf) % 8)
    eof = len(w.buf)
    sb = SIGNATURE + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0)
    sb += struct.pack("<HHI", leaf_k, w.internal_k, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    sb += struct.pack("<QQII", 0, oh, 1, 0) + struct.pack("<QQ", bt, hp)      # root symbol-table entry
    assert len(sb) == 96
    w.put(0, sb)
    with open(path, "wb") as f:
        f.write(bytes(w.buf))


# ---------------------------------------------------------------------------------------------
# reader
# ---------------------------------------------------------------------------------------------
class _Reader:
    def __init__(self, data: bytes):
        self.d = data
        base = 0
        while data[base:base + 8] != SIGNATURE:       # a user block may precede the superblock (512, 1024, ...)
            base = 512 if base == 0 else base * 2
            if base >= len(data):
                raise ValueError("not an HDF5 file")
        self.base = base
        ver = data[base + 8]
        if ver not in (0, 1):
            raise ValueError(f"hdf5 reader: superblock version {ver} is not supported (classic format only)")
        if data[base + 13] != 8 or data[base + 14] != 8:
            raise ValueError("hdf5 reader: only 8-byte offsets / lengths are supported")
        self.leaf_k, self.internal_k = struct.unpack_from("<HH", data, base + 16)
        off = base + 24 + (4 if ver == 1 else 0)
        self.base_addr, _, self.eof, _ = struct.unpack_from("<QQQQ", data, off)
        off += 32
        _, self.root_oh, cache, _ = struct.unpack_from("<QQII", data, off)
        self.gcol: dict[int, dict[int, bytes]] = {}

    def at(self, addr: int) -> int:
        return self.base + addr if self.base_addr == 0 else addr + self.base_addr

    # -- object headers ---------------------------------------------------------------------
    def messages(self, addr: int):
        d = self.d
        p = self.at(addr)
        ver = d[p]
        if ver != 1:
            raise ValueError("hdf5 reader: only version-1 object headers are supported")
        nmsg, _, size = struct.unpack_from("<HII", d, p + 2)
        blocks = [(p + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            q, left = blocks.pop(0)
            end = q + left
            while q + 8 <= end and len(out) < nmsg:
                mtype, msize, flags = struct.unpack_from("<HHB", d, q)
                body = d[q + 8:q + 8 + msize]
                q += 8 + msize
                if mtype == MSG_CONTINUATION:
                    caddr, clen = struct.unpack_from("<QQ", body, 0)
                    blocks.append((self.at(caddr), clen))
                out.append((mtype, body))
        return out

    def read_object(self, addr: int):
        msgs = self.messages(addr)
        attrs = {}
        for mtype, body in msgs:
            if mtype == MSG_ATTRIBUTE:
                k, v = self.attribute(body)
                attrs[k] = v
        sym = [b for t, b in msgs if t == MSG_SYMBOL_TABLE]
        if sym:
            g = Group()
            g.attrs = attrs
            bt, hp = struct.unpack_from("<QQ", sym[0], 0)
            for name, child in self.walk_btree(bt, hp):
                g.children[name] = self.read_object(child)
            return g
        shape = dtype = None
        layout = None
        for mtype, body in msgs:
            if mtype == MSG_DATASPACE:
                shape = self.dataspace(body)
            elif mtype == MSG_DATATYPE:
                dtype = self.datatype(body)
            elif mtype == MSG_LAYOUT:
                layout = body
        if shape is None or dtype is None or layout is None:
            raise ValueError("hdf5 reader: object is neither a group nor a simple data set")
        n = int(np.prod(shape, dtype=np.int64)) * dtype.itemsize
        if layout[0] == 3 and layout[1] in (0, 1):
            if layout[1] == 0:                          # compact: the data sits in the message
                size = struct.unpack_from("<H", layout, 2)[0]
                raw = layout[4:4 + size]
            else:
                daddr, _ = struct.unpack_from("<QQ", layout, 2)
                raw = b"" if daddr == UNDEF else self.d[self.at(daddr):self.at(daddr) + n]
        elif layout[0] in (1, 2) and layout[2] == 1:    # libhdf5 <= 1.6: version, rank, class, 5 reserved, address
            daddr = struct.unpack_from("<Q", layout, 8)[0]
            raw = b"" if daddr == UNDEF else self.d[self.at(daddr):self.at(daddr) + n]
        else:
            raise ValueError("hdf5 reader: only compact / contiguous data sets are supported (not chunked)")
        return np.frombuffer(raw, dtype=dtype, count=int(np.prod(shape, dtype=np.int64))).reshape(shape).copy()

    # -- groups ---------------------------------------------------------------------------------
    def heap_name(self, heap_addr: int, off: int) -> str:
        p = self.at(heap_addr)
        if self.d[p:p + 4] != b"HEAP":
            raise ValueError("hdf5 reader: bad local heap")
        _, _, data_addr = struct.unpack_from("<QQQ", self.d, p + 8)
        q = self.at(data_addr) + off
        return self.d[q:self.d.index(b"\0", q)].decode("utf8")

    def walk_btree(self, addr: int, heap: int):
        if addr == UNDEF:
            return
        p = self.at(addr)
        sig = self.d[p:p + 4]
        if sig == b"TREE":
            _, level, used = struct.unpack_from("<BBH", self.d, p + 4)
            q = p + 24
            for i in range(used):
                child = struct.unpack_from("<Q", self.d, q + 8 + 16 * i)[0]
                yield from self.walk_btree(child, heap)
        elif sig == b"SNOD":
            n = struct.unpack_from("<H", self.d, p + 6)[0]
            for i in range(n):
                off, oh = struct.unpack_from("<QQ", self.d, p + 8 + 40 * i)
                yield self.heap_name(heap, off), oh
        else:
            raise ValueError("hdf5 reader: bad group B-tree node")

    # -- messages -------------------------------------------------------------------------------
    @staticmethod
    def dataspace(body: bytes):
        ver, rank, flags = body[0], body[1], body[2]
        off = 8 if ver == 1 else 4
        return tuple(struct.unpack_from("<Q", body, off + 8 * i)[0] for i in range(rank))

    @staticmethod
    def datatype(body: bytes):
        cls, b0 = body[0] & 0x0F, body[1]
        size = struct.unpack_from("<I", body, 4)[0]
        order = ">" if (b0 & 1) else "<"
        if cls == 0:
            return np.dtype(f"{order}{'i' if b0 & 0x08 else 'u'}{size}")
        if cls == 1:
            return np.dtype(f"{order}f{size}")
        if cls == 3:
            return np.dtype(f"S{size}")
        if cls == 9:                                   # variable length (strings only)
            return np.dtype("O")
        raise ValueError(f"hdf5 reader: datatype class {cls} is not supported")

    def global_heap_object(self, addr: int, index: int) -> bytes:
        if addr not in self.gcol:
            p = self.at(addr)
            if self.d[p:p + 4] != b"GCOL":
                raise ValueError("hdf5 reader: bad global heap collection")
            size = struct.unpack_from("<Q", self.d, p + 8)[0]
            objs, q = {}, p + 16
            while q + 16 <= p + size:
                idx, _, osize = struct.unpack_from("<HH4xQ", self.d, q)
                if idx == 0:
                    break
                objs[idx] = self.d[q + 16:q + 16 + osize]
                q += 16 + osize + (-osize % 8)
            self.gcol[addr] = objs
        return self.gcol[addr][index]

    def attribute(self, body: bytes):
        ver = body[0]
        if ver not in (1, 2, 3):
            raise ValueError("hdf5 reader: unsupported attribute message version")
        nsz, tsz, ssz = struct.unpack_from("<HHH", body, 2)
        p = 8 + (1 if ver == 3 else 0)
        pad = (lambda n: n + (-n % 8)) if ver == 1 else (lambda n: n)
        name = body[p:p + nsz].split(b"\0")[0].decode("utf8")
        p += pad(nsz)
        tbody = body[p:p + tsz]
        p += pad(tsz)
        shape = self.dataspace(body[p:p + ssz]) if ssz >= 4 else ()
        p += pad(ssz)
        dt = self.datatype(tbody)
        count = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if dt == np.dtype("O"):                        # variable-length strings: (length, heap address, index)
            vals = []
            for i in range(count):
                _, gaddr, gidx = struct.unpack_from("<IQI", body, p + 16 * i)
                vals.append(self.global_heap_object(gaddr, gidx))
            arr = np.array(vals, dtype=object).reshape(shape) if shape else vals[0]
            return name, arr
        arr = np.frombuffer(body[p:p + count * dt.itemsize], dtype=dt, count=count)
        return name, (arr.reshape(shape).copy() if shape else arr[0])


def read_file(path) -> Group:
    with open(path, "rb") as f:
        data = f.read()
    r = _Reader(data)
    root = r.read_object(r.root_oh)
    if not isinstance(root, Group):
        raise ValueError("hdf5 reader: the root object is not a group")
    return root


def is_hdf5(path) -> bool:
    with open(path, "rb") as f:
        return f.read(8) == SIGNATURE


def as_text(v) -> str:
    """Attribute value (bytes, numpy bytes_, str) as text."""
    if isinstance(v, (bytes, np.bytes_)):
        return bytes(v).rstrip(b"\0").decode("utf8")
    return str(v)
