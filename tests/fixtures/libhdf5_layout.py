"""Hand-assembled HDF5 files in the layout libhdf5 1.8 / h5py (libver='earliest') produces, built from the HDF5 File
Format Specification 2.0 with nothing but ``struct`` -- deliberately NOT through the package's own writer
(``csrc/rt/hdf5_subset.cpp``), so that the reader is tested against bytes it did not produce.

What is reproduced of libhdf5's habits (each one a place where a reader written against its own writer can go wrong):

* superblock version 0, base address 0, free-space / driver-info addresses undefined, group leaf K = 4 and internal
  K = 16, root symbol-table entry with cache type 1 (B-tree and heap address in the scratch pad);
* version-1 object headers whose first chunk is too small for the attributes Keras adds afterwards: the header ends
  in an OBJECT HEADER CONTINUATION message (0x0010) and the attributes live in a continuation block elsewhere in the
  file, padded with NIL messages (type 0) exactly as the library leaves them;
* version-1 attribute messages: name, datatype and dataspace each padded to a multiple of 8 bytes, name size counting
  the terminating NUL;
* ``layer_names`` / ``weight_names`` as arrays of FIXED-LENGTH, NUL-padded strings (numpy ``S`` arrays: datatype class 3,
  padding type 1), including the zero-length array Keras writes for layers without weights;
* version-1 simple dataspaces WITH maximum dimensions (flag bit 0), version-3 contiguous layout messages, version-2
  fill-value and modification-time messages in dataset headers (a reader has to skip them);
* "old-style" groups: symbol-table message -> v1 B-tree node (``TREE``, 2K+1 keys) -> symbol-table nodes (``SNOD``, at
  most 2 * leaf K = 8 entries each, so 9+ links give several leaves) -> names in a local heap (``HEAP``) with a free
  block at the end; group entries carry cache type 1, dataset entries cache type 0;
* Python ``bytes`` attributes (``backend``, ``keras_version``: h5py maps a bytes scalar to a VARIABLE-LENGTH string) as
  class-9 datatypes whose characters sit in a global heap collection (``GCOL``, 4096 bytes, free-space object last) and
  whose attribute data is the 16-byte (length, collection address, object index) reference;
* datasets allocated AFTER all metadata of their group, 8-byte aligned, little-endian IEEE floats with the full
  12-byte floating-point property block.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
LEAF_K, INTERNAL_K = 4, 16


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _dtype_msg(dt: np.dtype) -> bytes:
    dt = np.dtype(dt)
    if dt.kind == "f":
        size = dt.itemsize
        exp_bits, mant_bits, bias = {4: (8, 23, 127), 8: (11, 52, 1023)}[size]
        # class 1, version 1; bits: little-endian, mantissa normalisation 2 (msb implied), sign bit position
        head = struct.pack("<BBBBI", 0x11, 0x20, size * 8 - 1, 0, size)
        props = struct.pack("<HHBBBBI", 0, size * 8, mant_bits, exp_bits, 0, mant_bits, bias)
        return head + props
    if dt.kind in "iu":
        head = struct.pack("<BBBBI", 0x10, 0x08 if dt.kind == "i" else 0x00, 0, 0, dt.itemsize)
        return head + struct.pack("<HH", 0, dt.itemsize * 8)
    if dt.kind == "S":
        # class 3, version 1; padding type 1 (NUL-padded, what h5py writes for numpy S), ASCII
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, dt.itemsize)
    raise TypeError(dt)


def _space_msg(shape) -> bytes:
    rank = len(shape)
    if rank == 0:
        return struct.pack("<BBBBI", 1, 0, 0, 0, 0)
    out = struct.pack("<BBBBI", 1, rank, 1, 0, 0)          # flag 1: maximum dimensions present
    out += b"".join(struct.pack("<Q", d) for d in shape)
    out += b"".join(struct.pack("<Q", d) for d in shape)
    return out


def _vlen_attr_msg(name: str, length: int, heap_addr: int, index: int) -> bytes:
    nm = name.encode() + b"\0"
    # class 9 (variable length), type = string, null-terminated, ASCII; descriptor size 16; base type: 1-byte string
    dt = struct.pack("<BBBBI", 0x19, 0x01, 0x00, 0, 16) + struct.pack("<BBBBI", 0x13, 0x00, 0, 0, 1)
    sp = _space_msg(())
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(sp))
    return body + _pad8(nm) + _pad8(dt) + _pad8(sp) + struct.pack("<IQI", length, heap_addr, index)


def _attr_msg(name: str, value) -> bytes:
    arr = np.asarray(value)
    if arr.dtype.kind == "U":
        arr = arr.astype("S")
    nm = name.encode() + b"\0"
    dt, sp = _dtype_msg(arr.dtype), _space_msg(arr.shape)
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(sp))
    return body + _pad8(nm) + _pad8(dt) + _pad8(sp) + arr.tobytes()


def _msg(mtype: int, data: bytes, flags: int = 0) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), flags) + data


class _Node:
    def __init__(self, is_group, value=None):
        self.is_group = is_group
        self.value = value
        self.attrs = []
        self.children = {}
        self.addr = None
        self.btree = self.heap = None


class LibHdf5File:
    """Collect groups / datasets / attributes, then ``tobytes()`` lays the file out."""

    def __init__(self):
        self.root = _Node(True)

    def _walk(self, path, create_groups=True):
        node = self.root
        for part in [p for p in path.split("/") if p]:
            if part not in node.children:
                node.children[part] = _Node(True)
            node = node.children[part]
        return node

    def group(self, path):
        return self._walk(path)

    def attr(self, path, name, value):
        self._walk(path).attrs.append((name, value))

    def dataset(self, path, value):
        parts = [p for p in path.split("/") if p]
        parent = self._walk("/".join(parts[:-1]))
        parent.children[parts[-1]] = _Node(False, np.ascontiguousarray(value))

    # ------------------------------------------------------------------ layout
    def _vlen(self, data: bytes):
        """Store ``data`` in the file's global heap collection (created on first use); returns (address, index)."""
        if self._gcol is None:
            self._gcol = self._alloc(4096)
            self._gcol_used, self._gcol_next = 16, 1
        size = 16 + len(_pad8(data))
        assert self._gcol_used + size + 16 <= 4096, "fixture builder: one global heap collection only"
        at = self._gcol + self._gcol_used
        self._put(at, struct.pack("<HHIQ", self._gcol_next, 1, 0, len(data)) + data)
        self._gcol_used += size
        idx, self._gcol_next = self._gcol_next, self._gcol_next + 1
        free = 4096 - self._gcol_used
        self._put(self._gcol, b"GCOL" + struct.pack("<B3xQ", 1, 4096))
        self._put(self._gcol + self._gcol_used, struct.pack("<HHIQ", 0, 0, 0, free))
        return self._gcol, idx

    def _attr_messages(self, node):
        out = []
        for n, v in node.attrs:
            if isinstance(v, bytes):             # what h5py does with a Python bytes scalar
                addr, idx = self._vlen(v)
                out.append(_msg(0x000C, _vlen_attr_msg(n, len(v), addr, idx)))
            else:
                out.append(_msg(0x000C, _attr_msg(n, v)))
        return out

    def tobytes(self) -> bytes:
        self.buf = bytearray(96)                 # superblock v0 is 96 bytes; the root object header follows it
        self._gcol = None
        self._emit_group(self.root)
        eof = len(self.buf)
        sb = b"\x89HDF\r\n\x1a\n" + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0)
        sb += struct.pack("<HHI", LEAF_K, INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
        sb += struct.pack("<QQII", 0, self.root.addr, 1, 0) + struct.pack("<QQ", self.root.btree, self.root.heap)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)

    def _alloc(self, n, align=8):
        while len(self.buf) % align:
            self.buf.append(0)
        a = len(self.buf)
        self.buf.extend(b"\0" * n)
        return a

    def _put(self, addr, data):
        self.buf[addr:addr + len(data)] = data

    def _object_header(self, first_len, has_later, spare=8):
        """Allocate a v1 object header whose first chunk holds ``first_len`` bytes of messages (+ ``spare`` bytes the
        library left unused, + a continuation message if attributes are added later).  Returns (address, finish);
        ``finish(first_msgs, later_msgs)`` writes it once everything the library allocates in between (B-tree, heap,
        children, raw data) has its address: ``later_msgs`` go to a continuation block at the current end of file."""
        first_chunk_size = first_len + spare + (24 if has_later else 0)
        addr = self._alloc(16 + first_chunk_size)

        def finish(first_msgs, later_msgs):
            chunk = b"".join(first_msgs)
            assert len(chunk) == first_len
            n = len(first_msgs)
            if spare:                                        # NIL message over the unused part of the first chunk
                chunk += struct.pack("<HHB3x", 0, spare - 8, 0) + b"\0" * (spare - 8)
                n += 1
            if has_later:
                cbody = b"".join(later_msgs)
                slack = 24                                   # continuation blocks are over-allocated and NIL-filled
                caddr = self._alloc(len(cbody) + slack)
                self._put(caddr, cbody + struct.pack("<HHB3x", 0, slack - 8, 0) + b"\0" * (slack - 8))
                chunk += _msg(0x0010, struct.pack("<QQ", caddr, len(cbody) + slack))
                n += 1 + len(later_msgs) + 1
            assert len(chunk) == first_chunk_size
            self._put(addr, struct.pack("<BBHII4x", 1, 0, n, 1, first_chunk_size) + chunk)

        return addr, finish

    def _emit_group(self, g: _Node):
        names = sorted(g.children)                           # symbol-table entries are kept in name order
        # a fresh group header has room for the symbol-table message (24 bytes) and little else
        g.addr, finish = self._object_header(24, bool(g.attrs), spare=16)
        # B-tree node: 2K + 1 keys and 2K children are always allocated
        g.btree = self._alloc(24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8)
        # local heap: header, then the data segment (empty name at offset 0, names 8-byte aligned, free block last)
        heap_data = bytearray(b"\0" * 8)
        offs = {}
        for n in names:
            offs[n] = len(heap_data)
            heap_data += _pad8(n.encode() + b"\0")
        free_off = len(heap_data)
        heap_data += struct.pack("<QQ", 1, 32) + b"\0" * 16   # free block: next = 1 (none), size 32
        g.heap = self._alloc(32)
        seg = self._alloc(len(heap_data))
        self._put(g.heap, b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free_off, seg))
        self._put(seg, bytes(heap_data))
        # symbol-table nodes are allocated when the first link is inserted, i.e. before the children's own objects
        leaves = [names[i:i + 2 * LEAF_K] for i in range(0, len(names), 2 * LEAF_K)] or [[]]
        assert len(leaves) <= 2 * INTERNAL_K, "fixture builder: one B-tree level only"
        leaf_addr = [self._alloc(8 + 2 * LEAF_K * 40) for _ in leaves]
        for n in names:                                      # children (depth first, like h5py creating them in order)
            c = g.children[n]
            if c.is_group:
                self._emit_group(c)
            else:
                self._emit_dataset(c)
        for la, leaf in zip(leaf_addr, leaves):
            ent = b""
            for n in leaf:
                c = g.children[n]
                if c.is_group:
                    ent += struct.pack("<QQII", offs[n], c.addr, 1, 0) + struct.pack("<QQ", c.btree, c.heap)
                else:
                    ent += struct.pack("<QQII", offs[n], c.addr, 0, 0) + b"\0" * 16
            self._put(la, b"SNOD" + struct.pack("<BBH", 1, 0, len(leaf)) + ent)
        keys = [0] + [offs[leaf[-1]] if leaf else 0 for leaf in leaves]
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(leaves) if names else 0, UNDEF, UNDEF)
        for i, la in enumerate(leaf_addr):
            node += struct.pack("<QQ", keys[i], la)
        node += struct.pack("<Q", keys[len(leaves)])
        self._put(g.btree, node)
        # the attributes were added last (Keras writes them after the weights): continuation block at the end
        finish([_msg(0x0011, struct.pack("<QQ", g.btree, g.heap))], self._attr_messages(g))

    def _emit_dataset(self, d: _Node):
        arr = d.value
        space = _msg(0x0001, _space_msg(arr.shape))
        dtype = _msg(0x0003, _dtype_msg(arr.dtype), flags=1)         # constant message
        fill = _msg(0x0005, struct.pack("<BBBB", 2, 2, 2, 1) + struct.pack("<I", 0))
        mtime = _msg(0x0012, struct.pack("<B3xI", 1, 1541548800))
        attr_msgs = self._attr_messages(d)
        layout_len = len(_msg(0x0008, struct.pack("<BBQQ", 3, 1, 0, 0)))
        d.addr, finish = self._object_header(len(space) + len(dtype) + len(fill) + layout_len + len(mtime), bool(attr_msgs))
        data = self._alloc(arr.nbytes) if arr.nbytes else UNDEF      # raw data after the header, 8-byte aligned
        if arr.nbytes:
            self._put(data, arr.tobytes())
        finish([space, dtype, fill, _msg(0x0008, struct.pack("<BBQQ", 3, 1, data, arr.nbytes)), mtime], attr_msgs)
