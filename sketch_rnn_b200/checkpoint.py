"""Checkpoint files under save/<dataset>/ (reference train.py:72-75, sample.py:59-64).

The reference saves with `tf.train.Saver(tf.all_variables())` of TensorFlow 0.x, i.e. the single-file
"V1" checkpoint: a LevelDB-style sorted table (tensorflow/core/lib/io/table) whose key "" maps to a
`SavedTensorSlices` protobuf holding the meta data (name, shape, dtype of every tensor) and whose other keys
(`EncodeTensorNameSlice(name, slice)`) map to `SavedTensorSlices{data: SavedSlice{name, slice,
data: TensorProto}}`.  This module reads and writes that container in pure Python so that
`save/kanji/model.ckpt-0` (absent from the reference mount) loads unchanged once it is
there, and so that files written here use the reference's format.

Validation status: TensorFlow cannot be installed here and no real V1 checkpoint is available, so the
format is restated from TensorFlow's published sources and checked by round trip only (tests/test_checkpoint.py).
The reader does not depend on the exact key encoding (it takes names and payloads from the protobuf values),
verifies block checksums (masked CRC-32C) and refuses compressed blocks with a clear error.

`write_checkpoint(path, {tf_variable_name: array})`, `read_checkpoint(path) -> {name: array}`.
An `.npz` container written by an earlier revision is still accepted on read.
"""
import io
import struct

import numpy as np

TF_TABLE_MAGIC = 0xdb4775248b80fb57
_BLOCK_SIZE = 256 * 1024
_RESTART_INTERVAL = 16
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_INT64 = 1, 2, 3, 9
_NP_OF_DT = {DT_FLOAT: np.float32, DT_DOUBLE: np.float64, DT_INT32: np.int32, DT_INT64: np.int64}

# ---- CRC-32C (Castagnoli), masked as LevelDB does ----------------------------------------------------
_CRC_TABLE = None


def _crc_table():
    global _CRC_TABLE
    if _CRC_TABLE is None:
        tbl = np.zeros(256, dtype=np.uint32)
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            tbl[i] = c
        _CRC_TABLE = [int(v) for v in tbl]
    return _CRC_TABLE


def crc32c(data, crc=0):
    try:  # fast path: the C implementation exported by libskrnn (host-only helper)
        from . import _lib
        lib = _lib.load()
        return int(lib.skrnn_crc32c(crc, bytes(data), len(data)))
    except Exception:
        pass
    tbl = _crc_table()
    c = crc ^ 0xFFFFFFFF
    for b in bytes(data):
        c = tbl[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def _mask(crc):
    return ((((crc >> 15) | (crc << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# ---- varints / protobuf wire format -----------------------------------------------------------------
def _varint(n):
    out = bytearray()
    n &= (1 << 64) - 1
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos):
    shift = res = 0
    while True:
        b = buf[pos]
        pos += 1
        res |= (b & 0x7F) << shift
        if not b & 0x80:
            return res, pos
        shift += 7


def _field(num, wire, payload):
    return _varint((num << 3) | wire) + payload


def _ld(num, payload):  # length-delimited
    return _field(num, 2, _varint(len(payload)) + payload)


def _parse(buf):
    """[(field_number, wire_type, value)] of one protobuf message (value: int or bytes)."""
    out, pos = [], 0
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        num, wire = key >> 3, key & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
        elif wire == 1:
            v, pos = buf[pos:pos + 8], pos + 8
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            v, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            v, pos = buf[pos:pos + 4], pos + 4
        else:
            raise IOError("unsupported protobuf wire type %d" % wire)
        out.append((num, wire, v))
    return out


def _shape_proto(shape):
    return b"".join(_ld(2, _field(1, 0, _varint(int(d)))) for d in shape)  # TensorShapeProto.dim{size}


def _parse_shape(buf):
    dims = []
    for num, _, v in _parse(buf):
        if num == 2:
            size = 0
            for n2, _, v2 in _parse(v):
                if n2 == 1:
                    size = v2
            dims.append(size)
    return tuple(dims)


# ---- OrderedCode pieces used by EncodeTensorNameSlice -------------------------------------------------
def _oc_num(v):
    b = b""
    while v > 0:
        b = bytes([v & 0xFF]) + b
        v >>= 8
    return bytes([len(b)]) + b


def _oc_string(s):
    return b"".join(b"\x00\xff" if c == 0 else (b"\xff\x00" if c == 255 else bytes([c])) for c in s) + b"\x00\x01"


def _oc_signed_small(v):
    assert -64 <= v < 64
    return bytes([0x80 ^ (v & 0xFF)])


def encode_tensor_name_slice(name, rank):
    """Key of the full slice of `name` (tensorflow/core/util/saved_tensor_slice_util.cc): 0, name,
    rank, then (start, length) = (0, -1) per dimension, all in OrderedCode."""
    key = _oc_num(0) + _oc_string(name.encode()) + _oc_num(rank)
    for _ in range(rank):
        key += _oc_signed_small(0) + _oc_signed_small(-1)
    return key


# ---- table writer -----------------------------------------------------------------------------------
class _BlockBuilder(object):
    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.count = 0
        self.last = b""

    def add(self, key, value):
        shared = 0
        if self.count % _RESTART_INTERVAL == 0 and self.count:
            self.restarts.append(len(self.buf))
        elif self.count:
            m = min(len(key), len(self.last))
            while shared < m and key[shared] == self.last[shared]:
                shared += 1
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value)) + key[shared:] + value
        self.last = key
        self.count += 1

    def finish(self):
        out = bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts) + struct.pack("<I", len(self.restarts))
        return out

    def size(self):
        return len(self.buf) + 4 * len(self.restarts) + 4


def _write_block(f, contents):
    off = f.tell()
    f.write(contents)
    f.write(b"\x00")  # kNoCompression
    f.write(struct.pack("<I", _mask(crc32c(contents + b"\x00"))))
    return off, len(contents)


def write_checkpoint(path, variables):
    names = sorted(variables)
    metas, entries = b"", []
    for n in names:
        a = np.asarray(variables[n])
        if a.dtype not in (np.float32, np.float64, np.int32, np.int64):
            a = a.astype(np.float32)
        dt = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.float64): DT_DOUBLE, np.dtype(np.int32): DT_INT32,
              np.dtype(np.int64): DT_INT64}[a.dtype]
        shape = _shape_proto(a.shape)
        slice_proto = b"".join(_ld(1, b"") for _ in a.shape)  # TensorSliceProto: one empty Extent per dim = full
        metas += _ld(1, _ld(1, n.encode()) + _ld(2, shape) + _field(3, 0, _varint(dt)) + _ld(4, slice_proto))
        if dt == DT_FLOAT:
            payload = _ld(5, a.astype("<f4").tobytes())            # TensorProto.float_val (packed)
        elif dt == DT_DOUBLE:
            payload = _ld(6, a.astype("<f8").tobytes())
        elif dt == DT_INT32:
            payload = _ld(7, b"".join(_varint(int(v)) for v in a.ravel()))
        else:
            payload = _ld(10, b"".join(_varint(int(v)) for v in a.ravel()))
        tensor = _field(1, 0, _varint(dt)) + _ld(2, shape) + payload
        saved_slice = _ld(1, n.encode()) + _ld(2, slice_proto) + _ld(3, tensor)
        entries.append((encode_tensor_name_slice(n, a.ndim), _ld(2, saved_slice)))
    # SavedTensorSliceMeta{tensor*, versions{producer}} under key ""
    meta = _ld(1, metas + _ld(2, _field(1, 0, _varint(0))))
    entries.append((b"", meta))
    entries.sort(key=lambda kv: kv[0])
    with open(path, "wb") as f:
        index, blk = [], _BlockBuilder()
        for k, v in entries:
            blk.add(k, v)
            if blk.size() >= _BLOCK_SIZE:
                off, size = _write_block(f, blk.finish())
                index.append((k, off, size))
                blk = _BlockBuilder()
        if blk.count:
            off, size = _write_block(f, blk.finish())
            index.append((blk.last, off, size))
        mi_off, mi_size = _write_block(f, _BlockBuilder().finish())  # empty metaindex
        ib = _BlockBuilder()
        for k, off, size in index:
            ib.add(k, _varint(off) + _varint(size))
        ix_off, ix_size = _write_block(f, ib.finish())
        footer = _varint(mi_off) + _varint(mi_size) + _varint(ix_off) + _varint(ix_size)
        f.write(footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", TF_TABLE_MAGIC))
    return path


# ---- table reader -----------------------------------------------------------------------------------
def _read_block(blob, off, size, verify=True):
    contents = blob[off:off + size]
    ctype = blob[off + size]
    if verify:
        want = struct.unpack("<I", blob[off + size + 1:off + size + 5])[0]
        if _mask(crc32c(contents + bytes([ctype]))) != want:
            raise IOError("checkpoint block at %d fails its CRC-32C check" % off)
    if ctype != 0:
        raise IOError("checkpoint block at %d is compressed (type %d); only uncompressed tables are supported" % (off, ctype))
    return contents


def _block_entries(block):
    nrest = struct.unpack("<I", block[-4:])[0]
    end = len(block) - 4 - 4 * nrest
    pos, key = 0, b""
    while pos < end:
        shared, pos = _read_varint(block, pos)
        unshared, pos = _read_varint(block, pos)
        vlen, pos = _read_varint(block, pos)
        key = key[:shared] + block[pos:pos + unshared]
        pos += unshared
        yield key, block[pos:pos + vlen]
        pos += vlen


def _tensor_from_proto(buf):
    dtype, shape, vals = DT_FLOAT, (), None
    for num, wire, v in _parse(buf):
        if num == 1:
            dtype = v
        elif num == 2:
            shape = _parse_shape(v)
        elif num == 4 and wire == 2:            # tensor_content
            vals = np.frombuffer(v, dtype=_NP_OF_DT.get(dtype, np.float32))
        elif num == 5:                          # float_val: packed or single
            vals = np.frombuffer(v, dtype="<f4") if wire == 2 else np.concatenate([vals if vals is not None else np.zeros(0, "<f4"), np.frombuffer(v, "<f4")])
        elif num == 6:
            vals = np.frombuffer(v, dtype="<f8") if wire == 2 else np.frombuffer(v, "<f8")
        elif num in (7, 10):
            if wire == 2:
                out, pos = [], 0
                while pos < len(v):
                    x, pos = _read_varint(v, pos)
                    out.append(x)
                vals = np.array(out, dtype=_NP_OF_DT[dtype])
            else:
                vals = np.array([v], dtype=_NP_OF_DT[dtype])
    n = int(np.prod(shape)) if shape else 1
    if vals is None:
        vals = np.zeros(n, dtype=_NP_OF_DT.get(dtype, np.float32))
    if vals.size == 1 and n > 1:
        vals = np.repeat(vals, n)  # TensorProto allows a single repeated value
    return np.array(vals).reshape(shape)


def _read_tf_table(blob):
    footer = blob[-48:]
    pos = 0
    _, pos = _read_varint(footer, pos)
    _, pos = _read_varint(footer, pos)
    ix_off, pos = _read_varint(footer, pos)
    ix_size, pos = _read_varint(footer, pos)
    out = {}
    for _, handle in _block_entries(_read_block(blob, ix_off, ix_size)):
        off, p = _read_varint(handle, 0)
        size, p = _read_varint(handle, p)
        for key, value in _block_entries(_read_block(blob, off, size)):
            if key == b"":
                continue  # meta entry: names / shapes are repeated inside every data entry
            for num, _, v in _parse(value):
                if num != 2:
                    continue
                name, tensor = None, None
                for n2, _, v2 in _parse(v):
                    if n2 == 1:
                        name = v2.decode()
                    elif n2 == 3:
                        tensor = _tensor_from_proto(v2)
                if name is not None and tensor is not None:
                    if name in out:
                        raise IOError("tensor %s is stored in several slices; partitioned variables are not supported" % name)
                    out[name] = tensor
    return out


def list_checkpoint(path):
    """{name: shape} of a checkpoint (cheap way to discover the variable names of a foreign file)."""
    return {k: tuple(v.shape) for k, v in read_checkpoint(path).items()}


def read_checkpoint(path):
    with open(path, "rb") as f:
        blob = f.read()
    if blob[:2] == b"PK":  # .npz container of an earlier revision
        with np.load(io.BytesIO(blob)) as z:
            return {k.replace("|", "/"): z[k] for k in z.files}
    if len(blob) >= 48 and struct.unpack("<Q", blob[-8:])[0] == TF_TABLE_MAGIC:
        return _read_tf_table(blob)
    raise IOError("%s is not a checkpoint this library can read" % path)
