"""Reader (and test writer) for TensorFlow "tensor bundle" checkpoints, without TensorFlow.

The reference restores the pretrained NSynth WaveNet encoder with ``tf.train.Saver(...).restore(sess,
'./nsynth/model/wavenet-ckpt/model.ckpt-200000')`` (methods.py:79-84): a V2 checkpoint, i.e. the pair
``<prefix>.index`` + ``<prefix>.data-00000-of-00001``.  This module reads that pair directly so that
``GatysNet.load_model`` here accepts the same ``--ckpt_path`` the reference does.

Formats (TensorFlow is a third-party dependency of the reference, not vendored; restated from its published
sources tensorflow/core/util/tensor_bundle/* and tensorflow/core/lib/io/{table,block,format}.cc, which are
LevelDB's table format):

* ``.index`` is an immutable sorted string table.  File = data blocks, a metaindex block, an index block and a
  48-byte footer (two varint BlockHandles {offset, size}, zero padding to 40 bytes, magic 0xdb4775248b80fb57 LE).
  Every block is followed by a 5-byte trailer {compression type (0 none, 1 snappy), masked CRC32C of block+type}.
  A block is a run of entries {varint shared, varint non_shared, varint value_len, key suffix, value} followed
  by the restart offsets (fixed32 each) and their count (fixed32).
* key "" -> BundleHeaderProto {1: num_shards, 2: endianness, 3: version}; every other key is a variable name ->
  BundleEntryProto {1: dtype, 2: TensorShapeProto, 3: shard_id, 4: offset, 5: size, 6: fixed32 masked crc32c,
  7: slices}.
* ``.data-SSSSS-of-NNNNN`` holds the raw little-endian tensor bytes at [offset, offset + size).

Variable names of the encoder (nsynth/wavenet/masked.py:141-145 under the scopes of model.py:93,106,113):
``ae_startconv/{W,biases}``, ``ae_dilatedconv_<i>/{W,biases}``, ``ae_res_<i>/{W,biases}``, i = 1..30; conv kernels
are stored [1, K, Cin, Cout], which is this package's [K][Cin][Cout] layout once the leading 1 is dropped.

NOT VERIFIED against a file written by TensorFlow itself (none is available offline); verified against the
format's own known answers (CRC32C check value, table magic, masking constant) and by write -> read round trips
(tests/test_tf_checkpoint.py).
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

TABLE_MAGIC = 0xdb4775248b80fb57
FOOTER_LEN = 48
BLOCK_TRAILER = 5
MASK_DELTA = 0xa282ead8

# tensorflow/core/framework/types.proto
DTYPES = {1: np.dtype('<f4'), 2: np.dtype('<f8'), 3: np.dtype('<i4'), 4: np.dtype('u1'), 5: np.dtype('<i2'),
          6: np.dtype('i1'), 9: np.dtype('<i8'), 10: np.dtype('?'), 17: np.dtype('<u2'), 19: np.dtype('<f2'),
          22: np.dtype('<u4'), 23: np.dtype('<u8')}
DT_OF = {np.dtype('float32'): 1, np.dtype('float64'): 2, np.dtype('int32'): 3, np.dtype('int64'): 9}


class CheckpointError(ValueError):
    pass


# ---- CRC32C (Castagnoli), masked as LevelDB / TF store it -----------------------------------------
def _make_table():
    poly = 0x82F63B78
    t = np.zeros(256, np.uint32)
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ (poly if c & 1 else 0)
        t[i] = c
    return t


_T0 = _make_table()
_T = None


def _slice_tables():
    """Slicing-by-8 tables."""
    global _T
    if _T is None:
        T = np.zeros((8, 256), np.uint32)
        T[0] = _T0
        for k in range(1, 8):
            T[k] = (T[k - 1] >> 8) ^ _T0[T[k - 1] & 0xFF]
        _T = [[int(v) for v in T[k]] for k in range(8)]
    return _T


def crc32c(data: bytes, crc: int = 0) -> int:
    """CRC-32C of `data` (check value: crc32c(b'123456789') == 0xE3069283)."""
    T = _slice_tables()
    t0, t1, t2, t3, t4, t5, t6, t7 = T
    c = crc ^ 0xFFFFFFFF
    mv = memoryview(data).cast('B')
    n = len(mv)
    n8 = n - (n % 8)
    if n8:
        words = np.frombuffer(mv[:n8], dtype='<u8').tolist()
        for w in words:
            w ^= c
            c = (t7[w & 0xFF] ^ t6[(w >> 8) & 0xFF] ^ t5[(w >> 16) & 0xFF] ^ t4[(w >> 24) & 0xFF] ^
                 t3[(w >> 32) & 0xFF] ^ t2[(w >> 40) & 0xFF] ^ t1[(w >> 48) & 0xFF] ^ t0[(w >> 56) & 0xFF])
    for b in mv[n8:]:
        c = t0[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def mask_crc(crc: int) -> int:
    return (((crc >> 15) | (crc << 17)) + MASK_DELTA) & 0xFFFFFFFF


def unmask_crc(masked: int) -> int:
    rot = (masked - MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ---- varints / protobuf wire format ------------------------------------------------------------------
def _get_varint(buf, pos: int) -> Tuple[int, int]:
    shift = result = 0
    while True:
        if pos >= len(buf):
            raise CheckpointError("truncated varint")
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 63:
            raise CheckpointError("varint too long")


def _put_varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _proto_fields(buf) -> Iterable[Tuple[int, int, object]]:
    """(field number, wire type, value) of one protobuf message; value is int or bytes."""
    pos = 0
    while pos < len(buf):
        tag, pos = _get_varint(buf, pos)
        fno, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _get_varint(buf, pos)
        elif wt == 1:
            v = struct.unpack_from('<Q', buf, pos)[0]
            pos += 8
        elif wt == 2:
            ln, pos = _get_varint(buf, pos)
            v = bytes(buf[pos:pos + ln])
            if len(v) != ln:
                raise CheckpointError("truncated length-delimited field")
            pos += ln
        elif wt == 5:
            v = struct.unpack_from('<I', buf, pos)[0]
            pos += 4
        else:
            raise CheckpointError(f"unsupported protobuf wire type {wt}")
        yield fno, wt, v


def _signed64(v: int) -> int:
    return v - (1 << 64) if v >= (1 << 63) else v


# ---- snappy (the table builder may compress blocks; the bundle writer does not, but be liberal) ------
def _snappy_decompress(buf: bytes) -> bytes:
    n, pos = _get_varint(buf, 0)
    out = bytearray()
    while pos < len(buf):
        tag = buf[pos]
        pos += 1
        kind = tag & 3
        if kind == 0:
            ln = tag >> 2
            if ln >= 60:
                nb = ln - 59
                ln = int.from_bytes(buf[pos:pos + nb], 'little')
                pos += nb
            ln += 1
            out += buf[pos:pos + ln]
            pos += ln
            continue
        if kind == 1:
            ln = ((tag >> 2) & 7) + 4
            off = ((tag >> 5) << 8) | buf[pos]
            pos += 1
        elif kind == 2:
            ln = (tag >> 2) + 1
            off = int.from_bytes(buf[pos:pos + 2], 'little')
            pos += 2
        else:
            ln = (tag >> 2) + 1
            off = int.from_bytes(buf[pos:pos + 4], 'little')
            pos += 4
        if off == 0 or off > len(out):
            raise CheckpointError("corrupt snappy block")
        for _ in range(ln):
            out.append(out[-off])
    if len(out) != n:
        raise CheckpointError("snappy length mismatch")
    return bytes(out)


# ---- the sorted string table (.index) ---------------------------------------------------------------------
def _read_block(buf: bytes, offset: int, size: int, verify: bool) -> bytes:
    end = offset + size + BLOCK_TRAILER
    if offset < 0 or end > len(buf):
        raise CheckpointError("block handle outside the file")
    body, ctype = buf[offset:offset + size], buf[offset + size]
    if verify:
        stored = struct.unpack_from('<I', buf, offset + size + 1)[0]
        if unmask_crc(stored) != crc32c(buf[offset:offset + size + 1]):
            raise CheckpointError(f"block checksum mismatch at offset {offset}")
    if ctype == 0:
        return body
    if ctype == 1:
        return _snappy_decompress(body)
    raise CheckpointError(f"unknown block compression type {ctype}")


def _block_entries(block: bytes) -> List[Tuple[bytes, bytes]]:
    if len(block) < 4:
        raise CheckpointError("block too short")
    n_restarts = struct.unpack_from('<I', block, len(block) - 4)[0]
    limit = len(block) - 4 - 4 * n_restarts
    if limit < 0:
        raise CheckpointError("bad restart array")
    out, pos, key = [], 0, b''
    while pos < limit:
        shared, pos = _get_varint(block, pos)
        non_shared, pos = _get_varint(block, pos)
        vlen, pos = _get_varint(block, pos)
        if shared > len(key) or pos + non_shared + vlen > limit:
            raise CheckpointError("corrupt block entry")
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        out.append((key, block[pos:pos + vlen]))
        pos += vlen
    return out


def read_table(path: str, verify: bool = True) -> Dict[bytes, bytes]:
    """All key -> value pairs of a LevelDB-format table file, in file order."""
    with open(path, 'rb') as f:
        buf = f.read()
    if len(buf) < FOOTER_LEN:
        raise CheckpointError(f"{path}: too short for a table footer")
    footer = buf[-FOOTER_LEN:]
    if struct.unpack_from('<Q', footer, 40)[0] != TABLE_MAGIC:
        raise CheckpointError(f"{path}: not a table file (bad magic)")
    pos = 0
    _mo, pos = _get_varint(footer, pos)
    _ms, pos = _get_varint(footer, pos)
    io, pos = _get_varint(footer, pos)
    isz, pos = _get_varint(footer, pos)
    out: Dict[bytes, bytes] = {}
    for _key, handle in _block_entries(_read_block(buf, io, isz, verify)):
        off, p = _get_varint(handle, 0)
        size, p = _get_varint(handle, p)
        for k, v in _block_entries(_read_block(buf, off, size, verify)):
            out[k] = v
    return out


# ---- the bundle ------------------------------------------------------------------------------------------
class BundleEntry:
    __slots__ = ('dtype', 'shape', 'shard_id', 'offset', 'size', 'crc32c', 'sliced')

    def __init__(self):
        self.dtype, self.shape, self.shard_id, self.offset, self.size, self.crc32c, self.sliced = 0, (), 0, 0, 0, None, False


def _parse_shape(buf: bytes) -> Tuple[int, ...]:
    dims = []
    for fno, wt, v in _proto_fields(buf):
        if fno == 2 and wt == 2:                       # repeated Dim dim = 2
            size = 0
            for f2, w2, v2 in _proto_fields(v):
                if f2 == 1 and w2 == 0:
                    size = _signed64(v2)
            dims.append(size)
        elif fno == 3 and wt == 0 and v:
            raise CheckpointError("tensor of unknown rank")
    return tuple(dims)


def parse_entry(buf: bytes) -> BundleEntry:
    e = BundleEntry()
    for fno, wt, v in _proto_fields(buf):
        if fno == 1 and wt == 0:
            e.dtype = v
        elif fno == 2 and wt == 2:
            e.shape = _parse_shape(v)
        elif fno == 3 and wt == 0:
            e.shard_id = v
        elif fno == 4 and wt == 0:
            e.offset = _signed64(v)
        elif fno == 5 and wt == 0:
            e.size = _signed64(v)
        elif fno == 6 and wt == 5:
            e.crc32c = v
        elif fno == 7:
            e.sliced = True
    return e


def parse_header(buf: bytes) -> Tuple[int, int]:
    """(num_shards, endianness) of the BundleHeaderProto; endianness 0 = little."""
    num_shards, endian = 1, 0
    for fno, wt, v in _proto_fields(buf):
        if fno == 1 and wt == 0:
            num_shards = v
        elif fno == 2 and wt == 0:
            endian = v
    return num_shards, endian


def is_bundle(prefix: str) -> bool:
    return os.path.exists(str(prefix) + '.index')


class BundleReader:
    """``BundleReader(prefix)[name]`` -> numpy array; ``.keys()`` lists the variables."""

    def __init__(self, prefix: str, verify: bool = True):
        self.prefix, self.verify = str(prefix), verify
        table = read_table(self.prefix + '.index', verify)
        if b'' not in table:
            raise CheckpointError(f"{self.prefix}.index has no bundle header entry")
        self.num_shards, endian = parse_header(table[b''])
        if endian != 0:
            raise CheckpointError("big-endian bundles are not supported")
        self.entries = {k.decode('utf-8'): parse_entry(v) for k, v in table.items() if k != b''}

    def keys(self) -> List[str]:
        return sorted(self.entries)

    def __contains__(self, name: str) -> bool:
        return name in self.entries

    def shape(self, name: str) -> Tuple[int, ...]:
        return self.entries[name].shape

    def _shard(self, shard_id: int) -> str:
        return f"{self.prefix}.data-{shard_id:05d}-of-{self.num_shards:05d}"

    def __getitem__(self, name: str) -> np.ndarray:
        if name not in self.entries:
            raise KeyError(f"{name!r} is not in checkpoint {self.prefix} ({len(self.entries)} variables)")
        e = self.entries[name]
        if e.sliced:
            raise CheckpointError(f"{name}: partitioned (sliced) variables are not supported")
        if e.dtype not in DTYPES:
            raise CheckpointError(f"{name}: unsupported dtype enum {e.dtype}")
        dt = DTYPES[e.dtype]
        count = int(np.prod(e.shape, dtype=np.int64)) if e.shape else 1
        if count * dt.itemsize != e.size:
            raise CheckpointError(f"{name}: shape {e.shape} x {dt} does not match {e.size} stored bytes")
        with open(self._shard(e.shard_id), 'rb') as f:
            f.seek(e.offset)
            raw = f.read(e.size)
        if len(raw) != e.size:
            raise CheckpointError(f"{name}: data shard truncated")
        if self.verify and e.crc32c is not None and unmask_crc(e.crc32c) != crc32c(raw):
            raise CheckpointError(f"{name}: tensor checksum mismatch")
        return np.frombuffer(raw, dtype=dt).reshape(e.shape).copy()


# ---- the encoder's variables (model.py:88-114, masked.py:141-145) -----------------------------------------
EMA_SUFFIX = '/ExponentialMovingAverage'


def encoder_variable_names(n_layers: int = 30) -> List[str]:
    names = ['ae_startconv/W', 'ae_startconv/biases']
    for i in range(1, n_layers + 1):
        names += [f'ae_dilatedconv_{i}/W', f'ae_dilatedconv_{i}/biases', f'ae_res_{i}/W', f'ae_res_{i}/biases']
    return names


def load_encoder_weights(prefix: str, n_layers: int = 30, ema: bool = False, verify: bool = True):
    """The NSynth encoder's weights in ast_set_weights layouts (synth.EncoderWeights).

    The reference restores the plain variable names (tf.train.Saver over tf.global_variables(), methods.py:80-84);
    ema=True reads the '<name>/ExponentialMovingAverage' shadows the NSynth checkpoint also carries."""
    from . import synth
    rd = BundleReader(prefix, verify)
    sfx = EMA_SUFFIX if ema else ''

    def get(name, shape):
        a = rd[name + sfx]
        if tuple(a.shape) != tuple(shape):
            raise CheckpointError(f"{name + sfx}: shape {tuple(a.shape)}, expected {tuple(shape)}")
        if a.dtype != np.float32:
            raise CheckpointError(f"{name + sfx}: dtype {a.dtype}, expected float32")
        return a

    C = 128
    W0 = get('ae_startconv/W', (1, 3, 1, C))[0]
    b0 = get('ae_startconv/biases', (C,))
    Wd = np.stack([get(f'ae_dilatedconv_{i}/W', (1, 3, C, C))[0] for i in range(1, n_layers + 1)])
    bd = np.stack([get(f'ae_dilatedconv_{i}/biases', (C,)) for i in range(1, n_layers + 1)])
    Wr = np.stack([get(f'ae_res_{i}/W', (1, 1, C, C))[0, 0] for i in range(1, n_layers + 1)])
    br = np.stack([get(f'ae_res_{i}/biases', (C,)) for i in range(1, n_layers + 1)])
    return synth.EncoderWeights(W0, b0, Wd, bd, Wr, br)


# ---- writer (tests, fixtures, exporting weights for the reference) ------------------------------------------
def _block(entries: List[Tuple[bytes, bytes]], restart_interval: int = 16) -> bytes:
    out, restarts, prev = bytearray(), [], b''
    for i, (k, v) in enumerate(entries):
        shared = 0
        if i % restart_interval == 0:
            restarts.append(len(out))
        else:
            while shared < min(len(prev), len(k)) and prev[shared] == k[shared]:
                shared += 1
        out += _put_varint(shared) + _put_varint(len(k) - shared) + _put_varint(len(v)) + k[shared:] + v
        prev = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        out += struct.pack('<I', r)
    out += struct.pack('<I', len(restarts))
    return bytes(out)


def _tag(fno: int, wt: int) -> bytes:
    return _put_varint((fno << 3) | wt)


def _entry_proto(dtype: int, shape, shard_id: int, offset: int, size: int, crc_masked: int) -> bytes:
    sh = bytearray()
    for d in shape:
        dim = _tag(1, 0) + _put_varint(int(d))
        sh += _tag(2, 2) + _put_varint(len(dim)) + dim
    out = _tag(1, 0) + _put_varint(dtype) + _tag(2, 2) + _put_varint(len(sh)) + bytes(sh)
    if shard_id:
        out += _tag(3, 0) + _put_varint(shard_id)
    if offset:
        out += _tag(4, 0) + _put_varint(offset)
    out += _tag(5, 0) + _put_varint(size) + _tag(6, 5) + struct.pack('<I', crc_masked)
    return out


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray], block_size: int = 4096) -> None:
    """Write {name: array} as a one-shard tensor bundle in the layout BundleWriter produces (sorted keys, header
    under the empty key, uncompressed blocks of about `block_size` bytes, restart interval 16)."""
    prefix = str(prefix)
    items, off = [], 0
    with open(prefix + '.data-00000-of-00001', 'wb') as f:
        for name in sorted(tensors):
            a = np.asarray(tensors[name])                                   # (ascontiguousarray would turn 0-d into 1-d)
            if a.dtype not in DT_OF:
                raise CheckpointError(f"{name}: dtype {a.dtype} not supported by the writer")
            raw = a.astype(a.dtype.newbyteorder('<'), copy=False).tobytes(order='C')
            f.write(raw)
            items.append((name.encode('utf-8'), _entry_proto(DT_OF[a.dtype], a.shape, 0, off, len(raw), mask_crc(crc32c(raw)))))
            off += len(raw)
    version = _tag(1, 0) + _put_varint(1)                                   # VersionDef.producer = 1
    header = _tag(1, 0) + _put_varint(1) + _tag(3, 2) + _put_varint(len(version)) + version
    entries = [(b'', header)] + items

    out = bytearray()

    def emit(block: bytes) -> bytes:
        handle = _put_varint(len(out)) + _put_varint(len(block))
        out.extend(block)
        out.append(0)                                                       # kNoCompression
        out.extend(struct.pack('<I', mask_crc(crc32c(block + b'\x00'))))
        return handle

    index, cur, cur_bytes = [], [], 0
    for k, v in entries:
        cur.append((k, v))
        cur_bytes += len(k) + len(v) + 3
        if cur_bytes >= block_size:
            index.append((cur[-1][0], emit(_block(cur))))
            cur, cur_bytes = [], 0
    if cur:
        index.append((cur[-1][0], emit(_block(cur))))
    meta_handle = emit(_block([]))
    index_handle = emit(_block(index, restart_interval=1))
    footer = meta_handle + index_handle
    footer += b'\x00' * (40 - len(footer)) + struct.pack('<Q', TABLE_MAGIC)
    out.extend(footer)
    with open(prefix + '.index', 'wb') as f:
        f.write(bytes(out))


def save_encoder_weights(prefix: str, w, ema: bool = False) -> None:
    """Inverse of load_encoder_weights: the encoder as a TF tensor bundle under the reference's variable names."""
    sfx = EMA_SUFFIX if ema else ''
    t = {'ae_startconv/W' + sfx: np.asarray(w.W0, np.float32)[None], 'ae_startconv/biases' + sfx: np.asarray(w.b0, np.float32)}
    for i in range(len(w.Wd)):
        t[f'ae_dilatedconv_{i + 1}/W' + sfx] = np.asarray(w.Wd[i], np.float32)[None]
        t[f'ae_dilatedconv_{i + 1}/biases' + sfx] = np.asarray(w.bd[i], np.float32)
        t[f'ae_res_{i + 1}/W' + sfx] = np.asarray(w.Wr[i], np.float32)[None, None]
        t[f'ae_res_{i + 1}/biases' + sfx] = np.asarray(w.br[i], np.float32)
    write_bundle(prefix, t)
