"""TensorFlow "V2" checkpoint (tensor bundle) ingestion without TensorFlow.

The reference restores its models with `tf.train.Saver` through tfsnippet's `CheckpointSaver`
(tfsnippet/scaffold/checkpoint.py:74-287; `restore_checkpoint = '…/checkpoint/checkpoint.dat-390000'`,
experiment/dmm/dmm_cifar10.py:1105-1126).  A Saver checkpoint with prefix P is

    P.index                   a LevelDB-format table (TF `table::Table`, uncompressed blocks) mapping
                              ""            -> BundleHeaderProto {num_shards, endianness, version}
                              variable name -> BundleEntryProto {dtype, shape, shard_id, offset, size, crc32c}
    P.data-0000S-of-0000N     the raw little-endian tensor bytes, addressed by (shard_id, offset, size)
    <dir>/checkpoint          text: model_checkpoint_path: "…"   (what tf.train.latest_checkpoint reads)

Variable names are the TF variable names without ":0" — exactly the names `apvae_load_weights` takes
(`q_net/level_0/residual/conv_0/kernel`, …, `p_net/xi`, `beta`, `get_log_Z/log_Z`).  This module reads that format (and writes
it, so that fixtures can be produced here, where TensorFlow cannot be installed); it is host-side only.

Format parity is UNPINNED: there is no TensorFlow in this image and the reference ships no checkpoint file, so the
reader is checked against this module's own writer, which follows the published table / tensor-bundle layout
(leveldb `table_format.md`; tensorflow/core/util/tensor_bundle, tensorflow/core/protobuf/tensor_bundle.proto).
"""
import os
import re
import struct
from collections import OrderedDict

import numpy as np

__all__ = ['crc32c', 'list_variables', 'load_checkpoint', 'save_checkpoint', 'latest_checkpoint',
           'model_weights_from_checkpoint', 'CheckpointError']

TABLE_MAGIC = 0xdb4775248b80fb57
FOOTER_LEN = 48
BLOCK_TRAILER = 5                      # 1 byte compression type + 4 bytes masked crc32c

# tensorflow/core/framework/types.proto
_DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8, 9: np.int64, 10: np.bool_,
           17: np.uint16, 19: np.float16, 22: np.uint32, 23: np.uint64}
_DT_STRING, _DT_BFLOAT16 = 7, 14
_DT_OF = {np.dtype(v): k for k, v in _DTYPES.items()}


class CheckpointError(ValueError):
    pass


# ------------------------------------------------------------------ crc32c (Castagnoli), masked as leveldb / TF do
def _make_table():
    t = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        t.append(c)
    return t


_CRC_TABLE = _make_table()


def crc32c(data, crc=0):
    crc ^= 0xFFFFFFFF
    t = _CRC_TABLE
    for b in bytes(data):
        crc = t[(crc ^ b) & 0xFF] ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def _mask(crc):
    return ((((crc >> 15) | (crc << 17)) & 0xFFFFFFFF) + 0xa282ead8) & 0xFFFFFFFF


def _unmask(m):
    rot = (m - 0xa282ead8) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ------------------------------------------------------------------ varints / protobuf wire format
def _get_varint(buf, pos):
    shift, out = 0, 0
    while True:
        if pos >= len(buf):
            raise CheckpointError('truncated varint')
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7
        if shift > 63:
            raise CheckpointError('varint too long')


def _put_varint(v):
    v &= (1 << 64) - 1
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _proto_fields(buf):
    """[(field number, wire type, value)] of one message; length-delimited values stay bytes."""
    pos, out = 0, []
    while pos < len(buf):
        key, pos = _get_varint(buf, pos)
        fno, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _get_varint(buf, pos)
        elif wt == 1:
            v = struct.unpack_from('<Q', buf, pos)[0]
            pos += 8
        elif wt == 2:
            n, pos = _get_varint(buf, pos)
            v = bytes(buf[pos:pos + n])
            if len(v) != n:
                raise CheckpointError('truncated protobuf field')
            pos += n
        elif wt == 5:
            v = struct.unpack_from('<I', buf, pos)[0]
            pos += 4
        else:
            raise CheckpointError('unsupported protobuf wire type %d' % wt)
        out.append((fno, wt, v))
    return out


def _signed64(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def _parse_entry(buf):
    """BundleEntryProto: 1 dtype, 2 shape{2: dim{1: size}}, 3 shard_id, 4 offset, 5 size, 6 crc32c (fixed32), 7 slices."""
    e = dict(dtype=0, shape=[], shard_id=0, offset=0, size=0, crc32c=None, slices=0)
    for fno, wt, v in _proto_fields(buf):
        if fno == 1:
            e['dtype'] = v
        elif fno == 2:
            for f2, _, v2 in _proto_fields(v):
                if f2 == 2:
                    size = 0
                    for f3, _, v3 in _proto_fields(v2):
                        if f3 == 1:
                            size = _signed64(v3)
                    e['shape'].append(size)
        elif fno == 3:
            e['shard_id'] = v
        elif fno == 4:
            e['offset'] = v
        elif fno == 5:
            e['size'] = v
        elif fno == 6:
            e['crc32c'] = v
        elif fno == 7:
            e['slices'] += 1
    return e


def _build_entry(dtype, shape, shard_id, offset, size, crc):
    dims = b''.join(b'\x12' + _put_varint(len(d)) + d for d in (b'\x08' + _put_varint(int(s)) for s in shape))
    out = b'\x08' + _put_varint(dtype) + b'\x12' + _put_varint(len(dims)) + dims
    if shard_id:
        out += b'\x18' + _put_varint(shard_id)
    if offset:
        out += b'\x20' + _put_varint(offset)
    out += b'\x28' + _put_varint(size) + b'\x35' + struct.pack('<I', crc)
    return out


# ------------------------------------------------------------------ the table (leveldb format)
def _read_block(buf, offset, size, verify):
    end = offset + size + BLOCK_TRAILER
    if end > len(buf):
        raise CheckpointError('block handle points outside the index file')
    body, ctype = buf[offset:offset + size], buf[offset + size]
    if verify:
        want = _unmask(struct.unpack_from('<I', buf, offset + size + 1)[0])
        if crc32c(buf[offset:offset + size + 1]) != want:
            raise CheckpointError('index block checksum mismatch')
    if ctype != 0:
        raise CheckpointError('compressed table blocks (type %d) are not supported; tf.train.Saver writes none' % ctype)
    return body


def _block_entries(block):
    if len(block) < 4:
        raise CheckpointError('malformed block')
    n_restarts = struct.unpack_from('<I', block, len(block) - 4)[0]
    limit = len(block) - 4 - 4 * n_restarts
    if limit < 0:
        raise CheckpointError('malformed block restart array')
    pos, key = 0, b''
    while pos < limit:
        shared, pos = _get_varint(block, pos)
        non_shared, pos = _get_varint(block, pos)
        vlen, pos = _get_varint(block, pos)
        if shared > len(key) or pos + non_shared + vlen > limit:
            raise CheckpointError('malformed block entry')
        key = key[:shared] + bytes(block[pos:pos + non_shared])
        pos += non_shared
        yield key, bytes(block[pos:pos + vlen])
        pos += vlen


def _read_table(path, verify=True):
    with open(path, 'rb') as f:
        buf = f.read()
    if len(buf) < FOOTER_LEN or struct.unpack_from('<Q', buf, len(buf) - 8)[0] != TABLE_MAGIC:
        raise CheckpointError('%s is not a TF checkpoint index (bad table magic)' % path)
    footer = buf[len(buf) - FOOTER_LEN:]
    _, p = _get_varint(footer, 0)          # metaindex handle (unused by the bundle)
    _, p = _get_varint(footer, p)
    ioff, p = _get_varint(footer, p)
    isize, p = _get_varint(footer, p)
    out = OrderedDict()
    for _, handle in _block_entries(_read_block(buf, ioff, isize, verify)):
        boff, q = _get_varint(handle, 0)
        bsize, q = _get_varint(handle, q)
        for k, v in _block_entries(_read_block(buf, boff, bsize, verify)):
            out[k] = v
    return out


class _BlockBuilder(object):
    def __init__(self, restart_interval):
        self.buf, self.restarts, self.count, self.last, self.ri = bytearray(), [0], 0, b'', restart_interval

    def add(self, key, value):
        shared = 0
        if self.count and self.count % self.ri == 0:
            self.restarts.append(len(self.buf))
        elif self.count:
            m = min(len(key), len(self.last))
            while shared < m and key[shared] == self.last[shared]:
                shared += 1
        self.buf += _put_varint(shared) + _put_varint(len(key) - shared) + _put_varint(len(value)) + key[shared:] + value
        self.last = key
        self.count += 1

    def finish(self):
        return bytes(self.buf) + b''.join(struct.pack('<I', r) for r in self.restarts) + struct.pack('<I', len(self.restarts))


def _write_table(path, items, block_size=4096):
    """items: [(key bytes, value bytes)] sorted by key."""
    out = bytearray()

    def emit(block):
        off = len(out)
        out.extend(block + b'\x00')
        out.extend(struct.pack('<I', _mask(crc32c(block + b'\x00'))))
        return _put_varint(off) + _put_varint(len(block))

    index = _BlockBuilder(1)
    bb = _BlockBuilder(16)
    for k, v in items:
        bb.add(k, v)
        if len(bb.buf) >= block_size:
            index.add(bb.last, emit(bb.finish()))
            bb = _BlockBuilder(16)
    if bb.count:
        index.add(bb.last, emit(bb.finish()))
    meta = emit(_BlockBuilder(1).finish())
    idx = emit(index.finish())
    footer = meta + idx
    out.extend(footer + b'\x00' * (40 - len(footer)) + struct.pack('<Q', TABLE_MAGIC))
    with open(path, 'wb') as f:
        f.write(bytes(out))


# ------------------------------------------------------------------ public API
def _index_path(prefix):
    p = prefix + '.index'
    if not os.path.exists(p):
        raise CheckpointError('no V2 checkpoint at %r (%s missing; V1 single-file checkpoints are not supported)' % (prefix, p))
    return p


def _entries(prefix, verify=True):
    table = _read_table(_index_path(prefix), verify)
    if b'' not in table:
        raise CheckpointError('bundle header missing')
    header = dict((f, v) for f, _, v in _proto_fields(table[b'']))
    if header.get(2, 0) != 0:
        raise CheckpointError('big-endian bundles are not supported')
    n_shards = header.get(1, 1)
    return n_shards, OrderedDict((k.decode('utf-8'), _parse_entry(v)) for k, v in table.items() if k != b'')


def list_variables(prefix):
    """[(name, shape tuple, numpy dtype or 'string'/'bfloat16')] like tf.train.list_variables."""
    _, ent = _entries(prefix)
    out = []
    for name, e in ent.items():
        dt = _DTYPES.get(e['dtype'], 'string' if e['dtype'] == _DT_STRING else 'bfloat16' if e['dtype'] == _DT_BFLOAT16 else None)
        out.append((name, tuple(e['shape']), dt))
    return out


def load_checkpoint(prefix, names=None, verify_data=False):
    """{variable name: ndarray} of a Saver checkpoint.  `names`: restrict to these (KeyError if one is missing);
    string tensors (tfsnippet's pickled-objects variable) and sliced entries are skipped unless asked for by name, in
    which case they raise.  `verify_data` also checks every tensor's crc32c (pure Python: ~1 MB/s)."""
    n_shards, ent = _entries(prefix)
    want = list(ent) if names is None else list(names)
    files, out = {}, OrderedDict()
    for name in want:
        if name not in ent:
            raise KeyError('variable %r is not in checkpoint %r' % (name, prefix))
        e = ent[name]
        unsupported = e['slices'] or e['dtype'] not in _DTYPES
        if unsupported:
            if names is None:
                continue
            raise CheckpointError('variable %r: dtype %d / sliced entries are not supported' % (name, e['dtype']))
        sid = e['shard_id']
        if sid not in files:
            files[sid] = np.memmap('%s.data-%05d-of-%05d' % (prefix, sid, n_shards), dtype=np.uint8, mode='r')
        raw = files[sid][e['offset']:e['offset'] + e['size']]
        dt = np.dtype(_DTYPES[e['dtype']])
        count = int(np.prod(e['shape'])) if e['shape'] else 1
        if len(raw) != e['size'] or count * dt.itemsize != e['size']:
            raise CheckpointError('variable %r: size %d does not match shape %r' % (name, e['size'], e['shape']))
        if verify_data and e['crc32c'] is not None and crc32c(raw.tobytes()) != _unmask(e['crc32c']):
            raise CheckpointError('variable %r: data checksum mismatch' % name)
        out[name] = np.frombuffer(raw.tobytes(), dtype=dt.newbyteorder('<')).astype(dt).reshape(e['shape'])
    return out


def save_checkpoint(prefix, arrays, write_state_file=True):
    """Writes {name: ndarray} as a one-shard V2 checkpoint (fixtures, export).  Returns the prefix."""
    d = os.path.dirname(prefix)
    if d:
        os.makedirs(d, exist_ok=True)
    items, offset = [], 0
    with open(prefix + '.data-00000-of-00001', 'wb') as f:
        for name in sorted(arrays, key=lambda s: s.encode('utf-8')):
            a = np.asarray(arrays[name])
            if a.dtype not in _DT_OF:
                raise CheckpointError('dtype %s cannot be written' % a.dtype)
            raw = np.ascontiguousarray(a.astype(a.dtype.newbyteorder('<'))).tobytes()
            f.write(raw)
            items.append((name.encode('utf-8'), _build_entry(_DT_OF[a.dtype], a.shape, 0, offset, len(raw), _mask(crc32c(raw)))))
            offset += len(raw)
    header = b'\x08\x01' + b'\x1a\x02\x08\x01'          # num_shards = 1, version { producer: 1 }
    _write_table(prefix + '.index', [(b'', header)] + items)
    if write_state_file:
        base = os.path.basename(prefix)
        with open(os.path.join(d, 'checkpoint'), 'w') as f:
            f.write('model_checkpoint_path: "%s"\nall_model_checkpoint_paths: "%s"\n' % (base, base))
    return prefix


def latest_checkpoint(checkpoint_dir):
    """tf.train.latest_checkpoint: the prefix named by <dir>/checkpoint, or None."""
    state = os.path.join(checkpoint_dir, 'checkpoint')
    if not os.path.isfile(state):
        return None
    with open(state) as f:
        m = re.search(r'^model_checkpoint_path:\s*"(.*)"\s*$', f.read(), re.M)
    if not m:
        return None
    p = m.group(1)
    p = p if os.path.isabs(p) else os.path.join(checkpoint_dir, p)
    return p if os.path.exists(p + '.index') else None


def model_weights_from_checkpoint(prefix, config):
    """(weights, scalars) for `Model(config, weights)` / `Model.set_scalars(**scalars)` out of a reference checkpoint:
    every variable of the path's schema (KeyError names the first missing one), and xi / beta / log_Z when the
    checkpoint holds them.  Names follow the reference's scoping rules: `p_net/xi` (tf.get_variable inside the
    @spt.global_reuse function p_net, dmm_cifar10.py:604), `beta` (a root-level tf.Variable, :860) and
    `get_log_Z/log_Z` (spt.ScheduledVariable('log_Z') = tf.get_variable inside the @spt.global_reuse function get_log_Z,
    :458-463, scaffold/scheduled_var.py:60; utils/reuse.py opens a variable scope named after the function); a bare
    `log_Z` is accepted as a fallback.  A scalar the checkpoint does not hold is left out of `scalars` with a warning —
    the caller's defaults (log_Z = 1.0, the variable's initial value) would otherwise shift adv_test_nll / pd_energy /
    kl_adv_and_gaussian silently."""
    import warnings
    from .model import weight_schema
    schema = weight_schema(config)
    w = load_checkpoint(prefix, [n for n, _ in schema])
    for name, shape in schema:
        if tuple(w[name].shape) != tuple(shape):
            raise CheckpointError('variable %r has shape %r in the checkpoint, the model expects %r'
                                  % (name, tuple(w[name].shape), tuple(shape)))
        w[name] = w[name].astype(np.float32)
    present = dict((n, s) for n, s, _ in list_variables(prefix))
    scalars = {}
    for key, names in (('xi_raw', ('p_net/xi',)), ('beta', ('beta',)), ('log_Z', ('get_log_Z/log_Z', 'log_Z'))):
        name = next((n for n in names if n in present), None)
        if name is None:
            warnings.warn('checkpoint %r holds no %s: %s keeps its initial value' % (prefix, ' / '.join(names), key))
            continue
        scalars[key] = float(np.asarray(load_checkpoint(prefix, [name])[name]).reshape(-1)[0])
    return w, scalars
