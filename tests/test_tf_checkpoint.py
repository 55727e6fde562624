"""TF tensor-bundle checkpoint reader (audio_style_transfer_b200/tf_checkpoint.py): the data format on the weights side
of the path (reference methods.py:79-84 restores './nsynth/model/wavenet-ckpt/model.ckpt-200000').  No TensorFlow here,
so: the format's known answers, write -> read round trips through every table feature the reader handles, corruption
is detected, and the committed fixture (tests/golden/tiny_bundle.*) still parses byte for byte."""
import struct

import numpy as np
import pytest

from audio_style_transfer_b200 import synth
from audio_style_transfer_b200 import tf_checkpoint as tfc


def test_crc32c_known_answers_and_masking():
    assert tfc.crc32c(b'') == 0
    assert tfc.crc32c(b'123456789') == 0xE3069283                       # CRC-32C check value
    assert tfc.crc32c(bytes(32)) == 0x8A9136AA                          # RFC 3720 B.4: 32 zero bytes
    assert tfc.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43                 # RFC 3720 B.4: 32 bytes of 0xFF
    assert tfc.crc32c(bytes(range(32))) == 0x46DD794E                   # RFC 3720 B.4: ascending bytes
    data = bytes(range(256)) * 5 + b'xyz'
    assert tfc.crc32c(data[100:], tfc.crc32c(data[:100])) == tfc.crc32c(data)      # incremental == one shot
    c = tfc.crc32c(b'foo')
    assert tfc.mask_crc(c) != c and tfc.unmask_crc(tfc.mask_crc(c)) == c
    assert tfc.mask_crc(tfc.mask_crc(c)) != c


def test_varint_and_proto_helpers():
    for v in (0, 1, 127, 128, 300, 2 ** 32, 2 ** 63 - 1):
        buf = tfc._put_varint(v)
        assert tfc._get_varint(buf, 0) == (v, len(buf))
    e = tfc.parse_entry(tfc._entry_proto(1, (1, 3, 128, 128), 0, 4096, 196608, 0xDEADBEEF))
    assert (e.dtype, e.shape, e.shard_id, e.offset, e.size, e.crc32c) == (1, (1, 3, 128, 128), 0, 4096, 196608, 0xDEADBEEF)
    with pytest.raises(tfc.CheckpointError):
        tfc._get_varint(b'\x80\x80', 0)


def test_encoder_round_trip_through_the_bundle_format(tmp_path):
    w = synth.make_weights(30, seed=3, bias_scale=0.1)
    prefix = tmp_path / 'model.ckpt-200000'
    tfc.save_encoder_weights(prefix, w)
    assert tfc.is_bundle(prefix)
    rd = tfc.BundleReader(prefix)
    assert rd.keys() == sorted(tfc.encoder_variable_names(30)) and len(rd.keys()) == 122
    assert rd.shape('ae_dilatedconv_7/W') == (1, 3, 128, 128) and rd.shape('ae_res_30/W') == (1, 1, 128, 128)
    w2 = tfc.load_encoder_weights(prefix)
    assert w2.digest() == w.digest()
    w10 = tfc.load_encoder_weights(prefix, n_layers=10)                 # TF prunes unused layers; so does the loader
    assert w10.n_layers == 10 and np.array_equal(w10.Wd, w.Wd[:10])
    with pytest.raises(KeyError):
        tfc.load_encoder_weights(prefix, ema=True)                      # no EMA shadows in this file


def test_ema_shadows_and_other_dtypes(tmp_path):
    w = synth.make_weights(2, seed=1)
    t = {}
    tfc.save_encoder_weights(tmp_path / 'a', w, ema=True)
    assert tfc.load_encoder_weights(tmp_path / 'a', n_layers=2, ema=True).digest() == w.digest()
    t = {'global_step': np.array(200000, np.int64), 'v/f64': np.linspace(0, 1, 7), 'v/i32': np.arange(6, dtype=np.int32).reshape(2, 3)}
    tfc.write_bundle(tmp_path / 'b', t)
    rd = tfc.BundleReader(tmp_path / 'b')
    assert rd['global_step'].shape == () and int(rd['global_step']) == 200000
    assert np.array_equal(rd['v/f64'], t['v/f64']) and np.array_equal(rd['v/i32'], t['v/i32'])


@pytest.mark.parametrize("block_size", [64, 4096, 1 << 20])
def test_table_blocks_restarts_and_prefix_compression(tmp_path, block_size):
    """Many keys sharing long prefixes: one block or dozens, restart points every 16 entries."""
    rng = np.random.default_rng(0)
    t = {f'scope/sub_{i // 7}/var_{i:04d}/W': rng.standard_normal((i % 5 + 1, 3)).astype(np.float32) for i in range(200)}
    tfc.write_bundle(tmp_path / 'm', t, block_size=block_size)
    table = tfc.read_table(str(tmp_path / 'm') + '.index')
    assert list(table)[0] == b'' and list(table)[1:] == sorted(k.encode() for k in t)
    rd = tfc.BundleReader(tmp_path / 'm')
    for k, v in t.items():
        assert np.array_equal(rd[k], v)


def test_corruption_is_detected(tmp_path):
    w = synth.make_weights(1, seed=0)
    prefix = str(tmp_path / 'c')
    tfc.save_encoder_weights(prefix, w)
    data = bytearray(open(prefix + '.data-00000-of-00001', 'rb').read())
    data[100] ^= 0x01
    open(prefix + '.data-00000-of-00001', 'wb').write(bytes(data))
    rd = tfc.BundleReader(prefix)
    with pytest.raises(tfc.CheckpointError, match='checksum'):
        rd['ae_dilatedconv_1/W']
    assert tfc.BundleReader(prefix, verify=False)['ae_dilatedconv_1/W'].shape == (1, 3, 128, 128)
    idx = bytearray(open(prefix + '.index', 'rb').read())
    idx[10] ^= 0x40
    open(prefix + '.index', 'wb').write(bytes(idx))
    with pytest.raises(tfc.CheckpointError):
        tfc.BundleReader(prefix)
    open(prefix + '.index', 'wb').write(bytes(idx[:-8]) + struct.pack('<Q', 1))
    with pytest.raises(tfc.CheckpointError, match='magic'):
        tfc.BundleReader(prefix)


def test_snappy_blocks_are_accepted():
    raw = b'abcdabcdabcdabcd' + b'x' * 70
    comp = tfc._put_varint(len(raw)) + bytes([3 << 2]) + b'abcd' + bytes([((12 - 1) << 2) | 2, 4, 0]) \
        + bytes([60 << 2, 69]) + b'x' * 70
    assert tfc._snappy_decompress(comp) == raw
    comp1 = tfc._put_varint(12) + bytes([3 << 2]) + b'abcd' + bytes([((8 - 4) << 2) | 1, 4])      # 1-byte-offset copy
    assert tfc._snappy_decompress(comp1) == b'abcd' * 3


def test_committed_fixture_parses(golden_dir):
    """A bundle written once and committed: pins the reader (and the writer's byte layout) against drift."""
    rd = tfc.BundleReader(golden_dir / 'tiny_bundle')
    assert rd.keys() == ['ae_startconv/W', 'ae_startconv/biases', 'global_step']
    w = synth.make_weights(1, seed=0)
    assert np.array_equal(rd['ae_startconv/W'][0], w.W0) and np.array_equal(rd['ae_startconv/biases'], w.b0)
    assert int(rd['global_step']) == 200000
    idx = open(str(golden_dir / 'tiny_bundle') + '.index', 'rb').read()
    assert idx[-8:] == bytes.fromhex('57fb808b247547db')                 # table magic, little endian


def test_gatysnet_load_model_takes_a_tf_checkpoint_prefix(tmp_path):
    """methods.py:79-84 through the host mirror, without a GPU: a recording stand-in for the step object."""
    from audio_style_transfer_b200 import methods
    w = synth.make_weights(10, seed=5)
    tfc.save_encoder_weights(tmp_path / 'model.ckpt-200000', w)

    class Rec:
        n_layers = 10

        def set_weights(self, ww):
            self.w = ww

    net = methods.GatysNet.__new__(methods.GatysNet)
    net.checkpoint_path, net.seed, net.graph = str(tmp_path / 'model.ckpt-200000'), 0, Rec()
    net.load_model(None)
    assert net.graph.w.digest() == w.digest()
