"""TF V2 checkpoint ingestion (apvae/checkpoint.py; reference: tfsnippet/scaffold/checkpoint.py:74-287 saves with
tf.train.Saver).  No TensorFlow here: the reader is exercised against the module's own writer, the crc32c against
the published check values (RFC 3720 / leveldb crc32c_test "StandardResults")."""
import os
import struct

import numpy as np
import pytest

from apvae import checkpoint as ck
from apvae.model import ExpConfig, weight_schema


def test_crc32c_known_answers():
    assert ck.crc32c(b'123456789') == 0xE3069283
    assert ck.crc32c(b'\x00' * 32) == 0x8a9136aa
    assert ck.crc32c(b'\xff' * 32) == 0x62a8ab43
    assert ck.crc32c(bytes(range(32))) == 0x46dd794e
    assert ck.crc32c(b'world', ck.crc32c(b'hello ')) == ck.crc32c(b'hello world')      # extend
    for v in (0, 1, 0xdeadbeef, 0xffffffff):
        assert ck._unmask(ck._mask(v)) == v and ck._mask(v) != v


def test_round_trip_with_a_multi_block_index(tmp_path):
    rs = np.random.RandomState(0)
    arrays = {'p_net/xi': np.float32(0.25).reshape(()), 'beta': np.array(-1.5, dtype=np.float32),
              'get_log_Z/log_Z': np.array(3.0, dtype=np.float32), 'global_step': np.array(390000, dtype=np.int64),
              'some/double': rs.normal(size=(3, 2)), 'flags': np.array([True, False])}
    for i in range(300):                       # > 4096 bytes of index entries: several data blocks + a real index block
        arrays['net/level_%d/residual/conv_0/kernel' % i] = rs.normal(size=(3, 3, 2, 4)).astype(np.float32)
    prefix = ck.save_checkpoint(str(tmp_path / 'checkpoint' / 'checkpoint.dat-390000'), arrays)
    assert os.path.getsize(prefix + '.index') > 3 * 4096
    got = ck.load_checkpoint(prefix, verify_data=True)
    assert set(got) == set(arrays)
    for k, v in arrays.items():
        assert got[k].dtype == np.asarray(v).dtype and got[k].shape == np.asarray(v).shape
        np.testing.assert_array_equal(got[k], v)
    lv = dict((n, (s, d)) for n, s, d in ck.list_variables(prefix))
    assert lv['net/level_7/residual/conv_0/kernel'] == ((3, 3, 2, 4), np.float32) and lv['p_net/xi'][0] == ()
    assert ck.latest_checkpoint(str(tmp_path / 'checkpoint')) == prefix
    assert ck.latest_checkpoint(str(tmp_path)) is None
    sub = ck.load_checkpoint(prefix, ['beta', 'net/level_299/residual/conv_0/kernel'])
    assert list(sub) == ['beta', 'net/level_299/residual/conv_0/kernel']
    with pytest.raises(KeyError):
        ck.load_checkpoint(prefix, ['missing/variable'])


def test_corruption_is_detected(tmp_path):
    arrays = {'a': np.arange(12, dtype=np.float32).reshape(3, 4), 'b': np.ones(5, dtype=np.float32)}
    prefix = ck.save_checkpoint(str(tmp_path / 'm.ckpt'), arrays)
    raw = bytearray(open(prefix + '.index', 'rb').read())
    raw[10] ^= 0x40                                               # inside the first data block
    open(prefix + '.index', 'wb').write(bytes(raw))
    with pytest.raises(ck.CheckpointError):
        ck.load_checkpoint(prefix)
    prefix = ck.save_checkpoint(str(tmp_path / 'n.ckpt'), arrays)
    data = bytearray(open(prefix + '.data-00000-of-00001', 'rb').read())
    data[3] ^= 1
    open(prefix + '.data-00000-of-00001', 'wb').write(bytes(data))
    ck.load_checkpoint(prefix)                                    # data checksums are opt-in
    with pytest.raises(ck.CheckpointError):
        ck.load_checkpoint(prefix, verify_data=True)
    with pytest.raises(ck.CheckpointError):
        ck.load_checkpoint(str(tmp_path / 'absent'))
    open(str(tmp_path / 'junk.index'), 'wb').write(b'\x00' * 100)
    with pytest.raises(ck.CheckpointError):
        ck.load_checkpoint(str(tmp_path / 'junk'))


def test_string_and_sliced_entries_are_skipped_unless_requested(tmp_path):
    """tfsnippet stores its pickled objects in a DT_STRING variable next to the weights (checkpoint.py:20,149-157)."""
    prefix = str(tmp_path / 'mixed')
    a = np.arange(6, dtype=np.float32)
    open(prefix + '.data-00000-of-00001', 'wb').write(a.tobytes() + b'pickle-bytes')
    entry_a = ck._build_entry(1, a.shape, 0, 0, a.nbytes, ck._mask(ck.crc32c(a.tobytes())))
    entry_s = ck._build_entry(7, (), 0, a.nbytes, 12, 0)
    ck._write_table(prefix + '.index', [(b'', b'\x08\x01'), (b'a', entry_a), (b'tfsnippet_checkpoint_pickle_variable_x', entry_s)])
    assert list(ck.load_checkpoint(prefix)) == ['a']
    assert dict((n, d) for n, _, d in ck.list_variables(prefix))['tfsnippet_checkpoint_pickle_variable_x'] == 'string'
    with pytest.raises(ck.CheckpointError):
        ck.load_checkpoint(prefix, ['tfsnippet_checkpoint_pickle_variable_x'])


def test_model_weights_from_a_reference_style_checkpoint(tmp_path):
    """The checkpoint the reference writes holds every graph variable (q_net, G_theta, D_psi, G_omega, D_kappa,
    optimizer slots, global_step): the loader keeps the path's schema and the three scalars."""
    import oracle as O
    cfg = ExpConfig('M', z_dim=8, use_flow=True, nf_layers=2)
    ocfg = O.OracleConfig('M', z_dim=8, use_flow=True, nf_layers=2)
    params = {k: v.astype(np.float32) for k, v in O.init_weights(ocfg).items()}
    extra = {'G_omega/level_0/kernel': np.zeros((8, 4), np.float32), 'q_net/level_0/residual/conv_0/kernel/Adam': np.zeros(1, np.float32),
             'global_step': np.array(7, dtype=np.int64), 'p_net/xi': np.array(0.4, np.float32), 'beta': np.array(-0.7, np.float32)}
    prefix = ck.save_checkpoint(str(tmp_path / 'ckpt' / 'checkpoint.dat-7'), dict(params, **extra))
    with pytest.warns(UserWarning, match='get_log_Z/log_Z'):              # no log_Z in the file: said, not silently defaulted
        w, scalars = ck.model_weights_from_checkpoint(prefix, cfg)
    assert [n for n, _ in weight_schema(cfg)] == list(w)
    for k in w:
        np.testing.assert_array_equal(w[k], params[k])
    assert scalars == {'xi_raw': pytest.approx(0.4), 'beta': pytest.approx(-0.7)}
    # log_Z under the name the reference's graph gives it: spt.ScheduledVariable('log_Z') inside the @spt.global_reuse
    # function get_log_Z (dmm_cifar10.py:458-463) -> 'get_log_Z/log_Z'; a bare 'log_Z' is the fallback
    import warnings
    for name in ('get_log_Z/log_Z', 'log_Z'):
        p2 = ck.save_checkpoint(str(tmp_path / name.replace('/', '_') / 'c'), dict(params, **extra, **{name: np.array(2.5, np.float32)}))
        with warnings.catch_warnings():
            warnings.simplefilter('error')
            _, sc = ck.model_weights_from_checkpoint(p2, cfg)
        assert sc == {'xi_raw': pytest.approx(0.4), 'beta': pytest.approx(-0.7), 'log_Z': pytest.approx(2.5)}
    both = dict(params, **extra, **{'get_log_Z/log_Z': np.array(2.5, np.float32), 'log_Z': np.array(9.0, np.float32)})
    assert ck.model_weights_from_checkpoint(ck.save_checkpoint(str(tmp_path / 'both' / 'c'), both), cfg)[1]['log_Z'] == pytest.approx(2.5)
    bad = dict(params)
    bad['D_psi/level_-1/kernel'] = np.zeros((64, 2), np.float32)
    with pytest.raises(ck.CheckpointError):
        ck.model_weights_from_checkpoint(ck.save_checkpoint(str(tmp_path / 'bad' / 'c'), bad), cfg)
    del bad['D_psi/level_-1/kernel']
    with pytest.raises(KeyError):
        ck.model_weights_from_checkpoint(ck.save_checkpoint(str(tmp_path / 'bad2' / 'c'), bad), cfg)
