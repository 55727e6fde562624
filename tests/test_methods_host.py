"""Host-side mirror of the reference's methods.py: CLI surface, run-directory naming, constants.
CPU only (no compute); the GPU behaviour of GatysNet is in test_gpu_methods.py."""
import numpy as np
import pytest
import torch

from audio_style_transfer_b200 import methods, mu_law


def test_cli_defaults_match_reference():
    """methods.py:246-267 of the reference, flag for flag."""
    a = methods.build_parser().parse_args(['pachelbel', 'organ'])
    assert (a.epochs, a.batch_size, a.sr, a.stack, a.cont_lyrs, a.style_lyrs) == (100, 16384, 16000, None, [29], None)
    assert (a.lambd, a.gamma, a.channels, a.cnt_channels, a.start, a.gatys) == (100.0, 0.0, 128, 128, 1.0, False)
    assert a.ckpt_path == './nsynth/model/wavenet-ckpt/model.ckpt-200000'
    assert (a.dir, a.outdir, a.logdir, a.cmt) == ('./data/src', './data/out', './log', None)


def test_cli_quirks_are_preserved():
    p = methods.build_parser()
    assert p.parse_args(['a', 'b', '--gatys']).gatys is True
    assert p.parse_args(['a', 'b', '--gatys', 'False']).gatys is True      # type=bool quirk (methods.py:261)
    a = p.parse_args(['a', 'b', '--cont_lyrs', '25', '--stack', '0', '--lambd', '100', '--gamma', '0', '--epochs', '100'])
    assert a.cont_lyrs == [25] and a.stack == 0


def test_run_directory_naming(tmp_path):
    """utils.gt_s_path (utils.py:18-64): sorted kwargs, abbreviations, list formatting, gatys prefix."""
    a = methods.build_parser().parse_args(['c', 's', '--stack', '0', '--cont_lyrs', '25', '--outdir', str(tmp_path)])
    d = methods.get_dir(str(tmp_path), a)
    name = d.split('/')[-1]
    assert name.startswith('ours_')
    for part in ('_btch_16384_', '_cnt_c_', '_style_s_', '_cont_lyrs_-25_', '_stk_0_', '_lbd_100.0_', '_sr16.0kHz_', '_ep_100_'):
        assert part in name, (part, name)
    assert 'outdir' not in name and 'ckpt' not in name


def test_late_and_variable_init():
    assert (16384 - (16384 // 4096) * 4000) // 2 == 192                   # methods.py:39


def test_mu_law_round_trip_and_floor():
    x = np.linspace(-0.99, 0.99, 1001).astype(np.float32)
    q = mu_law.mu_law_numpy(x)
    assert q.min() >= -128 and q.max() <= 127 and np.all(q == np.floor(q))
    y = mu_law.inv_mu_law_numpy(q)
    assert np.max(np.abs(y - x)) < 0.05
    assert mu_law.inv_mu_law_numpy(np.zeros(3))[0] == 0


def test_l2_normalize_matches_oracle():
    from oracle import step_np
    G = np.random.default_rng(0).standard_normal((5, 7, 7))
    assert np.allclose(methods.l2_normalize(G), step_np.l2_normalize(G))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_gatysnet_needs_a_gpu():
    with pytest.raises(RuntimeError):
        methods.GatysNet(savepath=None, batch_size=4096, stack=0, cont_lyr_ids=[9])


def test_length_must_be_divisible_by_dilations():
    with pytest.raises(AssertionError):                                   # masked.py:134
        methods.GatysNet.build(1000, [9], 0, 128)
