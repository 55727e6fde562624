"""GatysNet on the GPU: the L-BFGS-B loop of methods.py:140-181 driven through ast_step_host."""
import numpy as np
import pytest

from audio_style_transfer_b200 import synth
from audio_style_transfer_b200.mu_law import mu_law_numpy
from oracle import step_np

pytestmark = pytest.mark.gpu


def test_first_evaluation_matches_oracle_and_loss_decreases(tmp_path):
    from audio_style_transfer_b200.methods import GatysNet, l2_normalize
    T = 4096
    net = GatysNet(savepath=str(tmp_path), checkpoint_path=None, batch_size=T, stack=0, cont_lyr_ids=[9], precision='fp32', seed=0)
    net.load_model(None)
    content = synth.content_audio(T + 2 * 16000, seed=1)
    style = synth.style_audio(T, seed=2, chunks=2)
    w = synth.make_weights(10, seed=0)
    # targets the way GatysNet.run builds them (methods.py:192-212)
    st = int(1.0 * 16000 - net.late)
    aud = content[st:st + T]
    phi_c = net.get_embeds(None, aud)
    phi = l2_normalize(net.get_embeds(None, aud, is_content=False) + net.get_style_phi(None, style) - net.get_style_phi(None, content))
    # the oracle's targets from the same audio
    c_mu = mu_law_numpy(aud).astype(np.float64)
    Pc, Gc = step_np.embeds(c_mu, w, [9], list(range(10)), False)
    assert np.linalg.norm(phi_c - Pc) <= 1e-4 * np.linalg.norm(Pc)
    fun, last = net.define_loss('loss', phi, phi_c, 100.0, 0.0)
    x0 = net.x[0].copy()
    f0, g0 = fun(x0)
    lo, dx = step_np.step(x0, w, phi_c.astype(np.float64), phi.astype(np.float64), [9], list(range(10)), False, 100.0)
    assert abs(f0 - lo[0]) <= 1e-4 * abs(lo[0])
    assert np.linalg.norm(g0 - dx) <= 1e-3 * np.linalg.norm(dx)
    seen = []
    net.l_bfgs(None, phi_c, phi, epochs=1, lambd=100.0, gamma=0.0, callback=lambda ep, i, l: seen.append(l[0]))
    assert len(seen) >= 20 and seen[-1] < 0.5 * seen[0]
    assert (tmp_path / 'ep-0.wav').exists()


def test_cli_synthetic_run(tmp_path):
    from audio_style_transfer_b200 import methods
    out = methods.piece_work(methods.build_parser().parse_args(
        ['c', 's', '--synthetic', '--epochs', '1', '--batch_size', '4096', '--stack', '0', '--cont_lyrs', '9',
         '--outdir', str(tmp_path / 'out'), '--logdir', str(tmp_path / 'log')]))
    assert out.shape == (4096,) and np.all(np.isfinite(out))


def test_batched_target_precompute_equals_one_by_one_bit_for_bit(tmp_path):
    """the <= 1 + 5 + 5 forwards before the loop (methods.py:97-111,192-212) as ONE B > 1 ast_embeds call."""
    from audio_style_transfer_b200.methods import GatysNet
    T = 4096
    net = GatysNet(savepath=None, checkpoint_path=None, batch_size=T, stack=0, cont_lyr_ids=[9], seed=0)
    net.load_model(None)
    aud = synth.content_audio(T, seed=1)
    target = synth.style_audio(T, seed=2, chunks=5)
    source = synth.content_audio(3 * T + 100, seed=3)                 # 3 full chunks + a tail that is dropped
    phi_c, phi, phi_t, phi_s = net.precompute_targets(aud, source, target)
    one = lambda a, c: net.get_embeds(None, a, is_content=c)
    assert np.array_equal(phi_c, one(aud, True))
    assert np.array_equal(phi, one(aud, False))
    assert np.array_equal(phi_t, np.mean([one(target[i * T:(i + 1) * T], False) for i in range(5)], axis=0))
    assert np.array_equal(phi_s, np.mean([one(source[i * T:(i + 1) * T], False) for i in range(3)], axis=0))
    assert np.array_equal(net.get_style_phi(None, target), phi_t)


def test_missing_checkpoint_raises_like_saver_restore(tmp_path):
    """methods.py:79-84: Saver.restore raises on a missing checkpoint; random init needs an explicit opt-in."""
    from audio_style_transfer_b200.methods import GatysNet
    net = GatysNet(savepath=None, checkpoint_path=str(tmp_path / 'model.ckpt-200000'), batch_size=4096, stack=0, cont_lyr_ids=[9])
    with pytest.raises(FileNotFoundError):
        net.load_model(None)
    net.random_init = True
    net.load_model(None)
