"""Failure must be loud: a hand-off wait of the whole-stack kernels that expires invalidates the evaluation, and every
synchronising entry point has to say so (VERDICT r1 item 7 / ADVICE: ast_step_host used to return AST_OK).

Uses the -DAST_EXPERIMENTS build (libast_b200_exp.so) for its fault-injection hook: one tile stops releasing its
hand-off flags, its neighbours' bounded waits expire, the kernel terminates and raises the sticky error word."""
import numpy as np
import pytest

from audio_style_transfer_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def test_default_build_exports_no_debug_hooks():
    lib = _lib.load()
    for name in ("ast_debug_timeline", "ast_debug_skip_publish", "ast_debug_tl_layer", "ast_opt_debug_stamps", "ast_debug_skipw", "ast_debug_group_tiles"):
        assert not hasattr(lib, name), name


def test_expired_handoff_is_reported_by_step_host_and_check():
    from audio_style_transfer_b200.step import StyleStep
    T, NL = 2048, 12
    lib = _lib.load(experiments=True)
    st = StyleStep(T, [NL - 1], list(range(NL)), True, precision="bf16", experiments=True)
    st.set_weights(synth.make_weights(NL, 0, 0.1))
    rng = np.random.default_rng(0)
    x = rng.normal(0, 40, (1, T))
    Ec, Gn = st.embeds(x)
    st.set_targets(Ec * 0.9, Gn.roll(1, 1), 100.0)
    assert st.uses_stack_kernels
    lo, g = st.step_host(x)                       # healthy
    st.check()
    try:
        lib.ast_debug_skip_publish(5 + 1)         # tile 5 goes mute from publish 3 on
        with pytest.raises(_lib.AstError, match="timed out"):
            st.step_host(x)
        st.step(x)                                # asynchronous entry point: returns OK ...
        with pytest.raises(_lib.AstError, match="timed out"):
            st.check()                            # ... the sticky error is reported here
    finally:
        lib.ast_debug_skip_publish(0)
    st.check()                                    # cleared by the failing check
    lo2, g2 = st.step_host(x)                     # and the context is usable again
    assert np.array_equal(g, g2) and np.array_equal(lo, lo2)
    st.close()
