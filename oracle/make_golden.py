"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.npz from the fp64 numpy oracle.

    python -m oracle.make_golden [case ...]

The reference itself cannot produce fixtures here (TensorFlow 1.x is not installable; see the
header of oracle/step_np.py: PARITY UNPINNED), so these vectors pin the ORACLE against drift and let
the GPU tests check BASELINE.json sizes without trusting a single run.
"""
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from oracle import cases  # noqa: E402

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def regulariser_golden():
    """gamma term alone (oracle/stft_reg.py) on the BASELINE-size step input."""
    from oracle import stft_reg
    x = cases.build_inputs("t16384_cfg0")["x"]
    reg, g, raw = stft_reg.stft_l1(x)
    np.savez_compressed(OUT / "reg_t16384.npz", reg=np.float64(reg), raw=np.float64(raw), g=g[::16].astype(np.float64),
                        g_norm=np.float64(np.linalg.norm(g)), x_sum=np.float64(x.sum()))
    print(f"reg_t16384: reg={reg:.12g} |g|={np.linalg.norm(g):.6g}")


def main(names):
    OUT.mkdir(parents=True, exist_ok=True)
    if "reg_t16384" in names:
        regulariser_golden()
        names = [n for n in names if n != "reg_t16384"]
    for name in names:
        t0 = time.time()
        case = cases.build_inputs(name)
        Pc, Ps = cases.oracle_targets(case)
        losses, dx, aux = cases.oracle_step(case, Pc, Ps, return_aux=True)
        np.savez_compressed(OUT / f"{name}.npz", **cases.summary(case, Pc, Ps, losses, dx, aux))
        print(f"{name}: loss={losses[0]:.9g} Lc={losses[1]:.9g} Ls={losses[2]:.9g} |dx|={np.linalg.norm(dx):.6g} "
              f"({time.time() - t0:.1f}s)")


if __name__ == "__main__":
    main(sys.argv[1:] or list(cases.CASES) + ["reg_t16384"])
