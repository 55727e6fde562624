"""mu-law companding either side of the step (host side, numpy).

Mirrors utils.py:79-90 of the reference (``mu_law_numpy`` / ``inv_mu_law_numpy``): same names,
same argument meaning.  These run once per clip before/after the optimisation loop, not per step.
"""
import numpy as np


def mu_law_numpy(x, mu=255):
    """utils.py:79-82 -- sign(x) * log(1 + mu|x|) / log(1 + mu), scaled by 128 and floored."""
    x = np.asarray(x)
    out = np.sign(x) * np.log(1 + mu * np.abs(x)) / np.log(1 + mu)
    return np.floor(out * 128)


def inv_mu_law_numpy(x, mu=255.0):
    """utils.py:85-90 -- inverse companding; exact zeros stay zero."""
    x = np.array(x).astype(np.float32)
    out = (x + 0.5) * 2. / (mu + 1)
    out = np.sign(out) / mu * ((1 + mu) ** np.abs(out) - 1)
    return np.where(np.equal(x, 0), x, out)
