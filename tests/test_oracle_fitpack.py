"""oracle/fitpack_port.py against scipy's FITPACK run live (scipy is the unvendored dependency that does the math)."""
import warnings

import numpy as np
import pytest
from scipy.interpolate import splev, splprep

from oracle import fitpack_port as fp


def _curve(rng, m, noise):
    x = np.sort(rng.uniform(-0.2, 0.2, m))
    y = 0.05 + 0.3 * x * x + rng.normal(0, noise, m)
    z = 0.4 + 0.05 * (x / 0.25) ** 2 + rng.normal(0, noise, m)
    return np.stack([x, y, z], 1)


@pytest.mark.parametrize("m,noise,branch", [(30, 1e-3, "A"), (200, 1e-3, "A"), (900, 3e-3, "A"),
                                            (300, 2.5e-2, "B"), (400, 2e-2, "B")])
def test_parcur_matches_scipy(m, noise, branch):
    P = _curve(np.random.default_rng(m), m, noise)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        (tck, u), fpv, ier, _ = splprep([P[:, 0], P[:, 1], P[:, 2]], s=0.1, k=3, full_output=1)
    t, c, ierp = tck[0], np.array(tck[1]), ier
    T, C, FP, IER = fp.parcur(P, 0.1)
    assert (len(t) == 8) == (branch == "A")
    assert np.array_equal(t, T)                      # knot placement is bit-identical
    assert IER == ierp
    np.testing.assert_allclose(C, c, rtol=0, atol=1e-12)
    assert abs(FP - fpv) <= 1e-12
    np.testing.assert_allclose(fp.chord_parameter(P), u, rtol=0, atol=1e-15)
    uu = np.linspace(0, 1, 100)
    for der in (0, 1, 2):
        ref = np.array(splev(uu, tck, der=der))
        got = fp.splev(uu, T, C, der)
        np.testing.assert_allclose(got, ref, rtol=1e-11, atol=1e-11 * np.abs(ref).max())


def test_parcur_rejects_what_scipy_rejects():
    P = _curve(np.random.default_rng(1), 40, 1e-3)
    P[11] = P[10]                                   # duplicate point -> parameter not strictly increasing
    with pytest.raises(ValueError):
        splprep([P[:, 0], P[:, 1], P[:, 2]], s=0.1, k=3)
    with pytest.raises(ValueError):
        fp.parcur(P, 0.1)
    with pytest.raises(TypeError):
        fp.parcur(P[:3], 0.1)
