"""ORACLE (test infrastructure, not product code): restatement of the FITPACK routines behind `splprep`/`splev`.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs may import
this module.

Where the arithmetic lives: the reference calls `scipy.interpolate.splprep([x,y,z], s=0.1, k=3)` and
`splev(u, tck, der=0|1|2)` (`pkg/geometry_utils.py:78,84,148-149`).  SciPy is a third-party dependency that is
*not* vendored under `/root/reference` and is *not pinned* by it (absent from `requirements.txt:1-21`); the
version in this image is scipy 1.18.1 (`_fitpack` is FITPACK translated to C).  This file restates the published
algorithm (P. Dierckx, "Curve and Surface Fitting with Splines", 1993; routines `parcur`, `fppara`, `fpbspl`,
`fpgivs`, `fprota`, `fpback`, `fpknot`, `fpdisc`, `fprati`, `splev`, `splder`) for the call the reference
makes: idim=3, k=3, unit weights, task=0, chord-length parameter, nest=m+2k.

* Part 1: least-squares spline on a growing knot set, rows rotated into a banded triangle by Givens
  rotations.  With 8 knots this is one cubic Bezier (`ier=-2`), which is what plausible frames hit.
* Part 2: smoothing-parameter root find once the knot set is rich enough (`fp < s`).

Parity pin: `tests/test_oracle_fitpack.py` checks knots/coefficients/fp/ier against `scipy.interpolate.splprep`
run live (scipy is in the image on both machines) and against vectors the unmodified reference produced
(`tests/golden/`, made by `oracle/make_golden.py`).  Pure-Python loops: meant for <= a few hundred knots.
"""
from __future__ import annotations

import math

import numpy as np

K = 3
K1 = K + 1
K2 = K + 2
NMIN = 2 * K1
TOL = 1e-3
MAXIT = 20
CON1, CON9, CON4 = 0.1, 0.9, 0.04


def chord_parameter(points: np.ndarray) -> np.ndarray:
    """parcur with ipar=0: cumulative chord length, sequential summation, normalised to [0,1], u[m-1]=1."""
    m = points.shape[0]
    u = np.zeros(m)
    for i in range(1, m):
        d = points[i] - points[i - 1]
        u[i] = u[i - 1] + math.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
    if u[m - 1] <= 0.0:
        raise ValueError("Error on input data")
    u[1:] = u[1:] / u[m - 1]
    u[m - 1] = 1.0
    return u


def _bspl(t, x, l):
    """fpbspl: the K1 non-zero cubic B-splines at x for t[l] <= x < t[l+1] (0-based l)."""
    h = [1.0, 0.0, 0.0, 0.0]
    hh = [0.0, 0.0, 0.0]
    for j in range(1, K + 1):
        for i in range(j):
            hh[i] = h[i]
        h[0] = 0.0
        for i in range(j):
            li = l + 1 + i
            lj = li - j
            f = hh[i] / (t[li] - t[lj])
            h[i] = h[i] + f * (t[li] - x)
            h[i + 1] = f * (x - t[lj])
    return h


def _givens(piv, ww):
    store = abs(piv)
    if store >= ww:
        dd = store * math.sqrt(1.0 + (ww / piv) ** 2)
    else:
        dd = ww * math.sqrt(1.0 + (piv / ww) ** 2)
    return ww / dd, piv / dd, dd


def _back(a, z, n, bw):
    """fpback: solve the banded upper-triangular system a c = z; a[i, 0] is the diagonal."""
    c = np.zeros_like(z)
    c[n - 1] = z[n - 1] / a[n - 1, 0]
    for i in range(n - 2, -1, -1):
        i1 = min(bw - 1, n - 1 - i)
        store = z[i].copy()
        for l in range(1, i1 + 1):
            store = store - c[i + l] * a[i, l]
        c[i] = store / a[i, 0]
    return c


def _lsq(t, n, u, X):
    """Rows of the collocation matrix rotated into a banded triangle.  Returns a, z, q, fp, c."""
    m = u.shape[0]
    nk1 = n - K1
    a = np.zeros((nk1, K1))
    z = np.zeros((nk1, 3))
    q = np.zeros((m, K1))
    fp = 0.0
    l = K                       # 0-based interval index: t[l] <= u < t[l+1]
    for it in range(m):
        ui = u[it]
        xi = X[it].copy()
        while ui >= t[l + 1] and l != nk1 - 1:
            l += 1
        h = _bspl(t, ui, l)
        q[it] = h
        j = l - K1
        for i in range(K1):
            j += 1
            piv = h[i]
            if piv == 0.0:
                continue
            cs, sn, a[j, 0] = _givens(piv, a[j, 0])
            zj = z[j].copy()
            z[j] = cs * zj + sn * xi
            xi = cs * xi - sn * zj
            if i == K1 - 1:
                break
            i2 = 0
            for i1 in range(i + 1, K1):
                i2 += 1
                s1, s2 = h[i1], a[j, i2]
                a[j, i2] = cs * s2 + sn * s1
                h[i1] = cs * s1 - sn * s2
        fp += float(xi[0] * xi[0] + xi[1] * xi[1] + xi[2] * xi[2])
    c = _back(a, z, nk1, K1)
    return a, z, q, fp, c


def _interval_residuals(t, n, u, X, q, c, nrint):
    """Squared residuals per knot interval; a point on a knot is split half/half between its neighbours."""
    nk1 = n - K1
    fpint = np.zeros(nrint)
    fpart = 0.0
    i = 0
    l = K2 - 1                  # 0-based index of the next interior knot
    for it in range(u.shape[0]):
        new = False
        if u[it] >= t[l] and l <= nk1 - 1:
            new = True
            l += 1
        l0 = l - K2 + 1
        fac = q[it] @ c[l0:l0 + K1]
        d = fac - X[it]
        term = float(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
        fpart += term
        if new:
            store = term * 0.5
            fpint[i] = fpart - store
            i += 1
            fpart = store
    fpint[nrint - 1] = fpart
    return fpint


def _add_knot(u, t, n, fpint, nrdata, nrint):
    """fpknot with istart=1: split the interval with the largest residual that still holds data points."""
    fpmax = 0.0
    jbegin = 1
    number = -1
    for j in range(nrint):
        jpoint = nrdata[j]
        if fpmax < fpint[j] and jpoint != 0:
            fpmax, number, maxpt, maxbeg = fpint[j], j, jpoint, jbegin
        jbegin += jpoint + 1
    if number < 0:
        raise RuntimeError("fpknot: no interval left to split")
    ihalf = maxpt // 2 + 1
    nrx = maxbeg + ihalf        # 1-based index into u
    nxt = number + 1
    fpint = list(fpint)
    fpint.insert(nxt, 0.0)
    nrdata.insert(nxt, 0)
    t.insert(nxt + K, u[nrx - 1])   # FITPACK: t(next+k) (1-based) becomes the new knot, the rest shifts up
    nrdata[number] = ihalf - 1
    nrdata[nxt] = maxpt - ihalf
    fpint[number] = fpmax * nrdata[number] / maxpt
    fpint[nxt] = fpmax * nrdata[nxt] / maxpt
    return fpint


def _disc(t, n):
    """fpdisc: jumps of the K-th derivative of the B-splines at the interior knots, FITPACK scaling."""
    nk1 = n - K1
    nrint = nk1 - K
    fac = nrint / (t[nk1] - t[K1 - 1])
    b = np.zeros((nrint - 1, K2))
    h = [0.0] * (2 * K1)
    for l in range(K2, nk1 + 1):            # 1-based l
        lmk = l - K1
        for j in range(1, K1 + 1):
            ik = j + K1
            lj = l + j
            lk = lj - K2
            h[j - 1] = t[l - 1] - t[lk - 1]
            h[ik - 1] = t[l - 1] - t[lj - 1]
        lp = lmk
        for j in range(1, K2 + 1):
            jk = j
            prod = h[j - 1]
            for _ in range(K):
                jk += 1
                prod = prod * h[jk - 1] * fac
            lk = lp + K1
            b[lmk - 1, j - 1] = (t[lk - 1] - t[lp - 1]) / prod
            lp += 1
    return b


def _rati(p1, f1, p2, f2, p3, f3):
    if p3 > 0.0:
        h1 = f1 * (f2 - f3)
        h2 = f2 * (f3 - f1)
        h3 = f3 * (f1 - f2)
        p = -(p1 * p2 * h3 + p2 * p3 * h1 + p3 * p1 * h2) / (p1 * h1 + p2 * h2 + p3 * h3)
    else:
        p = (p1 * (f1 - f3) * f2 - p2 * (f2 - f3) * f1) / ((f1 - f2) * f3)
    if f2 < 0.0:
        p3, f3 = p2, f2
    else:
        p1, f1 = p2, f2
    return p, p1, f1, p3, f3


def _smooth(t, n, u, X, a, z, q, b, p):
    """Coefficients and residual of the spline minimising (LSQ residual) + (1/p) * (derivative jumps)."""
    nk1 = n - K1
    n8 = n - NMIN
    pinv = 1.0 / p
    c = z.copy()
    g = np.zeros((nk1, K2))
    g[:, :K1] = a
    for it in range(n8):
        h = list(b[it] * pinv)
        xi = np.zeros(3)
        for j in range(it, nk1):
            piv = h[0]
            cs, sn, g[j, 0] = _givens(piv, g[j, 0])
            cj = c[j].copy()
            c[j] = cs * cj + sn * xi
            xi = cs * xi - sn * cj
            if j == nk1 - 1:
                break
            i2 = K1
            if j + 1 > n8:
                i2 = nk1 - (j + 1)
            for i in range(i2):
                s1, s2 = h[i + 1], g[j, i + 1]
                g[j, i + 1] = cs * s2 + sn * s1
                h[i] = cs * s1 - sn * s2
            h[i2] = 0.0
    c = _back(g, c, nk1, K2)
    fp = 0.0
    l = K2 - 1
    for it in range(u.shape[0]):
        if u[it] >= t[l] and l <= nk1 - 1:
            l += 1
        l0 = l - K2 + 1
        d = q[it] @ c[l0:l0 + K1] - X[it]
        fp += float(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
    return c, fp


def parcur(points, s=0.1, max_knots=None):
    """`splprep([x,y,z], s=s, k=3)` for ordered (m,3) points.  Returns (t, c (3, n-4), fp, ier).

    Raises ValueError / TypeError where scipy would (`m <= k`, parameter not strictly increasing).
    `max_knots` lets a caller stop early in the many-knot regime (returns ier=99 then).
    """
    X = np.ascontiguousarray(points, dtype=np.float64)
    m = X.shape[0]
    if m <= K:
        raise TypeError("m > k must hold")
    u = chord_parameter(X)
    if np.any(u[1:] <= u[:-1]):
        raise ValueError("Error on input data")
    nest = m + 2 * K
    nmax = m + K1
    acc = TOL * s
    n = NMIN
    t = [0.0] * K1 + [1.0] * K1
    nrdata = [m - 2]
    fpold, nplus, fp0, ier = 0.0, 0, 0.0, 0
    for _ in range(m):
        if n == NMIN:
            ier = -2
        nrint = n - NMIN + 1
        ta = np.array(t)
        a, z, q, fp, c = _lsq(ta, n, u, X)
        if ier == -2:
            fp0 = fp
        fpms = fp - s
        if abs(fpms) < acc:
            return ta, c.T.copy(), fp, (ier if ier == -2 else 0)
        if fpms < 0.0:
            break
        if n == nmax:
            return ta, c.T.copy(), fp, -1
        if n == nest:
            return ta, c.T.copy(), fp, 1
        if max_knots is not None and n >= max_knots:
            return ta, c.T.copy(), fp, 99
        if ier == 0:
            npl1 = nplus * 2
            if fpold - fp > acc:
                npl1 = int(nplus * fpms / (fpold - fp))
            nplus = min(nplus * 2, max(npl1, nplus // 2, 1))
        else:
            nplus, ier = 1, 0
        fpold = fp
        fpint = _interval_residuals(ta, n, u, X, q, c, nrint)
        for _l in range(nplus):
            fpint = _add_knot(u, t, n, fpint, nrdata, nrint)
            n += 1
            nrint += 1
            if n == nmax:
                # interpolation knots: t[K1 + i] = u[2 + i]
                t = [0.0] * K1 + [float(u[2 + i]) for i in range(m - K1)] + [1.0] * K1
                break
            if n == nest:
                break
    if ier == -2:
        return ta, c.T.copy(), fp, -2
    # Part 2: smoothing spline on the final knot set
    nk1 = n - K1
    b = _disc(ta, n)
    p1, f1, p3, f3 = 0.0, fp0 - s, -1.0, fpms
    p = nk1 / float(a[:, 0].sum())
    ich1 = ich3 = False
    ier = 0
    for it in range(1, MAXIT + 1):
        c, fp = _smooth(ta, n, u, X, a, z, q, b, p)
        fpms = fp - s
        if abs(fpms) < acc:
            break
        if it == MAXIT:
            ier = 3
            break
        p2, f2 = p, fpms
        if not ich3:
            if (f2 - f3) <= acc:
                p3, f3 = p2, f2
                p = p * CON4
                if p <= p1:
                    p = p1 * CON9 + p2 * CON1
                continue
            if f2 < 0.0:
                ich3 = True
        if not ich1:
            if (f1 - f2) <= acc:
                p1, f1 = p2, f2
                p = p / CON4
                if p3 < 0.0:
                    continue
                if p >= p3:
                    p = p2 * CON1 + p3 * CON9
                continue
            if f2 > 0.0:
                ich1 = True
        if f2 >= f1 or f2 <= f3:
            ier = 2
            break
        p, p1, f1, p3, f3 = _rati(p1, f1, p2, f2, p3, f3)
    return ta, c.T.copy(), fp, ier


def splev(u, t, c, der=0):
    """Evaluate the cubic spline curve (or its `der`-th derivative, der <= 2) at parameters u -> (3, len(u))."""
    t = np.asarray(t, dtype=np.float64)
    c = np.asarray(c, dtype=np.float64)
    n = t.shape[0]
    k = K
    coef = c[:, : n - K1].copy()
    tt = t
    lo = 0
    for _ in range(der):            # splder: difference the coefficients, degree drops by one
        nc = coef.shape[1]
        new = np.zeros((coef.shape[0], nc - 1))
        for i in range(nc - 1):
            den = tt[lo + i + k + 1] - tt[lo + i + 1]
            new[:, i] = k * (coef[:, i + 1] - coef[:, i]) / den if den > 0 else 0.0
        coef = new
        lo += 1
        k -= 1
    tk = tt[lo: n - lo]              # knot vector of the derivative spline (degree k)
    nk = tk.shape[0]
    out = np.zeros((coef.shape[0], len(u)))
    l = k
    for idx, x in enumerate(np.asarray(u, dtype=np.float64)):
        x = min(max(x, tk[k]), tk[nk - k - 1])
        l = k
        while x >= tk[l + 1] and l != nk - k - 2:
            l += 1
        # de Boor basis of degree k
        h = [1.0] + [0.0] * k
        for j in range(1, k + 1):
            hh = h[:j]
            h[0] = 0.0
            for i in range(j):
                li = l + 1 + i
                lj = li - j
                f = hh[i] / (tk[li] - tk[lj])
                h[i] = h[i] + f * (tk[li] - x)
                h[i + 1] = f * (x - tk[lj])
        out[:, idx] = coef[:, l - k: l + 1] @ np.array(h)
    return out
