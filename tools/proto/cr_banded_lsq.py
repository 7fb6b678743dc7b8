"""Prototype (NumPy) of the block cyclic-reduction QR used by k_fit_knots for the banded least-squares solves.

Checks the index logic of the CUDA implementation against a dense solve: interval triangles (4 columns i..i+3) and
penalty rows (5 columns j..j+4) -> groups of two 4-column blocks -> log2 levels of `eliminate` -> back substitution.
"""
import numpy as np


def givens_rotate_in(R, Z, h, x):
    """Rotate row (h, x) into the upper-triangular R (n x n) with right-hand side Z (n x 3).  Returns the remainder x."""
    n = R.shape[0]
    h = h.copy(); x = x.copy()
    for j in range(n):
        if h[j] == 0.0:
            continue
        ww = R[j, j]
        dd = np.hypot(ww, h[j])
        cs, sn = ww / dd, h[j] / dd
        R[j, j] = dd
        for e in range(j + 1, n):
            a, b = h[e], R[j, e]
            R[j, e] = cs * b + sn * a
            h[e] = cs * a - sn * b
        for d in range(3):
            a, b = x[d], Z[j, d]
            Z[j, d] = cs * b + sn * a
            x[d] = cs * a - sn * b
        h[j] = 0.0
    return x


def build_group(g, tris, trhs, D, pinv, nrint, nk1):
    """Group g couples blocks g, g+1 (global columns 4g .. 4g+7)."""
    R = np.zeros((8, 8)); Z = np.zeros((8, 3))
    for t in range(4):
        i = 4 * g + t
        if i < nrint:
            for r in range(4):
                h = np.zeros(8); h[t + r:t + 4] = tris[i][r, r:]
                givens_rotate_in(R, Z, h, trhs[i][r])
        j = 4 * g + t
        if pinv > 0 and j < nrint - 1:
            h = np.zeros(8); h[t:t + 5] = D[j] * pinv
            givens_rotate_in(R, Z, h, np.zeros(3))
    for lc in range(4, 8):
        if 4 * g + lc >= nk1:
            h = np.zeros(8); h[lc] = 1.0
            givens_rotate_in(R, Z, h, np.zeros(3))
    return R, Z


def eliminate(GL, GR):
    """GL couples (c_l, c_b), GR couples (c_b, c_r).  Column order here: [c_b | c_l | c_r]."""
    (RL, ZL), (RR, ZR) = GL, GR
    P = np.zeros((4, 12)); PZ = np.zeros((4, 3))
    left = []
    rows = []
    for k in range(8):
        h = np.zeros(12); h[0:4] = RL[k, 4:8]; h[4:8] = RL[k, 0:4]
        rows.append((h, ZL[k].copy()))
    for k in range(8):
        h = np.zeros(12); h[0:4] = RR[k, 0:4]; h[8:12] = RR[k, 4:8]
        rows.append((h, ZR[k].copy()))
    for h, x in rows:                       # phase A: eliminate c_b
        h = h.copy(); x = x.copy()
        for j in range(4):
            if h[j] == 0.0:
                continue
            ww = P[j, j]
            dd = np.hypot(ww, h[j]); cs, sn = ww / dd, h[j] / dd
            P[j, j] = dd
            for e in range(j + 1, 12):
                a, b = h[e], P[j, e]
                P[j, e] = cs * b + sn * a; h[e] = cs * a - sn * b
            for d in range(3):
                a, b = x[d], PZ[j, d]
                PZ[j, d] = cs * b + sn * a; x[d] = cs * a - sn * b
            h[j] = 0.0
        left.append((h[4:12].copy(), x))
    R = np.zeros((8, 8)); Z = np.zeros((8, 3))
    for h, x in left:                       # phase B: compress onto (c_l, c_r)
        givens_rotate_in(R, Z, h, x)
    return (P, PZ), (R, Z)


def solve_cr(tris, trhs, D, pinv, nrint):
    nk1 = nrint + 3
    NB = (nk1 + 3) // 4
    L = 1
    while 4 * 2 ** L < nrint:      # every interval index below 4 * 2^L: each interval has its group
        L += 1
    assert 2 ** L + 1 >= NB
    G = [build_group(g, tris, trhs, D, pinv, nrint, nk1) for g in range(2 ** L)]
    E = {}
    for lvl in range(L):
        out = []
        for j in range(len(G) // 2):
            b = (2 * j + 1) * 2 ** lvl
            E[b], g = eliminate(G[2 * j], G[2 * j + 1])
            out.append(g)
        G = out
    (R, Z), = G
    c = np.zeros((4 * (2 ** L + 1), 3))
    top = np.zeros((8, 3))
    for j in range(7, -1, -1):
        top[j] = 0.0 if R[j, j] == 0 else (Z[j] - R[j, j + 1:] @ top[j + 1:]) / R[j, j]
    c[0:4] = top[0:4]; c[4 * 2 ** L:4 * 2 ** L + 4] = top[4:8]
    for lvl in range(L - 1, -1, -1):
        for j in range(2 ** (L - lvl - 1)):
            b = (2 * j + 1) * 2 ** lvl
            P, PZ = E[b]
            cl, cr = c[4 * (b - 2 ** lvl):4 * (b - 2 ** lvl) + 4], c[4 * (b + 2 ** lvl):4 * (b + 2 ** lvl) + 4]
            cb = np.zeros((4, 3))
            for q in range(3, -1, -1):
                acc = PZ[q] - P[q, q + 1:4] @ cb[q + 1:] - P[q, 4:8] @ cl - P[q, 8:12] @ cr
                cb[q] = 0.0 if P[q, q] == 0 else acc / P[q, q]
            c[4 * b:4 * b + 4] = cb
    return c[:nk1]


def main():
    rng = np.random.default_rng(0)
    for nrint, pinv in [(1, 0), (2, 0), (5, 0), (13, 0), (13, 0.3), (64, 0), (77, 2.0), (300, 0.01), (641, 0.5), (129, 0), (129, 0.7), (128, 0.7), (65, 0), (33, 1.0), (257, 0.2)]:
        nk1 = nrint + 3
        rows, rhs = [], []
        tris, trhs = [], []
        for i in range(nrint):
            npts = rng.integers(1, 6)
            A = np.zeros((npts, nk1)); A[:, i:i + 4] = rng.random((npts, 4))
            b = rng.normal(size=(npts, 3))
            rows.append(A); rhs.append(b)
            Q, R = np.linalg.qr(np.hstack([A[:, i:i + 4], b]), mode="complete")
            T = np.zeros((4, 4)); Zr = np.zeros((4, 3))
            k = min(4, R.shape[0])
            T[:k] = R[:k, :4]; Zr[:k] = R[:k, 4:]
            tris.append(np.triu(T)); trhs.append(Zr)
        if nrint < 4:            # make the tiny cases full rank
            A = np.eye(nk1); rows.append(A); rhs.append(rng.normal(size=(nk1, 3)))
            continue
        D = rng.normal(size=(max(nrint - 1, 0), 5))
        if pinv > 0:
            for j in range(nrint - 1):
                A = np.zeros((1, nk1)); A[0, j:j + 5] = D[j] * pinv
                rows.append(A); rhs.append(np.zeros((1, 3)))
        ref = np.linalg.lstsq(np.vstack(rows), np.vstack(rhs), rcond=None)[0]
        c = solve_cr(tris, trhs, D, pinv, nrint)
        print(nrint, pinv, "max abs diff", np.abs(c - ref).max(), "scale", np.abs(ref).max())


if __name__ == "__main__":
    main()
