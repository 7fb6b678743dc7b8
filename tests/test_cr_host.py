"""The cyclic-reduction least-squares solver of k_fit_knots (`cr_solve`, csrc/geometry.cu), compiled as host code.

The device function's text is taken verbatim from the .cu file (plus `givens_fast` and `rota`), wrapped with one-thread
stand-ins for threadIdx / __syncthreads, built with g++ and compared with a dense least-squares solve: random banded
problems at the sizes where the padding logic changes, and the real collocation problems of the golden frames."""
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import fitpack_port as fp
from oracle import geometry_port as gp
from rdp_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
CU = os.path.join(os.path.dirname(HERE), "robotic-discovery-platform_b200", "csrc", "geometry.cu")


def _between(s, a, b):
    i = s.index(a)
    return s[i:s.index(b, i)]


@pytest.fixture(scope="module")
def cr_host(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    src = open(CU).read()
    givens = _between(src, "__device__ __forceinline__ void givens_fast(", "__device__ __forceinline__ void prefetch_l1")
    rota = _between(src, "__device__ __forceinline__ void rota(", "// cubic B-spline basis at x")
    body = _between(src, "struct CrWs {", "__device__ __forceinline__ double block_sum1")
    d = tmp_path_factory.mktemp("cr_host")
    pre = open(os.path.join(HERE, "host_harness", "cr_host_pre.inc")).read()
    main = open(os.path.join(HERE, "host_harness", "cr_host_main.inc")).read()
    cpp = d / "cr_host.cpp"
    cpp.write_text(pre + givens + rota + body + main)
    exe = d / "cr_host"
    subprocess.run(["g++", "-O1", "-std=c++17", "-o", str(exe), str(cpp)], check=True)

    def run(packed, D, pinv):
        nrint = len(packed)
        text = f"{nrint} {pinv!r}\n" + "\n".join(" ".join(repr(float(v)) for v in p) for p in packed) + "\n" + \
            "\n".join(" ".join(repr(float(v)) for v in row) for row in D) + "\n"
        out = subprocess.run([str(exe)], input=text, capture_output=True, text=True, check=True).stdout
        return np.array([float(x) for x in out.split()]).reshape(nrint + 3, 3)

    return run


def _triangle(A4, b):
    """Rows over one interval's four columns -> packed 4x4 upper triangle (10) + right-hand side (12)."""
    R = np.linalg.qr(np.hstack([A4, b]), mode="complete")[1]
    T, Z = np.zeros((4, 4)), np.zeros((4, 3))
    k = min(4, R.shape[0])
    T[:k], Z[:k] = R[:k, :4], R[:k, 4:]
    T = np.triu(T)
    return np.concatenate([np.concatenate([T[r, r:] for r in range(4)]), Z.reshape(-1)])


def _dense(A, X, D, pinv):
    if pinv == 0:
        return np.linalg.lstsq(A, X, rcond=None)[0]
    n = A.shape[1]
    P = np.zeros((D.shape[0], n))
    for j in range(D.shape[0]):
        P[j, j:j + 5] = D[j] * pinv
    return np.linalg.lstsq(np.vstack([A, P]), np.vstack([X, np.zeros((D.shape[0], 3))]), rcond=None)[0]


@pytest.mark.parametrize("nrint,pinv", [(48, 0.0), (64, 0.0), (65, 0.0), (77, 2.0), (128, 0.7), (129, 0.0), (129, 0.7),
                                        (257, 0.2), (300, 0.01), (641, 0.5)])
def test_random_banded_problems(cr_host, nrint, pinv):
    rng = np.random.default_rng(nrint)
    nk1 = nrint + 3
    A_rows, X_rows, packed = [], [], []
    for i in range(nrint):
        npts = int(rng.integers(1, 6))
        A = np.zeros((npts, nk1))
        A[:, i:i + 4] = rng.random((npts, 4))
        b = rng.normal(size=(npts, 3))
        A_rows.append(A)
        X_rows.append(b)
        packed.append(_triangle(A[:, i:i + 4], b))
    D = rng.normal(size=(nrint - 1, 5))
    ref = _dense(np.vstack(A_rows), np.vstack(X_rows), D, pinv)
    assert np.abs(cr_host(packed, D, pinv) - ref).max() <= 1e-11 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("case", ["geom_f32_480x640_f552_knots136", "geom_f32_480x640_f555_knots167",
                                  "geom_f32_480x640_f60_knots2039"])
def test_collocation_problems_of_the_golden_frames(cr_host, manifest, golden, case):
    """Real data: 1..65 points per interval, rank-deficient interval triangles, entries that decay below 1e-150."""
    v, g = manifest["cases"][case], golden(case)
    f = synth.make_frame(v["seed"], v["H"], v["W"], v["fill"], v["depth"])
    pcd, pix = gp.deproject(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
    rows, _, _ = gp.select_edge(pcd, pix)
    X = pcd[gp.order_along_x(pcd, pix, rows)]
    u, t = fp.chord_parameter(X), g["t"]
    n = len(t)
    nrint, nk1 = n - 7, n - 4
    A, iv, l = np.zeros((len(u), nk1)), np.zeros(len(u), int), 3
    for i, ui in enumerate(u):
        while ui >= t[l + 1] and l != nk1 - 1:
            l += 1
        A[i, l - 3:l + 1] = fp._bspl(t, ui, l)
        iv[i] = l - 3
    packed = [_triangle(A[iv == i][:, i:i + 4], X[iv == i]) for i in range(nrint)]
    D = np.asarray(fp._disc(t, n))[:nrint - 1]
    for pinv in (0.0, 20.0, 1e-4):
        ref = _dense(A, X, D, pinv)
        c = cr_host(packed, D, pinv)
        assert not np.isnan(c).any(), (case, pinv)
        assert np.abs(c - ref).max() <= 1e-11, (case, pinv)
