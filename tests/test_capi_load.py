"""The C-ABI library builds, loads and exports every symbol include/rdp_b200.h declares (no GPU needed)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_match_header():
    import __graft_entry__
    __graft_entry__.build()
    from rdp_b200 import _lib
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "rdp_b200.h")).read()
    declared = set(re.findall(r"\b(rdp_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found in the header"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/rdp_b200.h but not exported"
    assert declared == set(_lib.EXPORTS)
    assert lib.rdp_abi_version() == _lib.ABI_VERSION == 3


def test_argument_errors_without_gpu():
    """Shape / pointer validation happens before any CUDA call, so it is testable on a CPU-only box."""
    import ctypes as C
    from rdp_b200 import _lib
    lib = _lib.load()
    n = C.c_size_t()
    assert lib.rdp_geom_workspace_bytes(0, 480, 640, C.byref(n)) == -1
    assert b"bad shape" in lib.rdp_last_error()
    assert lib.rdp_geom_workspace_bytes(64, 480, 640, C.byref(n)) == 0 and n.value > 64 * 480 * 640 * 9
    assert lib.rdp_unet_create(None, None) == -1
    out = _lib.CurvatureOut()
    assert lib.rdp_geom_curvature(None, None, 0, 1, 4, 4, None, 1.0, C.byref(out), None, 0, None) == -1


def test_product_code_never_imports_the_oracle():
    pkg_root = os.path.join(ROOT, "robotic-discovery-platform_b200")
    for d, _, files in os.walk(pkg_root):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(d, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f"{f} imports the oracle"


def test_cpu_inputs_are_rejected_not_emulated():
    import numpy as np
    import torch
    from pkg.geometry_utils import compute_curvature_profile
    from pkg.segmentation_model import UNet
    if torch.cuda.is_available():
        pytest.skip("CPU-only behaviour")
    with pytest.raises(RuntimeError):
        compute_curvature_profile(np.ones((8, 8), np.uint8), np.ones((8, 8), np.uint16), np.eye(3), 0.001)
    with pytest.raises(RuntimeError):
        UNet(3, 1).eval()(torch.zeros(1, 3, 32, 32))
