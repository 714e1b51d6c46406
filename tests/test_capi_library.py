"""CPU tests: the C-ABI library loads and exports every symbol include/b200reg.h declares; without a device the
product path fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from open3d_slam_private_b200 import build as b200_build
from open3d_slam_private_b200 import capi as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200reg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    b200_build.build()
    lib = ctypes.CDLL(B.lib_path())
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), s


def test_version_and_default_params():
    lib = B.load()
    assert lib.b200_version() >= 100
    p = B.default_icp_params()
    # ICPChainBase::setDefault + module defaults (ICP.cpp:100-113, MatchersImpl.h:83-86, OutlierFiltersImpl.h:137,
    # TransformationCheckersImpl.h:70,90-92)
    assert p.knn == 1 and np.isinf(p.maxDist) and p.use_trimmed == 1 and abs(p.trimmed_ratio - 0.85) < 1e-7
    assert p.max_iterations == 40 and p.use_differential == 1 and p.smooth_length == 3
    assert abs(p.min_diff_rot - 0.001) < 1e-9 and abs(p.min_diff_trans - 0.001) < 1e-9


def test_struct_sizes_match_header():
    # the ctypes mirrors must have the C layout of include/b200reg.h
    lib = B.load()
    assert ctypes.sizeof(B.IcpParams) == lib.b200_struct_size(0) == 24 * 4
    assert ctypes.sizeof(B.IcpResult) == lib.b200_struct_size(1)
    assert ctypes.sizeof(B.IndexOpts) == lib.b200_struct_size(2) == 16
    assert ctypes.sizeof(B.SurfaceNormalParams) == lib.b200_struct_size(3)
    assert ctypes.sizeof(B.O3dParams) == lib.b200_struct_size(4)
    assert ctypes.sizeof(B.O3dResult) == lib.b200_struct_size(5)


def test_no_cpu_fallback_without_device():
    if B.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(B.B200Error):
        B.Index(np.ones((4, 4), np.float32))
    with pytest.raises(B.B200Error):
        B.ICP()
    # straight through the C ABI as well
    h = ctypes.c_void_p()
    st = B.load().b200_icp_create(ctypes.byref(h), 0, ctypes.byref(B.default_icp_params()))
    assert st == B.B200_E_CUDA and h.value is None


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "open3d_slam_private_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                for pat in (r"import\s+oracle", r"from\s+oracle", r"oracle_py", r"liboracle", r"oracle/", r"orc_"):
                    assert not re.search(pat, txt), (pat, os.path.join(dirpath, f))
