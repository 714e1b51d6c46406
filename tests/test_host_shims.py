"""The host-side C++ shims (open3d_slam_private_b200/host): compiled against the interface mirror and the C ABI.
CPU: without a device every module fails loudly. GPU: modules run and are checked inside the self-test."""
import os
import subprocess

import pytest

from open3d_slam_private_b200 import build as b200_build
from open3d_slam_private_b200 import capi as B

HOST = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "open3d_slam_private_b200", "host")


def _build():
    b200_build.build()
    subprocess.check_call(["make", "-C", HOST, "-s"])
    return os.path.join(HOST, "selftest")


def test_host_shims_compile_and_fail_loudly_without_device():
    exe = _build()
    if B.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    out = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert out.returncode == 0, out.stdout
    assert "loud failure verified" in out.stdout


@pytest.mark.gpu
def test_host_shims_on_device():
    exe = _build()
    out = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert out.returncode == 0, out.stdout
    assert "selftest ok (device path)" in out.stdout
