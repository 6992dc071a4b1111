import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _has_gpu():
    try:
        from elevenrender_b200 import capi
        return capi.device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu():
    """The CUDA path must be present on a GPU box: no skip, no fallback."""
    from elevenrender_b200 import capi
    n = capi.device_count()
    assert n > 0, "no CUDA device"
    info = capi.device_info(0)
    assert info["is_compatible"], f"device {info['name']} is not sm_100"
    return info


@pytest.fixture(scope="session")
def ref():
    """the reference oracle (oracle/_ref/liber_ref.so): built here from /root/reference, travels prebuilt"""
    import refharness as R
    if not R.available():
        if os.path.isdir("/root/reference/src"):
            import subprocess
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"])
        else:
            pytest.fail("oracle/_ref/liber_ref.so missing and /root/reference absent: build it in the container first")
    return R
