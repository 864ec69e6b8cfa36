import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "tiled: keeps the temporally blocked smoother on the CPU test double")


@pytest.fixture(scope="session")
def R():
    import refharness
    return refharness


@pytest.fixture(scope="session")
def sim(R):
    """Backend bound to the CPU test double (tests/devsim): host logic + plain-kernel arithmetic."""
    if not os.path.exists(R.DEVSIM_LIB):
        subprocess.run(["bash", os.path.join(ROOT, "tests", "devsim", "build.sh")], check=True)
    return R.Backend(R.DEVSIM_LIB)


@pytest.fixture(scope="session")
def gpu(R):
    """Backend bound to the product library; fails loudly when it is missing or has no device."""
    B = R.Backend()
    assert B.lib.cfdnn_device_count() >= 1, "no CUDA device visible to libcfdnn_b200.so"
    return B


@pytest.fixture(scope="session")
def oracle(R):
    if not R.have_oracle():
        if os.path.isdir("/root/reference/src"):
            subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True)
        else:
            pytest.fail("oracle/_ref/ref_harness missing (it is built where /root/reference exists and "
                        "travels to the GPU box as a built file)")
    return R
