import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port_oracle():
    """oracle/libhbtoracle.so (plain-C restatement); built on demand."""
    from oracle import pyoracle
    if not pyoracle.available("port"):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
    return pyoracle.load("port")


@pytest.fixture(scope="session")
def ref_oracle():
    """oracle/_ref/libhbtref.so (the real reference); skip where it was never built."""
    from oracle import pyoracle
    if not pyoracle.available("reference"):
        if os.path.exists("/root/reference/hbt/src/perf_event/PerfEventsGroup.h"):
            import subprocess
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-j8"])
        else:
            pytest.skip("oracle/_ref/libhbtref.so not built and /root/reference absent")
    return pyoracle.load("reference")
