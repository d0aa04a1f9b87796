import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200")]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def hostemu_lib():
    """Kernel SOURCE compiled for the host (tests/host_emul) - test-only, lets the CPU suite exercise csrc/."""
    d = os.path.join(ROOT, "tests/host_emul")
    so = os.path.join(d, "libgo2_hostemu.so")
    srcs = [os.path.join(d, "go2_host_emul.cpp")] + [os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/csrc", f)
                                                      for f in os.listdir(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/csrc"))]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++14", "-fPIC", "-shared", "-o", so, srcs[0]])
    return so
