import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def fc():
    """the product package (directory name starts with a digit -> importlib)"""
    lib = os.path.join(ROOT, "4c_b200", "lib", "libfourc_b200.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "4c_b200", "csrc")], check=True)
    return importlib.import_module("4c_b200")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib
