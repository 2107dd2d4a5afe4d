"""pytest configuration: registers the ``gpu`` marker and makes sure the native pieces
(libptcore.so = the product, libpt_oracle.so / oracle/_ref = the checkers) are built.

``-m "not gpu"``: oracle vs golden vectors, host logic, C-ABI surface (no compute).
``-m gpu``     : parity of the CUDA path with the oracle / golden fixtures through the C ABI.
Nothing here (or in any test) reads /root/reference at run time; oracle/_ref is rebuilt
only when the reference happens to be present (the build container).
"""
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    import __graft_entry__ as ge
    ge.build()
    yield


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device here; run with gpurun")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from pt_testlib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref_cpu():
    """The reference's own code on host cores (oracle/_ref/libref_cpu.so), if built."""
    from pt_testlib import RefCpu
    try:
        return RefCpu()
    except FileNotFoundError:
        pytest.skip("oracle/_ref/libref_cpu.so not built (needs /root/reference at build time)")


@pytest.fixture(scope="session")
def core():
    from pt_testlib import pkg
    c = pkg().Core(0)
    yield c
    c.close()
