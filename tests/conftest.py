import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def hostemu_lib():
    """CPU build of the kernels' task functions (tests/hostemu) -- unit tests of
    the kernel logic only; the product library has no CPU path."""
    import __graft_entry__ as ge
    return ge.build_hostemu()


@pytest.fixture(scope="session")
def cuda_lib():
    import __graft_entry__ as ge
    lib = ge.LIB
    if not os.path.exists(lib):
        lib = ge.build_cuda()
    return lib


@pytest.fixture(scope="session")
def golden_index(hostemu_lib):
    """Index of the golden reference, built by the product's host-side index
    writer (clb_build_index; host code, present in both libraries)."""
    import columba_b200 as cb
    import golden_util as gu
    fa = gu.golden_fasta()
    base = os.path.join(gu.WORK, "g1")
    if not os.path.exists(base + ".rev.brt") or not os.path.exists(base + ".meta"):
        lib = hostemu_lib
        cb.build_index(fa, base, 4, lib_path=lib)
    return base
