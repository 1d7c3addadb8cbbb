"""C++ host layer (include/columba_b200.hpp) through a small C++ program:
ALL mode and BEST mode against the reference's golden outputs."""
import gzip
import os
import subprocess

import pytest

import columba_b200 as cb
import golden_util as gu
from test_best_mode import expected_best

SRC = os.path.join(gu.ROOT, "tests", "cpp", "host_cpp_test.cpp")


def build_and_run(lib, base, mode, value):
    exe = os.path.join(gu.WORK, "host_cpp_test_" + os.path.basename(lib).replace(".", "_"))
    libdir, libname = os.path.dirname(lib), os.path.basename(lib)[3:-3]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(SRC), os.path.getmtime(lib)):
        subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(gu.ROOT, "include"), SRC, "-L", libdir,
                               "-l" + libname, "-Wl,-rpath," + libdir, "-o", exe])
    fq = os.path.join(gu.WORK, "g1_edit.fq")
    if not os.path.exists(fq):
        with gzip.open(os.path.join(gu.GOLD, "g1_edit.fq.gz"), "rb") as i, open(fq, "wb") as o:
            o.write(i.read())
    p = subprocess.run([exe, base, fq, cb.SCHEMES_DIR, mode, str(value)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-500:]
    rows = []
    for line in p.stdout.splitlines():
        a = line.split("\t")
        rows.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5]))
    return sorted(rows)


def test_cpp_host_all_and_best_kernel_logic(golden_index, hostemu_lib):
    assert build_and_run(hostemu_lib, golden_index, "all", 4) == gu.expected("edit_k4_columba")[0]
    assert build_and_run(hostemu_lib, golden_index, "best", 95) == expected_best("best_I95")


@pytest.mark.gpu
def test_cpp_host_all_and_best_gpu(golden_index, cuda_lib):
    assert build_and_run(cuda_lib, golden_index, "all", 4) == gu.expected("edit_k4_columba")[0]
    assert build_and_run(cuda_lib, golden_index, "best", 95) == expected_best("best_I95")


def test_cpp_host_reports_missing_device_or_index(cuda_lib, tmp_path):
    """The product library has no CPU path: without a GPU (or without the
    index) the C++ layer throws with the library's message."""
    exe_rows = None
    try:
        exe_rows = build_and_run(cuda_lib, str(tmp_path / "nonexistent"), "all", 4)
    except AssertionError as e:
        assert "error:" in str(e)
    assert exe_rows is None
