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


def build_and_run(lib, base, mode, value, reads="g1_edit", x=0):
    exe = os.path.join(gu.WORK, "host_cpp_test_" + os.path.basename(lib).replace(".", "_"))
    libdir, libname = os.path.dirname(lib), os.path.basename(lib)[3:-3]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(SRC), os.path.getmtime(lib)):
        subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(gu.ROOT, "include"), SRC, "-L", libdir,
                               "-l" + libname, "-Wl,-rpath," + libdir, "-o", exe])
    fq = os.path.join(gu.WORK, reads + ".fq")
    if not os.path.exists(fq):
        with gzip.open(os.path.join(gu.GOLD, reads + ".fq.gz"), "rb") as i, open(fq, "wb") as o:
            o.write(i.read())
    p = subprocess.run([exe, base, fq, cb.SCHEMES_DIR, mode, str(value), str(x)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-500:]
    if mode == "sam":
        return p.stdout.splitlines(keepends=True)
    rows = []
    for line in p.stdout.splitlines():
        a = line.split("\t")
        rows.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5]))
    return sorted(rows)


def all_and_best(lib, golden_index):
    # SAM body of single-end ALL mode from the C++ layer, byte for byte
    from test_sam_records import expected_sam
    assert build_and_run(lib, golden_index, "sam", 4) == expected_sam("edit_k4_columba")
    assert build_and_run(lib, golden_index, "all", 4) == gu.expected("edit_k4_columba")[0]
    assert build_and_run(lib, golden_index, "best", 95) == expected_best("best_I95")
    assert build_and_run(lib, golden_index, "best", 95, x=2) == expected_best("best_I95_x2")
    # repeat-rich reference: many co-optimal alignments, exact stratum, strata after the best one
    g2 = gu.index_of({"fasta": "g2"}, lib)
    assert build_and_run(lib, g2, "best", 95, reads="g3") == expected_best("g3_best_I95")
    assert build_and_run(lib, g2, "best", 97, reads="g2", x=1) == expected_best("g2_best_I97_x1")


def test_cpp_host_all_and_best_kernel_logic(golden_index, hostemu_lib):
    all_and_best(hostemu_lib, golden_index)


@pytest.mark.gpu
def test_cpp_host_all_and_best_gpu(golden_index, cuda_lib):
    all_and_best(cuda_lib, golden_index)


def test_cpp_host_reports_missing_device_or_index(cuda_lib, tmp_path):
    """The product library has no CPU path: without a GPU (or without the
    index) the C++ layer throws with the library's message."""
    exe_rows = None
    try:
        exe_rows = build_and_run(cuda_lib, str(tmp_path / "nonexistent"), "all", 4)
    except AssertionError as e:
        assert "error:" in str(e)
    assert exe_rows is None
