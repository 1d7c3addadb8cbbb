"""Paired-end mode (include/columba_b200_pe.hpp, a C++ host layer over the C ABI) through a small C++ program,
against golden outputs of the unmodified reference run with -f R1 -F R2 -nI: every SAM record of every pair
(flags incl. proper-pair / discordant / primary / mate fields, positions, MAPQ, CIGAR, mate position, template
length, NM) and the five pair counters, in ALL and BEST mode, FR / RF / FF, with and without discordant pairs."""
import gzip
import json
import os
import re
import subprocess

import pytest

import columba_b200 as cb
import golden_util as gu

SRC = os.path.join(gu.ROOT, "tests", "cpp", "pe_cpp_test.cpp")
HDR = os.path.join(gu.ROOT, "include", "columba_b200_pe.hpp")
CASES = json.load(open(os.path.join(gu.GOLD, "pe_cases.json")))


def expected_pe(case):
    rows = []
    with gzip.open(os.path.join(gu.GOLD, "expected_pe", case + ".tsv.gz"), "rt") as f:
        for line in f:
            a = line.rstrip("\n").split("\t")
            rows.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5], a[6], int(a[7]), int(a[8]), int(a[9])))
    return rows, json.load(open(os.path.join(gu.GOLD, "expected_pe", case + ".json")))


def run_pe(lib, case):
    p = CASES[case]
    exe = os.path.join(gu.WORK, "pe_cpp_test_" + os.path.basename(lib).replace(".", "_"))
    libdir, libname = os.path.dirname(lib), os.path.basename(lib)[3:-3]
    newest = max(os.path.getmtime(x) for x in (SRC, HDR, lib, os.path.join(gu.ROOT, "include", "columba_b200.hpp")))
    if not os.path.exists(exe) or os.path.getmtime(exe) < newest:
        subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(gu.ROOT, "include"), SRC, "-L", libdir,
                               "-l" + libname, "-Wl,-rpath," + libdir, "-o", exe])
    base = gu.index_of({"fasta": p["set"]} if p["set"] != "g1" else {}, lib)
    fqs = []
    for k in (1, 2):
        fq = os.path.join(gu.WORK, "pe_%s_%d.fq" % (p["set"], k))
        if not os.path.exists(fq):
            with gzip.open(os.path.join(gu.GOLD, "pe_%s_%d.fq.gz" % (p["set"], k)), "rb") as i, open(fq, "wb") as o:
                o.write(i.read())
        fqs.append(fq)
    r = subprocess.run([exe, base, fqs[0], fqs[1], cb.SCHEMES_DIR, p["mode"], str(p["value"]), p["ori"],
                        str(p["max_frag"]), str(p["min_frag"]), str(p["disc"]), str(p["x"])] + (["infer"] if p.get("infer") else []),
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-500:]
    rows = []
    for line in r.stdout.splitlines():
        a = line.split("\t")
        rows.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5], a[6], int(a[7]), int(a[8]), int(a[9])))
    m = re.search(r"pairs (\d+) mapped (\d+) discordant (\d+) half (\d+) unpaired (\d+)", r.stderr)
    ctr = dict(zip(("pairs", "mapped", "discordant", "half", "unpaired"), map(int, m.groups())))
    return sorted(rows), ctr


def check(lib, case):
    got, ctr = run_pe(lib, case)
    exp, ectr = expected_pe(case)
    assert ctr == ectr
    assert len(got) == len(exp)
    assert got == exp


@pytest.mark.parametrize("case", sorted(CASES))
def test_paired_end_kernel_logic(case, golden_index, hostemu_lib):
    check(hostemu_lib, case)


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(CASES))
def test_paired_end_gpu(case, golden_index, cuda_lib):
    check(cuda_lib, case)
