"""Preparation stage (codes + bit planes of both strands, four characters per
instruction) against a character-by-character restatement; see
tests/cpp/prep_swar_test.cpp.  Kernel logic compiled for the host (CPU)."""
import os
import subprocess

import golden_util as gu


def test_prep_swar_matches_scalar_restatement():
    src = os.path.join(gu.ROOT, "tests", "cpp", "prep_swar_test.cpp")
    os.makedirs(gu.WORK, exist_ok=True)
    exe = os.path.join(gu.WORK, "prep_swar_test")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(gu.ROOT, "columba_b200", "csrc"), src,
                           "-o", exe])
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout.strip() == "OK", p.stdout[-300:]


def test_screen32_matches_64bit_window():
    """Screening pass on a 32-bit band (screenStep32) against the 64-bit window form, row by row, on 240 000
    candidate windows (tests/cpp/screen32_test.cpp)."""
    src = os.path.join(gu.ROOT, "tests", "cpp", "screen32_test.cpp")
    os.makedirs(gu.WORK, exist_ok=True)
    exe = os.path.join(gu.WORK, "screen32_test")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wno-unused-function", "-I",
                           os.path.join(gu.ROOT, "columba_b200", "csrc"), src, "-o", exe])
    p = subprocess.run([exe, "6000"], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout.strip() == "OK", p.stdout[-300:]
