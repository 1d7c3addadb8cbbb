"""Drop-in check: INTEGRATION.md's reference-side binding (tests/integration/b200strategy.h) compiled against the
reference's own headers and linked with the reference's own record-formatting objects (`make -C oracle dropin`,
build container only; the binaries travel in oracle/_ref like the reference binary).  The driver maps the golden
reads through the C ABI and writes the SAM body with the reference's TextOcc::generateSAMSingleEnd*: it must equal
the SAM body of `columba_dev -a all` byte for byte (tests/golden/expected_sam)."""
import gzip
import os
import subprocess

import pytest

import columba_b200 as cb
import golden_util as gu
from test_sam_records import CASES, expected_sam

REFDIR = os.path.join(gu.ROOT, "oracle", "_ref")


def dropin(binary, base, k, tmp_path):
    exe = os.path.join(REFDIR, binary)
    if os.path.isdir("/root/reference/src"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(gu.ROOT, "oracle"), "_ref/" + binary])
    if not os.path.exists(exe):
        pytest.skip("the drop-in binary is built where the reference sources are (it did not travel)")
    fq = os.path.join(gu.WORK, "g1_edit.fq")
    if not os.path.exists(fq):
        with gzip.open(os.path.join(gu.GOLD, "g1_edit.fq.gz"), "rb") as i, open(fq, "wb") as o:
            o.write(i.read())
    sam = str(tmp_path / "dropin.sam")
    p = subprocess.run([exe, base, fq, cb.SCHEMES_DIR, str(k), sam], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-500:]
    return open(sam).readlines()


@pytest.mark.parametrize("case", sorted(CASES))
def test_dropin_sam_is_byte_identical_kernel_logic(case, golden_index, hostemu_lib, tmp_path):
    assert dropin("columba_dropin_emu", golden_index, CASES[case], tmp_path) == expected_sam(case)


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(CASES))
def test_dropin_sam_is_byte_identical_gpu(case, golden_index, cuda_lib, tmp_path):
    assert dropin("columba_dropin_cuda", golden_index, CASES[case], tmp_path) == expected_sam(case)
