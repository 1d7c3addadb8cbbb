"""clb_find_sa_static -- the entry point that allocates no device memory -- against clb_find_sa, through the
C ABI.  Collected last on purpose."""
import numpy as np
import pytest

import columba_b200 as cb

pytestmark = pytest.mark.gpu


def test_find_sa_static_matches_find_sa_on_two_calls_in_a_row(golden_index, cuda_lib):
    ix = cb.Index(golden_index, 4, 10, 4, device=0, lib_path=cuda_lib)
    m = cb.Mapper(ix)
    try:
        n = ix.text_length
        rng = np.random.default_rng(11)
        # the second call is shorter than the first: what the first one left in the shared array must not show
        first = np.concatenate([[0, 1, 63, 64, n - 1], rng.integers(0, n, 200000)]).astype(np.uint32)
        second = rng.integers(0, n, 777).astype(np.uint32)
        for rows in (first, second):
            exp = m.find_sa(rows)
            got = m.find_sa_static(rows)
            assert got.dtype == exp.dtype and got.shape == exp.shape
            assert (got == exp).all()  # index work: bit-exact
        # the largest call the static array holds, and one row more
        full = rng.integers(0, n, cb.api.STATIC_MAX_ROWS).astype(np.uint32)
        assert (m.find_sa_static(full) == m.find_sa(full)).all()
        with pytest.raises(cb.ClbError) as e:
            m.find_sa_static(np.zeros(cb.api.STATIC_MAX_ROWS + 1, np.uint32))
        assert e.value.code == -5 and "CLB_STATIC_MAX_ROWS" in str(e.value)
        with pytest.raises(cb.ClbError):
            m.find_sa_static(np.array([n], np.uint32))  # row out of range, as clb_find_sa
        assert len(m.find_sa_static(np.zeros(0, np.uint32))) == 0
        # a second context on the same GPU uses the same array
        m2 = cb.Mapper(ix)
        assert (m2.find_sa_static(second) == m.find_sa(second)).all()
        m2.close()
    finally:
        m.close()
        ix.close()
