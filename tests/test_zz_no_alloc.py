"""GPU: lbm_reynolds_column_static -- the entry point that allocates nothing (static device scratch inside the library) -- against
lbm_reynolds_column through the C ABI.  No counterpart in the reference (include/lbm_b200.h says why it exists).  Named test_zz_*
so that pytest collects it last: a failure here must not hide the parity tests under `pytest -x`."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ROWS = 65536                                           # LBM_STATIC_MAX_ROWS, include/lbm_b200.h


@pytest.mark.parametrize("prec,inplace", [("dp", False), ("sp", True)])
def test_static_scratch_column_matches_the_allocating_entry_point(prec, inplace):
    """Two calls in a row with the lattice advanced in between (the shared scratch keeps the first call's contents: stale values
    would show), then a second context on the same GPU with fewer rows.  Tolerance: none -- the tests of this repository compare
    bits, and the two entry points run the same arithmetic in the same order."""
    from latticeboltzmann_b200 import Lattice
    nx, ny = 400, 2000
    with Lattice(nx, ny, prec, inplace=inplace) as lat:
        lat.InitializeArrays()
        uy_buf = np.full(nx, 7.0, dtype=lat.dtype)         # caller's buffers, reused by both calls
        fl_buf = np.full(nx, 5, dtype=np.int8)
        seen = []
        for steps in (200, 100):
            lat.run(steps)
            want_uy, want_fl = lat.reynolds_column()
            got_uy, got_fl = lat.reynolds_column_static(uy_buf, fl_buf)
            assert got_uy.tobytes() == want_uy.tobytes(), f"u_y of the centre column differs after {steps} more steps"
            assert np.array_equal(got_fl, want_fl)
            assert float(lat.reynolds_from_column(got_uy, got_fl)).hex() == float(lat.ComputeReynolds()).hex()
            seen.append(got_uy.copy())
        assert seen[0].tobytes() != seen[1].tobytes(), "the flow did not move between the two calls: the test would not see stale scratch"
        assert want_fl[0] == 0 and want_fl[nx // 2] == 1    # channel wall rows / fluid
        # a smaller lattice on the same GPU shares the array: the longer column of the call above is still in it
        with Lattice(96, 520, prec, inplace=inplace) as small:
            small.InitializeArrays()
            small.run(20)
            w_uy, w_fl = small.reynolds_column()
            g_uy, g_fl = small.reynolds_column_static()
            assert g_uy.tobytes() == w_uy.tobytes() and np.array_equal(g_fl, w_fl)
        # and the first context again, unchanged state: same answer as before
        again_uy, again_fl = lat.reynolds_column_static()
        assert again_uy.tobytes() == seen[1].tobytes() and np.array_equal(again_fl, want_fl)


def test_static_scratch_refuses_more_rows_than_it_holds():
    from latticeboltzmann_b200 import Lattice, LbmError
    from latticeboltzmann_b200 import _capi
    with Lattice(MAX_ROWS + 1, 8, "sp") as lat:            # 65537 x 8 sites: a few MB
        lat.InitializeArrays()
        with pytest.raises(LbmError) as ei:
            lat.reynolds_column_static()
        assert ei.value.code == _capi.ERR_ARG and "LBM_STATIC_MAX_ROWS" in str(ei.value)
        uy, fl = lat.reynolds_column()                     # the allocating entry point still serves it
        assert uy.shape == (MAX_ROWS + 1,)
    with Lattice(MAX_ROWS, 8, "sp") as lat:                # exactly the maximum is served
        lat.InitializeArrays()
        w_uy, w_fl = lat.reynolds_column()
        g_uy, g_fl = lat.reynolds_column_static()
        assert g_uy.tobytes() == w_uy.tobytes() and np.array_equal(g_fl, w_fl)
