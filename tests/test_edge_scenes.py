"""CPU checks of the edge-case scenes (tests/edge_scenes.py) and of the oracle on them; the GPU parity runs of the
same scenes are in test_gpu_zedge.py."""
import numpy as np
import pytest

import edge_scenes as E


def test_wide_keeps_end_points_and_shapes(tiny_trusti):
    for L in (1, 33, 70):
        sc = E.wide(tiny_trusti, L)
        assert sc.lnum == L and sc.dens_arr.shape == (sc.grid.ncell, L) and sc.lum_p_src.shape == (1, L)
        assert len(sc.kp_sca) == L and len(sc.npix) == L
        assert np.array_equal(sc.dens_arr[:, 0], tiny_trusti.dens_arr[:, 0])
        if L > 1:
            assert np.array_equal(sc.dens_arr[:, -1], tiny_trusti.dens_arr[:, -1])
            assert sc.lum_p_src[0, -1] == tiny_trusti.lum_p_src[0, -1]


def test_precalc_wavelengths_are_independent(tiny_galaxy):
    """In the precalculation the fate of a ray is geometric (rad_lim, beam tests: rt_routines.f90:2428-2462), so the
    field of a wavelength cannot depend on how many other wavelengths ride along: column 0 of a 33-wavelength run
    equals the 4-wavelength run bit for bit (one thread: same summation order)."""
    from dartray_b200 import abi
    from oracle import oracle as O
    out = []
    for sc in (tiny_galaxy, E.wide(tiny_galaxy, 33)):
        orc = O.Oracle(sc, nthreads=1)
        st = O.State(sc)
        c = orc.run_pass(abi.RTT_PRECALC_CELL, 0, st)
        out.append((c.intersections, st.u_fest[:, 0].copy()))
    assert out[0][0] == out[1][0] > 0
    assert out[0][1].max() > 0 and np.array_equal(out[0][1], out[1][1])


@pytest.mark.parametrize("fixture", ["tiny_trusti", "tiny_galaxy"])
def test_oracle_with_no_luminous_source(fixture, request):
    from dartray_b200 import abi
    from oracle import oracle as O
    sc = E.dark(request.getfixturevalue(fixture))
    orc = O.Oracle(sc, nthreads=2)
    st = O.State(sc)
    for rt in (abi.RTT_PRECALC_CELL, abi.RTT_DIR_CELL) + ((abi.RTT_PRECALC_SRC, abi.RTT_DIR_SRC) if sc.nsrc else ()):
        c = orc.run_pass(rt, 0, st)
        assert c.rays == 0 and c.intersections == 0 and c.as_dict()["sources"] == 0
    assert not st.u_final.any() and not st.u_fest.any()


def test_point_source_host_cell_is_dusty(tiny_galaxy):
    sc = E.with_point_source(tiny_galaxy, (1000.0, 500.0, 20.0))
    j = int(sc.cell_src[0])
    assert sc.nsrc == 1 and sc.src_cell[j] == 0 and sc.grid.cchild[j] == -1
    assert sc.dens_arr[j].min() > 0 and sc.lumcell[j].max() > 0     # Q4 / Q10 run with tau > 0, and the cell emits too
    assert np.all(np.abs(sc.ccoord_p_src[0] - sc.grid.ccoord[j]) < sc.grid.csize[j] / 2)
