"""GPU parity on edge cases of the path (through the C ABI, against the CPU oracle): wavelength counts that leave
lanes ragged or need several wavelengths per lane, both source kinds in one model, and no luminous source at all.
Same bar as test_gpu_parity.py: counters and the intersection checksum bit for bit, fields within 1e-5 relative."""
import numpy as np
import pytest

import edge_scenes as E
from test_gpu_parity import _engine, _oracle, assert_counters_equal, run_sequence

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("lnum", [1, 33, 40, 70])
def test_wavelength_counts(tiny_trusti, lnum):
    """lnum = 1 (31 idle lanes), 33 and 40 (two wavelengths per lane, second word ragged), 70 (three words, run by the
    four-per-lane instantiation of the ray march): precalc, direct light, i_obs, SED, maps, one scattering iteration.
    The intersection list itself is compared for lnum = 33 (the first count past one word)."""
    run_sequence(E.wide(tiny_trusti, lnum), record=4_000_000 if lnum == 33 else 0, iterations=1)


def test_point_source_in_emitting_dusty_grid(tiny_galaxy):
    """Emitting cells and a point source in one model; the source sits in a dusty cell of the disk, so the host-cell
    formulas (quirks Q4 and Q10: rt_routines.f90:2409-2416, 2519-2525, 2552) run with tau > 0."""
    sc = E.with_point_source(tiny_galaxy, (1000.0, 500.0, 20.0))
    assert sc.dens_arr[sc.cell_src[0]].min() > 0
    run_sequence(sc, record=3_000_000, iterations=1)


@pytest.mark.parametrize("fixture", ["tiny_trusti", "tiny_galaxy"])
def test_no_luminous_source(fixture, request):
    """Empty cell list (luminosity gate rt_routines.f90:1016 rejects every cell / the point source has zero
    luminosity): every pass returns at once with zero counters and leaves the accumulators untouched."""
    from dartray_b200 import abi
    from oracle import oracle as O
    sc = E.dark(request.getfixturevalue(fixture))
    eng = _engine(sc)
    orc, st = _oracle(sc)
    gst = O.State(sc)
    for rt in (abi.RTT_PRECALC_CELL, abi.RTT_PRECALC_SRC, abi.RTT_DIR_CELL, abi.RTT_DIR_SRC):
        if rt in (abi.RTT_PRECALC_SRC, abi.RTT_DIR_SRC) and sc.nsrc == 0:
            continue
        c_o = orc.run_pass(rt, 0, st)
        c_g = eng.rt_loop(rt, 0)
        assert_counters_equal(c_g, c_o)
        assert c_g.rays == 0 and c_g.intersections == 0
    eng.download_state(gst)
    assert not gst.u_final.any() and not gst.u_fest.any() and not gst.lum_lost.any()
    assert all(not p.any() for p in gst.scaspe)
    eng.close()
