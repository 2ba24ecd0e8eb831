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


def test_rounding_grid_makes_sums_order_independent():
    """The arithmetic behind the engine's option "deterministic" (DESIGN.md section 4), in numpy: values rounded to the
    grid q = 2^(e-52) with rq(v) = (v + C) - C, C = 1.5 * 2^e, add up exactly in fp64 as long as the sum stays below
    2^53 q, so any order of summation gives the same bits; without the rounding the orders disagree."""
    rng = np.random.default_rng(5)
    v = rng.random(200_000) * 10.0 ** rng.uniform(-9, 0, 200_000)        # 9 decades of dynamic range, sum ~ 2e4
    bound = 2.0 * v.sum()
    e = int(np.ceil(np.log2(bound)))
    C = 1.5 * 2.0 ** e
    q = 2.0 ** (e - 52)
    r = (v + C) - C
    assert np.all(np.abs(r - v) <= q / 2) and np.all(r / q == np.round(r / q))      # nearest grid point
    sums, raw = set(), set()
    for _ in range(6):
        p = rng.permutation(len(v))
        sums.add(float(np.add.reduce(r[p])))          # pairwise
        sums.add(float(np.cumsum(r[p])[-1]))          # sequential
        raw.add(float(np.cumsum(v[p])[-1]))
    assert len(sums) == 1 and len(raw) > 1
    assert abs(sums.pop() - v.sum()) <= q * np.sqrt(len(v))
    # scaling the inputs and the bound by a power of two scales the result exactly (the full-size linearity test)
    r2 = (2.0 * v + 2.0 * C) - 2.0 * C
    assert np.array_equal(r2, 2.0 * r)


@pytest.mark.parametrize("fixture", ["tiny_trusti", "small_magtar"])
def test_oracle_scattered_light_checksum_and_linearity(fixture, request):
    """Two pins of the oracle that need no second implementation (the same properties are checked on the GPU at the full
    C1 size, test_gpu_zedge.py::test_trusti_full_size_properties):
    - cross-formula checksum: the direct light scattered out of the rays (sum over cells and HEALPix directions of scaspe,
      process_scatt_rad rt_routines.f90:2638-2726 with sum_pix ffn = 1 :2258-2260) equals ksca_norm x the energy removed
      from them, sum_c u_final dens cs V (deposit :2499-2525) — valid for point sources in dust-free host cells;
    - linearity: twice the luminosity gives exactly twice the accumulators (one thread: fixed summation order)."""
    import copy
    from dartray_b200 import abi
    from oracle import oracle as O
    sc = request.getfixturevalue(fixture)
    assert sc.dens_arr[sc.cell_src[0]].max() == 0
    out = []
    for s in (sc, copy.copy(sc)):
        if out:
            s.lum_p_src = 2.0 * sc.lum_p_src
        orc = O.Oracle(s, nthreads=1)
        st = O.State(s)
        orc.run_pass(abi.RTT_PRECALC_SRC, 0, st)
        c = orc.run_pass(abi.RTT_DIR_SRC, 0, st)
        out.append((c, st))
    c, st = out[0]
    cs = abi.make_params(sc, abi.RTT_DIR_SRC).cs
    removed = (st.u_final * sc.dens_arr * (cs * sc.grid.csize ** 3)[:, None]).sum(axis=0)
    scattered = np.array([st.scaspe[l][:, :sc.npix_hp[l]].sum() for l in range(sc.lnum)])
    assert scattered.min() > 0
    assert np.allclose(scattered, sc.ksca_norm * removed, rtol=1e-10, atol=0.0)
    c2, st2 = out[1]
    assert c2.intersections == c.intersections and c2.isect_hash == c.isect_hash
    assert np.array_equal(st2.u_final, 2.0 * st.u_final) and np.array_equal(st2.lum_lost, 2.0 * st.lum_lost)
    for l in range(sc.lnum):
        assert np.array_equal(st2.scaspe[l], 2.0 * st.scaspe[l])
