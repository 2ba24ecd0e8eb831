"""GPU tests of the drop-in entry points that the full-sequence parity tests do not reach: `dre_run_pass` with HOST
arrays (upload -> pass -> download, the call a Fortran host makes), the separate small kernels (cell list,
find_cc_new2 locator, pix2vec) and the error behaviour of the C ABI (non-zero return + message, no exception)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from test_gpu_parity import _engine, _oracle, assert_counters_equal, assert_fields_close  # noqa: E402


def _compare_state(hs, st, scene, tag):
    assert_fields_close(hs.u_fest, st.u_fest, f"u_fest {tag}")
    assert_fields_close(hs.u_final, st.u_final, f"u_final {tag}")
    for l in range(scene.lnum):
        assert_fields_close(hs.scaspe[l], st.scaspe[l], f"scaspe[{l}] {tag}")
    assert_fields_close(hs.lum_lost, st.lum_lost, f"lum_lost {tag}")


@pytest.mark.parametrize("fixture", ["tiny_trusti", "tiny_galaxy"])
def test_host_array_call_matches_oracle(fixture, request):
    """dre_run_pass with caller-owned arrays: precalc -> direct light -> rt_prepare on the host arrays -> one scattering
    iteration; counters and intersection checksum bit for bit, fields <= 1e-5 relative."""
    from dartray_b200 import abi
    from oracle import oracle as O
    scene = request.getfixturevalue(fixture)
    eng = _engine(scene)
    orc, st = _oracle(scene)
    hs = O.State(scene)
    has_src, has_cell = scene.nsrc > 0, bool(np.any(scene.lumcell))
    seq = ([abi.RTT_PRECALC_CELL] if has_cell else []) + ([abi.RTT_PRECALC_SRC] if has_src else []) + \
          ([abi.RTT_DIR_CELL] if has_cell else []) + ([abi.RTT_DIR_SRC] if has_src else [])
    for rt in seq:
        c_o = orc.run_pass(rt, 0, st)
        c_g = eng.rt_loop_host(rt, 0, hs)
        assert_counters_equal(c_g, c_o)
        _compare_state(hs, st, scene, f"after host-array pass {rt}")
    # rt_prepare(rtt_scatt) (dartray_hub.f90:432-457) on both sets of host arrays
    assert not orc.prepare_scatt(1, hs)
    assert not orc.prepare_scatt(1, st)
    orc.set_sources(st.lumcell)
    eng.set_sources(hs.lumcell)
    c_o = orc.run_pass(abi.RTT_SCATT, 1, st)
    c_g = eng.rt_loop_host(abi.RTT_SCATT, 1, hs)
    assert_counters_equal(c_g, c_o)
    _compare_state(hs, st, scene, "after host-array scattering iteration 1")
    eng.close()


def test_cell_list_kernel(tiny_galaxy):
    """leaf && sum(lum) > 0 gate of rt_loop (rt_routines.f90:1016) as its own kernel."""
    from dartray_b200 import abi
    eng = _engine(tiny_galaxy)
    got = eng.build_source_list(abi.RTT_DIR_CELL)
    g = tiny_galaxy.grid
    want = np.nonzero((g.cchild == -1) & (tiny_galaxy.lumcell.sum(axis=1) > 0))[0]
    assert np.array_equal(np.sort(got), want)
    eng.close()


def test_locator_kernel_matches_oracle(small_trusti):
    """find_cc_new2 (rt_routines.f90:1962-2066) for arbitrary points: cell ids bit for bit, points outside -> -1."""
    eng = _engine(small_trusti)
    orc, _ = _oracle(small_trusti, nthreads=1)
    g = small_trusti.grid
    rng = np.random.default_rng(7)
    pts = (rng.random((4000, 3)) - 0.5) * g.modelsize * 1.1          # ~25 % outside the model
    leaves = np.nonzero(g.cchild == -1)[0].astype(np.int32)
    cc_old = leaves[rng.integers(0, len(leaves), len(pts))]
    want = orc.locate(pts, cc_old)
    got = eng.locate(pts, cc_old)
    assert np.array_equal(got, want)
    assert (want == -1).any() and (want >= 0).any()
    eng.close()


@pytest.mark.parametrize("nside", [1, 4, 64, 512])
def test_pix2vec_kernel_bit_exact(tiny_trusti, nside):
    """pix2vec = pix2ang_nest (healpix_routines.f90:30) + calc_cproj (rt_routines.f90:1945): the direction tables the
    march reads must be the oracle's doubles bit for bit."""
    eng = _engine(tiny_trusti)
    orc, _ = _oracle(tiny_trusti, nthreads=1)
    npix = 12 * nside * nside
    ip = np.unique(np.concatenate([np.arange(min(npix, 3000)), np.random.default_rng(1).integers(0, npix, 3000)])).astype(np.int32)
    _, _, v = eng.pix2vec(nside, ip)
    assert np.array_equal(v.view(np.int64), orc.pix2vec(nside, ip).view(np.int64))
    eng.close()


def test_error_behaviour(tiny_trusti):
    """Reference convention is print + STOP; the ABI returns non-zero and keeps a message."""
    from dartray_b200 import abi, engine
    eng = _engine(tiny_trusti)
    with pytest.raises(engine.EngineError, match="not configured"):
        eng.make_maps(32, 1.0, 0)                      # dre_set_params has not been called yet
    eng.set_params(abi.RTT_PRECALC_SRC)
    with pytest.raises(engine.EngineError, match="i_obs"):
        eng.make_maps(32, 1.0, 0)                      # no i_obs pass was run
    with pytest.raises(engine.EngineError, match="i_obs"):
        eng.calc_sed(1.0e4, 0)
    with pytest.raises(engine.EngineError):
        eng.rt_loop(99)                                # unknown rt_type
    with pytest.raises(engine.EngineError, match="units"):
        eng.rt_loop(abi.RTT_PRECALC_SRC); eng.rt_loop(abi.RTT_DIR_SRC); eng.rt_loop_iobs(abi.RTT_I_OBS_DIR_SRC)
        eng.make_maps(32, 1.0, 2)                      # W/m needs lambda_si
    # the engine stays usable after an error
    assert eng.make_maps(16, 1.0, 0).shape == (tiny_trusti.ndir, tiny_trusti.lnum, 16, 16)
    eng.close()
