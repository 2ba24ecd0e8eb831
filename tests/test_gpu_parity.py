"""GPU parity: the CUDA engine (through the C ABI) against the CPU oracle on the same inputs.
Bit-exact: intersection lists (cell, isel, psel bits), counters.  fp64 fields: <= 1e-5 relative
(north star tolerance; the engine's atomics only change the summation order)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _engine(scene, options=()):
    from dartray_b200 import engine
    eng = engine.Engine(scene)
    eng.set_option("hash", 1)   # intersection checksum (off by default: parity facility)
    for k, v in options:        # engine tuning options (dre_set_option): none of them may change a result
        eng.set_option(k, v)
    return eng


def _oracle(scene, nthreads=8):
    from oracle import oracle as O
    return O.Oracle(scene, nthreads=nthreads), O.State(scene)


def _sorted_records(r):
    return np.sort(r, order=["unit", "step", "ray", "k"])


def assert_fields_close(a, b, name, rtol=RTOL, atol=0.0):
    """atol (scalar or broadcastable array): extra absolute allowance, used only for the deterministic mode, whose
    accumulators are sums of values rounded to a known grid."""
    a = np.asarray(a); b = np.asarray(b)
    scale = np.abs(b).max()
    if scale == 0:
        assert np.abs(a).max() <= np.max(atol), name
        return
    # relative to each element, with an absolute floor far below any physically meaningful value
    err = np.maximum(np.abs(a - b) - atol, 0.0) / np.maximum(np.abs(b), 1e-12 * scale)
    assert err.max() <= rtol, f"{name}: max rel err {err.max():.3e}"


def assert_counters_equal(c1, c2, keys=("sources", "rays", "intersections", "deposits", "beam_tests", "splits", "merges",
                                        "sca_elems", "active_lambda", "steps", "isect_hash")):
    d1, d2 = c1.as_dict(), c2.as_dict()
    for k in keys:
        assert d1[k] == d2[k], f"counter {k}: engine {d1[k]} oracle {d2[k]}"


def assert_psel_av(eng, st, scene, where):
    """psel_av_arr(i0, iterations, :) (rt_routines.f90:1456-1463): the oracle keeps the per-source sums psel_sum / ipsel of the
    slice, the engine returns the averages; sources without a terminated ray stay 0 on both sides."""
    from dartray_b200 import abi
    got = np.zeros(scene.grid.ncell + scene.nsrc)
    eng.download(abi.ARR_PSEL_AV, got)
    want = np.where(st.ipsel > 0, st.psel_sum / np.maximum(st.ipsel, 1), 0.0)
    assert np.array_equal(got != 0, want != 0), f"psel_av {where}: different sources have an average path"
    assert_fields_close(got, want, f"psel_av {where}")


def run_sequence(scene, record=0, iterations=2, options=(), nthreads=8, maps=True):
    """precalc_src -> dir_src -> i_obs_dir_src -> scatt x iterations -> i_obs on both sides, comparing after each pass."""
    from dartray_b200 import abi
    eng = _engine(scene, options)
    orc, st = _oracle(scene, nthreads)
    from oracle import oracle as O
    gst = O.State(scene)
    has_src = scene.nsrc > 0
    has_cell = bool(np.any(scene.lumcell))
    passes = []
    if has_cell:
        passes.append(abi.RTT_PRECALC_CELL)
    if has_src:
        passes.append(abi.RTT_PRECALC_SRC)
    if has_cell:
        passes.append(abi.RTT_DIR_CELL)
    if has_src:
        passes.append(abi.RTT_DIR_SRC)
    for rt in passes:
        if record:
            eng.set_record(record); orc.set_record(record)
        c_o = orc.run_pass(rt, 0, st)
        c_g = eng.rt_loop(rt, 0)
        assert_counters_equal(c_g, c_o)
        if record:
            ro, rg = _sorted_records(orc.records()), _sorted_records(eng.records())
            assert len(ro) == len(rg) == c_o.intersections
            assert np.array_equal(ro, rg), f"intersection list differs in pass {rt}"
        eng.download_state(gst)
        assert_fields_close(gst.u_fest, st.u_fest, f"u_fest after pass {rt}")
        assert_fields_close(gst.u_final, st.u_final, f"u_final after pass {rt}")
        for l in range(scene.lnum):
            assert_fields_close(gst.scaspe[l], st.scaspe[l], f"scaspe[{l}] after pass {rt}")
        assert_fields_close(gst.lum_lost, st.lum_lost, f"lum_lost after pass {rt}")
    # one psel_av slice for the direct-light phase: cells from rtt_dir_cell, point sources from rtt_dir_src
    assert_psel_av(eng, st, scene, "after the direct-light passes")
    # direct-light i_obs
    for rt in ([abi.RTT_I_OBS_DIR_CELL] if has_cell else []) + ([abi.RTT_I_OBS_DIR_SRC] if has_src else []):
        c_o = orc.run_iobs(rt, st)
        c_g = eng.rt_loop_iobs(rt, gst)
        assert_counters_equal(c_g, c_o, keys=("sources", "rays", "intersections", "isect_hash"))
        assert_fields_close(gst.i_obs, st.i_obs, f"i_obs after pass {rt}")
        assert_fields_close(eng.calc_sed(1.0e4, 0), orc.calc_sed(st, 1.0e4, 0), f"SED (direct light) after pass {rt}")
        if maps:
            assert_fields_close(eng.make_maps(32, 1.5, 1), orc.make_maps(st, 32, 1.5, 1), f"maps (direct light) after pass {rt}")
        if len(scene.ccoord_obs):
            assert_fields_close(gst.i_obs_in, st.i_obs_in, f"i_obs_in after pass {rt}")
    # scattering iterations
    tot_g = np.zeros(scene.lnum)
    for it in range(1, iterations + 1):
        f_o = orc.prepare_scatt(it, st)
        f_g = eng.rt_prepare_scatt(it, tot_g)
        assert f_o == f_g
        assert_fields_close(tot_g, st.tot_rad_en, "tot_rad_en")
        if f_o:
            break
        orc.set_sources(st.lumcell)
        st.psel_sum[...] = 0; st.ipsel[...] = 0      # expand_psel_av_arr (dartray_hub.f90:453): a fresh slice per iteration
        if record:
            eng.set_record(record); orc.set_record(record)
        c_o = orc.run_pass(abi.RTT_SCATT, it, st)
        c_g = eng.rt_loop(abi.RTT_SCATT, it)
        assert_counters_equal(c_g, c_o)
        if record:
            ro, rg = _sorted_records(orc.records()), _sorted_records(eng.records())
            assert np.array_equal(ro, rg), f"intersection list differs in scattering iteration {it}"
        eng.download_state(gst)
        assert_fields_close(gst.u_final, st.u_final, f"u_final after scatt {it}")
        for l in range(scene.lnum):
            assert_fields_close(gst.scaspe[l], st.scaspe[l], f"scaspe[{l}] after scatt {it}")
            assert_fields_close(gst.scaspe_tot[l], st.scaspe_tot[l], f"scaspe_tot[{l}] after scatt {it}")
        assert_fields_close(gst.lum_lost, st.lum_lost, f"lum_lost after scatt {it}")
        assert_fields_close(gst.lumcell, st.lumcell, f"lumcell after scatt {it}")
        assert_psel_av(eng, st, scene, f"after scatt {it}")
    # scattered-light i_obs from scaspe_tot
    it += 1
    orc.prepare_scatt(it, st); eng.rt_prepare_scatt(it, tot_g)
    c_o = orc.run_iobs(abi.RTT_I_OBS, st)
    c_g = eng.rt_loop_iobs(abi.RTT_I_OBS, gst)
    assert_counters_equal(c_g, c_o, keys=("sources", "rays", "intersections", "isect_hash"))
    assert_fields_close(gst.i_obs, st.i_obs, "i_obs after rtt_i_obs")
    lam_si = np.asarray(scene.lambda_um) * 1e-6
    for units in (0, 1, 2):     # N1: calc_sed / sum_i_obs (sed_routines.f90:90-246) in the three i_obs unit systems
        assert_fields_close(eng.calc_sed(1.0e6, units, lam_si), orc.calc_sed(st, 1.0e6, units, lam_si), f"SED units {units}")
    # N2: make_maps / map_projection (visual_routines.f90:59-177, 570-772): external-observer maps of every direction
    for npx, fac, units in (((40, 1.0, 0), (24, 2.0, 2)) if maps else ()):
        assert_fields_close(eng.make_maps(npx, fac, units, lam_si), orc.make_maps(st, npx, fac, units, lam_si),
                            f"maps {npx} px, size factor {fac}, units {units}")
    if len(scene.ccoord_obs):
        assert_fields_close(gst.i_obs_in, st.i_obs_in, "i_obs_in after rtt_i_obs")
        # N2, internal observers: map_in_projection (visual_routines.f90:917-1182) on HEALPix spheres of nside 8 and 64
        for kp, units in (((3, 0), (6, 2)) if maps else ()):
            assert_fields_close(eng.make_maps_in(kp, units, lam_si), orc.make_maps_in(st, kp, units, lam_si),
                                f"internal-observer maps kp_maps {kp}, units {units}")
    eng.close()


def test_trusti_small_full_sequence(small_trusti):
    run_sequence(small_trusti, record=6_000_000, iterations=2)


def test_trusti_tiny_thick(tiny_trusti):
    run_sequence(tiny_trusti, record=4_000_000, iterations=3)


def test_magtar_small_full_sequence(small_magtar):
    """C2: point source on a cell corner (fix_ccoord_p_src), isotropic + kp 0 + kp 1 wavelengths, no wall."""
    run_sequence(small_magtar, record=12_000_000, iterations=2)


def test_galaxy_tiny_full_sequence(tiny_galaxy):
    """C3: emitting cells (rtt_precalc_cell / rtt_dir_cell / rtt_i_obs_dir_cell, quirk Q3), z wall."""
    run_sequence(tiny_galaxy, record=2_000_000, iterations=2)


def test_maps_at_reference_resolution(tiny_galaxy):
    """N2 at the smallest map the reference accepts (npixel_maps = 200, io_routines.f90:3207): 800 x 800 oversampled
    pixels, 8 lines of sight from face-on (theta = 0: the tilted-axis branch of calc_texture_array) to edge-on."""
    from dartray_b200 import abi
    eng = _engine(tiny_galaxy)
    orc, st = _oracle(tiny_galaxy)
    from oracle import oracle as O
    gst = O.State(tiny_galaxy)
    for rt in (abi.RTT_PRECALC_CELL, abi.RTT_DIR_CELL):
        orc.run_pass(rt, 0, st); eng.rt_loop(rt)
    orc.run_iobs(abi.RTT_I_OBS_DIR_CELL, st)
    eng.rt_loop_iobs(abi.RTT_I_OBS_DIR_CELL, gst)
    m_g = eng.make_maps(200, 1.0, 0)
    m_o = orc.make_maps(st, 200, 1.0, 0)
    assert m_g.shape == (tiny_galaxy.ndir, tiny_galaxy.lnum, 200, 200) and m_o.max() > 0
    assert_fields_close(m_g, m_o, "maps 200 px")
    eng.close()


def test_galaxy_small_counters(small_galaxy):
    """C3 at 3 levels: 2909 sources x 48 sectors; counters + order-independent intersection checksum
    (1.1e8 intersections: the record list is not kept)."""
    run_sequence(small_galaxy, record=0, iterations=1)


def test_kernel_exp_within_one_ulp(tiny_trusti):
    """The ray-march kernel's exp (constant-bank coefficients) against glibc/numpy exp: <= 1 ulp over the
    attenuation range, exact at 0, and finite / correct beyond the fast range."""
    eng = _engine(tiny_trusti)
    rng = np.random.default_rng(11)
    x = -np.concatenate([[0.0], 10.0 ** rng.uniform(-12, 2.8, 200000), [707.9, 708.0, 709.5, 745.0, 800.0]])
    y = eng.eval_exp(x)
    ref = np.exp(x)
    assert y[0] == 1.0
    big = ref > 1e-300
    ulp = np.abs(y[big] - ref[big]) / np.spacing(ref[big])
    assert ulp.max() <= 1.0, ulp.max()
    assert np.allclose(y[~big], ref[~big], rtol=1e-10, atol=5e-324)
    eng.close()


@pytest.mark.parametrize("kind", ["trusti", "galaxy"])
def test_two_gpus_match_one(kind):
    """(e) multi-GPU: 2 ranks over NCCL (units sharded cyclically, dre_allreduce = rt_mpi) against one GPU."""
    import os, subprocess, sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29517", os.path.join(root, "tests", "multigpu_check.py"), kind],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "multigpu_check ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_internal_observer_iobs(tiny_galaxy):
    """C4 (2DTO3D example: rt_algorithm 'i_obs' with an internal observer at (8000, 0, 0)): rt_loop_iobs rays
    towards internal observers (find_theta_phi_obs_in rt_routines.f90:3164-3217, ray_type_i_obs_in :1762-1769),
    direct and scattered light, on the galaxy-shaped grid."""
    import copy
    sc = copy.copy(tiny_galaxy)
    sc.ccoord_obs = np.array([[8000.0, 0.0, 0.0], [-3000.0, 5000.0, 700.0]])
    run_sequence(sc, record=0, iterations=1)


def run_iobs_only(sc, kp_maps, units=2):
    """rt_algorithm 'i_obs' with only_direct_rt (dartray_hub.f90:128-146): rt_loop_iobs for the point sources and the emitting cells,
    then the internal-observer maps.  Engine vs oracle."""
    from dartray_b200 import abi
    from oracle import oracle as O
    eng = _engine(sc)
    orc, st = _oracle(sc)
    gst = O.State(sc)
    c_o = orc.run_iobs(abi.RTT_I_OBS_DIR_CELL, st)
    c_g = eng.rt_loop_iobs(abi.RTT_I_OBS_DIR_CELL, gst)
    assert_counters_equal(c_g, c_o, keys=("sources", "rays", "intersections", "isect_hash"))
    assert c_o.rays == int((sc.lumcell.sum(axis=1) > 0).sum()) * len(sc.ccoord_obs)
    assert_fields_close(gst.i_obs_in, st.i_obs_in, "i_obs_in")
    lam_si = np.asarray(sc.lambda_um) * 1e-6
    m_g, m_o = eng.make_maps_in(kp_maps, units, lam_si), orc.make_maps_in(st, kp_maps, units, lam_si)
    assert m_o.max() > 0
    assert_fields_close(m_g, m_o, f"internal-observer maps kp_maps {kp_maps}")
    eng.close()


def test_c4_2dto3d_internal_observer():
    """C4 = 2DTO3D_GRIDS/input_2dto3d_example.in on the shipped 2-D tables (9 wavelengths 12-850 um, observer at (8000, 0, 0) pc),
    one level coarser than the example (max_lvl 3: 16 066 cells)."""
    from dartray_b200 import scenes
    sc = scenes.twod3d_scene(max_lvl=3)
    assert sc.lnum == 9 and sc.ndir == 0 and len(sc.ccoord_obs) == 1
    run_iobs_only(sc, kp_maps=6)


def test_c5_nbody_small_sequence():
    """C5 = NBODY_SPH_GRIDS/input_nbody_example1.in on the seeded synthetic particle set, reduced (max_lvl 3, 4 wavelengths,
    2 341 emitting leaves): the whole sequence of dartray_main, every leaf with particles emitting."""
    from dartray_b200 import scenes
    sc = scenes.nbody_scene(max_lvl=3, lnum=4, max_dlum=1e-3)
    run_sequence(sc, record=0, iterations=1)
