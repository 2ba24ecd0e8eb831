"""The C ABI called from a compiled C host with Fortran-shaped arguments (tests/abi_host.c, VERDICT r01 item 5 iv): column-major
arrays, an array of plane pointers for scaspe, dre_params by reference; precalc + direct light + rt_prepare(rtt_scatt) on the host
arrays + scattering iteration 1 through the drop-in call dre_run_pass.  No Python between the host program and the library; the
results it writes are compared with the CPU oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from test_abi import build_abi_host
from test_gpu_parity import assert_fields_close

pytestmark = pytest.mark.gpu


def dump_scene(sc, d):
    from dartray_b200 import abi
    g = sc.grid
    has_cell = int(bool(np.any(sc.lumcell)))
    np.array([g.ncell, g.max_lvl, g.base[0], g.base[1], sc.lnum, sc.ndir, sc.nsrc, has_cell], dtype=np.int32).tofile(f"{d}/header.i32")
    np.array([sc.conv_en_lim], dtype=np.float64).tofile(f"{d}/conv_en_lim.f64")
    for name, arr, dt in (("cchild.i32", g.cchild, np.int32), ("lvl.i32", g.lvl, np.int32), ("cindex.i64", g.cindex, np.int64),
                          ("ccoord.f64", g.ccoord, np.float64), ("csize.f64", g.csize, np.float64),
                          ("csize_arr.f64", g.csize_arr, np.float64), ("carea_arr.f64", g.carea_arr, np.float64),
                          ("cvol_arr.f64", g.cvol_arr, np.float64), ("dens_arr.f64", sc.dens_arr, np.float64),
                          ("ksca_norm.f64", sc.ksca_norm, np.float64), ("gsca.f64", sc.gsca, np.float64),
                          ("kp_sca.i32", sc.kp_sca, np.int32), ("npix.i32", sc.npix, np.int32), ("npix_hp.i32", sc.npix_hp, np.int32),
                          ("dir_i_out.f64", sc.dir_i_out, np.float64), ("lumcell.f64", sc.lumcell, np.float64),
                          ("ccoord_p_src.f64", sc.ccoord_p_src, np.float64), ("lum_p_src.f64", sc.lum_p_src, np.float64),
                          ("src_cell.i32", sc.src_cell, np.int32), ("cell_src.i32", sc.cell_src, np.int32)):
        np.ascontiguousarray(arr, dtype=dt).tofile(f"{d}/{name}")
    for fn, rt in (("params_direct.bin", abi.RTT_DIR_SRC), ("params_scatt.bin", abi.RTT_SCATT)):
        with open(f"{d}/{fn}", "wb") as f:
            f.write(bytes(abi.make_params(sc, rt)))


@pytest.mark.parametrize("kind", ["trusti", "galaxy"])
def test_c_host_matches_oracle(kind, tiny_trusti, tiny_galaxy, tmp_path):
    from dartray_b200 import abi
    from oracle import oracle as O
    sc = tiny_trusti if kind == "trusti" else tiny_galaxy
    exe = build_abi_host(tmp_path)
    dump_scene(sc, tmp_path)
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "abi_host ok" in r.stdout, r.stdout[-1000:] + r.stderr[-2000:]
    N, L = sc.grid.ncell, sc.lnum

    def out(name, shape):
        return np.fromfile(f"{tmp_path}/{name}", dtype=np.float64).reshape(shape)

    orc, st = O.Oracle(sc, nthreads=8), O.State(sc)
    has_src, has_cell = sc.nsrc > 0, bool(np.any(sc.lumcell))
    isect = 0
    for rt in ([abi.RTT_PRECALC_CELL] if has_cell else []) + ([abi.RTT_PRECALC_SRC] if has_src else []):
        isect += orc.run_pass(rt, 0, st).intersections
    assert_fields_close(out("u_fest.out", (N, L)), st.u_fest, "u_fest (C host)")
    for rt in ([abi.RTT_DIR_CELL] if has_cell else []) + ([abi.RTT_DIR_SRC] if has_src else []):
        isect += orc.run_pass(rt, 0, st).intersections
    assert_fields_close(out("u_final_direct.out", (N, L)), st.u_final, "u_final after the direct light (C host)")
    want = np.where(st.ipsel > 0, st.psel_sum / np.maximum(st.ipsel, 1), 0.0)
    assert_fields_close(out("psel_av_direct.out", (N + sc.nsrc,)), want, "psel_av of the direct light (C host)")
    assert not orc.prepare_scatt(1, st)
    assert_fields_close(out("lumcell.out", (N, L)), st.lumcell, "lumcell from the host-side rt_prepare (C host)")
    orc.set_sources(st.lumcell)
    st.psel_sum[...] = 0; st.ipsel[...] = 0
    isect += orc.run_pass(abi.RTT_SCATT, 1, st).intersections
    assert f"abi_host ok: {isect} intersections" in r.stdout, r.stdout
    assert_fields_close(out("u_final.out", (N, L)), st.u_final, "u_final after scattering iteration 1 (C host)")
    assert_fields_close(out("lum_lost.out", (L,)), st.lum_lost, "lum_lost (C host)")
    for l in range(L):
        assert_fields_close(out(f"scaspe_{l}.out", (N, int(sc.npix[l]))), st.scaspe[l], f"scaspe[{l}] (C host)")
    want = np.where(st.ipsel > 0, st.psel_sum / np.maximum(st.ipsel, 1), 0.0)
    assert_fields_close(out("psel_av.out", (N + sc.nsrc,)), want, "psel_av of scattering iteration 1 (C host)")
