"""ctypes binding of the CPU oracle (oracle/dartray_oracle.cpp).

TEST INFRASTRUCTURE: import only from tests/, __graft_entry__.smoke() and bench.py's CPU legs.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from dartray_b200 import abi  # noqa: E402

_LIB = os.path.join(_HERE, "_build", "liboracle.so")


def build(force=False):
    src = os.path.join(_HERE, "dartray_oracle.cpp")
    if force or not os.path.exists(_LIB) or (os.path.exists(src) and os.path.getmtime(_LIB) < os.path.getmtime(src)):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_create.restype = C.c_void_p
        _lib.orc_last_error.restype = C.c_char_p
        for n in ("orc_destroy", "orc_last_error", "orc_set_grid", "orc_set_opacity", "orc_set_sources", "orc_set_params",
                  "orc_set_record", "orc_get_records", "orc_run_pass", "orc_run_iobs", "orc_prepare_scatt",
                  "orc_get_lumcell", "orc_cindex_roundtrip", "orc_locate_points"):
            getattr(_lib, n).argtypes = None
    return _lib


class State:
    """Host copies of the accumulators of the path (u_fest_arr, u_final_arr, scaspe_arr, ...)."""

    def __init__(self, scene):
        N, L = scene.grid.ncell, scene.lnum
        self.u_fest = np.zeros((N, L))
        self.u_final = np.zeros((N, L))
        self.scaspe = [np.zeros((N, int(p))) for p in scene.npix]
        self.scaspe_prev = [np.zeros((N, int(p))) for p in scene.npix]
        self.scaspe_tot = [np.zeros((N, int(p))) for p in scene.npix]
        self.lum_lost = np.zeros(L)
        self.lum_lost_prev = np.zeros(L)
        self.psel_sum = np.zeros(N + scene.nsrc)
        self.ipsel = np.zeros(N + scene.nsrc, dtype=np.int64)
        self.tot_rad_en = np.zeros(L)
        self.i_obs = np.zeros((scene.ndir, N + scene.nsrc, L))
        nin = scene.ccoord_obs.shape[0]
        self.i_obs_in = np.zeros((max(nin, 1), N + scene.nsrc, L))
        self.lumcell = scene.lumcell.copy()


class Oracle:
    def __init__(self, scene, nthreads=1):
        self.scene = scene
        self.nthreads = nthreads
        self.l = lib()
        self.h = C.c_void_p(self.l.orc_create())
        g = scene.grid
        base = (C.c_int32 * 2)(*g.base)
        self._keep = [g.cchild, g.lvl, g.cindex, g.ccoord, g.csize]
        self.l.orc_set_grid(self.h, C.c_int32(g.ncell), C.c_int32(g.max_lvl), base, abi.iptr(g.cchild), abi.iptr(g.lvl),
                            abi.lptr(g.cindex), abi.dptr(g.ccoord), abi.dptr(g.csize), abi.dptr(g.csize_arr),
                            abi.dptr(g.carea_arr), abi.dptr(g.cvol_arr))
        self.l.orc_set_opacity(self.h, C.c_int32(scene.lnum), abi.dptr(scene.dens_arr), abi.dptr(scene.ksca_norm),
                               abi.dptr(scene.gsca), abi.iptr(scene.kp_sca), C.c_int32(scene.ndir), abi.dptr(scene.dir_i_out))
        self.set_sources(scene.lumcell)

    def __del__(self):
        try:
            self.l.orc_destroy(self.h)
        except Exception:
            pass

    def set_sources(self, lumcell):
        s = self.scene
        lumcell = np.ascontiguousarray(lumcell)
        self.l.orc_set_sources(self.h, abi.dptr(lumcell), C.c_int32(s.nsrc), abi.dptr(s.ccoord_p_src), abi.dptr(s.lum_p_src),
                               abi.iptr(s.src_cell), abi.iptr(s.cell_src))

    def set_record(self, n):
        self.l.orc_set_record(self.h, C.c_int64(n))

    def set_option(self, name, value):
        """"lock_cell" = 1: the reference's non-atomic vectorised update when one ray is in the cell (rt_routines.f90:2444,
        2497, 2649).  Timing legs only: it can lose an update, so the parity checks keep it off."""
        if self.l.orc_set_option(self.h, name.encode(), C.c_double(value)) != 0:
            raise ValueError(name)

    def records(self):
        n = C.c_int64(0)
        self.l.orc_get_records(self.h, None, C.c_int64(0), C.byref(n))
        out = np.zeros(n.value, dtype=abi.ISECT_DTYPE)
        if n.value:
            self.l.orc_get_records(self.h, out.ctypes.data_as(C.c_void_p), C.c_int64(n.value), C.byref(n))
        return out

    def run_pass(self, rt_type, iterations, st, rank=0, nranks=1, nthreads=None):
        p = abi.make_params(self.scene, rt_type)
        self.l.orc_set_params(self.h, C.byref(p))
        cnt = abi.Counters()
        precalc = rt_type in (abi.RTT_PRECALC_CELL, abi.RTT_PRECALC_SRC)
        accum = st.u_fest if precalc else st.u_final
        rc = self.l.orc_run_pass(self.h, C.c_int(rt_type), C.c_int(iterations), abi.dptr(st.u_fest), abi.dptr(accum),
                                 abi.plane_ptrs(st.scaspe), abi.plane_ptrs(st.scaspe_prev), abi.dptr(st.lum_lost),
                                 abi.dptr(st.psel_sum), abi.lptr(st.ipsel), C.c_int(nthreads or self.nthreads),
                                 C.c_int(rank), C.c_int(nranks), C.byref(cnt))
        if rc != 0:
            raise RuntimeError(self.l.orc_last_error(self.h).decode())
        return cnt

    # ---- N5: 2D mode -----------------------------------------------------------------------------------------------
    def fix_symmetry(self, rt_type, st):
        """fix_symmetry (rt_routines.f90:3605-3635) on the state arrays after the first loop of rt_loop_2D."""
        rc = self.l.orc_fix_symmetry(self.h, C.c_int(rt_type), abi.dptr(st.u_fest), abi.dptr(st.u_final), abi.plane_ptrs(st.scaspe),
                                     abi.dptr(st.lum_lost), abi.dptr(st.lum_lost_prev))
        if rc != 0:
            raise RuntimeError(self.l.orc_last_error(self.h).decode())

    def run_pass_2d(self, rt_type, iterations, st):
        """rt_loop_2D (rt_routines.f90:1061-1140): positive-octant sources, fix_symmetry, then the sources on the Cartesian planes.
        Returns the two pass counters."""
        out = []
        try:
            for i2d in (0, 1):
                self.set_option("i2d", i2d)
                out.append(self.run_pass(rt_type, iterations, st))
                if i2d == 0:
                    self.fix_symmetry(rt_type, st)
        finally:
            self.set_option("i2d", -1)
        return out

    def linked_cells(self, nc):
        out = np.zeros(7, dtype=np.int32)
        if self.l.orc_linked_cells(self.h, C.c_int32(int(nc)), abi.iptr(out)) != 0:
            raise RuntimeError("find_linked_cells failed")
        return out

    def scaspe_symmetry_index(self, kp, which):
        out = np.zeros(12 * 4 ** kp, dtype=np.int32)
        self.l.orc_scaspe_symmetry_index(C.c_int(kp), C.c_int(which), abi.iptr(out))
        return out

    def run_iobs(self, rt_type, st, nthreads=None):
        p = abi.make_params(self.scene, rt_type)
        self.l.orc_set_params(self.h, C.byref(p))
        cnt = abi.Counters()
        obs = np.ascontiguousarray(self.scene.ccoord_obs, dtype=np.float64)
        rc = self.l.orc_run_iobs(self.h, C.c_int(rt_type), abi.plane_ptrs(st.scaspe_tot), C.c_int32(obs.shape[0]),
                                 abi.dptr(obs), abi.dptr(st.i_obs), abi.dptr(st.i_obs_in),
                                 C.c_int(nthreads or self.nthreads), C.byref(cnt))
        if rc != 0:
            raise RuntimeError(self.l.orc_last_error(self.h).decode())
        return cnt

    def calc_sed(self, st, dist_obs, units, lambda_si=None):
        """calc_sed (sed_routines.f90:90) on st.i_obs; returns (ndir, L)."""
        sed = np.zeros((self.scene.ndir, self.scene.lnum))
        lam = np.ascontiguousarray(lambda_si if lambda_si is not None else np.zeros(self.scene.lnum), dtype=np.float64)
        self.l.orc_calc_sed(self.h, C.c_int32(self.scene.ndir), abi.dptr(st.i_obs), C.c_double(dist_obs), C.c_int(units),
                            abi.dptr(lam), abi.dptr(sed))
        return sed

    def make_maps(self, st, npixel_maps, map_size_factor, units, lambda_si=None):
        """make_maps / map_projection (visual_routines.f90:59, 570) on st.i_obs; returns (ndir, L, npix, npix) [MJy/sr],
        the memory image of map_arr_out(0:npix-1, 0:npix-1, 0:L-1, 0:ndir-1) (last index = X' pixel)."""
        sc = self.scene
        par = abi.make_params(sc, abi.RTT_I_OBS)          # modelsize
        self.l.orc_set_params(self.h, C.byref(par))
        maps = np.zeros((sc.ndir, sc.lnum, npixel_maps, npixel_maps))
        lam = np.ascontiguousarray(lambda_si if lambda_si is not None else np.zeros(sc.lnum), dtype=np.float64)
        self.l.orc_make_maps(self.h, C.c_int32(sc.ndir), abi.dptr(st.i_obs), C.c_int(npixel_maps), C.c_double(map_size_factor),
                             C.c_int(units), abi.dptr(lam), abi.dptr(maps))
        return maps

    def make_maps_in(self, st, kp_maps, units, lambda_si=None):
        """map_in_projection (visual_routines.f90:917) on st.i_obs_in for every internal observer; returns
        (ndir_in, L, 12*4**kp_maps) [MJy/sr] = the memory image of map_in_arr_out(pixel, lambda, observer)."""
        sc = self.scene
        obs = np.ascontiguousarray(sc.ccoord_obs, dtype=np.float64)
        npix = 12 * 4 ** kp_maps
        maps = np.zeros((obs.shape[0], sc.lnum, npix))
        lam = np.ascontiguousarray(lambda_si if lambda_si is not None else np.zeros(sc.lnum), dtype=np.float64)
        self.l.orc_make_maps_in(self.h, C.c_int32(obs.shape[0]), abi.dptr(obs), abi.dptr(st.i_obs_in), C.c_int(kp_maps),
                                C.c_int(units), abi.dptr(lam), abi.dptr(maps))
        return maps

    def neighbours(self, nside, ipix):
        out = np.zeros(8, dtype=np.int32)
        n = self.l.orc_neighbours(C.c_int(nside), C.c_int(int(ipix)), out.ctypes.data_as(C.c_void_p))
        return out[:n].copy()

    def map_texture(self, npixel_maps, map_size_factor, theta, phi, level):
        """calc_texture_array (visual_routines.f90:314): (prof2d[iy, ix], im0, im1, axes(3,3) = obs_vec, xi_dir, yi_dir)."""
        par = abi.make_params(self.scene, abi.RTT_I_OBS)   # modelsize
        self.l.orc_set_params(self.h, C.byref(par))
        axes = np.zeros((3, 3)); info = np.zeros(2, dtype=np.int32)
        self.l.orc_map_texture.restype = C.c_int
        nt = self.l.orc_map_texture(self.h, C.c_int(npixel_maps), C.c_double(map_size_factor), C.c_double(theta), C.c_double(phi),
                                    C.c_int(level), None, C.c_int(0), info.ctypes.data_as(C.c_void_p), abi.dptr(axes))
        prof = np.zeros((nt, nt))
        self.l.orc_map_texture(self.h, C.c_int(npixel_maps), C.c_double(map_size_factor), C.c_double(theta), C.c_double(phi),
                               C.c_int(level), abi.dptr(prof), C.c_int(nt * nt), info.ctypes.data_as(C.c_void_p), abi.dptr(axes))
        return prof, int(info[0]), int(info[1]), axes

    def dust_emission(self, kabs_stars, kabs_dust, lambda_si_dust, dl_stars, dl_dust, kext_ref, dens_ref, rf_stars, rf_dust=None,
                      iterations_dustem=0):
        """set_dust_emission 'eff' (sed_routines.f90:249-276, 377-430): (dens_stars (N, lnum_dust), t_dust (N))."""
        f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        ks, kd, lam, dls, dld, ref, rfs = map(f, (kabs_stars, kabs_dust, lambda_si_dust, dl_stars, dl_dust, dens_ref, rf_stars))
        rfd = None if rf_dust is None else f(rf_dust)
        N = self.scene.grid.ncell
        out = np.zeros((N, len(kd))); t = np.zeros(N)
        rc = self.l.orc_dust_emission(self.h, C.c_int32(len(ks)), C.c_int32(len(kd)), abi.dptr(ks), abi.dptr(kd), abi.dptr(lam), abi.dptr(dls),
                                      abi.dptr(dld), C.c_double(kext_ref), abi.dptr(ref), abi.dptr(rfs),
                                      abi.dptr(rfd) if rfd is not None else None, C.c_int32(iterations_dustem), abi.dptr(out), abi.dptr(t))
        if rc != 0:
            raise RuntimeError(self.l.orc_last_error(self.h).decode())
        return out, t

    def prepare_scatt(self, iterations, st):
        flag = C.c_int32(0)
        self.l.orc_prepare_scatt(self.h, C.c_int(iterations), C.c_double(self.scene.conv_en_lim), abi.dptr(st.u_fest),
                                 abi.dptr(st.u_final), abi.dptr(st.lum_lost_prev), abi.dptr(st.lum_lost),
                                 abi.plane_ptrs(st.scaspe), abi.plane_ptrs(st.scaspe_prev), abi.plane_ptrs(st.scaspe_tot),
                                 abi.dptr(st.tot_rad_en), C.byref(flag))
        self.l.orc_get_lumcell(self.h, abi.dptr(st.lumcell))
        return bool(flag.value)

    def pix2ang(self, nside, ipix):
        ipix = np.ascontiguousarray(ipix, dtype=np.int32)
        th = np.zeros(len(ipix)); ph = np.zeros(len(ipix))
        self.l.orc_pix2ang_nest(C.c_int(nside), C.c_int(len(ipix)), abi.iptr(ipix), abi.dptr(th), abi.dptr(ph))
        return th, ph

    def ang2pix(self, nside, th, ph):
        th = np.ascontiguousarray(th); ph = np.ascontiguousarray(ph)
        out = np.zeros(len(th), dtype=np.int32)
        self.l.orc_ang2pix_nest(C.c_int(nside), C.c_int(len(th)), abi.dptr(th), abi.dptr(ph), abi.iptr(out))
        return out

    def pix2vec(self, nside, ipix):
        ipix = np.ascontiguousarray(ipix, dtype=np.int32)
        v = np.zeros((len(ipix), 3))
        self.l.orc_pix2vec(C.c_int(nside), C.c_int(len(ipix)), abi.iptr(ipix), abi.dptr(v))
        return v

    def locate(self, xyz, cc_old):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64); cc_old = np.ascontiguousarray(cc_old, dtype=np.int32)
        out = np.zeros(len(cc_old), dtype=np.int32)
        p = abi.make_params(self.scene, abi.RTT_DIR_CELL)
        self.l.orc_set_params(self.h, C.byref(p))
        self.l.orc_locate_points(self.h, C.c_int32(len(cc_old)), abi.dptr(xyz), abi.iptr(cc_old), abi.iptr(out))
        return out

    def cindex_roundtrip(self, cell):
        a = C.c_int32(0); b = C.c_int32(0)
        self.l.orc_cindex_roundtrip(self.h, C.c_int32(int(cell)), C.byref(a), C.byref(b))
        return a.value, b.value
