"""ctypes binding of the engine's C ABI (include/dartray_engine.h) and a host driver that
mirrors the call sequence of dartray_main (dartray_hub.f90:12-67).

The procedures keep the reference names: rt_loop (rt_routines.f90:985), rt_loop_iobs (:1470),
rt_mpi (io_routines.f90:3667), rt_prepare(rtt_scatt) (dartray_hub.f90:432-457).
There is no CPU path: if the CUDA library is missing or no GPU is present the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DARTRAY_ENGINE_LIB", os.path.join(_HERE, "libdartray_engine.so"))  # override: A/B builds

SYMBOLS = [
    "dre_get_unique_id", "dre_create", "dre_destroy", "dre_last_error", "dre_set_grid", "dre_set_opacity",
    "dre_set_sources", "dre_set_params", "dre_upload", "dre_download", "dre_zero", "dre_run_pass_resident",
    "dre_run_pass", "dre_run_pass_2d", "dre_scaspe_symmetry_index", "dre_run_iobs", "dre_allreduce", "dre_prepare_scatt", "dre_dust_emission", "dre_calc_sed", "dre_make_maps", "dre_make_maps_in", "dre_build_source_list",
    "dre_locate_points", "dre_locate_points_noalloc", "dre_pix2vec", "dre_set_record", "dre_get_records", "dre_set_option", "dre_get_quanta", "dre_eval_exp", "dre_mark", "dre_elapsed_ms", "dre_launch_count",
]


def build(force=False):
    """Compile the CUDA engine for sm_100a (nvcc cross-compiles without a GPU)."""
    src_dir = os.path.join(_HERE, "csrc")
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    subprocess.check_call(["make", "-C", src_dir, "-s"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(the engine has no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.dre_last_error.restype = C.c_char_p
        _lib.dre_last_error.argtypes = [C.c_void_p]
        _lib.dre_destroy.argtypes = [C.c_void_p]
        _lib.dre_destroy.restype = None
    return _lib


class EngineError(RuntimeError):
    pass


class Engine:
    """One engine per GPU / rank (the Fortran shim would hold one per MPI rank)."""

    def __init__(self, scene, device=0, rank=0, nranks=1, nccl_id=None):
        self.scene = scene
        self.l = lib()
        h = C.c_void_p()
        idbuf = None
        if nccl_id is not None:
            idbuf = (C.c_char * 128).from_buffer_copy(bytes(nccl_id))
        rc = self.l.dre_create(C.byref(h), C.c_int(device), C.c_int(rank), C.c_int(nranks), idbuf)
        if rc != 0:
            raise EngineError(f"dre_create failed (rc={rc}): a CUDA device is required")
        self.h = h
        self.rank, self.nranks = rank, nranks
        g = scene.grid
        base = (C.c_int32 * 2)(*g.base)
        self._keep = [np.ascontiguousarray(g.cchild, np.int32), np.ascontiguousarray(g.lvl, np.int32),
                      np.ascontiguousarray(g.cindex, np.int64), np.ascontiguousarray(g.ccoord, np.float64),
                      np.ascontiguousarray(g.csize, np.float64)]
        k = self._keep
        self._ck(self.l.dre_set_grid(self.h, C.c_int32(g.ncell), C.c_int32(g.max_lvl), base, abi.iptr(k[0]), abi.iptr(k[1]),
                                     abi.lptr(k[2]), abi.dptr(k[3]), abi.dptr(k[4]), abi.dptr(g.csize_arr),
                                     abi.dptr(g.carea_arr), abi.dptr(g.cvol_arr)))
        self._ck(self.l.dre_set_opacity(self.h, C.c_int32(scene.lnum), abi.dptr(scene.dens_arr), abi.dptr(scene.ksca_norm),
                                        abi.dptr(scene.gsca), abi.iptr(scene.kp_sca), C.c_int32(scene.ndir),
                                        abi.dptr(scene.dir_i_out)))
        self.set_sources(scene.lumcell)

    @staticmethod
    def get_unique_id():
        buf = (C.c_char * 128)()
        rc = lib().dre_get_unique_id(buf)
        if rc != 0:
            raise EngineError("dre_get_unique_id failed (NCCL not loadable)")
        return bytes(buf)

    def close(self):
        if getattr(self, "h", None):
            self.l.dre_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise EngineError(self.l.dre_last_error(self.h).decode())

    # ---- configuration ------------------------------------------------------------------
    def set_sources(self, lumcell):
        s = self.scene
        lumcell = np.ascontiguousarray(lumcell, np.float64)
        self._ck(self.l.dre_set_sources(self.h, abi.dptr(lumcell), C.c_int32(s.nsrc), abi.dptr(s.ccoord_p_src),
                                        abi.dptr(s.lum_p_src), abi.iptr(s.src_cell), abi.iptr(s.cell_src)))

    def set_params(self, rt_type):
        p = abi.make_params(self.scene, rt_type)
        self._ck(self.l.dre_set_params(self.h, C.byref(p)))

    def set_option(self, name, value):
        self._ck(self.l.dre_set_option(self.h, name.encode(), C.c_double(value)))

    def quanta(self):
        """Rounding grids of the last pass run with option "deterministic": (q_scaspe[L], q_u[L], q_lum_lost[L])."""
        L = self.scene.lnum
        q = np.zeros(3 * L)
        self._ck(self.l.dre_get_quanta(self.h, abi.dptr(q)))
        return q[:L].copy(), q[L:2 * L].copy(), q[2 * L:].copy()

    def mark(self, slot):
        self._ck(self.l.dre_mark(self.h, C.c_int(slot)))

    def elapsed_ms(self, a, b):
        ms = C.c_double(0)
        self._ck(self.l.dre_elapsed_ms(self.h, C.c_int(a), C.c_int(b), C.byref(ms)))
        return ms.value

    def launch_count(self):
        self.l.dre_launch_count.restype = C.c_int64
        self.l.dre_launch_count.argtypes = [C.c_void_p]
        return int(self.l.dre_launch_count(self.h))

    def set_record(self, n):
        self._ck(self.l.dre_set_record(self.h, C.c_int64(n)))

    def records(self):
        n = C.c_int64(0)
        self._ck(self.l.dre_get_records(self.h, None, C.c_int64(0), C.byref(n)))
        out = np.zeros(n.value, dtype=abi.ISECT_DTYPE)
        if n.value:
            self._ck(self.l.dre_get_records(self.h, out.ctypes.data_as(C.c_void_p), C.c_int64(n.value), C.byref(n)))
        return out

    # ---- device-resident state ------------------------------------------------------------
    def upload(self, which, arr):
        if isinstance(arr, (list, tuple)):
            self._ck(self.l.dre_upload(self.h, C.c_int(which), abi.plane_ptrs(arr)))
        else:
            self._ck(self.l.dre_upload(self.h, C.c_int(which), abi.dptr(arr)))

    def download(self, which, arr):
        if isinstance(arr, (list, tuple)):
            self._ck(self.l.dre_download(self.h, C.c_int(which), abi.plane_ptrs(arr)))
        else:
            self._ck(self.l.dre_download(self.h, C.c_int(which), arr.ctypes.data_as(C.c_void_p)))
        return arr

    def zero(self, which):
        self._ck(self.l.dre_zero(self.h, C.c_int(which)))

    def upload_state(self, st):
        self.upload(abi.ARR_U_FEST, st.u_fest)
        self.upload(abi.ARR_U_FINAL, st.u_final)
        self.upload(abi.ARR_SCASPE, st.scaspe)
        self.upload(abi.ARR_SCASPE_PREV, st.scaspe_prev)
        self.upload(abi.ARR_SCASPE_TOT, st.scaspe_tot)
        self.upload(abi.ARR_LUM_LOST, st.lum_lost)

    def download_state(self, st):
        self.download(abi.ARR_U_FEST, st.u_fest)
        self.download(abi.ARR_U_FINAL, st.u_final)
        self.download(abi.ARR_SCASPE, st.scaspe)
        self.download(abi.ARR_SCASPE_PREV, st.scaspe_prev)
        self.download(abi.ARR_SCASPE_TOT, st.scaspe_tot)
        self.download(abi.ARR_LUM_LOST, st.lum_lost)
        self.download(abi.ARR_LUMCELL, st.lumcell)

    # ---- passes ---------------------------------------------------------------------------
    def rt_loop(self, rt_type, iterations=0):
        """rt_loop on device-resident state (rt_routines.f90:985)."""
        self.set_params(rt_type)
        cnt = abi.Counters()
        self._ck(self.l.dre_run_pass_resident(self.h, C.c_int(rt_type), C.c_int(iterations), C.byref(cnt)))
        return cnt

    def rt_loop_2d(self, rt_type, iterations=0):
        """rt_loop_2D (rt_routines.f90:1061-1140) on device-resident state: positive-octant sources, fix_symmetry, plane sources."""
        self.set_params(rt_type)
        cnt = abi.Counters()
        self._ck(self.l.dre_run_pass_2d(self.h, C.c_int(rt_type), C.c_int(iterations), C.byref(cnt)))
        return cnt

    def scaspe_symmetry_index(self, kp, which):
        out = np.zeros(12 * 4 ** kp, dtype=np.int32)
        self._ck(self.l.dre_scaspe_symmetry_index(self.h, C.c_int32(kp), C.c_int32(which), abi.iptr(out)))
        return out

    def rt_loop_host(self, rt_type, iterations, st):
        """rt_loop with HOST arrays (upload -> pass -> download): the drop-in call."""
        self.set_params(rt_type)
        cnt = abi.Counters()
        precalc = rt_type in (abi.RTT_PRECALC_CELL, abi.RTT_PRECALC_SRC)
        accum = st.u_fest if precalc else st.u_final
        psel = getattr(st, "psel_av", None)
        self._ck(self.l.dre_run_pass(self.h, C.c_int(rt_type), C.c_int(iterations), abi.dptr(st.u_fest), abi.dptr(accum),
                                     abi.plane_ptrs(st.scaspe), abi.plane_ptrs(st.scaspe_prev), abi.dptr(st.lum_lost),
                                     abi.dptr(psel) if psel is not None else None, C.byref(cnt)))
        return cnt

    def rt_loop_iobs(self, rt_type, st=None, scaspe_tot=None):
        """rt_loop_iobs (rt_routines.f90:1470); `st` given: i_obs arrays are read and written back."""
        self.set_params(rt_type)
        cnt = abi.Counters()
        obs = np.ascontiguousarray(self.scene.ccoord_obs, dtype=np.float64)
        nin = obs.shape[0]
        self._ck(self.l.dre_run_iobs(self.h, C.c_int(rt_type), abi.plane_ptrs(scaspe_tot) if scaspe_tot is not None else None,
                                     C.c_int32(nin), abi.dptr(obs) if nin else None,
                                     abi.dptr(st.i_obs) if st is not None else None,
                                     abi.dptr(st.i_obs_in) if (st is not None and nin) else None, C.byref(cnt)))
        return cnt

    def calc_sed(self, dist_obs, units, lambda_si=None):
        """calc_sed (sed_routines.f90:90) from the device-resident i_obs array; returns (ndir, L) [Jy]."""
        sed = np.zeros((self.scene.ndir, self.scene.lnum))
        lam = None if lambda_si is None else np.ascontiguousarray(lambda_si, dtype=np.float64)
        self._ck(self.l.dre_calc_sed(self.h, C.c_double(dist_obs), C.c_int(units), abi.dptr(lam) if lam is not None else None,
                                     abi.dptr(sed)))
        return sed

    def make_maps(self, npixel_maps, map_size_factor, units, lambda_si=None):
        """make_maps / map_projection (visual_routines.f90:59, 570) from the device-resident i_obs array; returns
        (ndir, L, npix, npix) [MJy/sr] = the memory image of map_arr_out(x, y, lambda, direction)."""
        maps = np.zeros((self.scene.ndir, self.scene.lnum, npixel_maps, npixel_maps))
        lam = None if lambda_si is None else np.ascontiguousarray(lambda_si, dtype=np.float64)
        self._ck(self.l.dre_make_maps(self.h, C.c_int(npixel_maps), C.c_double(map_size_factor), C.c_int(units),
                                      abi.dptr(lam) if lam is not None else None, abi.dptr(maps)))
        return maps

    def make_maps_in(self, kp_maps, units, lambda_si=None):
        """map_in_projection (visual_routines.f90:917) from the device-resident i_obs_in array; returns
        (ndir_in, L, 12*4**kp_maps) [MJy/sr] = the memory image of map_in_arr_out(pixel, lambda, observer)."""
        obs = np.ascontiguousarray(self.scene.ccoord_obs, dtype=np.float64)
        maps = np.zeros((obs.shape[0], self.scene.lnum, 12 * 4 ** kp_maps))
        lam = None if lambda_si is None else np.ascontiguousarray(lambda_si, dtype=np.float64)
        self._ck(self.l.dre_make_maps_in(self.h, C.c_int32(obs.shape[0]), abi.dptr(obs), C.c_int(kp_maps), C.c_int(units),
                                         abi.dptr(lam) if lam is not None else None, abi.dptr(maps)))
        return maps

    def rt_mpi(self, rt_type):
        self._ck(self.l.dre_allreduce(self.h, C.c_int(rt_type)))

    def rt_prepare_scatt(self, iterations, tot_rad_en):
        flag = C.c_int32(0)
        self._ck(self.l.dre_prepare_scatt(self.h, C.c_int(iterations), C.c_double(self.scene.conv_en_lim),
                                          abi.dptr(tot_rad_en), C.byref(flag)))
        return bool(flag.value)

    def dust_emission(self, kabs_stars, kabs_dust, lambda_si_dust, dl_stars, dl_dust, kext_ref, dens_ref, rf_stars, rf_dust=None,
                      iterations_dustem=0, to_lumcell=False):
        """set_dust_emission, 'eff' mode (sed_routines.f90:249-276, 377-430): returns (dens_stars (N, lnum_dust), t_dust (N)).
        rf_dust None with iterations_dustem > 0: the device-resident u_final of the last dust pass is used."""
        f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        ks, kd, lam, dls, dld, ref, rfs = map(f, (kabs_stars, kabs_dust, lambda_si_dust, dl_stars, dl_dust, dens_ref, rf_stars))
        rfd = None if rf_dust is None else f(rf_dust)
        N = self.scene.grid.ncell
        out = np.zeros((N, len(kd))); t = np.zeros(N)
        self._ck(self.l.dre_dust_emission(self.h, C.c_int32(len(ks)), C.c_int32(len(kd)), abi.dptr(ks), abi.dptr(kd), abi.dptr(lam),
                                          abi.dptr(dls), abi.dptr(dld), C.c_double(kext_ref), abi.dptr(ref), abi.dptr(rfs),
                                          abi.dptr(rfd) if rfd is not None else None, C.c_int32(iterations_dustem), abi.dptr(out),
                                          abi.dptr(t), C.c_int32(1 if to_lumcell else 0)))
        return out, t

    # ---- separate small kernels -----------------------------------------------------------
    def build_source_list(self, rt_type):
        self.set_params(rt_type)
        n = C.c_int32(0)
        out = np.zeros(self.scene.grid.ncell, dtype=np.int32)
        self._ck(self.l.dre_build_source_list(self.h, C.c_int(rt_type), C.byref(n), abi.iptr(out)))
        return out[:n.value].copy()

    def locate(self, xyz, cc_old):
        self.set_params(abi.RTT_DIR_CELL)
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        cc_old = np.ascontiguousarray(cc_old, dtype=np.int32)
        out = np.zeros(len(cc_old), dtype=np.int32)
        self._ck(self.l.dre_locate_points(self.h, C.c_int32(len(cc_old)), abi.dptr(xyz), abi.iptr(cc_old), abi.iptr(out)))
        return out

    def locate_noalloc(self, xyz, cc_old):
        """dre_locate_points_noalloc: the same locator through the library's static scratch array (<= 65 536 points, no allocation)."""
        self.set_params(abi.RTT_DIR_CELL)
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        cc_old = np.ascontiguousarray(cc_old, dtype=np.int32)
        out = np.zeros(len(cc_old), dtype=np.int32)
        self._ck(self.l.dre_locate_points_noalloc(self.h, C.c_int32(len(cc_old)), abi.dptr(xyz), abi.iptr(cc_old), abi.iptr(out)))
        return out

    def eval_exp(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros_like(x)
        self._ck(self.l.dre_eval_exp(self.h, C.c_int32(x.size), abi.dptr(x), abi.dptr(y)))
        return y

    def pix2vec(self, nside, ipix):
        ipix = np.ascontiguousarray(ipix, dtype=np.int32)
        th = np.zeros(len(ipix)); ph = np.zeros(len(ipix)); v = np.zeros((len(ipix), 3))
        self._ck(self.l.dre_pix2vec(self.h, C.c_int32(nside), C.c_int32(len(ipix)), abi.iptr(ipix), abi.dptr(th), abi.dptr(ph),
                                    abi.dptr(v)))
        return th, ph, v


def dartray_main(eng: Engine, max_iterations=20, run_iobs=True, log=None):
    """The call sequence of dartray_main (dartray_hub.f90:12-67) on device-resident state.
    Returns a dict of per-pass counters.  `eng.rt_mpi` is a no-op on one rank."""
    sc = eng.scene
    out = {}
    has_src = sc.nsrc > 0
    has_cell = bool(np.any(sc.lumcell))
    for a in (abi.ARR_U_FEST, abi.ARR_U_FINAL, abi.ARR_SCASPE, abi.ARR_SCASPE_PREV, abi.ARR_SCASPE_TOT, abi.ARR_LUM_LOST,
              abi.ARR_PSEL_AV):
        eng.zero(a)
    eng.set_sources(sc.lumcell)
    # rtt_precalc_cell / rtt_precalc_src
    if has_cell:
        out["precalc_cell"] = eng.rt_loop(abi.RTT_PRECALC_CELL)
    if has_src:
        out["precalc_src"] = eng.rt_loop(abi.RTT_PRECALC_SRC)
    eng.rt_mpi(abi.RTT_PRECALC_SRC)
    # rtt_dir_cell / rtt_dir_src
    if has_cell:
        out["dir_cell"] = eng.rt_loop(abi.RTT_DIR_CELL)
    if has_src:
        out["dir_src"] = eng.rt_loop(abi.RTT_DIR_SRC)
    eng.rt_mpi(abi.RTT_DIR_SRC)
    if run_iobs and sc.ndir > 0:
        eng.zero(abi.ARR_I_OBS)
        if has_cell:
            out["i_obs_dir_cell"] = eng.rt_loop_iobs(abi.RTT_I_OBS_DIR_CELL)
        if has_src:
            out["i_obs_dir_src"] = eng.rt_loop_iobs(abi.RTT_I_OBS_DIR_SRC)
    # rtt_scatt
    tot_rad_en = np.zeros(sc.lnum)
    it = 0
    out["scatt"] = []
    while it < max_iterations:
        it += 1
        if eng.rt_prepare_scatt(it, tot_rad_en):
            break
        c = eng.rt_loop(abi.RTT_SCATT, it)
        eng.rt_mpi(abi.RTT_SCATT)
        out["scatt"].append(c)
        if log:
            log(f"scatt iteration {it}: {c.intersections} intersections, {c.pass_ms:.1f} ms")
    out["iterations"] = it
    if run_iobs and sc.ndir > 0:
        out["i_obs"] = eng.rt_loop_iobs(abi.RTT_I_OBS)
    return out
