"""ctypes mirror of include/dartray_engine.h (structs and rt_type ids)."""
import ctypes as C

import numpy as np

RTT_PRECALC_CELL = 1
RTT_PRECALC_SRC = 2
RTT_DIR_CELL = 5
RTT_I_OBS_DIR_CELL = 6
RTT_DIR_SRC = 7
RTT_I_OBS_DIR_SRC = 8
RTT_SCATT = 10
RTT_I_OBS = 11

ARR_U_FEST, ARR_U_FINAL, ARR_SCASPE, ARR_SCASPE_PREV, ARR_SCASPE_TOT = 0, 1, 2, 3, 4
ARR_LUM_LOST, ARR_PSEL_AV, ARR_LUMCELL, ARR_I_OBS, ARR_I_OBS_IN = 5, 6, 7, 8, 9


class Params(C.Structure):
    _fields_ = [
        ("bm_par", C.c_int32), ("bm_par_max", C.c_int32), ("pabs_max", C.c_int32), ("print_psel_av", C.c_int32),
        ("rad_lim", C.c_double), ("en_lim", C.c_double), ("cs", C.c_double), ("modelsize", C.c_double),
        ("wall_on", C.c_int32 * 3), ("reserved0", C.c_int32), ("wall_coord", C.c_double * 6),
    ]


class Counters(C.Structure):
    _fields_ = [
        ("sources", C.c_int64), ("rays", C.c_int64), ("intersections", C.c_int64), ("deposits", C.c_int64),
        ("beam_tests", C.c_int64), ("splits", C.c_int64), ("merges", C.c_int64), ("sca_elems", C.c_int64),
        ("active_lambda", C.c_int64), ("steps", C.c_int64), ("kernel_ms", C.c_double), ("pass_ms", C.c_double),
        ("isect_hash", C.c_uint64), ("active_lambda_dep", C.c_int64), ("descent_loads", C.c_int64),
        ("launches", C.c_int64), ("trace_launches", C.c_int64),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


ISECT_DTYPE = np.dtype([("unit", "<i4"), ("step", "<i4"), ("ray", "<i4"), ("k", "<i4"), ("cell", "<i4"),
                        ("isel", "<i4"), ("psel", "<f8")], align=True)
assert ISECT_DTYPE.itemsize == 32


def make_params(scene, rt_type, print_psel_av=True):
    """Fill dre_params the way the Fortran shim would from the module variables."""
    p = Params()
    scatt = rt_type in (RTT_SCATT,)
    p.bm_par = scene.bm_par_sca if scatt else scene.bm_par  # dartray_hub.f90:442
    p.bm_par_max = scene.bm_par_max
    p.pabs_max = int(scene.grid.modelsize * 10)  # io_routines.f90:678
    p.print_psel_av = 1 if print_psel_av else 0
    p.rad_lim = scene.rad_lim
    p.en_lim = scene.en_lim()
    p.cs = scene.cs()
    p.modelsize = scene.grid.modelsize
    for a in range(3):
        p.wall_on[a] = int(scene.wall_on[a])
    for a in range(6):
        p.wall_coord[a] = float(scene.wall_coord[a])
    return p


def dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def iptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32)) if a is not None else None


def lptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64)) if a is not None else None


def plane_ptrs(planes):
    """array of L pointers to the (N,P(l)) planes (== Fortran scaspe_arr(l)%a(P(l),N))."""
    arr = (C.POINTER(C.c_double) * len(planes))()
    for i, p in enumerate(planes):
        arr[i] = dptr(p)
    return arr
