#!/usr/bin/env python
"""bench.py — ray-cell intersections/s of the DART-Ray ray-tracing hot path on B200.

One "step" = one full radiation-field solve of the workload: precalc pass, direct-light pass and
all scattering iterations up to convergence (the rt_loop calls of dartray_main,
dartray_hub.f90:12-67), on synthetic inputs shaped like the reference's example
(`--workload trustI`: TRUSTI_GRIDS/input_trustI_tau0.01_LD.in, 21 801 cells, 32 wavelengths).

  value   whole-job intersections/s, accumulators resident in HBM (device timed, max over ranks)
  e2e     the same solve through dre_run_pass with HOST (pinned) arrays: every pass uploads
          u_fest/u_final/scaspe/scaspe_prev/lum_lost and downloads the results
  roofline  algorithmic bytes of the deposit passes / device time of the ray-march kernel launches
  cpu_baseline  the CPU oracle (C++ restatement of the reference path, OpenMP over sources like
          rt_routines.f90:994) on all host cores: ONE WHOLE SOLVE of the same workload (no sampling)

`--impl reference` times that CPU path alone, whole solves of the same workload (the Fortran reference
cannot be built in this image: no Fortran compiler, MPI or HDF5).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def load_scene(name):
    from dartray_b200 import scenes
    if name == "trustI":
        return scenes.trusti_scene(), "TRUST-I slab tau=0.01 (input_trustI_tau0.01_LD.in): 21801 cells, 32 wavelengths, 1 point source"
    if name == "trustI_small":
        return scenes.trusti_scene(lnum=4, max_lvl=3), "TRUST-I slab reduced (5801 cells, 4 wavelengths)"
    if name == "galaxy":
        sc = scenes.galaxy_scene(max_lvl=3, min_lvl=2)
        return sc, ("galaxy-shaped grid, input_galaxy_example1.in parameters with max_lvl=3 (one level less than EX1 so that a "
                    f"step takes about a second): {sc.grid.ncell} cells, {int((sc.lumcell.sum(axis=1) > 0).sum())} emitting leaf cells, {sc.lnum} wavelengths")
    if name == "galaxy_full":
        sc = scenes.galaxy_scene()
        return sc, (f"galaxy-shaped grid at the EX1 depth (max_lvl=4): {sc.grid.ncell} cells, "
                    f"{int((sc.lumcell.sum(axis=1) > 0).sum())} emitting leaf cells, {sc.lnum} wavelengths")
    if name == "magtar":
        sc = scenes.magtar_scene()
        return sc, f"MAGTAR ellipsoidal shell (input_shell.in): {sc.grid.ncell} cells, {sc.lnum} wavelengths, 1 point source"
    if name == "nbody":
        sc = scenes.nbody_scene()
        return sc, (f"N-body/SPH-shaped grid (input_nbody_example1.in parameters, seeded synthetic particles): {sc.grid.ncell} cells, "
                    f"{int((sc.lumcell.sum(axis=1) > 0).sum())} emitting leaf cells, {sc.lnum} wavelengths")
    fn = getattr(scenes, name + "_scene", None)
    if fn is None:
        raise SystemExit(f"unknown workload {name}")
    sc = fn()
    return sc, f"{name}: {sc.grid.ncell} cells, {sc.lnum} wavelengths"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md).  Sampled in-process through NVML (the library
    nvidia-smi itself uses): spawning nvidia-smi every 200 ms steals host time from the rank that enqueues the solve, which shows
    up as load imbalance at 8 GPUs.  Falls back to the nvidia-smi command line when the NVML binding is missing."""

    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index = index
        self.rows = []           # (sm_mhz, sm_max_mhz, [active reason names])
        self.stop = threading.Event()
        self.th = None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.bits = [(pynvml.nvmlClocksEventReasonHwSlowdown, "hw_slowdown"), (pynvml.nvmlClocksEventReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                         (pynvml.nvmlClocksEventReasonSwThermalSlowdown, "sw_thermal_slowdown"), (pynvml.nvmlClocksEventReasonSwPowerCap, "sw_power_cap")]
        except Exception:
            self.nvml = None

    @staticmethod
    def _physical_index(i):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[i])
            except Exception:
                return i
        return i

    def _sample_nvml(self):
        n = self.nvml
        sm = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
        r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        self.rows.append((sm, self.sm_max, [name for bit, name in self.bits if r & bit]))
        if os.environ.get("BENCH_DIAG"):
            mem = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_MEM))
            pw = n.nvmlDeviceGetPowerUsage(self.h) / 1000.0
            print(f"[bench clocks] t={time.perf_counter():.3f} sm={sm:.0f} mem={mem:.0f} power_w={pw:.0f} reasons={r:#x}", file=sys.stderr)

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
            r = [x.strip() for x in out.split(",")]
            self.rows.append((float(r[0]), float(r[1]), [n for i, n in enumerate(self.NAMES) if len(r) > 2 + i and r[2 + i] == "Active"]))

    def _run(self):
        while not self.stop.is_set():
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self.stop.wait((0.05 if os.environ.get("BENCH_DIAG") else 0.25) if self.nvml is not None else 0.5)

    def __enter__(self):
        if os.environ.get("BENCH_NO_CLOCKS"):      # A/B switch for diagnostics: does the sampling disturb the solve?
            return self
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        if self.th is not None:
            self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({n for r in self.rows for n in r[2]})
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": self.rows[0][1], "reasons": reasons, "samples": len(self.rows),
                "source": "NVML (in-process)" if self.nvml is not None else "nvidia-smi"}


def algorithmic_bytes(c):
    """geometry 24 B ccoord + 4 B lvl per intersection + 4 B per cchild load; deposit
    8*La dens + 16*La u_final RMW + 16*sum P(lambda) scaspe RMW; beam test 8*La dens + 8*La u_fest."""
    beam_act = c.active_lambda - c.active_lambda_dep
    return 28 * c.intersections + 4 * c.descent_loads + 24 * c.active_lambda_dep + 16 * c.sca_elems + 16 * beam_act


def deposit_bytes(c):
    """The part of algorithmic_bytes that belongs to the deposit kernel k_deposit (the geometry -- cell records and tree
    descent, 28 B per intersection + 4 B per cchild load, < 1 % of the total -- is done by the march kernels, which are
    not inside the timed launches)."""
    beam_act = c.active_lambda - c.active_lambda_dep
    return 24 * c.active_lambda_dep + 16 * c.sca_elems + 16 * beam_act


def solve_resident(eng, scene, abi, max_iterations=30):
    """precalc -> direct -> scattering iterations on device-resident state; returns the pass counters."""
    diag = [] if os.environ.get("BENCH_DIAG") else None      # wall-clock segments of the solve (diagnostics only)
    t_seg = time.perf_counter()

    def seg(name):
        nonlocal t_seg
        if diag is not None:
            t = time.perf_counter(); diag.append((name, round((t - t_seg) * 1e3, 3))); t_seg = t

    for a in (abi.ARR_U_FEST, abi.ARR_U_FINAL, abi.ARR_SCASPE, abi.ARR_SCASPE_PREV, abi.ARR_SCASPE_TOT, abi.ARR_LUM_LOST,
              abi.ARR_PSEL_AV):
        eng.zero(a)
    seg("zero")
    eng.set_sources(scene.lumcell)
    seg("set_sources")
    has_src, has_cell = scene.nsrc > 0, bool(np.any(scene.lumcell))
    cs = []
    mpi_t = [0.0]
    _rt_mpi = eng.rt_mpi

    def rt_mpi(rt):
        seg("pass")
        t0 = time.perf_counter()
        _rt_mpi(rt)
        mpi_t[0] += time.perf_counter() - t0
        seg("rt_mpi")

    eng_rt_mpi = rt_mpi
    if has_cell:
        cs.append(("precalc_cell", eng.rt_loop(abi.RTT_PRECALC_CELL)))
    if has_src:
        cs.append(("precalc_src", eng.rt_loop(abi.RTT_PRECALC_SRC)))
    eng_rt_mpi(abi.RTT_PRECALC_SRC)
    if has_cell:
        cs.append(("dir_cell", eng.rt_loop(abi.RTT_DIR_CELL)))
    if has_src:
        cs.append(("dir_src", eng.rt_loop(abi.RTT_DIR_SRC)))
    eng_rt_mpi(abi.RTT_DIR_SRC)
    tot = np.zeros(scene.lnum)
    for it in range(1, max_iterations + 1):
        done = eng.rt_prepare_scatt(it, tot)
        seg("rt_prepare")
        if done:
            break
        cs.append((f"scatt{it}", eng.rt_loop(abi.RTT_SCATT, it)))
        eng_rt_mpi(abi.RTT_SCATT)
    solve_resident.last_mpi_ms = mpi_t[0] * 1e3
    if diag is not None:
        print(f"[bench diag] t={time.perf_counter():.3f} " + " ".join(f"{n}={v}" for n, v in diag), file=sys.stderr)
    return cs


def pinned(shape):
    import torch
    return torch.zeros(shape, dtype=torch.float64).pin_memory().numpy()


class HostState:
    """The Fortran host's arrays (u_fest_arr, u_final_arr, scaspe_arr, ...) in pinned memory."""

    def __init__(self, scene):
        N, L = scene.grid.ncell, scene.lnum
        self.u_fest, self.u_final = pinned((N, L)), pinned((N, L))
        self.scaspe = [pinned((N, int(p))) for p in scene.npix]
        self.scaspe_prev = [pinned((N, int(p))) for p in scene.npix]
        self.scaspe_tot = [pinned((N, int(p))) for p in scene.npix]
        self.lum_lost, self.lum_lost_prev = pinned((L,)), pinned((L,))
        self.lumcell = pinned((N, L))
        self.psel_av = pinned((N + scene.nsrc,))

    def nbytes_up(self, precalc, scatt):
        n = self.u_fest.nbytes
        if not precalc:
            n += self.u_final.nbytes + sum(p.nbytes for p in self.scaspe) + self.lum_lost.nbytes
            if scatt:
                n += sum(p.nbytes for p in self.scaspe_prev)
        return n

    def nbytes_down(self, precalc):
        if precalc:
            return self.u_fest.nbytes
        return self.u_final.nbytes + sum(p.nbytes for p in self.scaspe) + self.lum_lost.nbytes + self.psel_av.nbytes


def solve_host(eng, scene, abi, hs, max_iterations=30):
    """The same solve through the drop-in call dre_run_pass with host arrays; rt_prepare(rtt_scatt)
    (dartray_hub.f90:432-457) is done on the host arrays like the Fortran host does."""
    for a in (hs.u_fest, hs.u_final, hs.lum_lost, hs.lum_lost_prev, hs.lumcell, *hs.scaspe, *hs.scaspe_prev, *hs.scaspe_tot):
        a[...] = 0
    eng.set_sources(scene.lumcell)
    has_src, has_cell = scene.nsrc > 0, bool(np.any(scene.lumcell))
    up = down = 0
    isect = 0
    t_call = 0.0
    seq = ([abi.RTT_PRECALC_CELL] if has_cell else []) + ([abi.RTT_PRECALC_SRC] if has_src else []) + \
          ([abi.RTT_DIR_CELL] if has_cell else []) + ([abi.RTT_DIR_SRC] if has_src else [])
    for rt in seq:
        pre = rt in (abi.RTT_PRECALC_CELL, abi.RTT_PRECALC_SRC)
        t0 = time.perf_counter()
        c = eng.rt_loop_host(rt, 0, hs)
        t_call += time.perf_counter() - t0
        isect += c.intersections
        up += hs.nbytes_up(pre, False); down += hs.nbytes_down(pre)
    leaf = scene.grid.cchild == -1
    tot_rad_en = np.zeros(scene.lnum)
    nhp = scene.npix_hp
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1))   # the Fortran host runs these array loops under OpenMP
    for it in range(1, max_iterations + 1):
        hs.u_fest[...] = hs.u_final
        hs.lum_lost_prev[...] = hs.lum_lost

        def prep(l):
            sca = hs.scaspe[l]
            row = sca[:, :nhp[l]].sum(axis=1)
            if it == 1:
                hs.scaspe_tot[l][...] = sca
                tot_rad_en[l] = row.sum()
            else:
                hs.scaspe_tot[l] += sca
            hs.scaspe_prev[l][...] = sca
            hs.lumcell[:, l] = np.where(leaf, row, 0.0)
            sca[...] = 0

        list(pool.map(prep, range(scene.lnum)))
        tl = hs.lumcell.sum(axis=0)
        conv = bool(np.all(tl / tot_rad_en < scene.conv_en_lim)) if tot_rad_en.sum() > 0 else True
        if conv:
            break
        eng.set_sources(hs.lumcell)
        up += hs.lumcell.nbytes
        t0 = time.perf_counter()
        c = eng.rt_loop_host(abi.RTT_SCATT, it, hs)
        t_call += time.perf_counter() - t0
        isect += c.intersections
        up += hs.nbytes_up(False, True); down += hs.nbytes_down(False)
    pool.shutdown()
    return isect, up, down, t_call


def solve_shim(eng, scene, abi, hs, rank, max_iterations=30):
    """One solve the way the Fortran host drives the engine through fortran/rt_wrappers.f90 (the reference-facing C-ABI
    calls, host buffers on both ends): the luminosities go up from the host arrays (dre_set_sources), the passes run on
    device-resident accumulators (rt_loop -> dre_run_pass_resident, rt_mpi -> dre_allreduce, rt_prepare(rtt_scatt) ->
    dre_prepare_scatt), and the host arrays are refreshed where the reference reads them: make_output after the precalculation
    (u_fest), after the direct light (u_final, lum_lost, psel_av) and after the scattering iterations (u_final, every scaspe_tot
    plane, lum_lost, psel_av), on the main process only (io_routines.f90:1396: with no_communications only it writes files).
    Returns (intersections, h2d bytes, d2h bytes)."""
    for a in (abi.ARR_U_FEST, abi.ARR_U_FINAL, abi.ARR_SCASPE, abi.ARR_SCASPE_PREV, abi.ARR_SCASPE_TOT, abi.ARR_LUM_LOST, abi.ARR_PSEL_AV):
        eng.zero(a)                                            # create_scaspe / create_u_final ... of the reference
    hs.lumcell[...] = scene.lumcell
    eng.set_sources(hs.lumcell)
    up = hs.lumcell.nbytes + scene.lum_p_src.nbytes + scene.ccoord_p_src.nbytes + scene.src_cell.nbytes
    down = 0
    has_src, has_cell = scene.nsrc > 0, bool(np.any(scene.lumcell))
    isect = 0
    if has_cell:
        isect += eng.rt_loop(abi.RTT_PRECALC_CELL).intersections
    if has_src:
        isect += eng.rt_loop(abi.RTT_PRECALC_SRC).intersections
    eng.rt_mpi(abi.RTT_PRECALC_SRC)
    if rank == 0:                                              # make_output, rtt_output_part1
        eng.download(abi.ARR_U_FEST, hs.u_fest); down += hs.u_fest.nbytes
    if has_cell:
        isect += eng.rt_loop(abi.RTT_DIR_CELL).intersections
    if has_src:
        isect += eng.rt_loop(abi.RTT_DIR_SRC).intersections
    eng.rt_mpi(abi.RTT_DIR_SRC)
    if rank == 0:                                              # make_output, rtt_output_part2
        eng.download(abi.ARR_U_FINAL, hs.u_final); eng.download(abi.ARR_LUM_LOST, hs.lum_lost); eng.download(abi.ARR_PSEL_AV, hs.psel_av)
        down += hs.u_final.nbytes + hs.lum_lost.nbytes + hs.psel_av.nbytes
    tot = np.zeros(scene.lnum)
    for it in range(1, max_iterations + 1):
        if eng.rt_prepare_scatt(it, tot):
            break
        isect += eng.rt_loop(abi.RTT_SCATT, it).intersections
        eng.rt_mpi(abi.RTT_SCATT)
    if rank == 0:                                              # make_output, rtt_scatt (print_scaspe_tot, print_psel_av on)
        eng.download(abi.ARR_U_FINAL, hs.u_final); eng.download(abi.ARR_SCASPE_TOT, hs.scaspe_tot)
        eng.download(abi.ARR_LUM_LOST, hs.lum_lost); eng.download(abi.ARR_PSEL_AV, hs.psel_av)
        down += hs.u_final.nbytes + sum(p.nbytes for p in hs.scaspe_tot) + hs.lum_lost.nbytes + hs.psel_av.nbytes
    return isect, up, down


def cpu_solve(scene, abi, nthreads=None, max_iterations=30):
    """The whole solve on the host cores with the CPU oracle: the same call sequence as solve_resident
    (dartray_main, dartray_hub.f90:12-67).  Returns (intersections, seconds, per-pass list)."""
    from oracle import oracle as O
    nthreads = nthreads or os.cpu_count() or 1
    orc = O.Oracle(scene, nthreads=nthreads)
    orc.set_option("lock_cell", 1)    # the reference's own update scheme (vectorised when one ray is in the cell)
    st = O.State(scene)
    has_src, has_cell = scene.nsrc > 0, bool(np.any(scene.lumcell))
    passes = []
    t0 = time.perf_counter()
    seq = ([abi.RTT_PRECALC_CELL] if has_cell else []) + ([abi.RTT_PRECALC_SRC] if has_src else []) + \
          ([abi.RTT_DIR_CELL] if has_cell else []) + ([abi.RTT_DIR_SRC] if has_src else [])
    for rt in seq:
        t1 = time.perf_counter()
        c = orc.run_pass(rt, 0, st)
        passes.append((rt, 0, c.intersections, time.perf_counter() - t1))
    for it in range(1, max_iterations + 1):
        if orc.prepare_scatt(it, st):
            break
        orc.set_sources(st.lumcell)
        t1 = time.perf_counter()
        c = orc.run_pass(abi.RTT_SCATT, it, st)
        passes.append((abi.RTT_SCATT, it, c.intersections, time.perf_counter() - t1))
    dt = time.perf_counter() - t0
    return sum(p[2] for p in passes), dt, passes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="trustI")
    ap.add_argument("--no-also", action="store_true", help="skip the secondary workloads (galaxy-shaped grid, MAGTAR)")
    ap.add_argument("--also", default="galaxy,magtar,nbody", help="comma-separated secondary workloads")
    ap.add_argument("--prewarm-s", type=float, default=1.5, help="untimed solves for about this many seconds before the warm-up steps")
    ap.add_argument("--ref-budget-s", type=float, default=150.0, help="reference arm: stop starting new solves after this many seconds")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--option", action="append", default=[], help="engine option name=value")
    args = ap.parse_args()

    # NCCL's log goes to stderr so that stdout stays ONE JSON line; the level (NCCL_DEBUG) is left to the caller -- the driver
    # counts the communicator's ranks from the INFO lines; unset, NCCL prints nothing (NCCL_DEBUG=WARN puts a version banner on stdout)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from dartray_b200 import abi

    scene, desc = load_scene(args.workload)
    metric, unit = "ray-cell intersections/s (direct+scatter solve)", "intersections/s"
    config = {"workload": desc, "solve": "precalc + direct + scattering iterations to conv_en_lim=%g" % scene.conv_en_lim,
              "l2_policy": "L2 flushed (256 MB written) before every timed step; accumulator rows %.0f MB" %
                           (scene.grid.ncell * (scene.lnum + int(scene.npix.sum())) * 8 / 1e6),
              "prewarm": "untimed solves for about %.1f s before the --warmup steps (GPU wake-up)" % args.prewarm_s}

    if args.impl == "reference":
        if rank != 0:
            return
        ncores = os.cpu_count() or 1
        # One step = one WHOLE solve of the same workload on all host cores (no sampling).  The run is bounded in time, not
        # in work per step: at most `warmup` (capped at 1) untimed and `steps` timed solves, and no new solve is started
        # once --ref-budget-s seconds have gone by (C1: ~13 s per solve on 16 cores).
        t_start = time.perf_counter()
        n_warm = min(1, args.warmup)
        for _ in range(n_warm):
            cpu_solve(scene, abi, ncores)
        vals = []
        for i in range(max(1, args.steps)):
            if vals and time.perf_counter() - t_start > args.ref_budget_s:
                break
            n, dt, passes = cpu_solve(scene, abi, ncores)
            vals.append((n, dt))
        n = sum(v[0] for v in vals); t = sum(v[1] for v in vals)
        v = n / t
        sample = (f"whole solve (precalc + direct + {sum(1 for p in passes if p[0] == abi.RTT_SCATT)} scattering iterations), "
                  f"{vals[-1][0]} intersections per solve, {len(vals)} timed solves of {t / len(vals):.1f} s after {n_warm} warm-up")
        print(json.dumps({"impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": args.gpus, "steps": args.steps,
                          "steps_timed": len(vals), "warmup": args.warmup, "ms_per_step": 1e3 * t / len(vals), "higher_is_better": True,
                          "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": unit, "cores": ncores, "kind": "port", "sample": sample},
                          "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    from dartray_b200 import engine
    torch.cuda.set_device(local_rank)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        obj = [engine.Engine.get_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        nccl_id = obj[0]
    def new_id():
        obj = [engine.Engine.get_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        return obj[0]

    eng = engine.Engine(scene, device=local_rank, rank=rank, nranks=world, nccl_id=nccl_id)
    for o in args.option:
        k, v = o.split("=")
        eng.set_option(k, float(v))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def maxr(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sumr(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def flush_l2():
        flush_buf.zero_()
        torch.cuda.synchronize()

    def measure(e_, sc_, steps, warmup, prewarm_s=0.0):
        """`steps` timed solves (device time between two events on the engine's stream, max over ranks)."""
        # GPU wake-up before the warm-up steps: untimed solves for about --prewarm-s seconds.  The first solves of a process
        # run up to 25 % slower than the steady state on some boxes even after three warm-up solves (profiles/r02_sweeps.md);
        # every rank does the same number (the solves contain collectives)
        if prewarm_s > 0:
            t0 = time.perf_counter(); solve_resident(e_, sc_, abi); torch.cuda.synchronize()
            t1 = maxr(time.perf_counter() - t0)
            for _ in range(int(min(40, max(0, np.ceil(prewarm_s / max(t1, 1e-3)) - 1)))):
                solve_resident(e_, sc_, abi)
        for _ in range(warmup):
            solve_resident(e_, sc_, abi)
        barrier()
        l0 = e_.launch_count()
        step_ms, last = [], None
        for _ in range(steps):
            flush_l2()
            barrier()
            e_.mark(0)
            last = solve_resident(e_, sc_, abi)
            e_.mark(1)
            step_ms.append(maxr(e_.elapsed_ms(0, 1)))
            barrier()
        launches = e_.launch_count() - l0
        isect = sumr(float(sum(c.intersections for _, c in last)))
        deposits = sumr(float(sum(c.deposits for _, c in last)))
        ms = float(np.mean(step_ms))
        dep = [c for n, c in last if n.startswith("dir") or n.startswith("scatt")]
        abytes = sumr(float(sum(deposit_bytes(c) for c in dep)))
        abytes_all = sumr(float(sum(algorithmic_bytes(c) for c in dep)))
        kms = maxr(float(sum(c.kernel_ms for c in dep)))
        mpi_ms = maxr(getattr(solve_resident, "last_mpi_ms", 0.0))
        psum = float(sum(c.pass_ms for _, c in last))
        pass_minmax = [-maxr(-psum), maxr(psum)]     # fastest / slowest rank: device time inside the passes
        return {"step_ms": [round(x, 3) for x in step_ms], "pass_minmax": pass_minmax, "mpi_ms": mpi_ms, "ms": ms, "isect": isect, "deposits": deposits, "abytes": abytes, "abytes_all": abytes_all, "kms": kms, "last": last, "launches": launches,
                "trace_launches": int(sum(c.trace_launches for c in dep))}

    with ClockSampler(local_rank) as clk:
        m = measure(eng, scene, args.steps, args.warmup, args.prewarm_s)
    clocks = clk.summary()
    ms, isect, last = m["ms"], m["isect"], m["last"]
    value = isect / (ms * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))

    def roofline(mm):
        # algorithmic bytes of all ranks / (ranks x the slowest rank's ray-march kernel time): per-GPU GB/s
        ach = mm["abytes"] / world / (mm["kms"] * 1e-3) / 1e9 if mm["kms"] > 0 else 0.0
        return ach

    ach = roofline(m)
    def traffic_of(workload):
        """dram__bytes_read.sum + dram__bytes_write.sum of the captured k_trace launch of this workload
        (profiles/traffic.json, from an `ncu --set full` capture; per launch, like `achieved`)."""
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            ent = tj.get("workloads", {}).get(workload)
            return (ent["dram_bytes_per_launch"], ent) if ent else (None, None)
        except Exception:
            return None, None

    def roof_block(mm, workload):
        ach_ = roofline(mm)
        traffic, ent = traffic_of(workload)
        return {"bound": "hbm", "kernel": "k_deposit", "achieved": ach_, "peak": peak, "unit": "GB/s", "frac": ach_ / peak,
                "traffic": traffic, "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650",
                "algorithmic_bytes_per_step": mm["abytes"], "algorithmic_bytes_with_geometry": mm["abytes_all"],
                "kernel_ms_per_step": mm["kms"], "launches_per_step": mm["trace_launches"],
                "per": "GPU (algorithmic bytes of the whole job / n_gpus / slowest rank's kernel time)",
                "traffic_note": (ent or {}).get("launch"),
                "note": "achieved = ALGORITHMIC bytes of the deposits (SURVEY 8d: opacity row 8 B, u_final RMW 16 B per active wavelength, "
                        "scaspe RMW 16 B per element, beam test 16 B per active wavelength) / device time of the k_deposit launches (CUDA "
                        "events on the engine's stream around every launch); the rows are reduced in L2 (cp.reduce.async.bulk), so DRAM "
                        "`traffic` is far below it when the rows fit the 126-MB L2"}

    roof = roof_block(m, args.workload)
    line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": int(m["launches"]),
            "s_per_wavelength": ms * 1e-3 / scene.lnum, "intersections_per_step": isect, "deposits_per_step": m["deposits"],
            "passes": {n: {"ms": round(c.pass_ms, 3), "kernel_ms": round(c.kernel_ms, 3), "intersections": c.intersections,
                           "deposits": c.deposits, "rays": c.rays, "steps": c.steps} for n, c in last},
            "rt_mpi_ms_per_step": m["mpi_ms"], "rank_pass_ms_min_max": m["pass_minmax"], "ms_of_each_step": m["step_ms"], "ms_per_step_median": float(np.median(m["step_ms"])), "roofline": roof}

    # the smaller kernels of the path, once, on the state the last solve left on the device (device time of the pass from the
    # engine's CUDA events; host wall clock around the calls that return host arrays)
    if not args.no_e2e and scene.ndir > 0:
        prod = {}
        eng.zero(abi.ARR_I_OBS)
        has_src, has_cell = scene.nsrc > 0, bool(np.any(scene.lumcell))
        eng.set_sources(scene.lumcell)
        for nm, rt, on in (("i_obs_dir_cell", abi.RTT_I_OBS_DIR_CELL, has_cell), ("i_obs_dir_src", abi.RTT_I_OBS_DIR_SRC, has_src), ("i_obs", abi.RTT_I_OBS, True)):
            if on:
                c = eng.rt_loop_iobs(rt)
                prod[nm] = {"ms": round(c.pass_ms, 3), "kernel_ms": round(c.kernel_ms, 3), "rays": c.rays, "intersections": c.intersections}
        lam_si = np.asarray(scene.lambda_um) * 1e-6
        t0 = time.perf_counter(); eng.calc_sed(1.0e4, 2, lam_si); prod["calc_sed_ms"] = round((time.perf_counter() - t0) * 1e3, 3)
        t0 = time.perf_counter(); eng.make_maps(200, 1.0, 2, lam_si); prod["make_maps_200px_ms"] = round((time.perf_counter() - t0) * 1e3, 3)
        line["products"] = prod

    # the other workloads the metric is quoted on (north star: galaxy-shaped grid, MAGTAR shell), same engine options
    if not args.no_also:
        also = {}
        for w in [x for x in args.also.split(",") if x and x != args.workload]:
            sc2, desc2 = load_scene(w)
            e2 = engine.Engine(sc2, device=local_rank, rank=rank, nranks=world, nccl_id=None if world == 1 else new_id())
            for o in args.option:
                k, v = o.split("=")
                e2.set_option(k, float(v))
            n2 = 1 if w == "nbody" else 2          # a solve of the N-body grid takes ~25 s on one GPU
            # two warm-up solves: the first one sizes the unit batches with a probe batch, the second settles the buffers
            w2 = 1 if w == "nbody" else 2
            m2 = measure(e2, sc2, n2, w2)
            also[w] = {"workload": desc2, "value": m2["isect"] / (m2["ms"] * 1e-3), "ms_per_step": m2["ms"],
                       "s_per_wavelength": m2["ms"] * 1e-3 / sc2.lnum, "intersections_per_step": m2["isect"],
                       "rt_mpi_ms_per_step": m2["mpi_ms"], "ms_of_each_step": m2["step_ms"], "rank_pass_ms_min_max": m2["pass_minmax"], "roofline": roof_block(m2, w), "steps": n2, "warmup": w2}
            e2.close()
        line["also"] = also

    if not args.no_e2e:
        hs = HostState(scene)
        solve_shim(eng, scene, abi, hs, rank)  # warm-up
        barrier()
        n_e2e = max(3, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            isect_h, up, down = solve_shim(eng, scene, abi, hs, rank)
        barrier()
        dt = maxr(time.perf_counter() - t0) / n_e2e
        line["e2e"] = {"value": sumr(float(isect_h)) / dt, "unit": unit, "h2d_bytes_per_step": int(up), "d2h_bytes_per_step": int(sumr(float(down))),
                       "ms_per_step": dt * 1e3, "solves_timed": n_e2e,
                       "protocol": "fortran/rt_wrappers.f90: luminosities uploaded from the host arrays per solve (dre_set_sources), passes on "
                                   "device-resident accumulators, host arrays refreshed at the reference's three make_output points (u_fest; "
                                   "u_final + lum_lost + psel_av; u_final + every scaspe_tot plane + lum_lost + psel_av) by the main process",
                       "timing": "host wall clock around the whole solve (every C-ABI call, uploads and downloads inside), mean of the timed solves, max over ranks"}
        # the slower drop-in form that keeps the host arrays authoritative after EVERY pass (dre_run_pass with host arrays)
        solve_host(eng, scene, abi, hs)
        barrier()
        t0 = time.perf_counter()
        isect_h2, up2, down2, t_call = solve_host(eng, scene, abi, hs)
        barrier()
        dt2 = maxr(time.perf_counter() - t0)
        line["e2e"]["hostarrays_every_pass"] = {"ms_per_step": dt2 * 1e3, "ms_in_dre_run_pass": t_call * 1e3, "h2d_bytes_per_step": int(up2),
                                                "d2h_bytes_per_step": int(down2), "value": sumr(float(isect_h2)) / dt2}

    if rank == 0 and world == 1 and not args.no_cpu:
        ncores = os.cpu_count() or 1
        n_cpu, dt_cpu, passes_cpu = cpu_solve(scene, abi, ncores)     # one whole solve on the host cores, CPU only
        line["cpu_baseline"] = {"value": n_cpu / dt_cpu, "unit": unit, "cores": ncores, "kind": "port",
                                "sample": f"one whole solve of the same workload: {n_cpu} intersections in {dt_cpu:.1f} s",
                                "ms_per_step": dt_cpu * 1e3}
    if rank == 0:
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
