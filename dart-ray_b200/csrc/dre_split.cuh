// dre_split.cuh — the deposit passes (rtt_dir_cell, rtt_dir_src, rtt_scatt) as two kernels.
//
// Every decision that ends a ray of ray_tracing (rt_routines.f90:1618-1857) is geometric: the ray leaves the model
// (:1790), its beam is too coarse for the crossed cell (:2473: the energy test there only chooses between "split" and
// "gone" and zeroes intensities), or too fine (:2530).  The march through the grid therefore does not depend on the
// per-wavelength arithmetic, and the two are separated:
//
//   k_geom     one ray per THREAD (north star: "rays are bundled one per thread"): calc_psel :2307, the tabulated
//              neighbour step :1778-1841, find_cc_new2 (done by k_prep_rays), the beam tests of deposit :2435-2534.
//              It runs twice per step: pass 0 counts the events of every ray, a prefix sum places them, pass 1 writes
//              one 24-byte event per crossed cell with a non-zero path {cell, level, kind, length, psel}, the geometric
//              part of the ray's outcome (list entry fields) and the intersection checksum / records.
//   k_deposit  one ray per WARP, one wavelength per lane: walks the ray's events and does deposit :2378-2555 and
//              process_scatt_rad :2638-2726 -- exp, i_av, the energy test, u_final, the en_sca x phase-function row staged
//              in shared memory and added to the accumulator row with one TMA bulk reduction.  Events are known ahead,
//              so the opacity row of the next cell is always in flight while the current one is processed.
//
// The geometry costs 1/32 of the issue slots it took when every lane of the ray's warp evaluated it redundantly, and the
// deposit warps no longer wait for cell records or carry the geometry in registers.
#pragma once
#include <cooperative_groups.h>
#include "dre_kernels.cuh"

struct __align__(8) DreEvt { double length, psel; int cc; unsigned info; };   // info = clvl | kind << 4 | wall_hit << 6
enum { EV_DEP = 0, EV_COARSE = 1, EV_LOW = 2 };
// how a ray ended (geometry): written into the ray packet for k_deposit
enum { TERM_EXIT = 0,      // left the model (:1790): psel_av bookkeeping
       TERM_COARSE = 1,    // last event is EV_COARSE: split or gone, decided by the energy test
       TERM_LOW = 2,       // last event is EV_LOW: to the low list
       TERM_NONE = 3 };    // relaunch point outside the model (:1654) or a reference STOP condition: nothing to do

// Words of the ray packet (DreRay, dre_kernels.cuh) that k_geom pass 1 overwrites for k_deposit: the list-entry indices
// e0..e3 are dead once k_prep_rays has gathered the luminosities, prev_ray is dead once the march is done.
#define PKT_NEV 16      // int index: number of events
#define PKT_TERM 17
#define PKT_EVOFF 18
#define PKT_PSELT 10    // double index (prev_ray): psel of the terminal intersection

struct GeomCnt { unsigned long long isect, desc, hash; };

// The march of one ray.  WRITE = 0 counts events; WRITE = 1 writes them.  Returns the number of events.
template <int WRITE>
__device__ __forceinline__ int dre_geom_march(const DreDev& g, const DreAccum& A, const DreStep& S, int t, double* __restrict__ pk,
                                              DreEvt* __restrict__ ev, GeomCnt& gc) {
  const int* qi = reinterpret_cast<const int*>(pk);
  const int un = qi[0], nc = qi[1], gid = qi[2], step = qi[3], ray_ord = qi[4], nk = qi[5], dk = qi[6], dip = qi[7];
  const int cc_old = qi[9], isel_ray = qi[10], cc0 = qi[12], lvl0 = qi[13], loc_rc = qi[14], lvl_nc = qi[15];
  const double prev_ray = pk[10];
  const double src[3] = {pk[12], pk[13], pk[14]};
  const DreDir dir = *reinterpret_cast<const DreDir*>(pk + 16);
  const int rt = g.rt_type;
  int term = TERM_NONE, nev = 0;
  int o_isel_old = 0, o_cc = -1; double o_prev = 0, o_dplane = 0, psel_term = 0;
  // calc_cproj :1945-1953
  const double d[3] = {dir.st * dir.cp, dir.st * dir.sp, dir.ct};
  int cc, clvl, isel = 0, isel_old, cc_new, clvl_old = -1;
  double prev;
  bool flag_beam = true, go = true;
  if (cc_old == nc) { cc = nc; prev = 0.0; isel_old = 0; cc_new = -1; clvl = lvl_nc; }
  else {
    if (loc_rc == 1) go = false;                                        // relaunch point outside the model :1654
    else if (loc_rc == 2) { if (WRITE) atomicCAS(A.err, 0, 3); go = false; }
    cc_new = cc0; clvl = lvl0; cc = cc_new; prev = prev_ray; isel_old = isel_ray;
  }
  if (go) {
    int inc[3]; double inv[3];
    for (int a = 0; a < 3; a++) {
      const int v = (int)(d[a] / DRE_GLEPSILON);
      inc[a] = max(min(v, 1), -1);
      inv[a] = 1 / d[a];
    }
    const int nside = 1 << nk;
    const double beam_i = 4 * DRE_PI / (12 * (nside * nside));
    const int b2 = g.base[1];
    double pabs_arr[3] = {0, 0, 0};
    int kint = 0;
    const int nf0 = (inc[0] > 0) ? 1 : 0, nf1 = 2 + ((inc[1] > 0) ? 1 : 0), nf2 = 4 + ((inc[2] > 0) ? 1 : 0);
    for (;;) {
      // ---- calc_psel :2307-2351
      const DreCell* cp = &g.cell[cc];
      const double ccv[3] = {cp->c[0], cp->c[1], cp->c[2]};
      const unsigned e0 = cp->nbr[nf0], e1 = cp->nbr[nf1], e2 = cp->nbr[nf2];
      double pabs[3];
      const double hcs = g.csize_arr[clvl];
      for (int a = 0; a < 3; a++) pabs[a] = (double)g.par.pabs_max;
      if (clvl == clvl_old) {
        for (int a = 0; a < 3; a++) if (inc[a] != 0) pabs[a] = inv[a] * ccv[a] + pabs_arr[a];
      } else {
        for (int a = 0; a < 3; a++) {
          if (inc[a] != 0) pabs[a] = inv[a] * (ccv[a] - src[a] + 0.5 * (double)inc[a] * hcs);
          pabs_arr[a] = inv[a] * ((-src[a]) + 0.5 * (double)inc[a] * hcs);
        }
      }
      double psel;
      if (pabs[0] < pabs[2]) { if (pabs[0] < pabs[1]) { psel = pabs[0]; isel = 1; } else { psel = pabs[1]; isel = 2; } }
      else { if (pabs[2] < pabs[1]) { psel = pabs[2]; isel = 3; } else { psel = pabs[1]; isel = 2; } }
      gc.isect++;
      const unsigned e = (isel == 1) ? e0 : ((isel == 2) ? e1 : e2);
      if (WRITE && A.want_hash) {
        gc.hash += dre_isect_mix(gid, step, ray_ord, kint, cc, isel, (unsigned long long)__double_as_longlong(psel));
        if (A.rec_cap > 0) {
          const unsigned long long idx = atomicAdd(A.rec_n, 1ull);
          if ((long long)idx < A.rec_cap) {
            dre_isect q; q.unit = gid; q.step = step; q.ray = ray_ord; q.k = kint; q.cell = cc; q.isel = isel; q.psel = psel;
            A.rec[idx] = q;
          }
        }
      }
      kint++;
      double length = psel - prev;
      if (fabs(length) < (double)1E-5f * hcs) length = 0;
      if (length < 0) { if (WRITE) atomicCAS(A.err, 0, 1); term = TERM_NONE; break; }   // :1714-1727
      if (cc != nc && cc != cc_new) flag_beam = false;
      int kind = -1;
      if (length != 0.) {
        // ---- the beam tests of deposit :2473, :2494, :2530
        const double carea_f = g.carea_arr[clvl], p2 = psel * psel;
        if (!flag_beam && dre_beam_le(carea_f, p2, (double)g.par.bm_par, beam_i, false)) kind = EV_COARSE;
        else if (flag_beam || (nside <= 2 * 4) || dre_beam_le(carea_f, p2, (double)g.par.bm_par_max, beam_i, true)) kind = EV_DEP;
        else kind = EV_LOW;
        if (WRITE) {
          unsigned wall_hit = 0;                       // check_wall_hit :2572-2596
          if (kind == EV_COARSE && rt != DRE_RTT_DIR_SRC)
            for (int a = 0; a < 3; a++)
              if (g.par.wall_on[a] && (ccv[a] < g.par.wall_coord[2 * a] || ccv[a] > g.par.wall_coord[2 * a + 1])) wall_hit = 1;
          DreEvt q; q.length = length; q.psel = psel; q.cc = cc; q.info = (unsigned)clvl | ((unsigned)kind << 4) | (wall_hit << 6);
          ev[nev] = q;
        }
        nev++;
      }
      if (kind == EV_COARSE || kind == EV_LOW) {           // :1737-1758 (the split / gone choice is made by k_deposit)
        term = (kind == EV_COARSE) ? TERM_COARSE : TERM_LOW;
        o_isel_old = isel_old; o_cc = cc; psel_term = psel;
        const double dpl = (isel_old != 0) ? prev / ((isel_old == 1) ? inv[0] : ((isel_old == 2) ? inv[1] : inv[2])) : 0.;   // Q8, Q11
        if (kind == EV_COARSE) { o_dplane = (isel_old != 0) ? dpl : 0.; o_prev = (isel_old != 0) ? prev : 0.; }
        else { o_dplane = dpl; o_prev = prev; }
        break;
      }
      prev = psel; isel_old = isel; clvl_old = clvl;
      // ---- next cell :1778-1841 (tabulated per (cell, face); deeper digits from the exit point)
      const int isgn = (isel == 1) ? inc[0] : ((isel == 2) ? inc[1] : inc[2]);
      if (e == DRE_NBR_OUT) { term = TERM_EXIT; psel_term = psel; break; }   // left the model :1790-1801
      if (!(e >> 31)) { cc = (int)(e & 0x7ffffffu); clvl = (int)((e >> 27) & 15u); }
      else {
        double bco[3];
        for (int q = 0; q < 3; q++) bco[q] = (q + 1 != isel) ? psel / inv[q] + (src[q] - ccv[q] + 0.5 * hcs) : 0.0;   // :1819-1820
        int l = (int)(e & 0x7fffffffu); bool found = false;
        for (int i = clvl + 1; i <= g.max_lvl; i++) {
          const double cs = g.csize_arr[i];
          int v[3];
          for (int a = 0; a < 3; a++) {
            if (a + 1 == isel) v[a] = (isgn == 1) ? 0 : b2 - 1;   // :1807-1815
            else {                                // :1826-1838 ; fmod evaluated exactly with one fma
              const double qd = bco[a] / cs;
              v[a] = (int)qd;
              double n = trunc(qd);
              double rr = __fma_rn(-n, cs, bco[a]);
              if (bco[a] >= 0 ? (rr < 0) : (rr > 0)) { n += (bco[a] >= 0) ? -1.0 : 1.0; rr = __fma_rn(-n, cs, bco[a]); }
              if (v[a] == b2) v[a] = b2 - 1;
              bco[a] = rr;
            }
          }
          l += (v[2] * b2 + v[1]) * b2 + v[0];
          const int ch = __ldg(&g.cchild[l]);
          if (ch == -1) { cc = l; clvl = i; found = true; break; }
          l = ch;
        }
        if (!found) { if (WRITE) atomicCAS(A.err, 0, 2); term = TERM_NONE; break; }   // :1849-1852
      }
      gc.desc += (unsigned)clvl;
    }
  }
  if (WRITE) {
    int* qw = reinterpret_cast<int*>(pk);
    qw[PKT_NEV] = nev; qw[PKT_TERM] = term;
    pk[PKT_PSELT] = psel_term;
    // the geometric part of the list entry this ray leaves behind; k_deposit upgrades RT_GONE to RT_HIGH when the
    // energy test of a too-coarse beam keeps the ray alive, and writes the luminosities
    const int ty = (term == TERM_LOW) ? RT_LOW : RT_GONE;
    const DreList& out = S.OUT;
    out.unit[t] = un; out.dir_ipix[t] = dip; out.meta[t] = dre_meta(dk, nk, o_isel_old, ty);
    out.cc_old[t] = o_cc; out.dplane[t] = o_dplane; out.prev[t] = o_prev;
    S.fH[t] = 0; S.fW[t] = (term == TERM_LOW) ? 1 : 0;      // outcome flags of the list rebuild (k_deposit sets fH for a split)
  }
  return nev;
}

// Ray packets of a step (k_prep_rays) and pass 0 of the march together: the thread that resolves a ray's record walks it
// at once and leaves the number of its events in nev[t].  NT = threads per block of the calling kernel.
template <int NT>
__device__ __forceinline__ void ph_prep_geom(const DreAccum& A, const DreStep& S, int cur, int* __restrict__ nev) {
  const DreDev& g = c_g;
  __shared__ int s_e[4][NT], s_mode[NT], s_w[NT], s_nc[NT], s_dk[NT], s_dip[NT];
  if (S.ctl->abort) return;
  const int nt = S.ctl->nT;
  const DreList H = S.H[cur], W = S.W[cur];
  for (int t0 = blockIdx.x * NT; t0 < nt; t0 += gridDim.x * NT) {
    const int t = t0 + threadIdx.x;
    if (t < nt) dre_prep_record<MODE_DEP, NT>(g, S.u, H, W, S.tasks, t, S.pkts, S.pstride, s_e, s_mode, s_w, s_nc, s_dk, s_dip);
    __syncthreads();
    dre_prep_lum<NT>(g, A, H, W, t0, min(NT, nt - t0), S.pkts, S.pstride, s_e, s_mode, s_w, s_nc, s_dk, s_dip);
    if (t < nt) { GeomCnt gc; gc.isect = gc.desc = gc.hash = 0; nev[t] = dre_geom_march<0>(g, A, S, t, S.pkts + (size_t)t * S.pstride, nullptr, gc); }
    __syncthreads();
  }
}
#ifndef DRE_GEOM_MINB
#define DRE_GEOM_MINB 6     // (A/B on B200: 4 -> 93.6, 6 -> 92.5, 8 -> 93.1 ms per TRUST-I solve) resident 128-thread CTAs per SM the march kernels are compiled for (register cap 65536 / (128 * n))
#endif
__global__ void __launch_bounds__(128, DRE_GEOM_MINB) k_prep_geom(DreAccum A, DreStep S, int cur, int* __restrict__ nev) { ph_prep_geom<128>(A, S, cur, nev); }

// Pass 1 of the march: events written at evoff[t] (prefix sum of nev), or the whole step refused when they do not fit the
// event buffer.
__device__ __forceinline__ void ph_geom_write(const DreAccum& A, const DreStep& S, DreEvt* __restrict__ evt, const int* __restrict__ evoff, int capE) {
  const DreDev& g = c_g;
  DreCtl* c = S.ctl;
  const int nt = c->nT;
  if (nt <= 0 || c->abort) return;
  const int nE = evoff[nt];
  if (nE > capE || nE < 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) { c->needE = nE < 0 ? 0x7fffffff : nE; c->abort = 1; c->nT = 0; }
    return;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) c->needE = nE;
  GeomCnt gc; gc.isect = gc.desc = gc.hash = 0;
  unsigned long long rays = 0;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < nt; t += gridDim.x * blockDim.x) {
    double* pk = S.pkts + (size_t)t * S.pstride;
    dre_geom_march<1>(g, A, S, t, pk, evt + evoff[t], gc);
    reinterpret_cast<int*>(pk)[PKT_EVOFF] = evoff[t]; rays++;
  }
  for (int o = 16; o > 0; o >>= 1) {
    gc.isect += __shfl_xor_sync(0xffffffffu, gc.isect, o); gc.desc += __shfl_xor_sync(0xffffffffu, gc.desc, o);
    gc.hash += __shfl_xor_sync(0xffffffffu, gc.hash, o); rays += __shfl_xor_sync(0xffffffffu, rays, o);
  }
  if ((threadIdx.x & 31) == 0 && rays != 0) {
    atomicAdd(&A.counters[CNT_ISECT], gc.isect); atomicAdd(&A.counters[CNT_DESC], gc.desc);
    atomicAdd(&A.counters[CNT_HASH], gc.hash); atomicAdd(&A.counters[CNT_RAYS], rays);
  }
}
__global__ void __launch_bounds__(128, DRE_GEOM_MINB) k_geom_write(DreAccum A, DreStep S, DreEvt* __restrict__ evt, const int* __restrict__ evoff, int capE) {
  ph_geom_write(A, S, evt, evoff, capE);
}

// ---- the whole list side of a step in two cooperative launches (grid-wide barriers between the phases instead of kernel
// boundaries): a small step is bound by the latency of its chain of launches, not by work.
//   k_step_pre_coop : select, scans, extraction, sibling search, leader resolution, tasks, ray packets + march pass 0, event
//                     offsets, march pass 1      (then k_deposit, which needs its own shared-memory configuration)
//   k_step_post_coop: outcome scans, new list sizes, scans, rebuild, define_next_level
struct DreCoop { DreEvt* evt; int* nev; int* evoff; int capE; int L; };
__global__ void __launch_bounds__(SCAN_THREADS, 1) k_step_pre_coop(DreAccum A, DreStep S, int cur, DreCoop q) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  DreCtl* c = S.ctl;
  unsigned long long* D = S.scan_desc; const int ds = S.desc_stride;
  ScanJob j; j.n_const = 0; j.total_out = nullptr;
  ph_select(S, cur); grid.sync();
  j.in = S.selH; j.out = S.posH; j.n_ptr = &c->nH; j.slot = 0; dre_scan_job(j, c, D, ds, 0u, 0);
  j.in = S.selW; j.out = S.posW; j.n_ptr = &c->nW; j.slot = 1; dre_scan_job(j, c, D, ds, 1u, 0);
  grid.sync();
  ph_unit_ext(S, cur); grid.sync();
  j.in = S.extcnt; j.out = S.ext_seg; j.n_ptr = &c->U; j.slot = 2; j.total_out = &c->nExt; dre_scan_job(j, c, D, ds, 2u, 0); j.total_out = nullptr;
  grid.sync();
  ph_fill_ext(S, cur); grid.sync();
  ph_cand(S, cur); grid.sync();
  ph_resolve(S, cur); grid.sync();
  j.in = S.ntask; j.out = S.task_off; j.n_ptr = &c->nExt; j.slot = 3; dre_scan_job(j, c, D, ds, 3u, 0);
  j.in = S.nreco; j.out = S.reco_off; j.n_ptr = &c->nExt; j.slot = 4; dre_scan_job(j, c, D, ds, 4u, 0);
  grid.sync();
  ph_tasks(S, cur); grid.sync();
  ph_prep_geom<SCAN_THREADS>(A, S, cur, q.nev); grid.sync();
  j.in = q.nev; j.out = q.evoff; j.n_ptr = &c->nT; j.slot = 10; dre_scan_job(j, c, D, ds, 10u, 1);
  grid.sync();
  ph_geom_write(A, S, q.evt, q.evoff, q.capE);
}
__global__ void __launch_bounds__(SCAN_THREADS, 1) k_step_post_coop(DreStep S, int cur, DreCoop q) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  DreCtl* c = S.ctl;
  unsigned long long* D = S.scan_desc; const int ds = S.desc_stride;
  ScanJob j; j.n_const = 0; j.total_out = nullptr;
  j.in = S.fH; j.out = S.posOH; j.n_ptr = &c->nT; j.slot = 5; dre_scan_job(j, c, D, ds, 5u, 1);
  j.in = S.fW; j.out = S.posOW; j.n_ptr = &c->nT; j.slot = 6; dre_scan_job(j, c, D, ds, 6u, 1);
  grid.sync();
  ph_unit_new(S, cur); grid.sync();
  j.in = S.newHcnt; j.out = S.H[1 - cur].seg; j.n_ptr = &c->U; j.slot = 7; dre_scan_job(j, c, D, ds, 7u, 1);
  j.in = S.newWcnt; j.out = S.W[1 - cur].seg; j.n_ptr = &c->U; j.slot = 8; dre_scan_job(j, c, D, ds, 8u, 1);
  grid.sync();
  ph_build(S, cur, q.L); grid.sync();
  ph_next(S, cur);
}

// per-ray inputs of the deposit walk
struct DepIn {
  int nc, nk, fpix, type, lvl_nc, nev, term;
  double cs_nc, psel_term;
  DreDir fdir;
};

// Asynchronous copy of a cell's opacity row into the warp's look-ahead buffer.
template <int KL>
__device__ __forceinline__ void dre_fetch_dens(const DreDev& g, const WarpSmem& ws, int lane, int cc) {
#pragma unroll
  for (int k = 0; k < KL; k++) {
    const int l = lane + 32 * k;
    if (k < g.KLn && l < g.L) cp_async8(ws.pre + 8 + l, &g.dens[(size_t)cc * g.L + l]);
  }
  cp_async_commit();
}

// deposit :2362-2557 + process_scatt_rad :2638-2726 along the events of one ray.  Returns the outcome kind.
template <int KL, int DET>
__device__ __forceinline__ int deposit_ray(const DreDev& g, const DreAccum& A, WarpSmem& ws, int lane, const DepIn& r,
                                           const DreEvt* __restrict__ ev, double (&I)[KL], double (&lost)[KL], WarpStats& st,
                                           const LamK (&lk)[KL]) {
  const int L = g.L, nc = r.nc, rt = g.rt_type;
  const double cs_nc = r.cs_nc;
  const double norm = 4 * DRE_PI * cs_nc * cs_nc;
  unsigned actmask[KL]; int La = 0;
#pragma unroll
  for (int k = 0; k < KL; k++) {
    const int l = lane + 32 * k;
    I[k] = (k < g.KLn && l < L) ? I[k] / norm : 0.0;   // :1673
    actmask[k] = __ballot_sync(0xffffffffu, I[k] > 0);
    La += __popc(actmask[k]);
  }
  const int lvl_nc = r.lvl_nc;
  const double csize_i = g.csize_arr[lvl_nc], area_i = g.carea_arr[lvl_nc];
  const int nside = 1 << r.nk;
  const double beam_i = 4 * DRE_PI / (12 * (nside * nside));
  const bool src_pass = (rt == DRE_RTT_DIR_SRC);
  // u deposit factor area_i*beam_i/(cs*vol_f) per tree level, held by lane = level (:2499)
  const double kfl = (lane <= g.max_lvl) ? area_i * beam_i / (g.par.cs * g.cvol_arr[lane]) : 0.0;
  int nsca_ray = 0;   // sum of npix over the active wavelengths (event counter)
#pragma unroll
  for (int k = 0; k < KL; k++) if ((actmask[k] >> lane) & 1u) nsca_ray += lk[k].npix;
  for (int o2 = 16; o2 > 0; o2 >>= 1) nsca_ray += __shfl_xor_sync(0xffffffffu, nsca_ray, o2);
  const int nev = r.nev;
  // events: ev[k + 1] sits in registers while event k is processed, ev[k + 2] is in flight
  DreEvt eA, eB;
  eA.length = 0; eA.psel = 0; eA.cc = 0; eA.info = 0; eB = eA;
  if (nev > 0) eA = ev[0];
  if (nev > 1) eB = ev[1];
  DreRuns R; R.n = 0; R.total = 0; R.beg = ws.runs; R.len = ws.runs + DRE_MAXRUN; R.l0 = ws.runs + 2 * DRE_MAXRUN;
  if (ws.fwait) { mbar_wait(ws.mbar, ws.fphase); ws.fphase ^= 1u; ws.fwait = 0; }
  if (nev > 0) {
    bool same = ws.run_valid != 0;
#pragma unroll
    for (int k = 0; k < KL; k++) same = same && (ws.run_mask[k] == actmask[k]);
    if (same) { R.n = ws.run_n; R.total = ws.run_total; }
    else {
      dre_make_runs<KL>(g, actmask, R, lane);
#pragma unroll
      for (int k = 0; k < KL; k++) ws.run_mask[k] = actmask[k];
      ws.run_n = R.n; ws.run_total = R.total; ws.run_valid = 1;
    }
    if (r.fpix >= 0 && R.n > 0) {  // tabulated direction: TMA bulk copies of the active ranges of the row
      fence_async_smem();
      __syncwarp();
      if (lane == 0) {
        const double* src = g.ftab[r.nk] + (size_t)r.fpix * g.ROW;
        mbar_expect(ws.mbar, (unsigned)R.total * 8u);
        for (int q = 0; q < R.n; q++)
          bulk_load_part(ws.F + R.beg[q], src + R.beg[q], (unsigned)R.len[q] * 8u, ws.mbar);
      }
      ws.fwait = 1;
    } else if (r.fpix < 0) dre_phase<KL>(g, ws, lane, r.fdir, actmask);
  }
  cp_async_wait_all();       // (nothing of the previous ray is in flight, but keep the buffer protocol explicit)
  __syncwarp();
  if (nev > 0) dre_fetch_dens<KL>(g, ws, lane, eA.cc);
  unsigned rc_dep = 0, rc_beam = 0, rc_scaev = 0;
  int status = -1;      // OUT_HIGH / OUT_LOW / OUT_NONE + 10 (gone) / -1 (left the model or nothing)
  for (int kev = 0; kev < nev; kev++) {
    const DreEvt e = eA;
    eA = eB;
    if (kev + 2 < nev) eB = ev[kev + 2];
    cp_async_wait_all();
    __syncwarp();
    double dnv[KL];
#pragma unroll
    for (int k = 0; k < KL; k++) dnv[k] = (k < g.KLn && lane + 32 * k < L) ? ws.pre[8 + lane + 32 * k] : 0.0;
    __syncwarp();                               // every lane has read the buffer
    if (kev + 1 < nev) dre_fetch_dens<KL>(g, ws, lane, eA.cc);   // consumed at the next loop top
    const int cc = e.cc, clvl = (int)(e.info & 15u), kind = (int)((e.info >> 4) & 3u);
    const double length = e.length, psel = e.psel;
    if (kind == EV_LOW) { status = OUT_LOW; break; }
    // ---- deposit :2362-2557
    const double csize_f = g.csize_arr[clvl], vol_f = g.cvol_arr[clvl];
    const double kf = __shfl_sync(0xffffffffu, kfl, clvl);
    double tau[KL], ex[KL], iav[KL], dn[KL];
#pragma unroll
    for (int k = 0; k < KL; k++) {
      if (k >= g.KLn) { dn[k] = 0; tau[k] = 0; ex[k] = 1; iav[k] = I[k]; continue; }
      const bool act = (actmask[k] >> lane) & 1u;
      dn[k] = act ? dnv[k] : 0.0;
      tau[k] = dn[k] * length;
      ex[k] = dre_exp(-tau[k]);
      if (tau[k] < DRE_TINY || I[k] == 0) iav[k] = I[k];
      else if (cc != nc) iav[k] = I[k] / tau[k] * (1 - ex[k]);
      else if (!src_pass) {
        tau[k] = dn[k] * csize_i; ex[k] = dre_exp(-tau[k]);
        iav[k] = I[k] / (tau[k] * tau[k]) * (ex[k] + tau[k] - 1);
      } else iav[k] = I[k] / (dn[k] * length) * (1 - ex[k]);
    }
    if (kind == EV_COARSE) {
      rc_beam++;
      const double beam_f = g.carea_arr[clvl] / (psel * psel);
      bool pos = false;
#pragma unroll
      for (int k = 0; k < KL; k++) {
        if ((actmask[k] >> lane) & 1u) {
          const double u_beam = iav[k] * csize_f / g.par.cs / vol_f * area_i * beam_f;
          if (u_beam < g.par.en_lim * __ldg(&A.u_fest[(size_t)cc * L + lane + 32 * k])) {
            I[k] = 0;
            lost[k] = lost[k] + (DET ? dre_rq(iav[k] * beam_i * csize_i * csize_i, ws.cq[2 * dre_lq(L) + lane + 32 * k])
                                     : iav[k] * beam_i * csize_i * csize_i);
          }
          if (I[k] > 0) pos = true;
        }
      }
      const bool any_pos = __any_sync(0xffffffffu, pos);
      const bool wall_hit = (e.info >> 6) & 1u;
      status = (any_pos && !wall_hit) ? OUT_HIGH : OUT_NONE + 10;
      break;
    }
    rc_dep++;
    if (ws.fwait) { mbar_wait(ws.mbar, ws.fphase); ws.fphase ^= 1u; ws.fwait = 0; }
    bool nz = false;
    double* row = A.acc + (size_t)cc * g.ROW;
#pragma unroll
    for (int k = 0; k < KL; k++) {
      const int l = lane + 32 * k;
      if (k < g.KLn && l < L) {
        const bool act = (actmask[k] >> lane) & 1u;
        double en = 0;
        if (act) {
          atomicAdd(&row[l], DET ? dre_rq(iav[k] * length * kf, ws.cq[dre_lq(L) + l])
                                 : iav[k] * length * kf);                       // u_final(:,cc) += ... :2499, one coalesced fp64 reduction
          const double alb = lk[k].alb;
          if (cc != nc) en = (tau[k] > DRE_TINY) ? I[k] * (1.0 - ex[k]) * alb * area_i * beam_i
                                                 : I[k] * tau[k] * alb * area_i * beam_i;
          else en = (tau[k] > DRE_TINY) ? I[k] / (tau[k]) * (ex[k] + tau[k] - 1) * alb * area_i * beam_i
                                        : I[k] * tau[k] / 2 * alb * area_i * beam_i;
        }
        ws.en2[l] = en;
        if (en != 0) nz = true;
      }
    }
    const bool any_nz = __any_sync(0xffffffffu, nz);    // sum(en_sca) /= 0 (:2645), en_sca >= 0
    __syncwarp();
    if (any_nz) {
      rc_scaev++;
      if (A.use_bulk) {
        const double2* F2 = reinterpret_cast<const double2*>(ws.F);
        const unsigned* lam2 = reinterpret_cast<const unsigned*>(ws.lamof);
        const char* enb = reinterpret_cast<const char*>(ws.en2);
        const char* cqb = reinterpret_cast<const char*>(ws.cq);
        // One-slot stage buffer: a chunk is written only after every earlier bulk reduction has finished
        // reading it (cp.async.bulk.wait_group.read 0); ring_slots = 2 splits it into two halves that alternate.
        const bool two = A.ring_slots == 2;
        const int slot_len = two ? ((ws.stage_len >> 1) & ~1) : ws.stage_len;
        int pos = ws.spos, slot = ws.slot;
        for (int q = 0; q < R.n; q++) {
          const int rbeg = R.beg[q], rlen = R.len[q];
          for (int cb = 0; cb < rlen; ) {
            const int cl = min(rlen - cb, slot_len);
            if (pos + cl > slot_len) { slot = two ? (slot ^ 1) : 0; pos = 0; }
            if (pos == 0) { if (lane == 0) { if (two) bulk_wait_read1(); else bulk_wait_read(); } __syncwarp(); }
            double* sdst = ws.stage + slot * slot_len + pos;
            // en_sca(l) * phase function; a pair of elements never straddles two wavelengths (blocks are even-padded),
            // so one table look-up and one en_sca load serve both
            const int p0 = (rbeg + cb) >> 1, pn = cl >> 1;
            double2* S2 = reinterpret_cast<double2*>(sdst);
#pragma unroll kStageUnroll
            for (int j = lane; j < pn; j += 32) {
              const unsigned lm = lam2[p0 + j] & 0xffffu;
              const double2 f = F2[p0 + j];
              const double en = *reinterpret_cast<const double*>(enb + lm);
              double2 v;
              v.x = en * f.x; v.y = en * f.y;
              if (DET) {
                const double cq = *reinterpret_cast<const double*>(cqb + lm);
                v.x = dre_rq(v.x, cq); v.y = dre_rq(v.y, cq);
              }
              S2[j] = v;
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) bulk_red_add_f64(row + rbeg + cb, sdst, (unsigned)cl * 8u);
            pos += cl; cb += cl;
          }
        }
        ws.spos = pos; ws.slot = slot;
      } else {
        for (int j = g.off[0] + lane; j < g.ROW; j += 32) {
          const double en = ws.en2[ws.lamof[j] >> 3];
          if (en == 0) continue;          // inactive wavelength: its part of F was not loaded
          double v = en * ws.F[j];
          if (DET) v = dre_rq(v, ws.cq[ws.lamof[j] >> 3]);
          if (v != 0) atomicAdd(&row[j], v);
        }
      }
    }
    __syncwarp();
    // update ray_intensity :2543-2555
#pragma unroll
    for (int k = 0; k < KL; k++) {
      if (cc != nc) I[k] = I[k] * ex[k];
      else if (!(tau[k] < DRE_TINY)) I[k] = I[k] / (tau[k]) * (1 - ex[k]);
    }
  }
  cp_async_wait_all();       // a look-ahead copy may still be landing when the walk ended early
  __syncwarp();
  st.dep += rc_dep; st.beam += rc_beam;
  st.act += (unsigned long long)La * (rc_dep + rc_beam); st.actdep += (unsigned long long)La * rc_dep;
  st.sca += (unsigned long long)nsca_ray * rc_scaev;
  if (status == OUT_HIGH || status == OUT_LOW) {     // :1737-1758: list luminosities are I * 4 pi csize(nc)^2 (Q7)
    if (status == OUT_HIGH) st.split++; else st.merge++;
#pragma unroll
    for (int k = 0; k < KL; k++) I[k] = I[k] * norm;
    return status;
  }
  if (status == OUT_NONE + 10 || r.term == TERM_EXIT) {   // ras_gone :1746-1751, left the model :1790-1801
    st.npsel++; st.psel += DET ? dre_rq(r.psel_term, ws.cq[3 * dre_lq(L)]) : r.psel_term;
  }
  return OUT_NONE;
}

// ---- the deposit kernel: persistent, one ray per warp, rays handed out in chunks of consecutive tasks
template <int KL, int DET>
__global__ void __launch_bounds__(DRE_TRACE_THREADS, 1) k_deposit(DreAccum A, DreStep S, const DreEvt* __restrict__ evt, int chunk_max) {
  extern __shared__ __align__(128) unsigned char smem[];
  const DreDev& g = c_g;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nt = S.ctl->nT;
  if (nt <= 0 || S.ctl->abort) return;
  // a small step: the warps of the first CTAs cover every ray, the other CTAs leave before they set up shared memory
  if ((long long)blockIdx.x * (blockDim.x >> 5) >= nt) return;
  __shared__ unsigned long long s_cnt[8];
  __shared__ double s_lost[32 * DRE_KL];
  if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
  if (threadIdx.x < 32 * DRE_KL) s_lost[threadIdx.x] = 0;
  const double* __restrict__ pkts = S.pkts;
  const int pstride = S.pstride;
  const DreList out = S.OUT;
  int* __restrict__ counter = &S.ctl->trace_counter;
  const int chunk = max(1, min(chunk_max, nt / ((int)gridDim.x * (int)(blockDim.x >> 5) * 8)));
  WarpSmem ws;
  dre_warp_smem(g, smem, warp, lane, ws, true, A.stage_split, DET ? A.qc : nullptr);
  WarpStats st; memset(&st, 0, sizeof(st));
  double lost[KL];
#pragma unroll
  for (int k = 0; k < KL; k++) lost[k] = 0;
  const int L = g.L;
  LamK lk[KL];
#pragma unroll
  for (int k = 0; k < KL; k++) {
    const int l = lane + 32 * k;
    const bool ok = (k < g.KLn && l < L);
    lk[k].npix = ok ? g.npix[l] : 0; lk[k].alb = ok ? g.ksca[l] : 0.0;
  }
  const unsigned pbytes = (unsigned)pstride * 8u;
  unsigned pphase = 0;
  bool pk_inflight = false;
  int t = 0, t_end = 0;
  for (;;) {
    if (t >= t_end) {                         // next chunk of consecutive rays
      if (lane == 0) t = atomicAdd(counter, chunk);
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= nt) break;
      t_end = min(t + chunk, nt);
    }
    if (!pk_inflight) {                       // first ray of a chunk: its packet was not requested ahead
      if (lane == 0) bulk_load_row(ws.pkt, pkts + (size_t)t * pstride, pbytes, ws.pmbar);
    }
    mbar_wait(ws.pmbar, pphase); pphase ^= 1u;
    DepIn r;
    const int* qi = reinterpret_cast<const int*>(ws.pkt);
    r.nc = qi[1]; r.nk = qi[5];
    const int dk = qi[6], dip = qi[7];
    r.type = qi[8]; r.fpix = qi[11]; r.lvl_nc = qi[15];
    r.nev = qi[PKT_NEV]; r.term = qi[PKT_TERM];
    const int evoff = qi[PKT_EVOFF];
    r.psel_term = ws.pkt[PKT_PSELT]; r.cs_nc = ws.pkt[11];
    r.fdir.st = ws.pkt[16]; r.fdir.ct = ws.pkt[17]; r.fdir.sp = ws.pkt[18]; r.fdir.cp = ws.pkt[19];
    if (r.type == RT_RECO && r.nk <= 8 && r.fpix < 0) r.fdir = g.dirtab[r.nk][1];   // Q2: ipix = 1 for ray_type_reco
    double I[KL];
#pragma unroll
    for (int k = 0; k < KL; k++) {
      const int l = lane + 32 * k;
      I[k] = (k < g.KLn && l < L) ? ws.pkt[DRE_PKT_HDR + l] : 0.0;
    }
    __syncwarp();                             // every lane has read the packet: the buffer is free again
    pk_inflight = (t + 1 < t_end);
    if (pk_inflight && lane == 0) bulk_load_row(ws.pkt, pkts + (size_t)(t + 1) * pstride, pbytes, ws.pmbar);
    st.psel = 0; st.npsel = 0;
    const int okind = deposit_ray<KL, DET>(g, A, ws, lane, r, evt + evoff, I, lost, st, lk);
    if (lane == 0 && st.npsel > 0) {
      const int sid = (g.rt_type == DRE_RTT_DIR_SRC) ? g.N + g.src_cell[r.nc] : r.nc;
      atomicAdd(&A.psel_sum[sid], st.psel);
      atomicAdd((unsigned long long*)&A.ipsel[sid], (unsigned long long)st.npsel);
    }
    if (okind == OUT_HIGH && lane == 0) { out.meta[t] = (out.meta[t] & 0x00ffffffu) | ((unsigned)RT_HIGH << 24); S.fH[t] = 1; }
    if (okind != OUT_NONE) {
#pragma unroll
      for (int k = 0; k < KL; k++) {
        const int l = lane + 32 * k;
        if (k < g.KLn && l < L) out.lum[(size_t)t * L + l] = I[k];
      }
    }
    (void)dk; (void)dip;
    t++;
  }
  if (ws.fwait) { mbar_wait(ws.mbar, ws.fphase); ws.fwait = 0; }
  if (lane == 0) bulk_wait_all();
  // lum_lost and the event counters are reduced per CTA in shared memory, then added to the device totals once
  // (rays / intersections / descent loads / checksum are counted by k_geom)
#pragma unroll
  for (int k = 0; k < KL; k++) {
    const int l = lane + 32 * k;
    if (k < g.KLn && l < L && lost[k] != 0) atomicAdd(&s_lost[l], lost[k]);
  }
  if (lane == 0) {
    atomicAdd(&s_cnt[0], st.dep); atomicAdd(&s_cnt[1], st.beam); atomicAdd(&s_cnt[2], st.split); atomicAdd(&s_cnt[3], st.merge);
    atomicAdd(&s_cnt[4], st.sca); atomicAdd(&s_cnt[5], st.act); atomicAdd(&s_cnt[6], st.actdep);
  }
  __syncthreads();
  if (threadIdx.x < 7 && s_cnt[threadIdx.x] != 0) {
    const int idx = (threadIdx.x == 0) ? CNT_DEP : (threadIdx.x == 1) ? CNT_BEAM : (threadIdx.x == 2) ? CNT_SPLIT : (threadIdx.x == 3) ? CNT_MERGE
                  : (threadIdx.x == 4) ? CNT_SCA : (threadIdx.x == 5) ? CNT_ACT : CNT_ACTDEP;
    atomicAdd(&A.counters[idx], s_cnt[threadIdx.x]);
  }
  if (threadIdx.x < L && s_lost[threadIdx.x] != 0) atomicAdd(&A.lum_lost[threadIdx.x], s_lost[threadIdx.x]);
}
