// dartray_oracle.cpp — CPU restatement of the DART-Ray ray-tracing hot path.
//
// TEST INFRASTRUCTURE ONLY.  Imported/linked only by tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs.  The product (dart-ray_b200/)
// never calls it.
//
// PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this
// path and cannot be compiled here (Fortran 90 + MPI + HDF5, no
// Fortran compiler in the image).  This file is a function-by-function transcription
// with the reference file:line beside each block; its own fidelity is protected by the
// substitute pins in tests/ (HEALPix round trips, tree round trips, chord-length and
// optically-thin analytic cases, energy conservation).
//
// All paths below are under /root/reference.  Compile: g++ -O2 -ffp-contract=off -fopenmp
// (no FMA contraction: gfortran on baseline x86-64 emits none).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cstdio>
#include <vector>
#include <string>
#include <algorithm>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/dartray_engine.h"
#include "../include/dartray_hash.h"

namespace {

// ---- constants: smooth_grid_routines.f90:311-317, healpix_routines.f90:13 ---------
// halfpi=asin(1.) is a default-real (single precision) constant promoted to real64.
const double HALFPI = (double)1.57079637050628662109375f;  // float(pi/2)
const double PI = 2 * HALFPI;
const double TWOPI = 2 * PI;
const double TWOTHIRD = (double)(2.f / 3.f);
const double GLEPSILON = (double)1.0e-7f;   // rt_routines.f90:76
const int NSIDE_MIN = 4;                    // rt_routines.f90:60
const int NPIX_MAIN = 12;                   // rt_routines.f90:62
const int MAXL = 16;                        // max tree depth supported by the oracle

// ray types ray_list.f90:48-59 ; ray status rt_routines.f90:254-264
enum { RT_HIGH = 0, RT_LOW = 1, RT_RECO = 2, RT_GONE = 3, RT_IOBS = 4, RT_IOBS_IN = 5 };
enum { RAS_GO_HIGH = 0, RAS_GONE = 1, RAS_GO_LOW = 2, RAS_FIRST = 3, RAS_RELAUNCH = 4 };

// ---- HEALPix: healpix_routines.f90 -------------------------------------------------
int pix2x[1024], pix2y[1024], x2pix1[128], y2pix1[128];
bool hp_init_done = false;

void mk_pix2xy() {  // :145-175
  for (int kpix = 0; kpix < 1024; kpix++) {
    int jpix = kpix, ix = 0, iy = 0, ip = 1;
    while (jpix != 0) {
      int id = jpix % 2; jpix /= 2; ix = id * ip + ix;
      id = jpix % 2; jpix /= 2; iy = id * ip + iy;
      ip = 2 * ip;
    }
    pix2x[kpix] = ix; pix2y[kpix] = iy;
  }
}
void mk_xy2pix1() {  // :351-386
  for (int i = 0; i < 128; i++) {
    int j = i, k = 0, ip = 1;
    while (j != 0) { int id = j % 2; j /= 2; k = ip * id + k; ip *= 4; }
    x2pix1[i] = k; y2pix1[i] = 2 * k;
  }
}
void hp_init() { if (!hp_init_done) { mk_pix2xy(); mk_xy2pix1(); hp_init_done = true; } }

// pix2ang_nest :30-136 (nside <= 8192 branch)
void pix2ang_nest(int nside, int ipix, double& theta, double& phi) {
  static const int jrll[12] = {2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4};
  static const int jpll[12] = {1, 3, 5, 7, 0, 2, 4, 6, 1, 3, 5, 7};
  int npface = nside * nside;
  int nl4 = 4 * nside;
  int face_num = ipix / npface;
  int ipf = ipix % npface;
  double fn = (double)nside;
  double fact1 = 1.0 / (3.0 * fn * fn);
  double fact2 = 2.0 / (3.0 * fn);
  int ip_low = ipf & 1023;
  int ip_trunc = ipf / 1024;
  int ip_med = ip_trunc & 1023;
  int ip_hi = ip_trunc / 1024;
  int ix = 1024 * pix2x[ip_hi] + 32 * pix2x[ip_med] + pix2x[ip_low];
  int iy = 1024 * pix2y[ip_hi] + 32 * pix2y[ip_med] + pix2y[ip_low];
  int jrt = ix + iy;
  int jpt = ix - iy;
  int jr = jrll[face_num] * nside - jrt - 1;
  int nr, kshift; double z;
  if (jr < nside) {
    nr = jr; z = 1.0 - nr * fact1 * nr; kshift = 0;
  } else if (jr <= 3 * nside) {
    nr = nside; z = (2 * nside - jr) * fact2; kshift = (jr - nside) & 1;
  } else {
    nr = nl4 - jr; z = -1.0 + nr * fact1 * nr; kshift = 0;
  }
  theta = acos(z);
  int jp = (jpll[face_num] * nr + jpt + 1 + kshift) / 2;
  if (jp > nl4) jp -= nl4;
  if (jp < 1) jp += nl4;
  phi = (jp - (kshift + 1) * 0.5) * (HALFPI / nr);
}

// Fortran MODULO(a,p) for reals as gfortran expands it: fmod, then shift into the sign of p
inline double f_modulo(double a, double p) {
  double r = fmod(a, p);
  if (r != 0 && ((r < 0) != (p < 0))) r += p;
  return r;
}

// ang2pix_nest :238-348
int ang2pix_nest(int nside, double theta, double phi) {
  double z = cos(theta);
  double za = fabs(z);
  double tt = f_modulo(phi, TWOPI) / HALFPI;
  int face_num, ix, iy;
  if (za <= TWOTHIRD) {
    int jp = (int)(nside * (0.5 + tt - z * 0.75));
    int jm = (int)(nside * (0.5 + tt + z * 0.75));
    int ifp = jp / nside, ifm = jm / nside;
    if (ifp == ifm) face_num = (ifp & 3) + 4;
    else if (ifp < ifm) face_num = (ifp & 3);
    else face_num = (ifm & 3) + 8;
    ix = jm & (nside - 1);
    iy = nside - (jp & (nside - 1)) - 1;
  } else {
    int ntt = (int)tt;
    if (ntt >= 4) ntt = 3;
    double tp = tt - ntt;
    double tmp = sqrt(3.0 * (1.0 - za));
    int jp = (int)(nside * tp * tmp);
    int jm = (int)(nside * (1.0 - tp) * tmp);
    jp = std::min(nside - 1, jp);
    jm = std::min(nside - 1, jm);
    if (z >= 0) { face_num = ntt; ix = nside - jm - 1; iy = nside - jp - 1; }
    else { face_num = ntt + 8; ix = jp; iy = jm; }
  }
  int ix_low = ix & 127, iy_low = iy & 127;
  int ipf = x2pix1[ix_low] + y2pix1[iy_low] + (x2pix1[ix / 128] + y2pix1[iy / 128]) * 16384;
  return ipf + face_num * nside * nside;
}

// ---- ray list element: ray_list.f90:7-27 ---------------------------------------------
struct Ray {
  double theta, phi;
  std::vector<double> src_lum;
  int nside;
  double dplane, prev;
  int isel;
  int ray_type;
  int cc_old;
};

// ---- context ---------------------------------------------------------------------------
struct Ctx {
  // grid (smooth_grid_routines.f90:41-70)
  int N = 0, max_lvl = 0, base[2] = {0, 0};
  std::vector<int> cchild, lvl;
  std::vector<int64_t> cindex;
  std::vector<double> ccoord, csize, csize_arr, carea_arr, cvol_arr;
  int64_t basediv[2]; int basemask[2];
  // opacity
  int L = 0, ndir = 0;
  std::vector<double> dens_arr, ksca_norm, gsca, dir_i_out;
  std::vector<int> kp_sca, ik_sca, npix_hp_arr, npix_arr;
  int dim_npix_unique = 0;
  std::vector<int> kp_unique, npix_hp_unique, npix_unique;
  std::vector<std::vector<double>> sin_theta_sca, cos_theta_sca, sin_phi_sca, cos_phi_sca;
  // sources
  std::vector<double> lumcell, ccoord_p_src, lum_p_src;
  std::vector<int> src_cell, cell_src;
  int S = 0;
  dre_params par;
  // record
  int64_t max_records = 0;
  std::vector<dre_isect> records;
  std::string err;
  // lock_cell rt_routines.f90:1692-1693 (smooth_grid_routines.f90:191): occupancy counter per cell that selects the
  // reference's vectorised, non-atomic update when a single ray is inside the cell (:2444, :2497, :2649).  That path can lose
  // an update when a second ray enters the cell in the middle of it (the reference accepts this), so it is OFF for the parity
  // checks (every update atomic, exact) and switched ON by the timing legs of bench.py only, so that the CPU baseline runs
  // the reference's fast path (orc_set_option "lock_cell").
  int use_lock_cell = 0;
  int i2d = -1;      // 2D mode (rt_loop_2D rt_routines.f90:1061-1140): -1 off, 0 = positive-octant sources only, 1 = sources on the Cartesian planes only
  mutable std::vector<int> lock_cell;
};

// set_base smooth_grid_routines.f90:622-649
void set_base(Ctx& c) {
  for (int i = 0; i < 2; i++) {
    int j = 0; int64_t base3 = (int64_t)c.base[i] * c.base[i] * c.base[i];
    while (!(base3 < ((int64_t)1 << j))) j++;
    c.basediv[i] = (int64_t)1 << j;
    c.basemask[i] = (int)(c.basediv[i] - 1);
  }
}

inline int64_t ipow64(int64_t b, int e) { int64_t r = 1; for (int i = 0; i < e; i++) r *= b; return r; }

// cindex_to_ccindd smooth_grid_routines.f90:793-813 ; ccindd[axis 0..2][level 1..max_lvl]
void cindex_to_ccindd(const Ctx& c, int nc, int clvl, int ccindd[3][MAXL + 1]) {
  for (int i = 1; i <= clvl; i++) {
    int ib; int64_t j;
    if (i > 1) { ib = 1; j = ((c.cindex[nc] / (c.basediv[0] * ipow64(c.basediv[1], i - 2))) & c.basemask[1]) - 1; }
    else { ib = 0; j = ((c.cindex[nc] / ipow64(c.basediv[0], i - 1)) & c.basemask[0]) - 1; }
    int b = c.base[ib];
    ccindd[0][i] = (int)(j % b);
    ccindd[1][i] = (int)((j / b) % b);
    ccindd[2][i] = (int)((j / ((int64_t)b * b)) % b);
  }
}

// ccindd_to_cc smooth_grid_routines.f90:820-852
bool ccindd_to_cc(const Ctx& c, const int ccindd[3][MAXL + 1], int& cc, int& clvl) {
  int l = 1;
  for (int i = 1; i <= c.max_lvl; i++) {
    int b = c.base[i > 1 ? 1 : 0];
    int k = (ccindd[2][i] * b + ccindd[1][i]) * b + ccindd[0][i] + 1;
    l = l + k - 1;
    if (c.cchild[l] == -1) { cc = l; clvl = i; return true; }
    l = c.cchild[l];
  }
  return false;
}

// increment smooth_grid_routines.f90:751-769
inline void increment(const Ctx& c, int ilvl, int& val, int inc, int& ovr) {
  int b = c.base[ilvl > 1 ? 1 : 0];
  val = val + inc;
  ovr = inc;
  if (val >= b) val = 0;
  else if (val < 0) val = b - 1;
  else ovr = 0;
}

// (continued below)

// ---- per-thread state: the THREADPRIVATE list at rt_routines.f90:527-528, ray_list.f90:61 ----
struct Th {
  const Ctx* c;
  int rt_type, iterations;
  bool multi;                      // >1 thread: shared accumulators use atomics (lock_cell>1 branch)
  bool lock1 = false;              // lock_cell(cc) == 1 for the current intersection (only with Ctx::use_lock_cell)
  // accumulators (shared between threads)
  const double* u_fest; double* u_accum; double* const* scaspe; const double* const* scaspe_prev;
  const double* const* scaspe_tot;
  double src_ccoord[3];
  int nside;
  int ccindd_nc[3][MAXL + 1], ccindd[3][MAXL + 1];
  int clvl_old;
  double pabs_arr[3];
  double psel_av; int64_t ipsel_av;
  std::vector<double> lum_lost_temp, i_obs_temp;
  double vec_mod;
  std::vector<char> iq_a;
  std::vector<int> iq;
  std::vector<double> ray_intensity_dep;
  // per-thread scratch of deposit / ray_tracing (automatic arrays in the Fortran: no heap traffic in the hot loop)
  std::vector<double> s_tau, s_exp, s_iav, s_en, s_ri, s_sl;
  int lnum_a, lnum_a_old;
  std::vector<std::vector<double>> ffn_arr_mw, ffn_arr, ads_arr, scaspe_temp_arr;
  std::vector<Ray> high, low, ext;
  bool no_ray;
  // bookkeeping
  dre_counters cnt;
  std::vector<dre_isect> rec;
  int64_t max_rec;
  int unit, step, ray_ord;
};

struct LockGuard {   // lock_cell(cc) += 1 at the top of an intersection (:1692), -= 1 on every way out (:1743-1845)
  int* p;
  explicit LockGuard(int* q) : p(q) {
    if (p) {
#pragma omp atomic
      *p += 1;
    }
  }
  ~LockGuard() {
    if (p) {
#pragma omp atomic
      *p -= 1;
    }
  }
};
inline bool lock_is_one(const Ctx& c, int cc) {   // the test `lock_cell(cc) == 1` at the moment of the update
  if (!c.use_lock_cell) return false;
  int lc;
#pragma omp atomic read
  lc = c.lock_cell[cc];
  return lc == 1;
}
inline void atomic_add(bool multi, double* p, double v) {
  if (multi) {
#pragma omp atomic
    *p += v;
  } else {
    *p += v;
  }
}

// calc_cproj rt_routines.f90:1945-1953
inline void calc_cproj(double al, double dl, double cproj[3]) {
  cproj[0] = sin(dl) * cos(al);
  cproj[1] = sin(dl) * sin(al);
  cproj[2] = cos(dl);
}

// calc_outcube rt_routines.f90:2072-2085
inline int calc_outcube_sum(const double rel_vec[3], double cellsize) {
  int s = 0;
  for (int i = 0; i < 3; i++)
    if (fabs(rel_vec[i]) > (double)(1.0f + 1E-7f) * cellsize / 2.0) s++;  // (1+1E-7) is single precision
  return s;
}

// find_cc_new2 rt_routines.f90:1962-2066
bool find_cc_new2(Th& t, int cc_old, int& cc_new, const double cproj[3], double prev_ray, bool& out, int& clvl) {
  const Ctx& c = *t.c;
  out = false;
  double prev = prev_ray;
  double pcoord[3], rel_vec[3], norm_pcoord[3];
  for (int i = 0; i < 3; i++) pcoord[i] = t.src_ccoord[i] + prev * cproj[i];
  for (int i = 0; i < 3; i++)
    if (fabs(pcoord[i]) > c.par.modelsize / 2.) { out = true; return true; }
  for (int i = 0; i < 3; i++) rel_vec[i] = pcoord[i] - c.ccoord[3 * (size_t)cc_old + i];
  double cellsize = c.csize[cc_old];
  if (calc_outcube_sum(rel_vec, cellsize) == 0) {
    clvl = c.lvl[cc_old];
    cc_new = cc_old;
    cindex_to_ccindd(c, cc_new, clvl, t.ccindd);
    return true;
  }
  for (int i = 0; i < 3; i++) norm_pcoord[i] = pcoord[i] + c.par.modelsize / 2.;
  for (int i = 1; i <= c.max_lvl; i++) {
    int b = c.base[i > 1 ? 1 : 0];
    for (int a = 0; a < 3; a++) t.ccindd[a][i] = (int)(norm_pcoord[a] / c.csize_arr[i]);
    for (int a = 0; a < 3; a++) if (t.ccindd[a][i] == b) t.ccindd[a][i] = b - 1;
    for (int a = 0; a < 3; a++) norm_pcoord[a] = norm_pcoord[a] - t.ccindd[a][i] * c.csize_arr[i];
  }
  if (!ccindd_to_cc(c, t.ccindd, cc_new, clvl)) return false;  // 'cell not found in find_cc_new2'
  for (int i = 0; i < 3; i++) rel_vec[i] = pcoord[i] - c.ccoord[3 * (size_t)cc_new + i];
  cellsize = c.csize_arr[clvl];
  return calc_outcube_sum(rel_vec, cellsize) == 0;  // else 'host cell not found'
}

// calc_ads_arr rt_routines.f90:2288-2297
void calc_ads_arr(Th& t, double sin_dl, double cos_dl, double sin_al, double cos_al) {
  const Ctx& c = *t.c;
  for (int i = 0; i < c.dim_npix_unique; i++)
    for (int p = 0; p < c.npix_unique[i]; p++)
      t.ads_arr[i][p] = (c.cos_theta_sca[i][p] * cos_dl +
                         c.sin_theta_sca[i][p] * sin_dl * (c.cos_phi_sca[i][p] * cos_al + c.sin_phi_sca[i][p] * sin_al));
}

// calc_ffn_arr rt_routines.f90:2186-2281.  For nside<=256 the reference reads the
// precomputed table ads_arrNNN(:,ipix) built at :2129-2148 from pix2ang_nest(nside, ipix);
// the same values are produced here on the fly.
void calc_ffn_arr(Th& t, double al, double dl, int ipix) {
  const Ctx& c = *t.c;
  if (t.nside <= 256) {
    double theta, phi;
    pix2ang_nest(t.nside, ipix, theta, phi);
    calc_ads_arr(t, sin(theta), cos(theta), sin(phi), cos(phi));
  } else if (c.dim_npix_unique > 0) {
    calc_ads_arr(t, sin(dl), cos(dl), sin(al), cos(al));
  }
  int k = 0;
  for (int i = 0; i < c.L; i++) {
    if (t.iq_a[i]) {
      int ik_sca = c.ik_sca[i];
      if (ik_sca != -1) {
        double g = c.gsca[i];
        double cst = (1.0 - g * g) / c.npix_hp_arr[i];
        std::vector<double>& f = t.ffn_arr[ik_sca];
        int np = c.npix_unique[ik_sca];
        for (int p = 0; p < np; p++) f[p] = pow(1.0 + g * g - 2.0 * g * t.ads_arr[ik_sca][p], -1.5);
        for (int p = 0; p < np; p++) f[p] = cst * f[p];
        double tot_hg = 0;
        for (int p = 0; p < c.npix_hp_arr[i]; p++) tot_hg += f[p];
        for (int p = 0; p < np; p++) f[p] = f[p] / tot_hg;
        for (int p = 0; p < np; p++) t.ffn_arr_mw[k][p] = f[p];
      } else {
        // integer division 1/npix_hp_arr with npix_hp=1 (:2266)
        for (int p = 0; p < c.npix_arr[i]; p++) t.ffn_arr_mw[k][p] = (double)(1 / c.npix_hp_arr[i]);
      }
    }
    k++;
  }
}

// set_iq_list rt_routines.f90:1860-1893
void set_iq_list(Th& t, const std::vector<double>& ray_intensity, bool flag_beam) {
  const Ctx& c = *t.c;
  for (int i = 0; i < c.L; i++) if (ray_intensity[i] > 0) t.iq_a[i] = 1;
  t.lnum_a = 0;
  for (int i = 0; i < c.L; i++) t.lnum_a += t.iq_a[i] ? 1 : 0;
  if (flag_beam) t.lnum_a_old = (int)t.iq.size();
  if (t.lnum_a_old != t.lnum_a || flag_beam) {
    t.iq.clear();
    for (int i = 0; i < c.L; i++) if (t.iq_a[i]) t.iq.push_back(i);
    t.ray_intensity_dep.resize(t.lnum_a);
    for (int k = 0; k < t.lnum_a; k++) t.ray_intensity_dep[k] = ray_intensity[t.iq[k]];
    t.lnum_a_old = t.lnum_a;
  }
}

// calc_psel rt_routines.f90:2307-2351
void calc_psel(Th& t, const int incvec[3], double& psel, double pabs[3], int& isel, int clvl, int cc, const double cproj[3]) {
  const Ctx& c = *t.c;
  for (int a = 0; a < 3; a++) pabs[a] = (double)c.par.pabs_max;
  const double* cco = &c.ccoord[3 * (size_t)cc];
  if (clvl == t.clvl_old) {
    for (int a = 0; a < 3; a++) if (incvec[a] != 0) pabs[a] = cproj[a] * cco[a] + t.pabs_arr[a];
  } else {
    for (int a = 0; a < 3; a++)
      if (incvec[a] != 0) pabs[a] = cproj[a] * (cco[a] - t.src_ccoord[a] + 0.5 * (double)incvec[a] * c.csize_arr[clvl]);
    for (int a = 0; a < 3; a++) t.pabs_arr[a] = cproj[a] * ((-t.src_ccoord[a]) + 0.5 * (double)incvec[a] * c.csize_arr[clvl]);
  }
  if (pabs[0] < pabs[2]) {
    if (pabs[0] < pabs[1]) { psel = pabs[0]; isel = 1; } else { psel = pabs[1]; isel = 2; }
  } else {
    if (pabs[2] < pabs[1]) { psel = pabs[2]; isel = 3; } else { psel = pabs[1]; isel = 2; }
  }
}

// check_wall_hit rt_routines.f90:2572-2596
bool check_wall_hit(const Th& t, int cc) {
  const Ctx& c = *t.c;
  bool wall_hit = false;
  if (t.rt_type == DRE_RTT_DIR_SRC) return wall_hit;
  for (int a = 0; a < 3; a++)
    if (c.par.wall_on[a]) {
      double x = c.ccoord[3 * (size_t)cc + a];
      if (x < c.par.wall_coord[2 * a] || x > c.par.wall_coord[2 * a + 1]) wall_hit = true;
    }
  return wall_hit;
}

// process_scatt_rad rt_routines.f90:2638-2726 (no_communications, sequential_scattering)
void process_scatt_rad(Th& t, int cc, const std::vector<double>& en_sca) {
  const Ctx& c = *t.c;
  double s = 0;
  for (int iw = 0; iw < t.lnum_a; iw++) s += en_sca[iw];
  if (s == 0.) return;
  t.lock1 = lock_is_one(c, cc);
  for (int iw = 0; iw < t.lnum_a; iw++) {
    int il = t.iq[iw];
    double en_sca_out = en_sca[iw];
    int np = c.npix_arr[il];
    double* row = t.scaspe[il] + (size_t)np * cc;
    const std::vector<double>& f = t.ffn_arr_mw[il];
    if (t.lock1) for (int p = 0; p < np; p++) row[p] += en_sca_out * f[p];      // :2649-2656, vectorised
    else for (int p = 0; p < np; p++) atomic_add(t.multi, &row[p], en_sca_out * f[p]);
    t.cnt.sca_elems += np;
  }
}

// deposit rt_routines.f90:2362-2557 ; ray_intensity is the compact (iq) array
void deposit(Th& t, int cc, std::vector<double>& ray_intensity, double length, int nc, double psel,
             int& ray_status, bool flag_beam) {
  const Ctx& c = *t.c;
  const dre_params& P = c.par;
  const int La = t.lnum_a;
  const int L = c.L;
  int npix_beam = 12 * (t.nside * t.nside);
  double* tau_c = t.s_tau.data(); double* exp_tau_c = t.s_exp.data(); double* i_av = t.s_iav.data();
  for (int k = 0; k < La; k++) tau_c[k] = c.dens_arr[t.iq[k] + (size_t)L * cc] * length;
  for (int k = 0; k < La; k++) exp_tau_c[k] = exp(-tau_c[k]);
  int clvl = c.lvl[nc];
  double csize_i = c.csize_arr[clvl];
  double area_i = c.carea_arr[clvl];
  double beam_i = 4 * PI / npix_beam;
  clvl = c.lvl[cc];
  double csize_f = c.csize_arr[clvl];
  double vol_f = c.cvol_arr[clvl];
  double beam_f = c.carea_arr[clvl] / (psel * psel);
  const double TINY = (double)1E-008f;
  if (cc != nc) {
    for (int k = 0; k < La; k++) {
      if (tau_c[k] < TINY || ray_intensity[k] == 0) i_av[k] = ray_intensity[k];
      else i_av[k] = ray_intensity[k] / tau_c[k] * (1 - exp_tau_c[k]);
    }
  } else {
    for (int k = 0; k < La; k++) {
      if (tau_c[k] < TINY || ray_intensity[k] == 0) i_av[k] = ray_intensity[k];
      else if (t.rt_type != DRE_RTT_PRECALC_SRC && t.rt_type != DRE_RTT_DIR_SRC) {
        tau_c[k] = c.dens_arr[t.iq[k] + (size_t)L * cc] * csize_i;
        exp_tau_c[k] = exp(-tau_c[k]);
        i_av[k] = ray_intensity[k] / (tau_c[k] * tau_c[k]) * (exp_tau_c[k] + tau_c[k] - 1);
      } else {
        i_av[k] = ray_intensity[k] / (c.dens_arr[t.iq[k] + (size_t)L * cc] * length) * (1 - exp_tau_c[k]);
      }
    }
  }
  switch (t.rt_type) {
    case DRE_RTT_PRECALC_SRC:
    case DRE_RTT_PRECALC_CELL:
      if (psel > P.rad_lim * P.modelsize) { ray_status = RAS_GONE; return; }
      if ((beam_f / P.bm_par <= beam_i) && !flag_beam) { ray_status = RAS_GO_HIGH; return; }
      if ((beam_f / P.bm_par_max < beam_i) || (t.nside <= 2 * NSIDE_MIN) || flag_beam) {
        if (lock_is_one(c, cc)) for (int k = 0; k < La; k++) t.u_accum[t.iq[k] + (size_t)L * cc] += i_av[k] * length / P.cs / vol_f * area_i * beam_i;   // :2444
        else for (int k = 0; k < La; k++)
          atomic_add(t.multi, &t.u_accum[t.iq[k] + (size_t)L * cc], i_av[k] * length / P.cs / vol_f * area_i * beam_i);
        t.cnt.deposits++; t.cnt.active_lambda += La; t.cnt.active_lambda_dep += La;
      } else { ray_status = RAS_GO_LOW; return; }
      break;
    case DRE_RTT_DIR_SRC:
    case DRE_RTT_DIR_CELL:
    case DRE_RTT_SCATT:
      if ((beam_f / P.bm_par <= beam_i) && !flag_beam) {
        t.cnt.beam_tests++; t.cnt.active_lambda += La;
        bool any_pos = false;
        for (int k = 0; k < La; k++) {
          double u_beam = i_av[k] * csize_f / P.cs / vol_f * area_i * beam_f;
          if (u_beam < P.en_lim * t.u_fest[t.iq[k] + (size_t)L * cc]) {
            ray_intensity[k] = 0;
            t.lum_lost_temp[t.iq[k]] = t.lum_lost_temp[t.iq[k]] + i_av[k] * beam_i * csize_i * csize_i;
          }
        }
        for (int k = 0; k < La; k++) if (ray_intensity[k] > 0) any_pos = true;
        bool wall_hit = check_wall_hit(t, cc);
        if (any_pos && !wall_hit) { ray_status = RAS_GO_HIGH; return; }
        ray_status = RAS_GONE;
        return;
      } else if ((beam_f / P.bm_par_max < beam_i) || (t.nside <= 2 * NSIDE_MIN) || flag_beam) {
        if (lock_is_one(c, cc)) for (int k = 0; k < La; k++) t.u_accum[t.iq[k] + (size_t)L * cc] += i_av[k] * length / P.cs / vol_f * area_i * beam_i;   // :2497
        else for (int k = 0; k < La; k++)
          atomic_add(t.multi, &t.u_accum[t.iq[k] + (size_t)L * cc], i_av[k] * length / P.cs / vol_f * area_i * beam_i);
        t.cnt.deposits++; t.cnt.active_lambda += La; t.cnt.active_lambda_dep += La;
        std::vector<double>& en_sca = t.s_en;
        for (int k = 0; k < La; k++) en_sca[k] = 0.0;
        if (cc != nc) {
          for (int k = 0; k < La; k++) {
            if (tau_c[k] > TINY) en_sca[k] = ray_intensity[k] * (1.0 - exp_tau_c[k]) * c.ksca_norm[t.iq[k]] * area_i * beam_i;
            else en_sca[k] = ray_intensity[k] * tau_c[k] * c.ksca_norm[t.iq[k]] * area_i * beam_i;
          }
        } else {
          for (int k = 0; k < La; k++) {
            if (tau_c[k] > TINY) en_sca[k] = ray_intensity[k] / (tau_c[k]) * (exp_tau_c[k] + tau_c[k] - 1) * c.ksca_norm[t.iq[k]] * area_i * beam_i;
            else en_sca[k] = ray_intensity[k] * tau_c[k] / 2 * c.ksca_norm[t.iq[k]] * area_i * beam_i;
          }
        }
        process_scatt_rad(t, cc, en_sca);
      } else { ray_status = RAS_GO_LOW; return; }
      break;
    default: break;  // i_obs passes: attenuation only
  }
  // update ray_intensity :2543-2555
  if (cc != nc) {
    for (int k = 0; k < La; k++) ray_intensity[k] = ray_intensity[k] * exp_tau_c[k];
  } else {
    for (int k = 0; k < La; k++)
      if (!(tau_c[k] < TINY)) ray_intensity[k] = ray_intensity[k] / (tau_c[k]) * (1 - exp_tau_c[k]);
  }
}
// (continued below)

// add_to_high_ray_list ray_list.f90:463-521
void add_to_high_ray_list(Th& t, double theta, double phi, int nside, double prev, int isel_old, double dplane,
                          const std::vector<double>& src_lum, int cc, int ray_type) {
  Ray r;
  r.theta = theta; r.phi = phi; r.nside = nside;
  if (isel_old != 0) { r.dplane = dplane; r.prev = prev; } else { r.dplane = 0.; r.prev = 0.; }
  r.isel = isel_old; r.src_lum = src_lum; r.ray_type = ray_type; r.cc_old = cc;
  t.high.push_back(std::move(r));
}

// add_to_low_ray_list ray_list.f90:532-579 ; cproj is the INVERTED direction array (Q8)
void add_to_low_ray_list(Th& t, double theta, double phi, int nside, double prev, int isel_old, const double cproj[3],
                         const std::vector<double>& src_lum, int cc) {
  Ray r;
  r.theta = theta; r.phi = phi; r.nside = nside;
  // isel_old==0 would index cproj(0) (out of bounds) in the reference; unreachable because
  // flag_beam forbids go_low in the launch cell.  Guarded, not reproduced.
  r.dplane = (isel_old != 0) ? prev / cproj[isel_old - 1] : 0.;
  r.prev = prev; r.isel = isel_old; r.src_lum = src_lum; r.ray_type = RT_LOW; r.cc_old = cc;
  t.low.push_back(std::move(r));
}

inline void record_isect(Th& t, int k, int cc, int isel, double psel) {
  uint64_t bits; memcpy(&bits, &psel, 8);
  t.cnt.isect_hash += dre_isect_mix(t.unit, t.step, t.ray_ord, k, cc, isel, bits);
  if ((int64_t)t.rec.size() < t.max_rec) {
    dre_isect r; r.unit = t.unit; r.step = t.step; r.ray = t.ray_ord; r.k = k; r.cell = cc; r.isel = isel; r.psel = psel;
    t.rec.push_back(r);
  }
}

// ray_tracing rt_routines.f90:1618-1857.  Returns false on the reference's STOP conditions.
bool ray_tracing(Th& t, int nc, double al, double dl, int clvl, const std::vector<double>& src_lum, double dplane_ray,
                 double prev_ray, int isel_ray, int ray_type, int ipix, int cc_old) {
  const Ctx& c = *t.c;
  const int L = c.L;
  (void)dplane_ray;
  t.cnt.rays++;
  double cproj[3];
  calc_cproj(al, dl, cproj);
  int cc, isel = 0, isel_old, cc_new, ray_status;
  double prev;
  bool flag_beam;
  if (cc_old == nc) {
    cc = nc; prev = 0.0; ray_status = RAS_FIRST; isel_old = 0; flag_beam = true; t.clvl_old = -1;
    memcpy(t.ccindd, t.ccindd_nc, sizeof(t.ccindd)); cc_new = -1;
  } else {
    bool out;
    if (!find_cc_new2(t, cc_old, cc_new, cproj, prev_ray, out, clvl)) return false;
    prev = prev_ray;
    if (out) { t.ray_ord++; return true; }
    cc = cc_new; ray_status = RAS_RELAUNCH; isel_old = isel_ray; flag_beam = true; t.clvl_old = -1;
  }
  int incvec[3];
  for (int i = 0; i < 3; i++) incvec[i] = std::max(std::min((int)(cproj[i] / GLEPSILON), 1), -1);
  for (int i = 0; i < 3; i++) cproj[i] = 1 / cproj[i];
  std::vector<double>& ray_intensity = t.s_ri;
  for (int i = 0; i < L; i++) ray_intensity[i] = src_lum[i] / (4 * PI * c.csize[nc] * c.csize[nc]);
  std::fill(t.iq_a.begin(), t.iq_a.end(), 0);
  set_iq_list(t, ray_intensity, flag_beam);
  const bool dep_pass = (t.rt_type == DRE_RTT_DIR_SRC || t.rt_type == DRE_RTT_DIR_CELL || t.rt_type == DRE_RTT_SCATT);
  if (dep_pass) calc_ffn_arr(t, al, dl, ipix);
  const bool iobs_pass = (t.rt_type == DRE_RTT_I_OBS_DIR_CELL || t.rt_type == DRE_RTT_I_OBS_DIR_SRC || t.rt_type == DRE_RTT_I_OBS);
  int kint = 0;
  bool ok = true;
  for (;;) {
    LockGuard lg(c.use_lock_cell ? &c.lock_cell[cc] : nullptr);
    if (dep_pass) set_iq_list(t, ray_intensity, flag_beam);
    double psel, pabs[3];
    calc_psel(t, incvec, psel, pabs, isel, clvl, cc, cproj);
    t.cnt.intersections++;
    record_isect(t, kint++, cc, isel, psel);
    double length = psel - prev;
    if (fabs(length) < (double)1E-5f * c.csize_arr[clvl]) length = 0;
    if (length < 0) { ok = false; break; }  // 'what ? Something wrong in ray_tracing'
    if (cc != nc && cc != cc_new) flag_beam = false;
    if (length != 0.) {
      deposit(t, cc, t.ray_intensity_dep, length, nc, psel, ray_status, flag_beam);
      for (int k = 0; k < t.lnum_a; k++) ray_intensity[t.iq[k]] = t.ray_intensity_dep[k];
    }
    if (ray_status == RAS_GO_HIGH) {
      std::vector<double>& sl = t.s_sl;
      for (int i = 0; i < L; i++) sl[i] = ray_intensity[i] * (4 * PI * c.csize[nc] * c.csize[nc]);
      double dpl = (isel_old != 0) ? prev / cproj[isel_old - 1] : 0.;  // Q11: guard cproj(0)
      add_to_high_ray_list(t, dl, al, t.nside, prev, isel_old, dpl, sl, cc, RT_HIGH);
      t.cnt.splits++;
      break;
    } else if (ray_status == RAS_GONE) {
      t.ipsel_av++; t.psel_av += psel;
      break;
    } else if (ray_status == RAS_GO_LOW) {
      std::vector<double>& sl = t.s_sl;
      for (int i = 0; i < L; i++) sl[i] = ray_intensity[i] * (4 * PI * c.csize[nc] * c.csize[nc]);
      add_to_low_ray_list(t, dl, al, t.nside, prev, isel_old, cproj, sl, cc);
      t.cnt.merges++;
      break;
    }
    if (ray_type == RT_IOBS_IN) {
      if ((t.vec_mod >= prev) && (t.vec_mod <= psel)) {
        for (int i = 0; i < L; i++) t.i_obs_temp[i] = ray_intensity[i] * exp(c.dens_arr[i + (size_t)L * cc] * (psel - t.vec_mod));
        break;
      }
    }
    prev = psel; isel_old = isel; t.clvl_old = clvl;
    // next cell :1778-1841
    bool left = false;
    for (int i = 1; i <= 3; i++) {
      int a = i - 1;
      if (i == isel) {
        int k = incvec[a];
        for (int j = clvl; j >= 1; j--) { increment(c, j, t.ccindd[a][j], k, k); if (k == 0) break; }
        if (k != 0) {
          t.ipsel_av++; t.psel_av += psel;
          if (iobs_pass) for (int q = 0; q < L; q++) t.i_obs_temp[q] = ray_intensity[q];
          left = true; break;
        }
        for (int j = clvl + 1; j <= c.max_lvl; j++) {
          int b = c.base[j > 1 ? 1 : 0];
          if (incvec[a] == 1) t.ccindd[a][j] = 0;
          else if (incvec[a] == -1) t.ccindd[a][j] = b - 1;
        }
      } else {
        double cellsize = c.csize_arr[clvl];
        double b = psel / cproj[a] + (t.src_ccoord[a] - c.ccoord[3 * (size_t)cc + a] + 0.5 * cellsize);
        for (int j = clvl + 1; j <= c.max_lvl; j++) {
          int bb = c.base[j > 1 ? 1 : 0];
          cellsize = c.csize_arr[j];
          t.ccindd[a][j] = (int)(b / cellsize);
          if (t.ccindd[a][j] == bb) t.ccindd[a][j] = bb - 1;
          b = fmod(b, cellsize);
        }
      }
    }
    if (left) break;
    if (!ccindd_to_cc(c, t.ccindd, cc, clvl)) { ok = false; break; }  // 'Here you should never get'
    t.cnt.descent_loads += clvl;
  }
  t.ray_ord++;
  return ok;
}
// (continued below)

// extract_ray_list ray_list.f90:110-306
void extract_ray_list(Th& t, int nside) {
  const int L = t.c->L;
  t.no_ray = false;
  t.ext.clear();
  if (!t.high.empty()) {
    std::vector<Ray> keep;
    for (auto& r : t.high) {
      if (r.nside == nside / 2) t.ext.push_back(std::move(r)); else keep.push_back(std::move(r));
    }
    if (t.ext.empty()) { t.high.swap(keep); t.no_ray = true; return; }
    t.high.swap(keep);  // compress_high_list :361-403
  } else if (!t.low.empty()) {
    std::vector<Ray> keep;
    bool any = false;
    for (auto& r : t.low) if (r.nside == nside * 2) { any = true; break; }
    if (!any) { t.no_ray = true; return; }
    for (auto& r : t.low) {
      if (r.nside == nside * 2) t.ext.push_back(std::move(r)); else keep.push_back(std::move(r));
    }
    t.low.swap(keep);   // compress_low_list :406-451 (done at the end in the reference; order-independent)
    const int ni = (int)t.ext.size();
    std::vector<int> ipix_arr(ni);
    for (int i = 0; i < ni; i++) ipix_arr[i] = ang2pix_nest(nside, t.ext[i].theta, t.ext[i].phi);
    for (int i = 0; i < ni; i++) {
      int ic_arr[5] = {-1, -1, -1, -1, -1};
      if (t.ext[i].ray_type == RT_GONE) continue;
      double dplane_ray = t.ext[i].dplane;
      int isel_ray = t.ext[i].isel;
      int ipix = ipix_arr[i];
      double theta, phi;
      pix2ang_nest(nside, ipix, theta, phi);
      int icount = 1;
      ic_arr[icount] = i;
      for (int j = i + 1; j < ni; j++) {
        double dplane_ray_add = t.ext[j].dplane;
        int isel_ray_add = t.ext[j].isel;
        int ipix_add = ipix_arr[j];
        if ((ipix_add == ipix) && (isel_ray == isel_ray_add) &&
            (fabs(dplane_ray - dplane_ray_add) / fabs(dplane_ray) < (double)0.001f)) {
          icount++;
          ic_arr[icount] = j;
        }
        if (icount == 4) break;
      }
      if (icount == 4) {
        t.ext[i].nside = nside; t.ext[i].theta = theta; t.ext[i].phi = phi;
        for (int iw = 0; iw < L; iw++) {
          for (int ic = 2; ic <= 4; ic++) t.ext[i].src_lum[iw] = t.ext[i].src_lum[iw] + t.ext[ic_arr[ic]].src_lum[iw];
          t.ext[i].src_lum[iw] = t.ext[i].src_lum[iw] / 4;
        }
        double s = 0;
        for (int ic = 1; ic <= 4; ic++) s += t.ext[ic_arr[ic]].prev;
        t.ext[i].prev = s / 4;
        for (int ic = 2; ic <= 4; ic++) t.ext[ic_arr[ic]].ray_type = RT_GONE;
      } else {
        for (int j = 1; j <= icount; j++) {
          Ray& e = t.ext[ic_arr[j]];
          add_to_high_ray_list(t, e.theta, e.phi, nside, e.prev, e.isel, e.dplane, e.src_lum, e.cc_old, RT_RECO);
          e.ray_type = RT_GONE;
        }
      }
    }
  }
}

// assign_src_lum rt_routines.f90:3043-3071 (the dead code after :3071 is not restated)
void assign_src_lum(Th& t, double theta, double phi, std::vector<double>& src_lum) {
  const Ctx& c = *t.c;
  std::vector<int> ip_arr(std::max(1, c.dim_npix_unique));
  for (int i = 0; i < c.dim_npix_unique; i++) ip_arr[i] = ang2pix_nest(1 << c.kp_unique[i], theta, phi);
  if (t.rt_type == DRE_RTT_SCATT || t.rt_type == DRE_RTT_I_OBS) {
    for (int il = 0; il < c.L; il++) {
      int ik = c.ik_sca[il];
      int ip = (ik != -1) ? ip_arr[ik] : 0;
      src_lum[il] = t.scaspe_temp_arr[il][ip] * c.npix_hp_arr[il];
    }
  }
}

// src_lum_value rt_routines.f90:1897-1919
void src_lum_value(const Th& t, int id, std::vector<double>& out) {
  const Ctx& c = *t.c;
  const int L = c.L;
  int rt = t.rt_type;
  if (rt == DRE_RTT_PRECALC_SRC || rt == DRE_RTT_DIR_SRC || rt == DRE_RTT_I_OBS_DIR_SRC) {
    if (c.S > 0 && c.src_cell[id] != -1) for (int l = 0; l < L; l++) out[l] = c.lum_p_src[l + (size_t)L * c.src_cell[id]];
    else for (int l = 0; l < L; l++) out[l] = 0;
  } else {
    for (int l = 0; l < L; l++) out[l] = c.lumcell[l + (size_t)L * id];
  }
}

// calc_src_ccoord rt_routines.f90:1922-1939
void calc_src_ccoord(Th& t, int id) {
  const Ctx& c = *t.c;
  int rt = t.rt_type;
  if (rt == DRE_RTT_PRECALC_SRC || rt == DRE_RTT_DIR_SRC || rt == DRE_RTT_I_OBS_DIR_SRC)
    for (int a = 0; a < 3; a++) t.src_ccoord[a] = c.ccoord_p_src[3 * (size_t)c.src_cell[id] + a];
  else
    for (int a = 0; a < 3; a++) t.src_ccoord[a] = c.ccoord[3 * (size_t)id + a];
}

// main_dir_loop rt_routines.f90:1343-1467.  `sector_ok(unit)` implements the engine's
// (source, sector) work-unit sharding for the multi-rank tests; the reference shards by
// cell (rt_routines.f90:1005) — sums over ranks are identical either way.
bool main_dir_loop(Th& t, int src_id, int src_ord, const std::vector<double>& src_lum_in, double* lum_lost,
                   int rank, int nranks) {
  const Ctx& c = *t.c;
  const int L = c.L;
  std::vector<double> src_lum = src_lum_in, src_lum_ray(L);
  for (int idir = 0; idir < NPIX_MAIN; idir++) {
    for (int idir2 = 0; idir2 < 4; idir2++) {
      t.unit = src_ord * 48 + idir * 4 + idir2;
      if (t.unit % nranks != rank) continue;
      t.step = 0;
      t.nside = NSIDE_MIN;
      std::fill(t.lum_lost_temp.begin(), t.lum_lost_temp.end(), 0.0);
      // create_high_ray_list ray_list.f90:71-106
      {
        int nside_low = t.nside / 2;
        int ipix = (int)(idir * nside_low * nside_low + idir2 * nside_low * nside_low / 4.);
        Ray r;
        pix2ang_nest(nside_low, ipix, r.theta, r.phi);
        r.nside = nside_low; r.dplane = 0; r.prev = 0; r.isel = 0; r.src_lum = src_lum; r.ray_type = RT_HIGH; r.cc_old = src_id;
        t.high.clear(); t.low.clear();
        t.high.push_back(std::move(r));
      }
      for (;;) {
        extract_ray_list(t, t.nside);
        int next = t.no_ray ? 0 : (int)t.ext.size();
        t.ray_ord = 0;
        for (int i_ext = 0; i_ext < next; i_ext++) {
          Ray& e = t.ext[i_ext];
          int nside_ray = e.nside;
          double theta = e.theta, phi = e.phi;
          src_lum_ray = e.src_lum;
          double dplane_ray = e.dplane, prev_ray = e.prev;
          int isel_ray = e.isel, ray_type = e.ray_type, cc_old = e.cc_old;
          int i0 = 0, i1 = -1;
          if (ray_type == RT_HIGH) { i0 = ang2pix_nest(nside_ray, theta, phi); i0 = i0 * 4; i1 = i0 + 3; }
          if (ray_type == RT_RECO) { i0 = 1; i1 = 1; }
          if (ray_type == RT_LOW) { i0 = ang2pix_nest(t.nside, theta, phi); i1 = i0; }
          if (ray_type == RT_GONE) continue;
          for (int ipix = i0; ipix <= i1; ipix++) {
            if (ray_type == RT_HIGH) pix2ang_nest(t.nside, ipix, theta, phi);
            if (t.rt_type == DRE_RTT_SCATT && cc_old == src_id) {
              assign_src_lum(t, theta, phi, src_lum);
              src_lum_ray = src_lum;
            }
            int clvl = c.lvl[src_id];
            if (!ray_tracing(t, src_id, phi, theta, clvl, src_lum_ray, dplane_ray, prev_ray, isel_ray, ray_type, ipix, cc_old))
              return false;
          }
        }
        t.step++;
        t.cnt.steps++;
        // define_next_level rt_routines.f90:2745-2770
        if (!t.high.empty()) t.nside = t.nside * 2;
        else if (!t.low.empty()) t.nside = t.nside / 2;
        else break;
      }
      if (t.rt_type == DRE_RTT_DIR_SRC || t.rt_type == DRE_RTT_DIR_CELL || t.rt_type == DRE_RTT_SCATT)
        for (int i = 0; i < L; i++) atomic_add(t.multi, &lum_lost[i], t.lum_lost_temp[i]);
    }
  }
  return true;
}

void th_init(Th& t, const Ctx& c, int rt_type, int iterations, bool multi) {
  t.c = &c; t.rt_type = rt_type; t.iterations = iterations; t.multi = multi;
  const int L = c.L;
  t.lum_lost_temp.assign(L, 0.0); t.i_obs_temp.assign(L, 0.0);
  t.iq_a.assign(L, 0); t.iq.clear(); t.ray_intensity_dep.clear();
  t.s_tau.assign(L, 0.0); t.s_exp.assign(L, 0.0); t.s_iav.assign(L, 0.0); t.s_en.assign(L, 0.0); t.s_ri.assign(L, 0.0); t.s_sl.assign(L, 0.0);
  t.lnum_a = t.lnum_a_old = 0;
  t.ffn_arr_mw.resize(L); t.scaspe_temp_arr.resize(L);
  for (int l = 0; l < L; l++) { t.ffn_arr_mw[l].assign(c.npix_arr[l], 0.0); t.scaspe_temp_arr[l].assign(c.npix_arr[l], 0.0); }
  t.ffn_arr.resize(c.dim_npix_unique); t.ads_arr.resize(c.dim_npix_unique);
  for (int i = 0; i < c.dim_npix_unique; i++) { t.ffn_arr[i].assign(c.npix_unique[i], 0.0); t.ads_arr[i].assign(c.npix_unique[i], 0.0); }
  memset(&t.cnt, 0, sizeof(t.cnt));
  t.max_rec = c.max_records; t.rec.clear();
  t.vec_mod = 0; t.psel_av = 0; t.ipsel_av = 0; t.clvl_old = -1;
  memset(t.ccindd, 0, sizeof(t.ccindd)); memset(t.ccindd_nc, 0, sizeof(t.ccindd_nc));
  t.unit = t.step = t.ray_ord = 0;
}

void cnt_add(dre_counters& a, const dre_counters& b) {
  a.sources += b.sources; a.rays += b.rays; a.intersections += b.intersections; a.deposits += b.deposits;
  a.beam_tests += b.beam_tests; a.splits += b.splits; a.merges += b.merges; a.sca_elems += b.sca_elems;
  a.active_lambda += b.active_lambda; a.steps += b.steps; a.isect_hash += b.isect_hash;
  a.active_lambda_dep += b.active_lambda_dep; a.descent_loads += b.descent_loads;
}

}  // namespace
// (continued below)

struct orc_ctx { Ctx c; };

extern "C" {

orc_ctx* orc_create() { hp_init(); orc_ctx* o = new orc_ctx(); memset(&o->c.par, 0, sizeof(dre_params)); return o; }
void orc_destroy(orc_ctx* o) { delete o; }
const char* orc_last_error(orc_ctx* o) { return o->c.err.c_str(); }

int orc_set_grid(orc_ctx* o, int32_t N, int32_t max_lvl, const int32_t base[2], const int32_t* cchild, const int32_t* lvl,
                 const int64_t* cindex, const double* ccoord, const double* csize, const double* csize_arr,
                 const double* carea_arr, const double* cvol_arr) {
  Ctx& c = o->c;
  if (max_lvl > MAXL) { c.err = "max_lvl too large"; return 1; }
  c.N = N; c.max_lvl = max_lvl; c.base[0] = base[0]; c.base[1] = base[1];
  c.cchild.assign(cchild, cchild + N); c.lvl.assign(lvl, lvl + N); c.cindex.assign(cindex, cindex + N);
  c.ccoord.assign(ccoord, ccoord + 3 * (size_t)N); c.csize.assign(csize, csize + N);
  c.csize_arr.assign(csize_arr, csize_arr + max_lvl + 1);
  c.carea_arr.assign(carea_arr, carea_arr + max_lvl + 1);
  c.cvol_arr.assign(cvol_arr, cvol_arr + max_lvl + 1);
  set_base(c);
  return 0;
}

// set_npix_arr class bookkeeping rt_routines.f90:5289-5311 and create_scaspe :2827-2848
int orc_set_opacity(orc_ctx* o, int32_t L, const double* dens_arr, const double* ksca_norm, const double* gsca,
                    const int32_t* kp_sca, int32_t ndir, const double* dir_i_out) {
  Ctx& c = o->c;
  c.L = L; c.ndir = ndir;
  c.dens_arr.assign(dens_arr, dens_arr + (size_t)L * c.N);
  c.ksca_norm.assign(ksca_norm, ksca_norm + L); c.gsca.assign(gsca, gsca + L); c.kp_sca.assign(kp_sca, kp_sca + L);
  c.dir_i_out.assign(dir_i_out, dir_i_out + 2 * (size_t)ndir);
  c.npix_hp_arr.resize(L); c.npix_arr.resize(L); c.ik_sca.assign(L, -1);
  int kmax = -1;
  for (int l = 0; l < L; l++) {
    c.npix_hp_arr[l] = (kp_sca[l] < 0) ? 1 : 12 * (1 << (2 * kp_sca[l]));
    c.npix_arr[l] = c.npix_hp_arr[l] + ndir;
    kmax = std::max(kmax, kp_sca[l]);
  }
  c.kp_unique.clear(); c.npix_hp_unique.clear(); c.npix_unique.clear();
  for (int k = 0; k <= kmax; k++) {
    bool inc = false;
    for (int l = 0; l < L; l++) if (kp_sca[l] == k) inc = true;
    if (inc) {
      for (int l = 0; l < L; l++) if (kp_sca[l] == k) c.ik_sca[l] = (int)c.kp_unique.size();
      c.kp_unique.push_back(k); c.npix_hp_unique.push_back(12 * (1 << (2 * k))); c.npix_unique.push_back(12 * (1 << (2 * k)) + ndir);
    }
  }
  c.dim_npix_unique = (int)c.kp_unique.size();
  c.sin_theta_sca.assign(c.dim_npix_unique, {}); c.cos_theta_sca = c.sin_theta_sca; c.sin_phi_sca = c.sin_theta_sca; c.cos_phi_sca = c.sin_theta_sca;
  for (int i = 0; i < c.dim_npix_unique; i++) {
    int np = c.npix_unique[i], nhp = c.npix_hp_unique[i];
    std::vector<double> th(np), ph(np);
    for (int j = 0; j < nhp; j++) pix2ang_nest(1 << c.kp_unique[i], j, th[j], ph[j]);
    for (int j = 0; j < ndir; j++) { th[nhp + j] = dir_i_out[j]; ph[nhp + j] = dir_i_out[ndir + j]; }
    c.sin_theta_sca[i].resize(np); c.cos_theta_sca[i].resize(np); c.sin_phi_sca[i].resize(np); c.cos_phi_sca[i].resize(np);
    for (int j = 0; j < np; j++) {
      c.sin_theta_sca[i][j] = sin(th[j]); c.cos_theta_sca[i][j] = cos(th[j]);
      c.sin_phi_sca[i][j] = sin(ph[j]); c.cos_phi_sca[i][j] = cos(ph[j]);
    }
  }
  return 0;
}

int orc_set_sources(orc_ctx* o, const double* lumcell, int32_t S, const double* ccoord_p_src, const double* lum_p_src,
                    const int32_t* src_cell, const int32_t* cell_src) {
  Ctx& c = o->c;
  if (lumcell) c.lumcell.assign(lumcell, lumcell + (size_t)c.L * c.N); else c.lumcell.assign((size_t)c.L * c.N, 0.0);
  c.S = S;
  if (S > 0) {
    c.ccoord_p_src.assign(ccoord_p_src, ccoord_p_src + 3 * (size_t)S);
    c.lum_p_src.assign(lum_p_src, lum_p_src + (size_t)c.L * S);
    c.src_cell.assign(src_cell, src_cell + c.N); c.cell_src.assign(cell_src, cell_src + S);
  } else { c.src_cell.assign(c.N, -1); c.cell_src.clear(); }
  return 0;
}

int orc_set_params(orc_ctx* o, const dre_params* p) { o->c.par = *p; return 0; }
int orc_set_record(orc_ctx* o, int64_t n) { o->c.max_records = n; return 0; }
int orc_set_option(orc_ctx* o, const char* name, double value) {
  if (!strcmp(name, "lock_cell")) { o->c.use_lock_cell = value != 0; o->c.lock_cell.assign(o->c.N, 0); return 0; }
  if (!strcmp(name, "i2d")) { o->c.i2d = (int)value; return 0; }
  return 1;
}
int orc_get_records(orc_ctx* o, dre_isect* out, int64_t cap, int64_t* n) {
  int64_t m = std::min<int64_t>(cap, (int64_t)o->c.records.size());
  if (out && m > 0) memcpy(out, o->c.records.data(), m * sizeof(dre_isect));
  *n = (int64_t)o->c.records.size();
  return 0;
}

// skip_cell_2D_rt_loop rt_routines.f90:1174-1187
bool skip_cell_2D_rt_loop(const Ctx& c, int i) {
  if (c.i2d < 0) return false;
  const double* cc = &c.ccoord[3 * (size_t)i];
  if (c.i2d == 0) return (cc[0] <= 0.) || (cc[1] <= 0.) || (cc[2] <= 0.);      // only positive cells in the first loop
  return (cc[0] != 0.) && (cc[1] != 0.) && (cc[2] != 0.);                     // only cells on the Cartesian planes in the second
}

// rt_loop rt_routines.f90:985-1058.  psel_sum/ipsel (N+S): per-source sums (the reference stores
// psel_av/ipsel_av at :1459-1460; the ratio is taken by the caller so sharded runs can be added).
int orc_run_pass(orc_ctx* o, int rt_type, int iterations, const double* u_fest, double* u_accum, double* const* scaspe,
                 const double* const* scaspe_prev, double* lum_lost, double* psel_sum, int64_t* ipsel, int nthreads,
                 int rank, int nranks, dre_counters* out) {
  Ctx& c = o->c;
  const int L = c.L, N = c.N;
  if (nthreads < 1) nthreads = 1;
  c.records.clear();
  dre_counters tot; memset(&tot, 0, sizeof(tot));
  std::vector<double> dummy_lost(L, 0.0);
  if (!lum_lost) lum_lost = dummy_lost.data();
  // ordinal of each source among processed sources (unit numbering)
  std::vector<int> src_ord(N, -1);
  {
    Th t0; th_init(t0, c, rt_type, iterations, false);
    std::vector<double> sl(L);
    int n = 0;
    for (int i = 0; i < N; i++) {
      src_lum_value(t0, i, sl);
      double s = 0; for (int l = 0; l < L; l++) s += sl[l];
      if (c.cchild[i] != -1 || s == 0) continue;
      if (skip_cell_2D_rt_loop(c, i)) continue;
      src_ord[i] = n++;
    }
    tot.sources = n;
  }
  int chunk = std::max(1, N / (nthreads * nthreads));
  int failed = 0;
  auto t_start = std::chrono::steady_clock::now();
#pragma omp parallel num_threads(nthreads)
  {
    Th t; th_init(t, c, rt_type, iterations, nthreads > 1);
    t.u_fest = u_fest; t.u_accum = u_accum; t.scaspe = scaspe; t.scaspe_prev = scaspe_prev; t.scaspe_tot = nullptr;
    std::vector<double> src_lum(L);
#pragma omp for schedule(dynamic, chunk)
    for (int i = 0; i < N; i++) {
      if (failed) continue;
      if (src_ord[i] < 0) continue;
      src_lum_value(t, i, src_lum);
      int src_id = i;
      t.ipsel_av = 0; t.psel_av = 0;
      calc_src_ccoord(t, i);
      int sid = (rt_type == DRE_RTT_DIR_SRC) ? N + c.src_cell[i] : i;
      cindex_to_ccindd(c, src_id, c.lvl[src_id], t.ccindd_nc);
      if (rt_type == DRE_RTT_SCATT)  // assign_scaspe_temp_arr :1161-1163
        for (int il = 0; il < L; il++)
          for (int p = 0; p < c.npix_arr[il]; p++) t.scaspe_temp_arr[il][p] = scaspe_prev[il][p + (size_t)c.npix_arr[il] * i];
      if (!main_dir_loop(t, src_id, src_ord[i], src_lum, lum_lost, rank, nranks)) {
#pragma omp atomic write
        failed = 1;
      }
      if (rt_type == DRE_RTT_DIR_SRC || rt_type == DRE_RTT_DIR_CELL || rt_type == DRE_RTT_SCATT) {
        if (psel_sum) psel_sum[sid] += t.psel_av;
        if (ipsel) ipsel[sid] += t.ipsel_av;
      }
    }
#pragma omp critical
    {
      cnt_add(tot, t.cnt);
      for (auto& r : t.rec) if ((int64_t)c.records.size() < c.max_records) c.records.push_back(r);
    }
  }
  auto t_end = std::chrono::steady_clock::now();
  tot.pass_ms = std::chrono::duration<double, std::milli>(t_end - t_start).count();
  tot.kernel_ms = tot.pass_ms;
  if (out) *out = tot;
  if (failed) { c.err = "ray_tracing: reference STOP condition hit (negative length / cell not found)"; return 2; }
  return 0;
}

// rt_loop_iobs rt_routines.f90:1470-1603
int orc_run_iobs(orc_ctx* o, int rt_type, const double* const* scaspe_tot, int32_t ndir_in, const double* ccoord_obs,
                 double* i_obs, double* i_obs_in, int nthreads, dre_counters* out) {
  Ctx& c = o->c;
  const int L = c.L, N = c.N, S = c.S;
  if (nthreads < 1) nthreads = 1;
  c.records.clear();
  dre_counters tot; memset(&tot, 0, sizeof(tot));
  if (rt_type == DRE_RTT_I_OBS) {  // calc_total_luminosity_sca case(rtt_i_obs) :760-771
    c.lumcell.assign((size_t)L * N, 0.0);
    for (int i = 0; i < N; i++) {
      if (c.cchild[i] != -1) continue;
      for (int j = 0; j < L; j++) {
        double s = 0;
        for (int p = 0; p < c.npix_arr[j]; p++) s += scaspe_tot[j][p + (size_t)c.npix_arr[j] * i];
        c.lumcell[j + (size_t)L * i] = s;
      }
    }
  }
  std::vector<int> src_ord(N, -1);
  {
    Th t0; th_init(t0, c, rt_type, 0, false);
    std::vector<double> sl(L); int n = 0;
    for (int i = 0; i < N; i++) {
      src_lum_value(t0, i, sl);
      double s = 0; for (int l = 0; l < L; l++) s += sl[l];
      if (c.cchild[i] != -1 || s == 0) continue;
      src_ord[i] = n++;
    }
    tot.sources = n;
  }
  const int ndir = c.ndir;
  const int nunit = ndir + ndir_in;
  int chunk = std::max(1, N / (nthreads * nthreads));
  int failed = 0;
  auto t_start = std::chrono::steady_clock::now();
#pragma omp parallel num_threads(nthreads)
  {
    Th t; th_init(t, c, rt_type, 0, nthreads > 1);
    t.u_fest = nullptr; t.u_accum = nullptr; t.scaspe = nullptr; t.scaspe_prev = nullptr; t.scaspe_tot = scaspe_tot;
    std::vector<double> src_lum_ray(L), src_lum(L);
#pragma omp for schedule(dynamic, chunk)
    for (int i = 0; i < N; i++) {
      if (failed || src_ord[i] < 0) continue;
      src_lum_value(t, i, src_lum_ray);
      int src_id = i;
      calc_src_ccoord(t, i);
      int sid = (rt_type == DRE_RTT_I_OBS_DIR_SRC) ? N + c.src_cell[i] : i;
      cindex_to_ccindd(c, src_id, c.lvl[src_id], t.ccindd_nc);
      if (rt_type == DRE_RTT_I_OBS)
        for (int il = 0; il < L; il++)
          for (int p = 0; p < c.npix_arr[il]; p++) t.scaspe_temp_arr[il][p] = scaspe_tot[il][p + (size_t)c.npix_arr[il] * i];
      t.nside = NSIDE_MIN;  // not used by i_obs deposits
      for (int j = 0; j < ndir; j++) {
        std::fill(t.i_obs_temp.begin(), t.i_obs_temp.end(), 0.0);
        double theta = c.dir_i_out[j], phi = c.dir_i_out[ndir + j];
        if (rt_type == DRE_RTT_I_OBS)
          for (int il = 0; il < L; il++) src_lum_ray[il] = t.scaspe_temp_arr[il][c.npix_hp_arr[il] + j] * c.npix_hp_arr[il];
        int clvl = c.lvl[src_id];
        t.unit = src_ord[i] * nunit + j; t.step = 0; t.ray_ord = 0;
        if (!ray_tracing(t, src_id, phi, theta, clvl, src_lum_ray, 0., 0., 0, RT_IOBS, 0, src_id)) {
#pragma omp atomic write
          failed = 1;
        }
        double* dst = i_obs + (size_t)L * (sid + (size_t)(N + S) * j);
        for (int l = 0; l < L; l++) dst[l] += t.i_obs_temp[l];
      }
      for (int j = 0; j < ndir_in; j++) {
        std::fill(t.i_obs_temp.begin(), t.i_obs_temp.end(), 0.0);
        // find_theta_phi_obs_in rt_routines.f90:3164-3217
        double vec[3], theta, phi;
        for (int a = 0; a < 3; a++) vec[a] = ccoord_obs[3 * (size_t)j + a] - t.src_ccoord[a];
        t.vec_mod = sqrt(vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2]);
        const double E8 = (double)1E-8f;
        if (t.vec_mod > E8) {
          double theta_ratio = (0 * vec[0] + 0 * vec[1] + 1 * vec[2]) / t.vec_mod;
          if (fabs(theta_ratio - 1.0) < E8) theta_ratio = 1.0;
          if (fabs(theta_ratio + 1.0) < E8) theta_ratio = -1.0;
          theta = acos(theta_ratio);
          double vec_xy = sin(theta), phi_ratio;
          if (vec_xy > E8) phi_ratio = vec[0] / (t.vec_mod * vec_xy); else phi_ratio = 1.;
          if (fabs(phi_ratio - 1.0) < E8) phi_ratio = 1.0;
          if (fabs(phi_ratio + 1.0) < E8) phi_ratio = -1.0;
          phi = acos(phi_ratio);
          if (vec[1] < 0) phi = 2 * PI - phi;
        } else { theta = 0; phi = 0; t.vec_mod = 0; }
        if (rt_type == DRE_RTT_I_OBS) { assign_src_lum(t, theta, phi, src_lum); src_lum_ray = src_lum; }
        int clvl = c.lvl[src_id];
        t.unit = src_ord[i] * nunit + ndir + j; t.step = 0; t.ray_ord = 0;
        if (!ray_tracing(t, src_id, phi, theta, clvl, src_lum_ray, 0., 0., 0, RT_IOBS_IN, 0, src_id)) {
#pragma omp atomic write
          failed = 1;
        }
        double* dst = i_obs_in + (size_t)L * (sid + (size_t)(N + S) * j);
        for (int l = 0; l < L; l++) dst[l] += t.i_obs_temp[l];
      }
    }
#pragma omp critical
    {
      cnt_add(tot, t.cnt);
      for (auto& r : t.rec) if ((int64_t)c.records.size() < c.max_records) c.records.push_back(r);
    }
  }
  auto t_end = std::chrono::steady_clock::now();
  tot.pass_ms = std::chrono::duration<double, std::milli>(t_end - t_start).count();
  tot.kernel_ms = tot.pass_ms;
  if (out) *out = tot;
  if (failed) { c.err = "ray_tracing: reference STOP condition hit"; return 2; }
  return 0;
}

// rt_prepare(rtt_scatt) dartray_hub.f90:432-457 + calc_total_luminosity_sca rt_routines.f90:663-758
// (no_communications + sequential_scattering).  lumcell is stored in the context for the next pass.
// ---- N4: set_dust_emission, dust_heating_type 'eff' (sed_routines.f90:249-276) ---------------------------------------
namespace dustem {
const double parsec = (double)3.08567758E16f, cspeed = (double)2.99792458E8f;       // smooth_grid_routines.f90:373-377
const double hplanck = (double)6.62606957E-34f, kboltz = (double)1.3806488E-23f;   // :440-443
// bplanck sed_routines.f90:939-948
double bplanck(double t_source, double la) {
  double a1 = 2 * hplanck * pow(cspeed, 2.) / pow(la, 5.);
  double a2 = exp(hplanck * cspeed / (la * kboltz * t_source)) - 1;
  return a1 * pow(a2, -1.);
}
// abs_en_diff :917-936 (em_int is the caller's work array)
double abs_en_diff(double t_dust, const double* kabs_arr_planck, double abs_en, std::vector<double>& em_int, const double* lambda_si_dust,
                   const double* delta_lambda_bin_dust) {
  const int lnum = (int)em_int.size();
  for (int i = 0; i < lnum; i++) em_int[i] = (kabs_arr_planck[i] * bplanck(t_dust, lambda_si_dust[i]));
  double em_en = 0;
  for (int i = 0; i < lnum; i++) em_en += em_int[i] * delta_lambda_bin_dust[i];
  em_en = 4 * PI * em_en;
  return em_en - abs_en;
}
// ZBRENT_tdust :993-1078.  t0 = 0, t1 = 1E4 (sed_routines.f90:9-11); EPS = 3.E-16 (default real); TOL is implicitly typed and
// never assigned in the reference (undefined there; 0 here).  Returns false on the reference's two STOP conditions.
bool zbrent_tdust(const double* kabs_arr_planck, double abs_en, int lnum, const double* lambda_si_dust, const double* dl_dust, double& root) {
  const int ITMAX = 100; const double EPS = (double)3.E-16f, TOL = 0.0;
  std::vector<double> em_int(lnum);
  double A = 0.0, B = (double)1E4f, C = 0, D = 0, E = 0;
  double FA = abs_en_diff(A, kabs_arr_planck, abs_en, em_int, lambda_si_dust, dl_dust);
  double FB = abs_en_diff(B, kabs_arr_planck, abs_en, em_int, lambda_si_dust, dl_dust);
  if (FB * FA > 0.) return false;
  double FC = FB;
  for (int ITER = 1; ITER <= ITMAX; ITER++) {
    if (FB * FC > 0.) { C = A; FC = FA; D = B - A; E = D; }
    if (fabs(FC) < fabs(FB)) { A = B; B = C; C = A; FA = FB; FB = FC; FC = FA; }
    double TOL1 = 2. * EPS * fabs(B) + 0.5 * TOL;
    double XM = .5 * (C - B);
    if (fabs(XM) <= TOL1 || FB == 0.) { root = B; return true; }
    if (fabs(E) >= TOL1 && fabs(FA) > fabs(FB)) {
      double P, Q, R, S = FB / FA;
      if (A == C) { P = 2. * XM * S; Q = 1. - S; }
      else { Q = FA / FC; R = FB / FC; P = S * (2. * XM * Q * (Q - R) - (B - A) * (R - 1.)); Q = (Q - 1.) * (R - 1.) * (S - 1.); }
      if (P > 0.) Q = -Q;
      P = fabs(P);
      if (2. * P < std::min(3. * XM * Q - fabs(TOL1 * Q), fabs(E * Q))) { E = D; D = P / Q; }
      else { D = XM; E = D; }
    } else { D = XM; E = D; }
    A = B; FA = FB;
    if (fabs(D) > TOL1) B = B + D; else B = B + copysign(TOL1, XM);
    FB = abs_en_diff(B, kabs_arr_planck, abs_en, em_int, lambda_si_dust, dl_dust);
  }
  return false;
}
}  // namespace dustem

// calc_dens_dustem sed_routines.f90:377-430 + set_units_dens_stars_arr :904-914.  rf_dust may be null when iterations_dustem == 0.
int orc_dust_emission(orc_ctx* o, int32_t lnum_stars, int32_t lnum_dust, const double* kabs_stars, const double* kabs_dust,
                      const double* lambda_si_dust, const double* dl_stars, const double* dl_dust, double kext_ref, const double* dens_ref,
                      const double* rf_stars, const double* rf_dust, int32_t iterations_dustem, double* dens_stars, double* t_dust_out) {
  Ctx& c = o->c;
  int bad = 0;
  for (int i = 0; i < c.N; i++) {
    for (int j = 0; j < lnum_dust; j++) dens_stars[(size_t)lnum_dust * i + j] = 0;
    if (t_dust_out) t_dust_out[i] = 0;
    if (c.cchild[i] != -1 || dens_ref[i] == 0) continue;
    // calc_t_dust_equil :846-874
    double abs_en = 0;
    { double s = 0; for (int l = 0; l < lnum_stars; l++) s += (kabs_stars[l] * rf_stars[(size_t)lnum_stars * i + l]) * dl_stars[l]; abs_en = s; }
    { double s = 0; if (iterations_dustem > 0) for (int l = 0; l < lnum_dust; l++) s += (kabs_dust[l] * rf_dust[(size_t)lnum_dust * i + l]) * dl_dust[l]; abs_en = abs_en + s; }
    if (abs_en < 0) { bad = 1; continue; }
    double t_dust = 0;
    if (!dustem::zbrent_tdust(kabs_dust, abs_en, lnum_dust, lambda_si_dust, dl_dust, t_dust)) { bad = 2; continue; }
    for (int j = 0; j < lnum_dust; j++)
      dens_stars[(size_t)lnum_dust * i + j] = 4 * PI * kabs_dust[j] / kext_ref * dens_ref[i] * dustem::bplanck(t_dust, lambda_si_dust[j]) * pow(dustem::parsec, 2.);
    if (t_dust_out) t_dust_out[i] = t_dust;
  }
  if (bad) { c.err = bad == 1 ? "calc_t_dust_equil: negative absorbed energy" : "zbrent_tdust: root not bracketed / too many iterations"; return 2; }
  return 0;
}

int orc_prepare_scatt(orc_ctx* o, int iterations, double conv_en_lim, double* u_fest, const double* u_final,
                      double* lum_lost_prev, const double* lum_lost, double* const* scaspe, double* const* scaspe_prev,
                      double* const* scaspe_tot, double* tot_rad_en, int32_t* cnflag) {
  Ctx& c = o->c;
  const int L = c.L, N = c.N;
  memcpy(u_fest, u_final, sizeof(double) * (size_t)L * N);
  if (lum_lost_prev && lum_lost) memcpy(lum_lost_prev, lum_lost, sizeof(double) * L);
  if (iterations == 1)  // create_scaspe_tot rt_routines.f90:3012-3040: scaspe_tot_arr = scaspe_arr
    for (int l = 0; l < L; l++) memcpy(scaspe_tot[l], scaspe[l], sizeof(double) * (size_t)c.npix_arr[l] * N);
  for (int l = 0; l < L; l++) memcpy(scaspe_prev[l], scaspe[l], sizeof(double) * (size_t)c.npix_arr[l] * N);
  if (iterations == 1) {
    for (int i = 0; i < L; i++) {
      double s = 0;   // sum(scaspe_arr(k)%a(0:npix_hp-1,:)) : column-major order
      for (int cc = 0; cc < N; cc++) for (int p = 0; p < c.npix_hp_arr[i]; p++) s += scaspe[i][p + (size_t)c.npix_arr[i] * cc];
      tot_rad_en[i] = s;
    }
  }
  c.lumcell.assign((size_t)L * N, 0.0);
  for (int i = 0; i < N; i++) {
    if (c.cchild[i] != -1) continue;
    for (int j = 0; j < L; j++) {
      double s = 0;
      for (int p = 0; p < c.npix_hp_arr[j]; p++) s += scaspe[j][p + (size_t)c.npix_arr[j] * i];
      c.lumcell[j + (size_t)L * i] = s;
    }
  }
  double sum_tot = 0; for (int i = 0; i < L; i++) sum_tot += tot_rad_en[i];
  int flag;
  if (sum_tot > 0) {
    flag = 1;
    for (int i = 0; i < L; i++) {
      double tl = 0; for (int cc = 0; cc < N; cc++) tl += c.lumcell[i + (size_t)L * cc];
      if (!(tl / tot_rad_en[i] < conv_en_lim)) flag = 0;
    }
  } else flag = 1;
  *cnflag = flag;
  if (iterations > 1)
    for (int l = 0; l < L; l++) { size_t n = (size_t)c.npix_arr[l] * N; for (size_t k = 0; k < n; k++) scaspe_tot[l][k] += scaspe[l][k]; }
  for (int l = 0; l < L; l++) memset(scaspe[l], 0, sizeof(double) * (size_t)c.npix_arr[l] * N);
  return 0;
}

// calc_sed / sum_i_obs sed_routines.f90:90-146, 212-246: sed(l, j) = sum_c i_obs(l, c, j) * conv(l) * (csize(c)/dist_obs)^2
// (+ the point sources with the size of their host cell), sequential sums in cell order.
// units: 0 'erg/s/Hz/pc^2/sr', 1 'W/Hz/pc^2/sr', 2 'W/m/pc^2/sr' (needs lambda_si).  i_obs is (L, N+S, ndir) Fortran order.
int orc_calc_sed(orc_ctx* o, int32_t ndir, const double* i_obs, double dist_obs, int units, const double* lambda_si,
                 double* sed /*(L, ndir) Fortran order*/) {
  Ctx& c = o->c;
  const int L = c.L, N = c.N, S = c.S;
  const size_t NS = (size_t)N + S;
  const double parsec = (double)3.08567758e16f, cspeed = (double)2.99792458e8f;
  for (int l = 0; l < L; l++) {
    for (int j = 0; j < ndir; j++) {
      double s = 0, sp = 0;
      for (size_t q = 0; q < NS; q++) {
        double v = i_obs[((size_t)j * NS + q) * L + l];
        if (units == 0) { v = v * 1e-7; v = v * 1e26 / (parsec * parsec); }
        else if (units == 1) { v = v * 1e26 / (parsec * parsec); }
        else { v = v * (lambda_si[l] * lambda_si[l]) / cspeed; v = v * 1e26 / (parsec * parsec); }
        const double cs = (q < (size_t)N) ? c.csize[q] : c.csize[c.cell_src[q - N]];
        const double w = (cs / dist_obs) * (cs / dist_obs);
        if (q < (size_t)N) s += v * w; else sp += v * w;
      }
      sed[(size_t)j * L + l] = (S > 0) ? s + sp : s;
    }
  }
  return 0;
}

int orc_get_lumcell(orc_ctx* o, double* out) { memcpy(out, o->c.lumcell.data(), sizeof(double) * o->c.lumcell.size()); return 0; }

void orc_pix2ang_nest(int nside, int n, const int32_t* ipix, double* theta, double* phi) {
  hp_init();
  for (int i = 0; i < n; i++) pix2ang_nest(nside, ipix[i], theta[i], phi[i]);
}
void orc_ang2pix_nest(int nside, int n, const double* theta, const double* phi, int32_t* ipix) {
  hp_init();
  for (int i = 0; i < n; i++) ipix[i] = ang2pix_nest(nside, theta[i], phi[i]);
}
void orc_pix2vec(int nside, int n, const int32_t* ipix, double* vec) {
  hp_init();
  for (int i = 0; i < n; i++) { double th, ph; pix2ang_nest(nside, ipix[i], th, ph); calc_cproj(ph, th, vec + 3 * (size_t)i); }
}
int orc_cindex_roundtrip(orc_ctx* o, int32_t cell, int32_t* cell_out, int32_t* lvl_out) {
  int cc[3][MAXL + 1]; memset(cc, 0, sizeof(cc));
  cindex_to_ccindd(o->c, cell, o->c.lvl[cell], cc);
  int a, b; bool ok = ccindd_to_cc(o->c, cc, a, b);
  *cell_out = ok ? a : -1; *lvl_out = ok ? b : -1; return 0;
}
// find_cc_new2 for arbitrary points: src=(0,0,0), prev=1, cproj=point
int orc_locate_points(orc_ctx* o, int32_t npts, const double* xyz, const int32_t* cc_old, int32_t* cell_out) {
  Th t; th_init(t, o->c, DRE_RTT_DIR_CELL, 0, false);
  t.src_ccoord[0] = t.src_ccoord[1] = t.src_ccoord[2] = 0;
  for (int i = 0; i < npts; i++) {
    int cc_new = -1, clvl = 0; bool out = false;
    bool ok = find_cc_new2(t, cc_old[i], cc_new, xyz + 3 * (size_t)i, 1.0, out, clvl);
    cell_out[i] = (out || !ok) ? -1 : cc_new;
  }
  return 0;
}

}  // extern "C"

// =====================================================================================================
// External-observer surface-brightness maps: make_maps / map_projection (visual_routines.f90:59-177, 251-284,
// 287-309, 314-453, 456-566, 570-772, 801-863) + convert_maps_to_MJy_sr (:180-246).
// All wavelengths are mapped (ind_out_maps = every index, iq_maps_id(il) = il: io_routines.f90:2439-2445).
// Literal restatement; three things a reader of the Fortran would otherwise miss:
//  * `map_section = map_arr_hd_old(ix0:ix1, iy0:iy1, il)` (:689) assigns a 5x5 section to the 4x4 array
//    `map_section(0:ns-1,0:ns-1)`; gfortran scalarises over the explicit-shape left-hand side, so the filter
//    window is [ix-2, ix+1] x [iy-2, iy+1].
//  * the splat at :639 writes (2*(delta_p/2)+1)^2 pixels from a (delta_p+1)^2 texture section; identical
//    when delta_p is even (it is, for the symmetric textures built at :378-450); the left-hand extents are used.
//  * default-real literals: dtheta = 1E-3, the 1E-6 tests, cellsize = csize*0.99, (1+1E-7) are single precision.
namespace {

struct MapTex { int nt = 0, im0 = 0, im1 = 0; std::vector<double> prof; };   // prof2d(ix,iy) = prof[ix + nt*iy]

double scal_prod3(const double* a, const double* b) { double s = 0.0; for (int i = 0; i < 3; i++) s = s + a[i] * b[i]; return s; }

// calc_ray_inters :479-566 ; returns through flag_ray/p1/p2
void calc_ray_inters(int j, const double* rc, double cellsize, const double* ro, const double* obs_vec, int& flag_ray, double* p1, double* p2) {
  const int isel_ray = j / 2;
  const double dx = (j == 0 || j == 2 || j == 4) ? cellsize / 2. : -cellsize / 2.;
  double po[3];
  for (int i = 0; i < 3; i++) po[i] = rc[i] + ((i == isel_ray) ? 1.0 : 0.0) * dx;
  double num = 0, den = 0;
  for (int i = 0; i < 3; i++) { const double n = (i == isel_ray) ? 1.0 : 0.0; num += (po[i] - ro[i]) * n; den += obs_vec[i] * n; }
  if (den == 0.) return;
  const double dist = num / den;
  double pcoord[3], rel_vec[3];
  for (int i = 0; i < 3; i++) { pcoord[i] = ro[i] + dist * obs_vec[i]; rel_vec[i] = pcoord[i] - rc[i]; }
  int out = 0;
  for (int i = 0; i < 3; i++) if ((fabs(rel_vec[i]) > (double)(1.0f + 1E-7f) * cellsize / 2.) && i != isel_ray) out++;
  if (out > 0) return;
  if (flag_ray == 0) { for (int i = 0; i < 3; i++) p1[i] = pcoord[i]; flag_ray = 1; }
  else if (flag_ray == 1) { for (int i = 0; i < 3; i++) p2[i] = pcoord[i]; flag_ray = 2; }
}

// find_ray_cell_intersections :456-475
double ray_cell_length(const double* rc, double cellsize, const double* ro, const double* obs_vec) {
  int flag_ray = 0; double p1[3] = {0, 0, 0}, p2[3] = {0, 0, 0};
  for (int j = 0; j <= 5; j++) {
    calc_ray_inters(j, rc, cellsize, ro, obs_vec, flag_ray, p1, p2);
    if (flag_ray == 2) {
      double s = 0; for (int i = 0; i < 3; i++) s += (p2[i] - p1[i]) * (p2[i] - p1[i]);
      return pow(s, 0.5);
    }
  }
  return 0.0;
}

struct MapGeom { int npix, npix_hd; double size_map, pixel_size, pixel_size_hd, area_pixel, area_pixel_hd; };
MapGeom map_geom(const Ctx& c, int npixel_maps, double map_size_factor) {   // create_map_arrays :252-271
  MapGeom m; m.npix = npixel_maps; m.npix_hd = 4 * npixel_maps;
  m.size_map = map_size_factor * c.par.modelsize;
  m.pixel_size = m.size_map / m.npix; m.pixel_size_hd = m.size_map / m.npix_hd;
  m.area_pixel = m.pixel_size * m.pixel_size; m.area_pixel_hd = m.pixel_size_hd * m.pixel_size_hd;
  return m;
}

// create_texture_array :287-309 + calc_texture_array :314-453
void calc_texture_array(const Ctx& c, const MapGeom& m, double theta, double phi, double* obs_vec, double* xi_dir, double* yi_dir,
                        std::vector<MapTex>& tex) {
  double dtheta = (double)1E-3f;
  const double z_dir[3] = {0, 0, 1};
  obs_vec[0] = sin(theta) * cos(phi); obs_vec[1] = sin(theta) * sin(phi); obs_vec[2] = cos(theta);
  double z_zi = scal_prod3(obs_vec, z_dir);
  if (fabs(fabs(z_zi) - 1.) > (double)1E-6f) {
    for (int i = 0; i < 3; i++) yi_dir[i] = z_dir[i] - z_zi * obs_vec[i];
  } else {
    if (fabs(theta - PI) < (double)1E-6f) dtheta = -dtheta;
    obs_vec[0] = sin(theta + dtheta) * cos(phi); obs_vec[1] = sin(theta + dtheta) * sin(phi); obs_vec[2] = cos(theta + dtheta);
    z_zi = scal_prod3(obs_vec, z_dir);
    for (int i = 0; i < 3; i++) yi_dir[i] = z_dir[i] - z_zi * obs_vec[i];
    z_zi = scal_prod3(yi_dir, z_dir);
    for (int i = 0; i < 3; i++) yi_dir[i] = yi_dir[i] - z_zi * z_dir[i];
    obs_vec[0] = sin(theta - dtheta) * cos(phi); obs_vec[1] = sin(theta - dtheta) * sin(phi); obs_vec[2] = cos(theta - dtheta);
  }
  { double s = 0; for (int i = 0; i < 3; i++) s += yi_dir[i] * yi_dir[i]; const double n = sqrt(s); for (int i = 0; i < 3; i++) yi_dir[i] = yi_dir[i] / n; }
  xi_dir[0] = yi_dir[1] * obs_vec[2] - yi_dir[2] * obs_vec[1];          // cross_prod(yi_dir, obs_vec, xi_dir) :897-914
  xi_dir[1] = -(yi_dir[0] * obs_vec[2] - yi_dir[2] * obs_vec[0]);
  xi_dir[2] = yi_dir[0] * obs_vec[1] - yi_dir[1] * obs_vec[0];
  tex.assign(c.max_lvl + 1, MapTex());
  const double rc[3] = {0, 0, 0};
  for (int i = 1; i <= c.max_lvl; i++) {
    int den = c.base[0]; for (int k = 0; k < i - 1; k++) den *= c.base[1];
    const double lside = c.csize[0] / den;
    const int nt = (int)lround(m.npix_hd / m.size_map * lside * (double)2.5f);
    MapTex& t = tex[i]; t.nt = nt; t.prof.assign((size_t)nt * nt, 0.0);
    std::vector<double> xpp(nt);
    for (int ip = 0; ip < nt; ip++) {
      if (nt % 2 == 0) xpp[ip] = ip * m.pixel_size_hd - nt * m.pixel_size_hd / 2;
      else xpp[ip] = ip * m.pixel_size_hd - nt * m.pixel_size_hd / 2 + m.pixel_size_hd / 2;
    }
    const double cellsize = c.csize_arr[i] * (double)0.99f;
    for (int ix = 0; ix < nt; ix++)
      for (int iy = 0; iy < nt; iy++) {
        double ro[3];
        for (int a = 0; a < 3; a++) ro[a] = rc[a] + obs_vec[a] * c.par.modelsize + xpp[ix] * xi_dir[a] + xpp[iy] * yi_dir[a];
        const double length = ray_cell_length(rc, cellsize, ro, obs_vec);
        if (length > 0) t.prof[ix + (size_t)nt * iy] = length;
      }
    double tot = 0; for (size_t k = 0; k < t.prof.size(); k++) tot += t.prof[k];     // sum() in array element order
    for (size_t k = 0; k < t.prof.size(); k++) t.prof[k] = t.prof[k] / tot;
    t.im0 = 0; t.im1 = nt - 1;
    auto rowcol = [&](int ix) {
      double sr = 0, sc = 0;
      for (int k = 0; k < nt; k++) { sr += t.prof[ix + (size_t)nt * k]; sc += t.prof[k + (size_t)nt * ix]; }
      return sr > 0 || sc > 0;
    };
    for (int ix = 1; ix <= nt - 1; ix++) if (rowcol(ix)) { t.im0 = ix - 1; break; }
    for (int ix = nt - 1; ix >= 1; ix--) if (rowcol(ix)) { t.im1 = ix + 1; break; }
  }
}

// value_locate (sed_routines.f90:1662-1693, reverse_order = .FALSE.): ig0
int value_locate0(double val, const std::vector<double>& a) {
  const int num = (int)a.size();
  if (val <= a[0]) return 0;
  if (val >= a[num - 1]) return num - 2;
  for (int i = 0; i < num - 1; i++) if (val > a[i] && val <= a[i + 1]) return i;
  return 0;
}
// find_projected_point :839-863
void find_projected_point(const double* r, const double* obs_vec, const double* xi_dir, const double* yi_dir,
                          const std::vector<double>& xpp, int& ix_p, int& iy_p) {
  const double vec_n = scal_prod3(obs_vec, r);
  double ri[3]; for (int i = 0; i < 3; i++) ri[i] = r[i] - vec_n * obs_vec[i];
  ix_p = value_locate0(scal_prod3(ri, xi_dir), xpp);
  iy_p = value_locate0(scal_prod3(ri, yi_dir), xpp);
}
// median_section :801-836 on the 16 values of the window (any correct ascending sort gives the same array)
double median_section16(double* v, bool zero_value) {
  const int ntot = 16;
  std::sort(v, v + ntot);
  int i = 0; while (i < ntot && v[i] == 0) i++;
  int iel = 0;
  if (i < ntot && !(zero_value && i > (ntot - 1) / 2)) iel = (ntot - i) / 2 + i;
  return v[iel];
}


// ---- internal-observer maps: map_in_projection (visual_routines.f90:917-1076), calc_cube_texture_sphere (:1080-1182) ----
// neighbours of a NESTED pixel.  The reference's neighbours_nest (healpix_routines.f90:416-852) is the HEALPix library
// routine; the flood fill below only uses the SET of neighbours (the order changes nothing but the order of a sum), so the
// set is produced here from the library's face-adjacency tables ((x, y, face) arithmetic) instead of transcribing the
// 440-line case analysis; tests pin it with the nside = 1 table the reference hard-codes (:466-477) and with symmetry /
// adjacency invariants.  Returns the number of distinct neighbours (8, or 7 at the eight 3-face corners).
void nest2xyf(int nside, int ipix, int& ix, int& iy, int& face) {
  const int nsq = nside * nside;
  face = ipix / nsq;
  const int ipf = ipix % nsq;
  const int ip_low = ipf & 1023, ip_trunc = ipf / 1024, ip_med = ip_trunc & 1023, ip_hi = ip_trunc / 1024;   // pix2xy_nest :947-1010
  ix = 1024 * pix2x[ip_hi] + 32 * pix2x[ip_med] + pix2x[ip_low];
  iy = 1024 * pix2y[ip_hi] + 32 * pix2y[ip_med] + pix2y[ip_low];
}
int xyf2nest(int nside, int ix, int iy, int face) {      // xy2pix_nest :868-927
  return x2pix1[ix & 127] + y2pix1[iy & 127] + (x2pix1[ix / 128] + y2pix1[iy / 128]) * 16384 + face * nside * nside;
}
int neighbours_set(int nside, int ipix, int* out) {
  static const int xo[8] = {-1, -1, 0, 1, 1, 1, 0, -1}, yo[8] = {0, 1, 1, 1, 0, -1, -1, -1};
  static const int facearray[9][12] = {
      {8, 9, 10, 11, -1, -1, -1, -1, 10, 11, 8, 9},   // S
      {5, 6, 7, 4, 8, 9, 10, 11, 9, 10, 11, 8},       // SE
      {-1, -1, -1, -1, 5, 6, 7, 4, -1, -1, -1, -1},   // E
      {4, 5, 6, 7, 11, 8, 9, 10, 11, 8, 9, 10},       // SW
      {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11},         // centre
      {1, 2, 3, 0, 0, 1, 2, 3, 5, 6, 7, 4},           // NE
      {-1, -1, -1, -1, 7, 4, 5, 6, -1, -1, -1, -1},   // W
      {3, 0, 1, 2, 3, 0, 1, 2, 4, 5, 6, 7},           // NW
      {2, 3, 0, 1, -1, -1, -1, -1, 0, 1, 2, 3}};      // N
  static const int swaparray[9][3] = {{0, 0, 3}, {0, 0, 6}, {0, 0, 0}, {0, 0, 5}, {0, 0, 0}, {5, 0, 0}, {0, 0, 0}, {6, 0, 0}, {3, 0, 0}};
  int ix, iy, face; nest2xyf(nside, ipix, ix, iy, face);
  int n = 0;
  for (int i = 0; i < 8; i++) {
    int x = ix + xo[i], y = iy + yo[i], nb = 4;
    if (x < 0) { x += nside; nb -= 1; } else if (x >= nside) { x -= nside; nb += 1; }
    if (y < 0) { y += nside; nb -= 3; } else if (y >= nside) { y -= nside; nb += 3; }
    const int f = facearray[nb][face];
    if (f < 0) continue;
    const int bits = swaparray[nb][face >> 2];
    if (bits & 1) x = nside - x - 1;
    if (bits & 2) y = nside - y - 1;
    if (bits & 4) std::swap(x, y);
    const int p = xyf2nest(nside, x, y, f);
    bool dup = (p == ipix);
    for (int k = 0; k < n; k++) if (out[k] == p) dup = true;
    if (!dup) out[n++] = p;
  }
  return n;
}

// find_theta_phi_obs_in rt_routines.f90:3164-3217 with (ro, rc) = (a, b): angles of a - b
void theta_phi_of(const double* a, const double* b, double& theta, double& phi) {
  double vec[3]; for (int i = 0; i < 3; i++) vec[i] = a[i] - b[i];
  const double vec_mod = sqrt(vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2]);
  const double E8 = (double)1E-8f;
  if (vec_mod > E8) {
    double theta_ratio = (0 * vec[0] + 0 * vec[1] + 1 * vec[2]) / vec_mod;
    if (fabs(theta_ratio - 1.0) < E8) theta_ratio = 1.0;
    if (fabs(theta_ratio + 1.0) < E8) theta_ratio = -1.0;
    theta = acos(theta_ratio);
    const double vec_xy = sin(theta);
    double phi_ratio = (vec_xy > E8) ? vec[0] / (vec_mod * vec_xy) : 1.;
    if (fabs(phi_ratio - 1.0) < E8) phi_ratio = 1.0;
    if (fabs(phi_ratio + 1.0) < E8) phi_ratio = -1.0;
    phi = acos(phi_ratio);
    if (vec[1] < 0) phi = 2 * PI - phi;
  } else { theta = 0; phi = 0; }
}

// calc_cube_texture_sphere :1080-1182: flood fill over the HEALPix neighbours from the pixel of the cell centre while the
// pixel-centre lines of sight cross the cube; mproj = normalised chord lengths, iq_px = the pixels found (n_px)
void calc_cube_texture_sphere(double cellsize, const double* rc, const double* ro, int nside_map, std::vector<double>& mproj,
                              std::vector<int>& iq_px) {
  iq_px.clear();
  double theta, phi; theta_phi_of(rc, ro, theta, phi);
  const int ip = ang2pix_nest(nside_map, theta, phi);
  double obs_vec[3] = {sin(theta) * cos(phi), sin(theta) * sin(phi), cos(theta)};
  double length = ray_cell_length(rc, cellsize, ro, obs_vec);
  iq_px.push_back(ip);
  if (length > 0) mproj[ip] = length; else { mproj[ip] = 1; return; }
  std::vector<int> neighbour_list, pix_list;
  int nb[8]; int nn = neighbours_set(nside_map, ip, nb);
  neighbour_list.assign(nb, nb + nn);
  while (!neighbour_list.empty()) {
    pix_list.swap(neighbour_list); neighbour_list.clear();
    for (int ipix : pix_list) {
      if (mproj[ipix] > 0) continue;
      pix2ang_nest(nside_map, ipix, theta, phi);
      obs_vec[0] = sin(theta) * cos(phi); obs_vec[1] = sin(theta) * sin(phi); obs_vec[2] = cos(theta);
      length = ray_cell_length(rc, cellsize, ro, obs_vec);
      if (length > 0) {
        mproj[ipix] = length; iq_px.push_back(ipix);
        nn = neighbours_set(nside_map, ipix, nb);
        neighbour_list.insert(neighbour_list.end(), nb, nb + nn);
      }
    }
  }
  double s = 0; for (int p : iq_px) s += mproj[p];
  for (int p : iq_px) mproj[p] = mproj[p] / s;
}

}  // namespace

extern "C" {

// Textures of one line of sight (for tests): prof of level `level` (nt*nt, ix fastest), returns nt; info = {im0, im1}
int orc_map_texture(orc_ctx* o, int npixel_maps, double map_size_factor, double theta, double phi, int level,
                    double* prof, int cap, int32_t* info, double* axes /*obs_vec, xi_dir, yi_dir: 9*/) {
  const Ctx& c = o->c;
  const MapGeom m = map_geom(c, npixel_maps, map_size_factor);
  std::vector<MapTex> tex;
  calc_texture_array(c, m, theta, phi, axes, axes + 3, axes + 6, tex);
  if (level < 1 || level > c.max_lvl) return -1;
  const MapTex& t = tex[level];
  if (prof && (size_t)cap >= t.prof.size()) memcpy(prof, t.prof.data(), t.prof.size() * 8);
  if (info) { info[0] = t.im0; info[1] = t.im1; }
  return t.nt;
}

// make_maps :59-126 + 169-173.  i_obs (L, N+S, ndir), maps (npix, npix, L, ndir) Fortran order, units as orc_calc_sed.
int orc_make_maps(orc_ctx* o, int32_t ndir, const double* i_obs, int npixel_maps, double map_size_factor, int units,
                  const double* lambda_si, double* maps) {
  const Ctx& c = o->c;
  const int L = c.L, N = c.N, S = c.S;
  const size_t NS = (size_t)N + S;
  const MapGeom m = map_geom(c, npixel_maps, map_size_factor);
  const int nh = m.npix_hd, np = m.npix;
  const size_t plane_hd = (size_t)nh * nh, plane = (size_t)np * np;
  std::vector<double> hd(plane_hd * L), hd_old(plane_hd * L), tmp(plane * L), map(plane * L);
  std::vector<double> xpp_hd(nh), xpp(np);
  for (int ip = 0; ip < nh; ip++) xpp_hd[ip] = ip * m.pixel_size_hd - nh * m.pixel_size_hd / 2;     // :593-596
  for (int ip = 0; ip < np; ip++) xpp[ip] = ip * m.pixel_size - np * m.pixel_size / 2;              // :747-750
  const double parsec = (double)3.08567758e16f, cspeed = (double)2.99792458e8f;
  for (int idir = 0; idir < ndir; idir++) {
    const double theta = c.dir_i_out[idir], phi = c.dir_i_out[ndir + idir];      // dir_i_out(idir,1), (idir,2)
    double obs_vec[3], xi_dir[3], yi_dir[3];
    std::vector<MapTex> tex;
    calc_texture_array(c, m, theta, phi, obs_vec, xi_dir, yi_dir, tex);
    const double* io = i_obs + (size_t)idir * NS * L;
    std::fill(map.begin(), map.end(), 0.0);
    std::vector<double> flux_old(L), flux_new(L);
    for (int ilevel = 1; ilevel <= c.max_lvl; ilevel++) {       // map_projection :601-737
      std::fill(hd.begin(), hd.end(), 0.0); std::fill(tmp.begin(), tmp.end(), 0.0);
      const MapTex& t = tex[ilevel];
      for (int i = 0; i < N; i++) {
        if (c.cchild[i] != -1 || c.lvl[i] != ilevel) continue;
        double s = 0; for (int l = 0; l < L; l++) s += io[(size_t)i * L + l];
        if (s == 0) continue;
        int ix_p, iy_p;
        find_projected_point(&c.ccoord[3 * (size_t)i], obs_vec, xi_dir, yi_dir, xpp_hd, ix_p, iy_p);
        const int delta_p = t.im1 - t.im0;
        const int ix0 = ix_p - delta_p / 2, ix1 = ix_p + delta_p / 2, iy0 = iy_p - delta_p / 2, iy1 = iy_p + delta_p / 2;
        if (ix0 < 0 || ix1 > nh - 1 || iy0 < 0 || iy1 > nh - 1) continue;
        for (int il = 0; il < L; il++) {
          const double v = io[(size_t)i * L + il];
          for (int b = 0; b <= iy1 - iy0; b++)
            for (int a = 0; a <= ix1 - ix0; a++) {
              // im1 can reach nt for textures narrower than ~3 pixels (maps far below the npixel_maps >= 200 rule,
              // io_routines.f90:3207); the Fortran then reads past prof2d — guarded, the element counts as 0
              if (t.im0 + a >= t.nt || t.im0 + b >= t.nt) continue;
              hd[(size_t)il * plane_hd + (size_t)(iy0 + b) * nh + ix0 + a] +=
                  v * t.prof[(t.im0 + a) + (size_t)t.nt * (t.im0 + b)] * (c.csize[i] * c.csize[i]) / m.area_pixel_hd;
            }
        }
      }
      for (int il = 0; il < L; il++) { double s = 0; for (size_t k = 0; k < plane_hd; k++) s += hd[il * plane_hd + k]; flux_old[il] = s * m.area_pixel_hd; }
      hd_old = hd; std::fill(hd.begin(), hd.end(), 0.0);
      for (int ix = 0; ix < nh; ix++)
        for (int iy = 0; iy < nh; iy++) {
          double s = 0; for (int il = 0; il < L; il++) s += hd_old[il * plane_hd + (size_t)iy * nh + ix];
          const bool zero_value = (s == 0);
          const int ix0 = ix - 2, ix1 = ix + 2, iy0 = iy - 2, iy1 = iy + 2;
          if (ix0 < 0 || iy0 < 0 || ix1 > nh - 1 || iy1 > nh - 1) {
            for (int il = 0; il < L; il++) hd[il * plane_hd + (size_t)iy * nh + ix] = hd_old[il * plane_hd + (size_t)iy * nh + ix];
            continue;
          }
          for (int il = 0; il < L; il++) {
            double v[16];
            for (int b = 0; b < 4; b++) for (int a = 0; a < 4; a++) v[b * 4 + a] = hd_old[il * plane_hd + (size_t)(iy0 + b) * nh + ix0 + a];
            hd[il * plane_hd + (size_t)iy * nh + ix] = median_section16(v, zero_value);
          }
        }
      for (int ix = 0; ix < np; ix++)           // rebinning :702-720 (the all-zero shortcut at :707 gives the same zeros)
        for (int iy = 0; iy < np; iy++)
          for (int il = 0; il < L; il++) {
            double s = 0;
            for (int b = 0; b < 4; b++) for (int a = 0; a < 4; a++) s += hd[il * plane_hd + (size_t)(iy * 4 + b) * nh + ix * 4 + a];
            tmp[il * plane + (size_t)iy * np + ix] = s / 16;
          }
      for (int il = 0; il < L; il++) {
        double s = 0; for (size_t k = 0; k < plane; k++) s += tmp[il * plane + k];
        flux_new[il] = s * m.area_pixel;
        if (flux_new[il] > 0) for (size_t k = 0; k < plane; k++) map[il * plane + k] = map[il * plane + k] + tmp[il * plane + k] * flux_old[il] / flux_new[il];
      }
    }
    for (int i = 0; i < S; i++) {               // point sources :742-770
      const int ihost = c.cell_src[i];
      int ix_p, iy_p;
      find_projected_point(&c.ccoord_p_src[3 * (size_t)i], obs_vec, xi_dir, yi_dir, xpp, ix_p, iy_p);
      for (int il = 0; il < L; il++)
        map[il * plane + (size_t)iy_p * np + ix_p] += io[((size_t)N + i) * L + il] * (c.csize[ihost] * c.csize[ihost]) / m.area_pixel;
    }
    for (int il = 0; il < L; il++)               // map_arr_out(:,:,kall,idir) = map_arr ; convert_maps_to_MJy_sr :180-246
      for (size_t k = 0; k < plane; k++) {
        double v = map[il * plane + k];
        if (units == 0) { v = v * 1e-7; v = v * 1e20 / (parsec * parsec); }
        else if (units == 1) { v = v * 1e20 / (parsec * parsec); }
        else { v = v * (lambda_si[il] * lambda_si[il]) / cspeed; v = v * 1e20 / (parsec * parsec); }
        maps[((size_t)idir * L + il) * plane + k] = v;
      }
  }
  return 0;
}


// neighbours of a NESTED pixel as a set (see neighbours_set); returns the count
int orc_neighbours(int nside, int ipix, int32_t* out) { hp_init(); int nb[8]; const int n = neighbours_set(nside, ipix, nb); for (int i = 0; i < n; i++) out[i] = nb[i]; return n; }

// map_in_projection :917-1076 for every internal observer + convert_maps_to_MJy_sr.  i_obs_in (L, N+S, ndir_in),
// maps_in = map_in_arr_out(0:npix_maps-1, 0:L-1, 0:ndir_in-1) Fortran order, npix_maps = 12 * 4**kp_maps.
int orc_make_maps_in(orc_ctx* o, int32_t ndir_in, const double* ccoord_obs, const double* i_obs_in, int kp_maps, int units,
                     const double* lambda_si, double* maps_in) {
  hp_init();
  const Ctx& c = o->c;
  const int L = c.L, N = c.N, S = c.S;
  const size_t NS = (size_t)N + S;
  const int nside_map = 1 << kp_maps, npix = 12 * nside_map * nside_map;
  const double omega_hp = 4 * PI / npix;
  const double parsec = (double)3.08567758e16f, cspeed = (double)2.99792458e8f;
  std::vector<double> mproj(npix, 0.0), map((size_t)npix * L);
  std::vector<int> iq_px;
  for (int idir = 0; idir < ndir_in; idir++) {
    const double* ro = ccoord_obs + 3 * (size_t)idir;
    const double* io = i_obs_in + (size_t)idir * NS * L;
    std::fill(map.begin(), map.end(), 0.0);
    for (int i = 0; i < N; i++) {
      if (c.cchild[i] != -1) continue;
      double s = 0; for (int l = 0; l < L; l++) s += io[(size_t)i * L + l];
      if (s == 0) continue;
      const double* rc = &c.ccoord[3 * (size_t)i];
      double rel_vec[3]; for (int a = 0; a < 3; a++) rel_vec[a] = rc[a] - ro[a];
      double dist2 = 0; for (int a = 0; a < 3; a++) dist2 += rel_vec[a] * rel_vec[a];
      const double cellsize = c.csize[i];
      const double omega_cell = (dist2 != 0) ? cellsize * cellsize / dist2 : 4 * PI;
      const double omega_ratio = omega_cell / omega_hp;
      if (omega_ratio < 1) {                                    // case 1: one pixel holds the whole cell
        double theta, phi; theta_phi_of(rc, ro, theta, phi);
        const int ip = ang2pix_nest(nside_map, theta, phi);
        for (int il = 0; il < L; il++) map[(size_t)il * npix + ip] += io[(size_t)i * L + il] * omega_ratio;
        continue;
      }
      if (calc_outcube_sum(rel_vec, cellsize) == 0) {           // case 2: the observer is inside the cell
        for (int il = 0; il < L; il++) { const double v = io[(size_t)i * L + il] / 2.; for (int p = 0; p < npix; p++) map[(size_t)il * npix + p] += v; }
        continue;
      }
      calc_cube_texture_sphere(cellsize, rc, ro, nside_map, mproj, iq_px);     // case 3
      for (int il = 0; il < L; il++)
        for (int p : iq_px) map[(size_t)il * npix + p] += io[(size_t)i * L + il] * mproj[p] * omega_ratio;
      for (int p : iq_px) mproj[p] = 0;
    }
    for (int i = 0; i < S; i++) {                                // point sources :1038-1073
      const double* rc = &c.ccoord_p_src[3 * (size_t)i];
      double dist2 = 0; for (int a = 0; a < 3; a++) dist2 += (rc[a] - ro[a]) * (rc[a] - ro[a]);
      if (dist2 == 0) continue;
      const double cellsize = c.csize[c.cell_src[i]];
      const double omega_ratio = cellsize * cellsize / dist2 / omega_hp;
      double theta, phi; theta_phi_of(rc, ro, theta, phi);
      const int ip = ang2pix_nest(nside_map, theta, phi);
      for (int il = 0; il < L; il++) map[(size_t)il * npix + ip] += io[((size_t)N + i) * L + il] * omega_ratio;
    }
    for (int il = 0; il < L; il++)
      for (int p = 0; p < npix; p++) {
        double v = map[(size_t)il * npix + p];
        if (units == 0) { v = v * 1e-7; v = v * 1e20 / (parsec * parsec); }
        else if (units == 1) { v = v * 1e20 / (parsec * parsec); }
        else { v = v * (lambda_si[il] * lambda_si[il]) / cspeed; v = v * 1e20 / (parsec * parsec); }
        maps_in[((size_t)idir * L + il) * npix + p] = v;
      }
  }
  return 0;
}

}  // extern "C"

// =========================================================================================================================
// N5: the 2D mode.  rt_loop_2D (rt_routines.f90:1061-1140) = rt_loop with the source filter skip_cell_2D_rt_loop, run twice
// (i2d = 0, 1); after the first run fix_symmetry (:3605-3635) adds what the seven mirror images of every positive cell received
// and copies the result back to them.  The reference's scratch arrays are sized by the HEALPix part of the scaspe rows only and
// assigned from whole rows (:3720-3750), which conforms only when there are no observer directions (dartray_main_2D computes no
// i_obs, dartray_hub.f90:69): tot_ndir_scaspe = 0 is required here.
// =========================================================================================================================
namespace sym2d {

// find_linked_cells rt_routines.f90:3364-3578: the cells at the mirrored tree coordinates, in the order X, Y, Z, XY, XZ, YZ, XYZ
bool find_linked_cells(const Ctx& c, int nc, int link_list[7]) {
  const int nlevel = c.lvl[nc];
  int ccindd[3][MAXL + 1];
  memset(ccindd, 0, sizeof(ccindd));
  cindex_to_ccindd(c, nc, nlevel, ccindd);
  static const int flip[7][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 1, 0}, {1, 0, 1}, {0, 1, 1}, {1, 1, 1}};
  for (int s = 0; s < 7; s++) {
    int nw[3][MAXL + 1];
    memcpy(nw, ccindd, sizeof(nw));
    for (int i = 1; i <= nlevel; i++) {
      const int b = c.base[i > 1 ? 1 : 0];
      for (int a = 0; a < 3; a++) if (flip[s][a]) nw[a][i] = b - 1 - ccindd[a][i];
    }
    // ccindd_to_cc descends all max_lvl levels; the digits below nlevel are zero as in the reference's (3, nlevel) arrays
    int cc = -1, clvl = -1;
    int l = 1; bool found = false;
    for (int i = 1; i <= nlevel; i++) {
      const int b = c.base[i > 1 ? 1 : 0];
      l += (nw[2][i] * b + nw[1][i]) * b + nw[0][i];
      if (c.cchild[l] == -1) { cc = l; clvl = i; found = true; break; }
      l = c.cchild[l];
    }
    if (!found || clvl != nlevel) return false;      // 'something wrong here: find_linked_cells'
    link_list[s] = cc;
  }
  return true;
}

// minloc(abs(arr - x0), mask) - 1: first index of the smallest masked value (Fortran MINLOC)
int minloc_masked(const std::vector<double>& arr, double x0, const std::vector<char>* mask) {
  int best = -1; double bv = 0;
  for (int i = 0; i < (int)arr.size(); i++) {
    if (mask && !(*mask)[i]) continue;
    const double v = fabs(arr[i] - x0);
    if (best < 0 || v < bv) { best = i; bv = v; }
  }
  return best;
}

struct Idx { std::vector<int> ix, iy, iz, ixy, ixz, iyz, ixyz; };

// calc_scaspe_indices rt_routines.f90:4083-4217 for one resolution class: theta / phi of the HEALPix centres of nside 2^kp
Idx calc_scaspe_indices(int kp) {
  const int n = 12 * (1 << (2 * kp));
  std::vector<double> th(n), ph(n);
  for (int j = 0; j < n; j++) pix2ang_nest(1 << kp, j, th[j], ph[j]);
  Idx q;
  q.ix.assign(n, 0); q.iy.assign(n, 0); q.iz.assign(n, 0); q.ixy.assign(n, 0); q.ixz.assign(n, 0); q.iyz.assign(n, 0); q.ixyz.assign(n, 0);
  std::vector<char> mask(n);
  auto mask_eq = [&](const std::vector<double>& a, double v) { for (int j = 0; j < n; j++) mask[j] = (a[j] == v); };
  const double pi = PI;
  for (int i = 0; i < n; i++) {
    double phi0, theta0; int is;
    // X symmetry
    if (ph[i] >= 0 && ph[i] <= pi / 2) { phi0 = pi - ph[i]; mask_eq(th, th[i]); is = minloc_masked(ph, phi0, &mask); q.ix[i] = is; q.ix[is] = i; }
    if (ph[i] > pi && ph[i] <= (double)(3.f / 2.f) * pi) { phi0 = 2 * pi - (ph[i] - pi); mask_eq(th, th[i]); is = minloc_masked(ph, phi0, &mask); q.ix[i] = is; q.ix[is] = i; }
    // Y symmetry
    if (ph[i] >= 0 && ph[i] <= pi) { phi0 = 2 * pi - ph[i]; if (phi0 == 2 * pi) phi0 = 0; mask_eq(th, th[i]); is = minloc_masked(ph, phi0, &mask); q.iy[i] = is; q.iy[is] = i; }
    // Z symmetry
    if (th[i] >= 0 && th[i] <= pi / 2) { theta0 = pi - th[i]; mask_eq(ph, ph[i]); is = minloc_masked(th, theta0, &mask); q.iz[i] = is; q.iz[is] = i; }
    // XY symmetry
    if (ph[i] >= 0 && ph[i] < pi) { phi0 = ph[i] + pi; mask_eq(th, th[i]); is = minloc_masked(ph, phi0, &mask); q.ixy[i] = is; q.ixy[is] = i; }
    // XZ symmetry
    if (ph[i] >= 0 && ph[i] <= pi / 2) {
      phi0 = pi - ph[i]; theta0 = pi - th[i]; is = minloc_masked(th, theta0, nullptr); theta0 = th[is];
      mask_eq(th, theta0); is = minloc_masked(ph, phi0, &mask); q.ixz[i] = is; q.ixz[is] = i;
    }
    if (ph[i] > pi && ph[i] <= (double)(3.f / 2.f) * pi) {
      phi0 = 2 * pi - (ph[i] - pi); theta0 = pi - th[i]; is = minloc_masked(th, theta0, nullptr); theta0 = th[is];
      mask_eq(th, theta0); is = minloc_masked(ph, phi0, &mask); q.ixz[i] = is; q.ixz[is] = i;
    }
    // YZ symmetry
    if (ph[i] >= 0 && ph[i] <= pi) {
      phi0 = 2 * pi - ph[i]; if (phi0 == 2 * pi) phi0 = 0; theta0 = pi - th[i]; is = minloc_masked(th, theta0, nullptr); theta0 = th[is];
      mask_eq(th, theta0); is = minloc_masked(ph, phi0, &mask); q.iyz[i] = is; q.iyz[is] = i;
    }
    // XYZ symmetry
    if (ph[i] >= 0 && ph[i] < pi) {
      phi0 = ph[i] + pi; theta0 = pi - th[i]; is = minloc_masked(th, theta0, nullptr); theta0 = th[is];
      mask_eq(th, theta0); is = minloc_masked(ph, phi0, &mask); q.ixyz[i] = is; q.ixyz[is] = i;
    }
  }
  return q;
}

}  // namespace sym2d

extern "C" {

// the index arrays of calc_scaspe_indices for resolution class kp; which = 0..6 = X, Y, Z, XY, XZ, YZ, XYZ
int orc_scaspe_symmetry_index(int kp, int which, int32_t* out) {
  hp_init();
  const sym2d::Idx q = sym2d::calc_scaspe_indices(kp);
  const std::vector<int>* v[7] = {&q.ix, &q.iy, &q.iz, &q.ixy, &q.ixz, &q.iyz, &q.ixyz};
  for (size_t i = 0; i < v[which]->size(); i++) out[i] = (*v[which])[i];
  return (int)v[which]->size();
}
int orc_linked_cells(orc_ctx* o, int32_t nc, int32_t* link_list) {
  int ll[7];
  if (!sym2d::find_linked_cells(o->c, nc, ll)) return 1;
  for (int i = 0; i < 7; i++) link_list[i] = ll[i];
  return 0;
}

// fix_symmetry rt_routines.f90:3605-3635 with fix_symmetry_part1 (:3638-3670, rtt_precalc_cell), part2 (:3673-3810,
// rtt_dir_cell with iterations_dustem <= 1) and part3 (:3813-4025, rtt_scatt; the sequential_scattering branches), then
// lum_lost = (lum_lost - lum_lost_prev) * 8 + lum_lost_prev.  Arrays as in orc_run_pass (planes (P(l), N) pixel-fastest).
// Isotropic wavelengths in part2 (:3789-3799): the reference adds scaspe(:, cc) seven times with a cc that the loop does not
// set -- it is whatever the code above left in it, link_list(6) of this cell once any wavelength has been handled; that value is
// used here for every wavelength (it is undefined in the reference when the very first wavelength of the first cell is isotropic).
int orc_fix_symmetry(orc_ctx* o, int rt_type, double* u_fest, double* u_final, double* const* scaspe, double* lum_lost,
                     const double* lum_lost_prev) {
  Ctx& c = o->c;
  const int L = c.L, N = c.N;
  if (c.ndir != 0) { c.err = "fix_symmetry: the 2D mode needs tot_ndir_scaspe = 0"; return 3; }
  hp_init();
  std::vector<sym2d::Idx> idx(c.dim_npix_unique);
  for (int i = 0; i < c.dim_npix_unique; i++) idx[i] = sym2d::calc_scaspe_indices(c.kp_unique[i]);
  for (int i = 0; i < N; i++) {
    if (c.cchild[i] != -1) continue;
    const double* cc3 = &c.ccoord[3 * (size_t)i];
    if (cc3[0] < 0. || cc3[1] < 0. || cc3[2] < 0.) continue;
    int link[7];
    if (!sym2d::find_linked_cells(c, i, link)) { c.err = "something wrong here: find_linked_cells"; return 2; }
    for (int j = 0; j < L; j++) {
      if (rt_type == DRE_RTT_PRECALC_CELL) {                                   // part1
        double u_arr[7], s = 0;
        for (int q = 0; q < 7; q++) u_arr[q] = u_fest[j + (size_t)L * link[q]];
        for (int q = 0; q < 7; q++) s += u_arr[q];
        u_fest[j + (size_t)L * i] = u_fest[j + (size_t)L * i] + s;
        for (int q = 0; q < 7; q++) u_fest[j + (size_t)L * link[q]] = u_fest[j + (size_t)L * i];
        continue;
      }
      double u_arr[7], s = 0;
      if (rt_type == DRE_RTT_DIR_CELL) for (int q = 0; q < 7; q++) u_arr[q] = u_final[j + (size_t)L * link[q]];
      else for (int q = 0; q < 7; q++) {                                       // part3: delta over u_fest, negatives removed
        u_arr[q] = u_final[j + (size_t)L * link[q]] - u_fest[j + (size_t)L * link[q]];
        if (u_arr[q] < 0) u_arr[q] = 0;
      }
      for (int q = 0; q < 7; q++) s += u_arr[q];
      u_final[j + (size_t)L * i] = u_final[j + (size_t)L * i] + s;
      for (int q = 0; q < 7; q++) u_final[j + (size_t)L * link[q]] = u_final[j + (size_t)L * i];
      // scaspe
      const int ik = c.ik_sca[j], np = c.npix_arr[j];
      double* P = scaspe[j];
      if (ik != -1) {
        const sym2d::Idx& X = idx[ik];
        const std::vector<int>* perm[7] = {&X.ix, &X.iy, &X.iz, &X.ixy, &X.ixz, &X.iyz, &X.ixyz};
        std::vector<std::vector<double>> sx(7, std::vector<double>(np));
        for (int q = 0; q < 7; q++) {
          const double* src = P + (size_t)np * link[q];
          std::vector<double> tmp(src, src + np);
          for (int p = 0; p < np; p++) sx[q][p] = tmp[(*perm[q])[p]];
        }
        double* own = P + (size_t)np * i;
        for (int p = 0; p < np; p++) own[p] = own[p] + sx[0][p] + sx[1][p] + sx[2][p] + sx[3][p] + sx[4][p] + sx[5][p] + sx[6][p];
        std::vector<double> fix(own, own + np);
        for (int q = 0; q < 7; q++) {
          double* dst = P + (size_t)np * link[q];
          for (int p = 0; p < np; p++) dst[p] = fix[(*perm[q])[p]];
        }
      } else {
        double* own = P + (size_t)np * i;
        if (rt_type == DRE_RTT_DIR_CELL) {                                     // part2 quirk: cc is not set by the loop
          const int ccq = link[6];
          for (int icc = 0; icc < 7; icc++) for (int p = 0; p < np; p++) own[p] = own[p] + P[(size_t)np * ccq + p];
        } else {
          for (int icc = 0; icc < 7; icc++) for (int p = 0; p < np; p++) own[p] = own[p] + P[(size_t)np * link[icc] + p];
        }
        for (int icc = 0; icc < 7; icc++) for (int p = 0; p < np; p++) P[(size_t)np * link[icc] + p] = own[p];
      }
    }
  }
  if (lum_lost) for (int l = 0; l < L; l++) lum_lost[l] = (lum_lost[l] - lum_lost_prev[l]) * 8 + lum_lost_prev[l];
  return 0;
}

}  // extern "C"
