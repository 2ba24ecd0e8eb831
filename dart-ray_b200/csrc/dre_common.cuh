// dre_common.cuh — device-side data model of the B200 ray-tracing engine.
//
// All reference citations are file:line under /root/reference (DART-Ray).  Geometry code in
// this translation unit is compiled with -fmad=false: gfortran on baseline x86-64 emits no FMA
// and the intersection lists must be bit-exact.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/dartray_engine.h"
#include "../../include/dartray_hash.h"

#ifndef DRE_RARE
#define DRE_RARE __forceinline__   // measured: __noinline__ on the rare paths costs 25% (call ABI spills), see profiles/
#endif
#define DRE_MAXLVL 8        // tree levels packed 8 bits each into one 64-bit word per axis
#define DRE_MAXK 13         // log2 of the largest HEALPix nside (8192, healpix_routines.f90:9)
#define DRE_KL 4            // wavelengths per lane: lnum <= 32*DRE_KL
#ifndef DRE_TRACE_THREADS
#define DRE_TRACE_THREADS 512   // threads per CTA of the ray-march kernel (register budget = 65536 / threads)
#endif
#define DRE_MAXCLASS 4      // scattering-resolution classes kp = 0..3 (set_npix_arr rt_routines.f90:5289)
#define DRE_NSIDE_MIN_K 2   // nside_min = 4, rt_routines.f90:60

// ray types (ray_list.f90:48-59) and list-entry "kind" of a task outcome
enum { RT_HIGH = 0, RT_LOW = 1, RT_RECO = 2, RT_GONE = 3, RT_IOBS = 4, RT_IOBS_IN = 5 };
enum { OUT_NONE = 0, OUT_HIGH = 1, OUT_LOW = 2 };
enum { MODE_PRECALC = 0, MODE_DEP = 1, MODE_IOBS = 2 };

// single-precision constants promoted to double (smooth_grid_routines.f90:311-317, rt_routines.f90:76)
#define DRE_HALFPI ((double)1.57079637050628662109375f)
#define DRE_PI (2.0 * DRE_HALFPI)
#define DRE_GLEPSILON ((double)1.0e-7f)
#define DRE_TINY ((double)1E-008f)

// One direction of the HEALPix table: (sin theta, cos theta, sin phi, cos phi) of pix2ang_nest
// (healpix_routines.f90:30-136), evaluated on the host with libm so that calc_cproj
// (rt_routines.f90:1945-1953) is reproduced bit for bit.
struct __align__(32) DreDir { double st, ct, sp, cp; };

// Per-cell traversal record (64 B, one L2 line half): centre + the six face neighbours.
// nbr[2*axis + (inc>0)] is the result of the reference's neighbour step restricted to the levels
// <= lvl(cell) -- `increment` (smooth_grid_routines.f90:751-769) followed by the tree descent
// `ccindd_to_cc` (:820-852), which for those levels depends on the integer tree coordinates only:
//   DRE_NBR_OUT               the ray leaves the model (carry out of level 1, rt_routines.f90:1790)
//   bit 31 clear              leaf reached at level (e >> 27) & 15 <= lvl(cell), id = e & 0x7ffffff
//   bit 31 set                the same-level neighbour is subdivided; e & 0x7fffffff = its first child:
//                             the descent continues with digits taken from the exit point (:1807-1838)
// Built on the device by k_build_nbr at dre_set_grid.
#define DRE_NBR_OUT 0xffffffffu
struct __align__(64) DreCell { double c[3]; unsigned nbr[6]; int lvl; int pad; double csize; };   // csize = csize(cell), not csize_arr(lvl)

// Static inputs in HBM (SoA).  Passed to kernels by value (lives in the constant bank).
struct DreDev {
  // grid: smooth_grid_routines.f90:41-70
  int N, max_lvl, base[2];
  long long basediv[2];
  int basemask[2];
  int baseshift[2];         // log2(basediv)
  const int* cchild;        // [N]
  const unsigned char* lvl; // [N]
  const long long* cindex;  // [N]
  const double* ccoord;     // [N][3]
  const double* csize;      // [N]
  const DreCell* cell;      // [N] traversal records
  double csize_arr[DRE_MAXLVL + 1], carea_arr[DRE_MAXLVL + 1], cvol_arr[DRE_MAXLVL + 1];
  // opacity / scattering classes
  int L, KLn;               // wavelengths, ceil(L/32)
  int ndir;                 // tot_ndir_scaspe
  const double* dens;       // [N][L]
  const double* ksca;       // [L] ksca_arr_norm
  const double* gsca;       // [L]
  const int* ik_sca;        // [L] class index or -1
  const int* npix_hp;       // [L]
  const int* npix;          // [L] npix_hp + ndir
  const int* off;           // [L] offset of wavelength l inside an accumulator row
  int off_c[32 * DRE_KL], npix_c[32 * DRE_KL];   // the same tables in the constant bank (warp-uniform lookups)
  const unsigned short* lamof; // [ROW] scaspe element -> byte offset 8*l of its wavelength in the per-warp en_sca array
  int ROW;                  // doubles per accumulator row: [0,L) u_final, then scaspe blocks (even-padded)
  int nclass;
  int class_kp[DRE_MAXCLASS], class_np[DRE_MAXCLASS], class_nhp[DRE_MAXCLASS];
  const DreDir* class_dir[DRE_MAXCLASS];   // [class_np] scaspe directions (create_scaspe rt_routines.f90:2827-2848)
  // direction tables per log2(nside)
  const DreDir* dirtab[DRE_MAXK + 1];
  int dir_kmax;             // largest k with a table
  // phase-function rows per direction (calc_ffn_arr for every pixel of nside 2^k), accumulator-row layout
  const double* ftab[DRE_MAXK + 1];
  // sources
  int S;
  const double* lumcell;    // [N][L]
  const double* ccoord_p_src; // [S][3]
  const double* lum_p_src;  // [S][L]
  const int* src_cell;      // [N]
  // params
  dre_params par;
  int rt_type;
  int i2d;                  // 2D mode source filter (skip_cell_2D_rt_loop rt_routines.f90:1174-1187): 0 off, 1 = first loop, 2 = second loop
};

// The static inputs of the current pass live in constant memory (one engine per device at a time).
__constant__ DreDev c_g;

__device__ __forceinline__ int dre_base(const DreDev& g, int lev) { return g.base[lev > 1 ? 1 : 0]; }

// ---- packed tree coordinates: digit of level j (1-based) lives in bits [8(j-1), 8j) -----
__device__ __forceinline__ int cx_get(unsigned long long c, int j) { return (int)((c >> (8 * (j - 1))) & 0xffull); }
__device__ __forceinline__ unsigned long long cx_set(unsigned long long c, int j, int v) {
  const int s = 8 * (j - 1);
  return (c & ~(0xffull << s)) | ((unsigned long long)(unsigned)v << s);
}

__device__ __forceinline__ long long dre_ipow64(long long b, int e) { long long r = 1; for (int i = 0; i < e; i++) r *= b; return r; }

// cindex_to_ccindd smooth_grid_routines.f90:793-813 (basediv are powers of two: divisions become shifts)
__device__ DRE_RARE void dre_decode(const DreDev& g, int nc, int clvl, unsigned long long c[3]) {
  unsigned long long c0 = 0, c1 = 0, c2 = 0;
  const unsigned long long ci = (unsigned long long)g.cindex[nc];
  for (int i = 1; i <= clvl; i++) {
    int ib, j;
    if (i > 1) { ib = 1; j = (int)((ci >> (g.baseshift[0] + g.baseshift[1] * (i - 2))) & (unsigned)g.basemask[1]) - 1; }
    else { ib = 0; j = (int)(ci & (unsigned)g.basemask[0]) - 1; }
    const int b = g.base[ib];
    c0 = cx_set(c0, i, j % b);
    c1 = cx_set(c1, i, (j / b) % b);
    c2 = cx_set(c2, i, (j / (b * b)) % b);
  }
  c[0] = c0; c[1] = c1; c[2] = c2;
}

// ccindd_to_cc smooth_grid_routines.f90:820-852 ; returns false if no leaf was found
__device__ __forceinline__ bool dre_descend(const DreDev& g, const unsigned long long c[3], int& cc, int& clvl) {
  int l = 1;
  for (int i = 1; i <= g.max_lvl; i++) {
    const int b = dre_base(g, i);
    l += (cx_get(c[2], i) * b + cx_get(c[1], i)) * b + cx_get(c[0], i);
    const int ch = __ldg(&g.cchild[l]);
    if (ch == -1) { cc = l; clvl = i; return true; }
    l = ch;
  }
  return false;
}

// calc_outcube rt_routines.f90:2072-2085 (sum of the three flags)
__device__ __forceinline__ int dre_outcube(const double r[3], double cellsize) {
  int s = 0;
  const double lim = (double)(1.0f + 1E-7f) * cellsize / 2.0;
  for (int i = 0; i < 3; i++) if (fabs(r[i]) > lim) s++;
  return s;
}

// find_cc_new2 rt_routines.f90:1962-2066.  Returns 0 ok, 1 outside the model, 2 reference STOP.
__device__ DRE_RARE int dre_locate(const DreDev& g, const double src[3], const double d[3], double prev,
                                          int cc_old, int& cc_new, int& clvl) {
  double p[3], rel[3], q[3];
  for (int i = 0; i < 3; i++) p[i] = src[i] + prev * d[i];
  for (int i = 0; i < 3; i++) if (fabs(p[i]) > g.par.modelsize / 2.) return 1;
  const DreCell* co = &g.cell[cc_old];
  for (int i = 0; i < 3; i++) rel[i] = p[i] - co->c[i];
  const int lold = co->lvl;
  if (dre_outcube(rel, co->csize) == 0) { clvl = lold; cc_new = cc_old; return 0; }   // csize(cc_old), :1991
  for (int i = 0; i < 3; i++) q[i] = p[i] + g.par.modelsize / 2.;
  int l = 1; bool found = false;
  for (int i = 1; i <= g.max_lvl; i++) {
    const int b = dre_base(g, i);
    const double cs = g.csize_arr[i];
    int v[3];
    for (int a = 0; a < 3; a++) {
      v[a] = (int)(q[a] / cs);
      if (v[a] == b) v[a] = b - 1;
      q[a] = q[a] - v[a] * cs;
    }
    l += (v[2] * b + v[1]) * b + v[0];
    const int ch = __ldg(&g.cchild[l]);
    if (ch == -1) { cc_new = l; clvl = i; found = true; break; }
    l = ch;
  }
  if (!found) return 2;
  for (int i = 0; i < 3; i++) rel[i] = p[i] - g.cell[cc_new].c[i];
  return dre_outcube(rel, g.csize_arr[clvl]) == 0 ? 0 : 2;
}

// ---- SoA ray list (ray_list.f90:7-27).  A direction is always a HEALPix pixel centre, so
// (theta,phi) is stored as (dir_k = log2 nside, dir_ipix); ang2pix_nest at a coarser nside is
// the NESTED parent shift (tests/ pin this against the oracle's ang2pix_nest).
struct DreList {
  int* unit;          // owning work unit (index inside the batch)
  int* dir_ipix;
  unsigned* meta;     // dir_k | ns_k<<8 | isel<<16 | type<<24
  int* cc_old;
  double* dplane;
  double* prev;
  double* lum;        // [cap][L]
  int* seg;           // [U+1] offsets per unit (lists only)
  int n, cap;
};
__device__ __host__ __forceinline__ unsigned dre_meta(int dir_k, int ns_k, int isel, int type) {
  return (unsigned)dir_k | ((unsigned)ns_k << 8) | ((unsigned)isel << 16) | ((unsigned)type << 24);
}
#define META_DIRK(m) ((int)((m) & 0xff))
#define META_NSK(m) ((int)(((m) >> 8) & 0xff))
#define META_ISEL(m) ((int)(((m) >> 16) & 0xff))
#define META_TYPE(m) ((int)(((m) >> 24) & 0xff))

// One ray_tracing call of the wavefront (main_dir_loop rt_routines.f90:1378-1431).
struct DreTask {
  int unit;      // batch-local unit
  int e0;        // source entry: index into the old HIGH list (kind 0,1) or LOW list (kind 2)
  int e1, e2, e3;// merged siblings (kind 2)
  int kind;      // 0 high child (child = sub), 1 reco, 2 low (merged)
  int sub;       // child 0..3
  int pad;
};

struct DreUnits {
  int* nc;        // [U] source cell
  int* gid;       // [U] global unit id = source_ordinal*48 + sector (record/hash key)
  int* nk;        // [U] log2 of the current nside
  int* step;      // [U]
  int* mode;      // [U] 0 done, 1 extract from HIGH, 2 extract from LOW
  int* task_seg;  // [U+1]
  int U;
};

struct DreAccum {
  double* acc;          // [N][ROW] u_final + scaspe (dir/scatt passes)
  double* u_fest;       // [N][L]  (read in dep passes, accumulated in precalc passes)
  const double* prev;   // [N][ROW] scaspe_prev (scatt) / scaspe_tot (i_obs)
  double* lum_lost;     // [L]
  double* psel_sum;     // [N+S]
  long long* ipsel;     // [N+S]
  unsigned long long* counters; // [16]
  dre_isect* rec; long long rec_cap; unsigned long long* rec_n;
  int* err;             // first reference STOP condition hit
  int stage_split;      // the deposit row is staged and reduced in this many chunks (smem vs occupancy)
  int want_hash;        // accumulate the intersection checksum (parity runs)
  int ring_slots;       // stage buffer used as 1 slot (wait for all reads when it wraps) or as a 2-slot ring
  int use_bulk;         // 1: TMA bulk reduce of the whole row; 0: per-lane red.global.add.f64
  // Order-independent accumulation (option "deterministic"): every contribution is first rounded to a multiple of a
  // per-wavelength quantum q = 2^(e-52), 2^e >= an a-priori bound of the sum, so that all partial sums are exact in
  // fp64 whatever the order of the reductions.  qc holds the rounding constants C = 1.5 * 2^52 * q (rq(v) = (v + C) - C):
  // [0,Lq) scaspe, [Lq,2Lq) u_final / u_fest, [2Lq,3Lq) lum_lost, [3Lq] psel_sum; Lq = L rounded up to even.
  int det;
  const double* qc;
};
__host__ __device__ inline int dre_lq(int L) { return L + (L & 1); }
enum { CNT_RAYS = 0, CNT_ISECT, CNT_DEP, CNT_BEAM, CNT_SPLIT, CNT_MERGE, CNT_SCA, CNT_ACT, CNT_HASH, CNT_N, CNT_ACTDEP, CNT_DESC };
