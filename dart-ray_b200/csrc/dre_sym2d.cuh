// dre_sym2d.cuh — N5, the 2D mode: fix_symmetry (rt_routines.f90:3605-3635) on the device-resident accumulators.
//
// rt_loop_2D (:1061-1140) traces only the sources of the positive octant (skip_cell_2D_rt_loop :1174-1187, applied by the cell-list
// kernel), then every leaf cell with non-negative coordinates receives what its seven mirror images got and hands the sum back to
// them: fix_symmetry_part1 (u_fest, :3638-3670), part2 (u_final + scaspe after the direct light, :3673-3810), part3 (the same on the
// increments of a scattering iteration, :3813-4025, sequential-scattering branches).  A scaspe row is mirrored by permuting its
// HEALPix pixels (calc_scaspe_indices :4083-4217).  One CTA per cell; the orbits of different cells are disjoint.
#pragma once
#include "dre_common.cuh"

struct DreSym {
  const int* perm[DRE_MAXCLASS];   // [7][nhp] per resolution class: pixel -> mirrored pixel for X, Y, Z, XY, XZ, YZ, XYZ
  double* rows;                    // [N][ROW] u_final + scaspe
  double* u_fest;                  // [N][L]
  int part;                        // 1, 2, 3
};

// find_linked_cells :3364-3578: the leaf at the mirrored tree coordinates (digit -> base - 1 - digit along the flipped axes)
__device__ inline int dre_linked_cell(const DreDev& g, int nc, int fx, int fy, int fz) {
  const int nlevel = g.lvl[nc];
  unsigned long long c[3];
  dre_decode(g, nc, nlevel, c);
  int l = 1;
  for (int i = 1; i <= nlevel; i++) {
    const int b = dre_base(g, i);
    int d0 = cx_get(c[0], i), d1 = cx_get(c[1], i), d2 = cx_get(c[2], i);
    if (fx) d0 = b - 1 - d0;
    if (fy) d1 = b - 1 - d1;
    if (fz) d2 = b - 1 - d2;
    l += (d2 * b + d1) * b + d0;
    const int ch = g.cchild[l];
    if (ch == -1) return (i == nlevel) ? l : -1;
    l = ch;
  }
  return -1;
}

#define SYM_MAXPIX 192     // 12 * 4^2: kp_sca <= 2 (dre_set_opacity)
__global__ void __launch_bounds__(128) k_fix_symmetry(DreSym q, int* __restrict__ err) {
  const DreDev& g = c_g;
  const int i = blockIdx.x;
  if (g.cchild[i] != -1) return;
  const double* cc3 = g.ccoord + 3 * (size_t)i;
  if (cc3[0] < 0. || cc3[1] < 0. || cc3[2] < 0.) return;
  __shared__ int link[7];
  __shared__ double fix[SYM_MAXPIX];
  if (threadIdx.x < 7) {
    const int s = threadIdx.x;
    const int fx = (s == 0 || s == 3 || s == 4 || s == 6), fy = (s == 1 || s == 3 || s == 5 || s == 6), fz = (s == 2 || s == 4 || s == 5 || s == 6);
    link[s] = dre_linked_cell(g, i, fx, fy, fz);
    if (link[s] < 0) atomicCAS(err, 0, 9);
  }
  __syncthreads();
  for (int s = 0; s < 7; s++) if (link[s] < 0) return;
  const int L = g.L, ROW = g.ROW;
  // ---- radiation field: u(i) += sum over the links (part3: of their increments over u_fest), then copied to the links
  if (threadIdx.x < L) {
    const int l = threadIdx.x;
    double* U = (q.part == 1) ? q.u_fest : q.rows;
    const size_t st = (q.part == 1) ? (size_t)L : (size_t)ROW;
    double ua[7], s = 0;
    for (int k = 0; k < 7; k++) {
      ua[k] = U[st * link[k] + l];
      if (q.part == 3) { ua[k] = ua[k] - q.u_fest[(size_t)L * link[k] + l]; if (ua[k] < 0) ua[k] = 0; }   // remove_negative
    }
    for (int k = 0; k < 7; k++) s += ua[k];
    const double v = U[st * i + l] + s;
    U[st * i + l] = v;
    for (int k = 0; k < 7; k++) U[st * link[k] + l] = v;
  }
  if (q.part == 1) return;
  __syncthreads();
  // ---- scattered-luminosity rows, one wavelength at a time
  for (int l = 0; l < L; l++) {
    const int off = g.off[l], np = g.npix[l], ik = g.ik_sca[l];
    double* own = q.rows + (size_t)ROW * i + off;
    if (ik != -1) {
      const int* P = q.perm[ik];
      for (int p = threadIdx.x; p < np; p += blockDim.x) {
        double v = own[p];
        for (int k = 0; k < 7; k++) v = v + q.rows[(size_t)ROW * link[k] + off + P[k * np + p]];
        fix[p] = v;
      }
      __syncthreads();
      for (int p = threadIdx.x; p < np; p += blockDim.x) own[p] = fix[p];
      for (int k = 0; k < 7; k++) {                 // in the reference's order: a link that is the cell itself is overwritten last
        __syncthreads();
        double* dst = q.rows + (size_t)ROW * link[k] + off;
        for (int p = threadIdx.x; p < np; p += blockDim.x) dst[p] = fix[P[k * np + p]];
      }
      __syncthreads();
    } else if (threadIdx.x == 0) {
      // isotropic wavelength (one element).  part2 adds the row of `cc` seven times where cc is not set by the loop
      // (:3789-3793): it holds link_list(6) of this cell once any wavelength has been handled -- that value is used.
      for (int p = 0; p < np; p++) {
        for (int k = 0; k < 7; k++) own[p] = own[p] + q.rows[(size_t)ROW * ((q.part == 2) ? link[6] : link[k]) + off + p];
        for (int k = 0; k < 7; k++) q.rows[(size_t)ROW * link[k] + off + p] = own[p];
      }
    }
  }
}
// lum_lost = (lum_lost - lum_lost_prev) * 8 + lum_lost_prev (:3627)
__global__ void k_lum_lost_x8(double* __restrict__ ll, const double* __restrict__ lp, int L) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l < L) ll[l] = (ll[l] - lp[l]) * 8 + lp[l];
}
