// dre_maps.cuh — external-observer surface-brightness maps from the device-resident i_obs array:
// make_maps / map_projection of DART-Ray (visual_routines.f90:59-177, 570-772) as CUDA kernels.
//
// Per line of sight and per subdivision level the reference (1) splats every leaf cell of that level onto a 4x
// oversampled map with the level's normalised projected-cube texture, (2) median-filters the oversampled map
// over a 4x4 window, (3) rebins 4x4 -> 1, (4) renormalises to the flux before filtering and adds the level to
// the output map; point sources are added to one pixel each.  The textures (a few hundred ray-cube
// intersections per level, libm sin/cos/pow) are built on the host in dre_engine.cu; everything that scales
// with cells x pixels x wavelengths runs here.  Map layout = the Fortran arrays: x fastest, then y, then wavelength.
#pragma once
#include "dre_common.cuh"

struct DreMapGeom {
  int npix, npix_hd;                 // npixel_maps, npixel_maps_hd = 4 * npixel_maps (hd_xfactor, visual_routines.f90:40)
  double area_pixel, area_pixel_hd;  // :268-269
  double obs_vec[3], xi_dir[3], yi_dir[3];   // calc_texture_array :352-376
};
struct DreMapTex { const double* prof; int nt, im0, im1; };   // prof2d(ix, iy) = prof[ix + nt * iy]

// value_locate (sed_routines.f90:1662-1693, ascending): i with a[i] < val <= a[i+1], clamped at both ends.
// The table holds the reference's own floating-point pixel edges, so the index is the one its linear search finds.
__device__ __forceinline__ int dre_value_locate(double val, const double* __restrict__ a, int num) {
  if (val <= a[0]) return 0;
  if (val >= a[num - 1]) return num - 2;
  int lo = 0, hi = num - 1;          // a[lo] < val <= a[hi]
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (val > a[mid]) lo = mid; else hi = mid; }
  return lo;
}
// find_projected_point :839-863
__device__ __forceinline__ void dre_project_point(const DreMapGeom& m, const double* r, const double* __restrict__ xpp, int num, int& ix_p, int& iy_p) {
  double vec_n = 0.0;
  for (int i = 0; i < 3; i++) vec_n = vec_n + m.obs_vec[i] * r[i];
  double ri[3];
  for (int i = 0; i < 3; i++) ri[i] = r[i] - vec_n * m.obs_vec[i];
  double xi_p = 0.0, yi_p = 0.0;
  for (int i = 0; i < 3; i++) xi_p = xi_p + ri[i] * m.xi_dir[i];
  for (int i = 0; i < 3; i++) yi_p = yi_p + ri[i] * m.yi_dir[i];
  ix_p = dre_value_locate(xi_p, xpp, num);
  iy_p = dre_value_locate(yi_p, xpp, num);
}

// Leaf cells of one level with a non-zero i_obs (:619): compacted so that the splat grid has no idle blocks.
__global__ void k_map_select(const double* __restrict__ io /*(L, N+S) of this direction*/, int level, int* __restrict__ list, int* __restrict__ n) {
  const DreDev& g = c_g;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  if (g.cchild[i] != -1 || g.lvl[i] != level) return;
  double s = 0;
  for (int l = 0; l < g.L; l++) s += io[(size_t)i * g.L + l];
  if (s == 0) return;
  list[atomicAdd(n, 1)] = i;
}

// Splat (:622-640): one block per selected cell; threads run over the texture section (x fastest), wavelengths inside.
__global__ void k_map_splat(DreMapGeom m, DreMapTex t, const double* __restrict__ io, const double* __restrict__ xpp_hd,
                            const int* __restrict__ list, double* __restrict__ hd) {
  const DreDev& g = c_g;
  const int i = list[blockIdx.x];
  __shared__ int s_ix0, s_iy0, s_ok;
  if (threadIdx.x == 0) {
    const double r[3] = {g.ccoord[3 * (size_t)i], g.ccoord[3 * (size_t)i + 1], g.ccoord[3 * (size_t)i + 2]};
    int ix_p, iy_p;
    dre_project_point(m, r, xpp_hd, m.npix_hd, ix_p, iy_p);
    const int delta_p = t.im1 - t.im0;
    const int ix0 = ix_p - delta_p / 2, ix1 = ix_p + delta_p / 2, iy0 = iy_p - delta_p / 2, iy1 = iy_p + delta_p / 2;
    s_ix0 = ix0; s_iy0 = iy0;
    s_ok = !(ix0 < 0 || ix1 > m.npix_hd - 1 || iy0 < 0 || iy1 > m.npix_hd - 1);     // :635
  }
  __syncthreads();
  if (!s_ok) return;
  const int w = 2 * ((t.im1 - t.im0) / 2) + 1;       // extents of the left-hand side section ix0:ix1
  const double cs = g.csize[i];
  const double cs2 = cs * cs;
  const size_t plane = (size_t)m.npix_hd * m.npix_hd;
  for (int k = threadIdx.x; k < w * w; k += blockDim.x) {
    const int a = k % w, b = k / w;
    // im1 can be nt (a texture narrower than ~3 pixels, below the reference's npixel_maps >= 200 rule): the reference
    // then reads past prof2d; the index is guarded here and the element counts as 0
    if (t.im0 + a >= t.nt || t.im0 + b >= t.nt) continue;
    const double p = t.prof[(t.im0 + a) + (size_t)t.nt * (t.im0 + b)];
    if (p == 0) continue;                                // adds exactly zero
    double* dst = hd + (size_t)(s_iy0 + b) * m.npix_hd + (s_ix0 + a);
    for (int il = 0; il < g.L; il++) {
      const double v = io[(size_t)i * g.L + il] * p * cs2 / m.area_pixel_hd;
      if (v != 0) atomicAdd(dst + il * plane, v);
    }
  }
}

// sum over a plane (x, y) per wavelength, times `scale` (tot_flux_old :659, tot_flux_new :727): one block per wavelength,
// fixed-order strided partial sums + tree.
__global__ void k_map_flux(const double* __restrict__ map, size_t plane, double scale, double* __restrict__ out) {
  __shared__ double sh[512];
  const double* p = map + blockIdx.x * plane;
  double s = 0;
  for (size_t k = threadIdx.x; k < plane; k += blockDim.x) s += p[k];
  sh[threadIdx.x] = s; __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) out[blockIdx.x] = sh[0] * scale;
}

// Median smoothing (:670-694, median_section :801-836).  The window is [ix-2, ix+1] x [iy-2, iy+1]: the reference assigns
// the 5x5 section (ix-2:ix+2, iy-2:iy+2) to a 4x4 array and gfortran scalarises over the 4x4 left-hand side.
__global__ void k_map_median(const double* __restrict__ old, double* __restrict__ out, int nh, int L) {
  const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y;
  if (ix >= nh) return;
  const size_t plane = (size_t)nh * nh, at = (size_t)iy * nh + ix;
  double s = 0;
  for (int il = 0; il < L; il++) s += old[il * plane + at];
  const bool zero_value = (s == 0);
  const bool edge = (ix - 2 < 0 || iy - 2 < 0 || ix + 2 > nh - 1 || iy + 2 > nh - 1);
  for (int il = 0; il < L; il++) {
    const double* p = old + il * plane;
    if (edge) { out[il * plane + at] = p[at]; continue; }
    double v[16];
#pragma unroll
    for (int b = 0; b < 4; b++)
#pragma unroll
      for (int a = 0; a < 4; a++) v[b * 4 + a] = p[(size_t)(iy - 2 + b) * nh + (ix - 2 + a)];
    // ascending sort of 16 values (odd-even transposition network, fully unrolled: stays in registers)
#pragma unroll
    for (int r = 0; r < 16; r++) {
#pragma unroll
      for (int k = (r & 1); k + 1 < 16; k += 2) {
        const double lo = fmin(v[k], v[k + 1]), hi = fmax(v[k], v[k + 1]);
        v[k] = lo; v[k + 1] = hi;
      }
    }
    int i = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) i += (v[k] == 0) ? 1 : 0;      // values are >= 0: the zeros lead the sorted array
    int iel = 0;
    if (i < 16 && !(zero_value && i > 7)) iel = (16 - i) / 2 + i;
    double res = v[0];
#pragma unroll
    for (int k = 0; k < 16; k++) if (k == iel) res = v[k];
    out[il * plane + at] = res;
  }
}

// Rebinning (:702-720): 4x4 -> 1, summed in array-element order (x fastest), / hd_xfactor**2
__global__ void k_map_rebin(const double* __restrict__ hd, double* __restrict__ tmp, int np, int L) {
  const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y, il = blockIdx.z;
  if (ix >= np) return;
  const int nh = 4 * np;
  const double* p = hd + (size_t)il * nh * nh;
  double s = 0;
  for (int b = 0; b < 4; b++) for (int a = 0; a < 4; a++) s += p[(size_t)(iy * 4 + b) * nh + ix * 4 + a];
  tmp[((size_t)il * np + iy) * np + ix] = s / 16;
}

// Renormalisation (:731-735): map += tmp * tot_flux_old / tot_flux_new where tot_flux_new > 0
__global__ void k_map_accum(double* __restrict__ map, const double* __restrict__ tmp, const double* __restrict__ flux_old,
                            const double* __restrict__ flux_new, size_t plane, int L) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int il = blockIdx.y;
  if (k >= plane) return;
  if (flux_new[il] > 0) map[il * plane + k] = map[il * plane + k] + tmp[il * plane + k] * flux_old[il] / flux_new[il];
}

// Point sources (:753-766): one pixel each, host-cell area
__global__ void k_map_psrc(DreMapGeom m, const double* __restrict__ io, const double* __restrict__ xpp, const int* __restrict__ cell_src,
                           double* __restrict__ map) {
  const DreDev& g = c_g;
  const int i = blockIdx.x, il = threadIdx.x;
  if (il >= g.L) return;
  const double r[3] = {g.ccoord_p_src[3 * (size_t)i], g.ccoord_p_src[3 * (size_t)i + 1], g.ccoord_p_src[3 * (size_t)i + 2]};
  int ix_p, iy_p;
  dre_project_point(m, r, xpp, m.npix, ix_p, iy_p);
  const double cs = g.csize[cell_src[i]];
  atomicAdd(&map[((size_t)il * m.npix + iy_p) * m.npix + ix_p], io[((size_t)g.N + i) * g.L + il] * (cs * cs) / m.area_pixel);
}

// convert_maps_to_MJy_sr :180-246 while copying into map_arr_out(:,:,l,idir)
__global__ void k_map_convert(const double* __restrict__ map, double* __restrict__ out, size_t plane, int units, const double* __restrict__ lambda_si) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int il = blockIdx.y;
  if (k >= plane) return;
  const double parsec = (double)3.08567758e16f, cspeed = (double)2.99792458e8f;
  double v = map[il * plane + k];
  if (units == 0) { v = v * 1e-7; v = v * 1e20 / (parsec * parsec); }
  else if (units == 1) { v = v * 1e20 / (parsec * parsec); }
  else { v = v * (lambda_si[il] * lambda_si[il]) / cspeed; v = v * 1e20 / (parsec * parsec); }
  out[il * plane + k] = v;
}

// =====================================================================================================
// Internal-observer maps: map_in_projection (visual_routines.f90:917-1076) + calc_cube_texture_sphere (:1080-1182).
// A cell is projected on the observer's HEALPix sphere (nside 2^kp_maps): (1) smaller than a pixel -> its pixel,
// (2) containing the observer -> uniform, (3) otherwise a flood fill over the pixel neighbours from the pixel of the
// cell centre while the pixel-centre lines of sight cross the cube, weighted by the normalised chord lengths.
// The host classifies every leaf cell and evaluates the centre direction (acos/sin/cos/ang2pix_nest with libm, as
// find_theta_phi_obs_in does); the flood fill and all accumulation run here.
struct DreMapInCell { int cell, kind, ip, pad; double ratio, len0; };   // kind 1/2/3 as above; len0 = chord of the centre line of sight

// (x, y, face) <-> NESTED pixel (pix2xy_nest / xy2pix_nest, healpix_routines.f90:868-1010): bit de-interleaving
__device__ __forceinline__ int dre_compact_bits(unsigned v) {
  v &= 0x55555555u; v = (v | (v >> 1)) & 0x33333333u; v = (v | (v >> 2)) & 0x0f0f0f0fu; v = (v | (v >> 4)) & 0x00ff00ffu; v = (v | (v >> 8)) & 0x0000ffffu;
  return (int)v;
}
__device__ __forceinline__ unsigned dre_spread_bits(unsigned v) {
  v &= 0x0000ffffu; v = (v | (v << 8)) & 0x00ff00ffu; v = (v | (v << 4)) & 0x0f0f0f0fu; v = (v | (v << 2)) & 0x33333333u; v = (v | (v << 1)) & 0x55555555u;
  return v;
}
__constant__ signed char c_nb_face[9][12] = {
    {8, 9, 10, 11, -1, -1, -1, -1, 10, 11, 8, 9}, {5, 6, 7, 4, 8, 9, 10, 11, 9, 10, 11, 8}, {-1, -1, -1, -1, 5, 6, 7, 4, -1, -1, -1, -1},
    {4, 5, 6, 7, 11, 8, 9, 10, 11, 8, 9, 10},     {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11},   {1, 2, 3, 0, 0, 1, 2, 3, 5, 6, 7, 4},
    {-1, -1, -1, -1, 7, 4, 5, 6, -1, -1, -1, -1}, {3, 0, 1, 2, 3, 0, 1, 2, 4, 5, 6, 7},     {2, 3, 0, 1, -1, -1, -1, -1, 0, 1, 2, 3}};
__constant__ signed char c_nb_swap[9][3] = {{0, 0, 3}, {0, 0, 6}, {0, 0, 0}, {0, 0, 5}, {0, 0, 0}, {5, 0, 0}, {0, 0, 0}, {6, 0, 0}, {3, 0, 0}};
// the i-th of the 8 neighbour slots of a pixel (neighbours_nest healpix_routines.f90:416-852 as a set), or -1
__device__ __forceinline__ int dre_neighbour(int kp, int ipix, int i) {
  const int nside = 1 << kp;
  const int face = ipix >> (2 * kp);
  const unsigned ipf = (unsigned)ipix & ((1u << (2 * kp)) - 1u);
  const int ix = dre_compact_bits(ipf), iy = dre_compact_bits(ipf >> 1);
  const int xo = (i == 0 || i == 1 || i == 7) ? -1 : ((i == 3 || i == 4 || i == 5) ? 1 : 0);
  const int yo = (i == 1 || i == 2 || i == 3) ? 1 : ((i == 5 || i == 6 || i == 7) ? -1 : 0);
  int x = ix + xo, y = iy + yo, nb = 4;
  if (x < 0) { x += nside; nb -= 1; } else if (x >= nside) { x -= nside; nb += 1; }
  if (y < 0) { y += nside; nb -= 3; } else if (y >= nside) { y -= nside; nb += 3; }
  const int f = c_nb_face[nb][face];
  if (f < 0) return -1;
  const int bits = c_nb_swap[nb][face >> 2];
  if (bits & 1) x = nside - x - 1;
  if (bits & 2) y = nside - y - 1;
  if (bits & 4) { const int t = x; x = y; y = t; }
  return (int)(dre_spread_bits((unsigned)x) | (dre_spread_bits((unsigned)y) << 1)) + (f << (2 * kp));
}

// find_ray_cell_intersections / calc_ray_inters (:456-566): chord of the line ro + t*dir through the cube (rc, cellsize)
__device__ __forceinline__ double dre_cube_chord(const double* rc, double cellsize, const double* ro, const double* dir) {
  const double half = cellsize / 2., tol = (double)(1.0f + 1E-7f) * cellsize / 2.;
  double h0[3] = {0, 0, 0}, h1[3] = {0, 0, 0}; int nhit = 0;
  for (int j = 0; j < 6 && nhit < 2; j++) {
    const int ax = j >> 1;
    const double dx = (j & 1) ? -half : half;
    const double den = dir[ax];
    if (den == 0.) continue;
    const double dist = ((rc[ax] + dx) - ro[ax]) / den;
    double pc[3]; bool inside = true;
    for (int i = 0; i < 3; i++) { pc[i] = ro[i] + dist * dir[i]; if (i != ax && fabs(pc[i] - rc[i]) > tol) inside = false; }
    if (!inside) continue;
    if (nhit == 0) { h0[0] = pc[0]; h0[1] = pc[1]; h0[2] = pc[2]; } else { h1[0] = pc[0]; h1[1] = pc[1]; h1[2] = pc[2]; }
    nhit++;
  }
  if (nhit < 2) return 0.0;
  double s = 0;
  for (int i = 0; i < 3; i++) { const double d = h1[i] - h0[i]; s += d * d; }
  return sqrt(s);
}

// Cases 1 and 2 (and the point sources, passed as kind 1 with cell = N + i): one thread per (entry, wavelength)
__global__ void k_mapin_simple(const DreMapInCell* __restrict__ cells, int n, const double* __restrict__ io, int npix, int L,
                               double* __restrict__ map, double* __restrict__ unif) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n * L) return;
  const DreMapInCell c = cells[k / L];
  const int il = k % L;
  const double v = io[(size_t)c.cell * L + il];
  if (v == 0) return;
  if (c.kind == 1) atomicAdd(&map[(size_t)il * npix + c.ip], v * c.ratio);      // :993, :1067
  else if (c.kind == 2) atomicAdd(&unif[il], v / 2.);                               // :1007 (added to every pixel by k_mapin_finish)
}

// Case 3: one block per cell (dynamic fetch).  Per-block scratch: mark[npix] (token of the cell that last touched the pixel:
// never cleared), two frontier lists, the found pixels and their chords.
__global__ void k_mapin_flood(const DreMapInCell* __restrict__ cells, int n, const double* __restrict__ io, const double* __restrict__ obs,
                              int kp, int npix, int L, const DreDir* __restrict__ dirtab, int* __restrict__ mark_all, int* __restrict__ list_all,
                              double* __restrict__ len_all, int* __restrict__ counter, double* __restrict__ map, int token0) {
  const DreDev& g = c_g;
  int* mark = mark_all + (size_t)blockIdx.x * npix;
  int* cur = list_all + (size_t)blockIdx.x * 3 * npix;
  int* nxt = cur + npix;
  int* fpix = cur + 2 * (size_t)npix;
  double* flen = len_all + (size_t)blockIdx.x * npix;
  __shared__ int s_k, s_ncur, s_nnext, s_nfound;
  __shared__ double s_red[32];
  __shared__ double s_total;
  const double ro[3] = {obs[0], obs[1], obs[2]};
  int token = token0;      // marks of earlier launches (other observers) hold smaller tokens
  for (;;) {
    if (threadIdx.x == 0) s_k = atomicAdd(counter, 1);
    __syncthreads();
    const int k = s_k;
    if (k >= n) break;
    token++;
    const DreMapInCell c = cells[k];
    const double rc[3] = {g.ccoord[3 * (size_t)c.cell], g.ccoord[3 * (size_t)c.cell + 1], g.ccoord[3 * (size_t)c.cell + 2]};
    const double cellsize = g.csize[c.cell];
    bool any = false;
    for (int il = 0; il < L; il++) any = any || (io[(size_t)c.cell * L + il] != 0);      // sum(i_obs_in(:, i, idir)) == 0 (:955), values >= 0
    if (!any) { __syncthreads(); continue; }
    if (threadIdx.x == 0) {
      mark[c.ip] = token; fpix[0] = c.ip; s_nfound = 1; s_nnext = 0; s_ncur = 0;
      if (c.len0 > 0) {
        flen[0] = c.len0;
        for (int i = 0; i < 8; i++) {
          const int q = dre_neighbour(kp, c.ip, i);
          if (q >= 0 && mark[q] != token) { mark[q] = token; cur[s_ncur++] = q; }
        }
      } else flen[0] = 1;       // no intersection despite omega_ratio >= 1: the centre pixel alone, weight 1 (:1125)
    }
    __syncthreads();
    while (s_ncur > 0) {
      const int ncur = s_ncur;
      for (int t = threadIdx.x; t < ncur; t += blockDim.x) {
        const int p = cur[t];
        const DreDir d = dirtab[p];
        const double dir[3] = {d.st * d.cp, d.st * d.sp, d.ct};
        const double len = dre_cube_chord(rc, cellsize, ro, dir);
        if (len > 0) {
          const int slot = atomicAdd(&s_nfound, 1);
          fpix[slot] = p; flen[slot] = len;
          for (int i = 0; i < 8; i++) {
            const int q = dre_neighbour(kp, p, i);
            if (q >= 0 && atomicExch(&mark[q], token) != token) nxt[atomicAdd(&s_nnext, 1)] = q;
          }
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) { s_ncur = s_nnext; s_nnext = 0; }
      int* tmp = cur; cur = nxt; nxt = tmp;
      __syncthreads();
    }
    const int nf = s_nfound;
    double total = 1.0;
    if (c.len0 > 0) {            // normalisation (:1178)
      double s = 0;
      for (int t = threadIdx.x; t < nf; t += blockDim.x) s += flen[t];
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
      __syncthreads();
      if (threadIdx.x == 0) { double a = 0; for (int w = 0; w < (int)(blockDim.x >> 5); w++) a += s_red[w]; s_total = a; }
      __syncthreads();
      total = s_total;
    }
    for (int t = threadIdx.x; t < nf; t += blockDim.x) {
      const double w = (c.len0 > 0) ? flen[t] / total : flen[t];
      for (int il = 0; il < L; il++) {
        const double v = io[(size_t)c.cell * L + il] * w * c.ratio;           // :1018
        if (v != 0) atomicAdd(&map[(size_t)il * npix + fpix[t]], v);
      }
    }
    __syncthreads();
  }
}

// uniform part (case 2) + convert_maps_to_MJy_sr into map_in_arr_out(:, l, idir)
__global__ void k_mapin_finish(const double* __restrict__ map, const double* __restrict__ unif, double* __restrict__ out, int npix, int units,
                               const double* __restrict__ lambda_si) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x, il = blockIdx.y;
  if (p >= npix) return;
  const double parsec = (double)3.08567758e16f, cspeed = (double)2.99792458e8f;
  double v = map[(size_t)il * npix + p] + unif[il];
  if (units == 0) { v = v * 1e-7; v = v * 1e20 / (parsec * parsec); }
  else if (units == 1) { v = v * 1e20 / (parsec * parsec); }
  else { v = v * (lambda_si[il] * lambda_si[il]) / cspeed; v = v * 1e20 / (parsec * parsec); }
  out[(size_t)il * npix + p] = v;
}
