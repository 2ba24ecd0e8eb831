// dre_host.h — host-side helpers of the engine: growable device buffers and the host HEALPix
// routines used to build the direction tables (pix2ang_nest healpix_routines.f90:30-136,
// ang2pix_nest :238-348) with the same libm the Fortran host uses.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "dre_common.cuh"

struct DreError { std::string msg; };
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) throw DreError{std::string(#call) + ": " + cudaGetErrorString(e_)}; } while (0)
#define DRE_REQUIRE(cond, text) do { if (!(cond)) throw DreError{std::string(text)}; } while (0)

template <class T> struct DBuf {
  T* p = nullptr; size_t cap = 0;
  DBuf() {}
  DBuf(const DBuf&) = delete; DBuf& operator=(const DBuf&) = delete;
  ~DBuf() { if (p) cudaFree(p); }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  // grow to at least n elements; contents are NOT preserved
  void ensure(size_t n, double slack = 1.25) {
    if (n <= cap) return;
    if (p) { cudaFree(p); p = nullptr; }
    size_t want = (size_t)(n * slack) + 64;
    CK(cudaMalloc(&p, want * sizeof(T)));
    cap = want;
  }
  // grow to at least n elements, keeping the first `keep` ones
  void grow_keep(size_t n, size_t keep, cudaStream_t s, double slack = 1.25) {
    if (n <= cap) return;
    const size_t want = (size_t)(n * slack) + 64;
    T* q = nullptr;
    CK(cudaMalloc(&q, want * sizeof(T)));
    if (p && keep) CK(cudaMemcpyAsync(q, p, std::min(keep, cap) * sizeof(T), cudaMemcpyDeviceToDevice, s));
    if (p) { CK(cudaStreamSynchronize(s)); cudaFree(p); }
    p = q; cap = want;
  }
  void upload(const T* h, size_t n, cudaStream_t s) { ensure(n, 1.0); CK(cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, s)); }
  void zero(size_t n, cudaStream_t s) { ensure(n, 1.0); CK(cudaMemsetAsync(p, 0, n * sizeof(T), s)); }
};

// device SoA list with its own storage
struct ListStore {
  DBuf<int> unit, dir_ipix, cc_old, seg; DBuf<unsigned> meta; DBuf<double> dplane, prev, lum;
  int n = 0;
  void ensure(size_t cap, int L, size_t nunits) {
    unit.ensure(cap); dir_ipix.ensure(cap); cc_old.ensure(cap); meta.ensure(cap); dplane.ensure(cap); prev.ensure(cap);
    lum.ensure(cap * (size_t)L); seg.ensure(nunits + 1, 1.0);
  }
  // grow to `cap` entries keeping the first `keep` (the segment array is sized by the unit count and kept as it is)
  void grow_keep(size_t cap, size_t keep, int L, cudaStream_t s) {
    unit.grow_keep(cap, keep, s); dir_ipix.grow_keep(cap, keep, s); cc_old.grow_keep(cap, keep, s); meta.grow_keep(cap, keep, s);
    dplane.grow_keep(cap, keep, s); prev.grow_keep(cap, keep, s); lum.grow_keep(cap * (size_t)L, keep * (size_t)L, s);
  }
  DreList view() {
    DreList l; l.unit = unit.p; l.dir_ipix = dir_ipix.p; l.meta = meta.p; l.cc_old = cc_old.p; l.dplane = dplane.p;
    l.prev = prev.p; l.lum = lum.p; l.seg = seg.p; l.n = n; l.cap = (int)unit.cap; return l;
  }
};

namespace hp {
// single-precision pi constants of the reference (smooth_grid_routines.f90:311-317, healpix_routines.f90:13)
static const double HALFPI = (double)1.57079637050628662109375f;
static const double PI_ = 2 * HALFPI, TWOPI = 2 * PI_;
static const double TWOTHIRD = (double)(2.f / 3.f);

// de-interleave the bits of a NESTED in-face index (what the pix2x/pix2y tables of mk_pix2xy encode)
static inline void ipf2xy(int ipf, int& ix, int& iy) {
  ix = iy = 0;
  for (int b = 0; b < 15; b++) { ix |= ((ipf >> (2 * b)) & 1) << b; iy |= ((ipf >> (2 * b + 1)) & 1) << b; }
}
static inline int xy2ipf(int ix, int iy) {
  int r = 0;
  for (int b = 0; b < 15; b++) r |= (((ix >> b) & 1) << (2 * b)) | (((iy >> b) & 1) << (2 * b + 1));
  return r;
}
// pix2ang_nest: returns z = cos(theta) as computed by the reference, theta = acos(z), phi
static inline void pix2ang_nest(int nside, int ipix, double& theta, double& phi) {
  static const int jrll[12] = {2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4};
  static const int jpll[12] = {1, 3, 5, 7, 0, 2, 4, 6, 1, 3, 5, 7};
  const int npface = nside * nside, nl4 = 4 * nside;
  const int face = ipix / npface, ipf = ipix % npface;
  const double fn = (double)nside, fact1 = 1.0 / (3.0 * fn * fn), fact2 = 2.0 / (3.0 * fn);
  int ix, iy; ipf2xy(ipf, ix, iy);
  const int jrt = ix + iy, jpt = ix - iy;
  const int jr = jrll[face] * nside - jrt - 1;
  int nr, kshift; double z;
  if (jr < nside) { nr = jr; z = 1.0 - nr * fact1 * nr; kshift = 0; }
  else if (jr <= 3 * nside) { nr = nside; z = (2 * nside - jr) * fact2; kshift = (jr - nside) & 1; }
  else { nr = nl4 - jr; z = -1.0 + nr * fact1 * nr; kshift = 0; }
  theta = acos(z);
  int jp = (jpll[face] * nr + jpt + 1 + kshift) / 2;
  if (jp > nl4) jp -= nl4;
  if (jp < 1) jp += nl4;
  phi = (jp - (kshift + 1) * 0.5) * (HALFPI / nr);
}
static inline double f_modulo(double a, double p) { double r = fmod(a, p); if (r != 0 && ((r < 0) != (p < 0))) r += p; return r; }
static inline int ang2pix_nest(int nside, double theta, double phi) {
  const double z = cos(theta), za = fabs(z);
  const double tt = f_modulo(phi, TWOPI) / HALFPI;
  int face, ix, iy;
  if (za <= TWOTHIRD) {
    const int jp = (int)(nside * (0.5 + tt - z * 0.75)), jm = (int)(nside * (0.5 + tt + z * 0.75));
    const int ifp = jp / nside, ifm = jm / nside;
    if (ifp == ifm) face = (ifp & 3) + 4; else if (ifp < ifm) face = (ifp & 3); else face = (ifm & 3) + 8;
    ix = jm & (nside - 1); iy = nside - (jp & (nside - 1)) - 1;
  } else {
    int ntt = (int)tt; if (ntt >= 4) ntt = 3;
    const double tp = tt - ntt, tmp = sqrt(3.0 * (1.0 - za));
    int jp = (int)(nside * tp * tmp), jm = (int)(nside * (1.0 - tp) * tmp);
    if (jp > nside - 1) jp = nside - 1;
    if (jm > nside - 1) jm = nside - 1;
    if (z >= 0) { face = ntt; ix = nside - jm - 1; iy = nside - jp - 1; } else { face = ntt + 8; ix = jp; iy = jm; }
  }
  return xy2ipf(ix, iy) + face * nside * nside;
}
static inline DreDir make_dir(double theta, double phi) { DreDir d; d.st = sin(theta); d.ct = cos(theta); d.sp = sin(phi); d.cp = cos(phi); return d; }
}  // namespace hp
