// dre_dust.cuh — N4: the dust-emission source term of the dust-heating iterations, effective-grain mode
// (set_dust_emission / calc_dens_dustem sed_routines.f90:249-276, 377-430), one thread per cell:
//   absorbed energy   abs_en = sum kabs(l) rf_stars(l,c) dlambda_stars(l) + sum kabs(l) rf_dust(l,c) dlambda_dust(l)   (calc_t_dust_equil :846-874)
//   equilibrium T     4 pi sum kabs(l) B(T, lambda_l) dlambda_dust(l) = abs_en, Brent's method on [t0, t1] = [0, 1e4] K     (zbrent_tdust :993-1078)
//   emissivity        dens_stars(l,c) = 4 pi kabs(l) / kext_ref * dens_ref(c) * B(T, lambda_l) * parsec^2                  (:411, set_units_dens_stars_arr :904-914)
// so that dartray_dust (dartray_hub.f90:210-253) re-enters the ray tracing without bringing u_final to the host: rf_dust can be read
// from the device-resident u_final rows and the result left on the device as lumcell = dens_stars * csize^3 (calc_total_luminosity :605-660).
#pragma once
#include "dre_common.cuh"

struct DreDust {
  int N, ls, ld;                 // cells, stellar-emission wavelengths, dust-emission wavelengths
  const double* kabs_s;          // [ls] kabs_arr(0:lnum_stars-1)
  const double* kabs_d;          // [ld] kabs_arr(i0:lnum_tot-1)
  const double* lam_d;           // [ld] lambda_arr_SI(i0:)
  const double* dl_s;            // [ls] delta_lambda_bin_stars
  const double* dl_d;            // [ld] delta_lambda_bin_dust
  const double* dens_ref;        // [N]
  const int* cchild;             // [N]
  const double* csize;           // [N]
  const double* rf_s;            // [N][ls] u_final_uv_opt after convert_ufield_ifield (W/m/m^2)
  const double* rf_d;            // [N][ld] u_final_arr/parsec^3*cspeed, or nullptr
  const double* rows; int ROW;   // device-resident u_final (first ld doubles of every accumulator row) when rf_d == nullptr
  double kext_ref;
  int iterations_dustem;
  double* dens_stars;            // [N][ld] out
  double* t_dust;                // [N] out (0 where no dust)
  double* lumcell;               // [N][ld] out or nullptr
  int* err;
};

// mks constants of the reference: real64 parameters initialised from default-real literals (smooth_grid_routines.f90:373-443)
#define DUST_PARSEC ((double)3.08567758E16f)
#define DUST_CSPEED ((double)2.99792458E8f)
#define DUST_HPLANCK ((double)6.62606957E-34f)
#define DUST_KBOLTZ ((double)1.3806488E-23f)

// bplanck sed_routines.f90:939-948
__host__ __device__ inline double dust_bplanck(double t, double la) {
  const double a1 = 2 * DUST_HPLANCK * pow(DUST_CSPEED, 2.) / pow(la, 5.);
  const double a2 = exp(DUST_HPLANCK * DUST_CSPEED / (la * DUST_KBOLTZ * t)) - 1;
  return a1 * (1.0 / a2);
}
// abs_en_diff :917-936
__host__ __device__ inline double dust_en_diff(double t, const double* kabs_d, const double* lam_d, const double* dl_d, int ld, double abs_en, double pi) {
  double s = 0;
  for (int i = 0; i < ld; i++) s += (kabs_d[i] * dust_bplanck(t, lam_d[i])) * dl_d[i];
  return 4 * pi * s - abs_en;
}
// zbrent_tdust :993-1078 (Numerical Recipes ZBRENT; EPS = 3e-16 as a default-real literal; TOL is an implicitly typed local
// that is never assigned in the reference: taken as 0).  rc != 0: root not bracketed / too many iterations (the reference STOPs).
__host__ __device__ inline double dust_zbrent(const double* kabs_d, const double* lam_d, const double* dl_d, int ld, double abs_en, double pi, int* rc) {
  const double EPS = (double)3.E-16f, TOL = 0.0;
  double a = 0.0, b = (double)1E4f, c = 0, d = 0, e = 0;
  double fa = dust_en_diff(a, kabs_d, lam_d, dl_d, ld, abs_en, pi), fb = dust_en_diff(b, kabs_d, lam_d, dl_d, ld, abs_en, pi);
  if (fb * fa > 0.) { *rc = 1; return 0.0; }
  double fc = fb;
  for (int iter = 1; iter <= 100; iter++) {
    if (fb * fc > 0.) { c = a; fc = fa; d = b - a; e = d; }
    if (fabs(fc) < fabs(fb)) { a = b; b = c; c = a; fa = fb; fb = fc; fc = fa; }
    const double tol1 = 2. * EPS * fabs(b) + 0.5 * TOL;
    const double xm = .5 * (c - b);
    if (fabs(xm) <= tol1 || fb == 0.) return b;
    if (fabs(e) >= tol1 && fabs(fa) > fabs(fb)) {
      double p, q, r;
      const double s = fb / fa;
      if (a == c) { p = 2. * xm * s; q = 1. - s; }
      else { q = fa / fc; r = fb / fc; p = s * (2. * xm * q * (q - r) - (b - a) * (r - 1.)); q = (q - 1.) * (r - 1.) * (s - 1.); }
      if (p > 0.) q = -q;
      p = fabs(p);
      if (2. * p < fmin(3. * xm * q - fabs(tol1 * q), fabs(e * q))) { e = d; d = p / q; }
      else { d = xm; e = d; }
    } else { d = xm; e = d; }
    a = b; fa = fb;
    if (fabs(d) > tol1) b = b + d; else b = b + copysign(tol1, xm);
    fb = dust_en_diff(b, kabs_d, lam_d, dl_d, ld, abs_en, pi);
  }
  *rc = 2;
  return b;
}

__global__ void k_dust_emission(DreDust q) {
  extern __shared__ double sh[];           // kabs_d | lam_d | dl_d (read by every Brent iteration of every thread)
  double* s_k = sh; double* s_l = sh + q.ld; double* s_d = sh + 2 * q.ld;
  for (int i = threadIdx.x; i < q.ld; i += blockDim.x) { s_k[i] = q.kabs_d[i]; s_l[i] = q.lam_d[i]; s_d[i] = q.dl_d[i]; }
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= q.N) return;
  double* out = q.dens_stars + (size_t)c * q.ld;
  double t = 0.0;
  const bool on = (q.cchild[c] == -1) && (q.dens_ref[c] != 0);      // :401
  if (on) {
    double abs_en = 0;                                                // calc_t_dust_equil :855-863
    { double s = 0; for (int l = 0; l < q.ls; l++) s += (q.kabs_s[l] * q.rf_s[(size_t)c * q.ls + l]) * q.dl_s[l]; abs_en = s; }
    if (q.iterations_dustem > 0) {
      double s = 0;
      for (int l = 0; l < q.ld; l++) {
        const double rf = q.rf_d ? q.rf_d[(size_t)c * q.ld + l] : q.rows[(size_t)c * q.ROW + l] / pow(DUST_PARSEC, 3.) * DUST_CSPEED;
        s += (s_k[l] * rf) * s_d[l];
      }
      abs_en = abs_en + s;
    }
    if (abs_en < 0) { atomicCAS(q.err, 0, 6); return; }
    int rc = 0;
    t = dust_zbrent(s_k, s_l, s_d, q.ld, abs_en, DRE_PI, &rc);
    if (rc != 0) { atomicCAS(q.err, 0, 6 + rc); return; }
  }
  for (int l = 0; l < q.ld; l++) {
    double v = 0.0;
    if (on) v = 4 * DRE_PI * s_k[l] / q.kext_ref * q.dens_ref[c] * dust_bplanck(t, s_l[l]) * pow(DUST_PARSEC, 2.);
    out[l] = v;
    if (q.lumcell) { const double cs = q.csize[c]; q.lumcell[(size_t)c * q.ld + l] = v * cs * cs * cs; }
  }
  if (q.t_dust) q.t_dust[c] = t;
}
