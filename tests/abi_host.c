/* abi_host.c -- a plain C host of the engine's C ABI (include/dartray_engine.h), no Python in the call path.
 *
 * It builds the arguments the way the Fortran wrappers of fortran/rt_wrappers.f90 do: column-major arrays exactly as the
 * Fortran host holds them (u_fest_arr(lnum, tot_ncell), ccoord(3, tot_ncell), ...), an ARRAY OF POINTERS to the ragged
 * planes scaspe_arr(k)%a(npix_arr(k), tot_ncell), dre_params by reference, and runs
 *     precalc -> direct light -> rt_prepare(rtt_scatt) on the host arrays -> scattering iteration 1
 * through the drop-in call dre_run_pass (host arrays in, host arrays out), like rt_loop_hostarrays.
 * Inputs: raw little-endian arrays dumped by tests/test_gpu_abi_host.py into argv[1]; outputs: *.out files there, which
 * the test compares with the CPU oracle.
 *
 *   gcc -O1 -Iinclude tests/abi_host.c -Ldart-ray_b200 -ldartray_engine -Wl,-rpath,$PWD/dart-ray_b200 -o abi_host
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "dartray_engine.h"

static const char* g_dir;

static void* rd(const char* name, size_t want_bytes) {
  char path[1024];
  snprintf(path, sizeof path, "%s/%s", g_dir, name);
  FILE* f = fopen(path, "rb");
  if (!f) { fprintf(stderr, "abi_host: cannot open %s\n", path); exit(2); }
  fseek(f, 0, SEEK_END);
  size_t n = (size_t)ftell(f);
  fseek(f, 0, SEEK_SET);
  if (want_bytes && n != want_bytes) { fprintf(stderr, "abi_host: %s has %zu bytes, expected %zu\n", name, n, want_bytes); exit(2); }
  void* p = malloc(n ? n : 1);
  if (fread(p, 1, n, f) != n) { fprintf(stderr, "abi_host: short read of %s\n", name); exit(2); }
  fclose(f);
  return p;
}
static void wr(const char* name, const void* p, size_t bytes) {
  char path[1024];
  snprintf(path, sizeof path, "%s/%s", g_dir, name);
  FILE* f = fopen(path, "wb");
  if (!f || fwrite(p, 1, bytes, f) != bytes) { fprintf(stderr, "abi_host: cannot write %s\n", path); exit(2); }
  fclose(f);
}
#define CHECK(call) do { int rc_ = (call); if (rc_ != 0) { fprintf(stderr, "abi_host: %s failed (rc %d): %s\n", #call, rc_, dre_last_error(e)); return 3; } } while (0)

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: abi_host <scene directory>\n"); return 1; }
  g_dir = argv[1];
  /* scalars: N, max_lvl, base(2), lnum, tot_ndir_scaspe, tot_p_src, has_cell */
  int32_t* hdr = (int32_t*)rd("header.i32", 8 * sizeof(int32_t));
  const int32_t N = hdr[0], max_lvl = hdr[1], base[2] = {hdr[2], hdr[3]}, L = hdr[4], ndir = hdr[5], S = hdr[6], has_cell = hdr[7];
  double* conv = (double*)rd("conv_en_lim.f64", sizeof(double));
  const size_t NL = (size_t)N * L;
  int32_t* cchild = (int32_t*)rd("cchild.i32", (size_t)N * 4);
  int32_t* lvl = (int32_t*)rd("lvl.i32", (size_t)N * 4);
  int64_t* cindex = (int64_t*)rd("cindex.i64", (size_t)N * 8);
  double* ccoord = (double*)rd("ccoord.f64", (size_t)N * 24);               /* ccoord(3, N) */
  double* csize = (double*)rd("csize.f64", (size_t)N * 8);
  double* csize_arr = (double*)rd("csize_arr.f64", (size_t)(max_lvl + 1) * 8);
  double* carea_arr = (double*)rd("carea_arr.f64", (size_t)(max_lvl + 1) * 8);
  double* cvol_arr = (double*)rd("cvol_arr.f64", (size_t)(max_lvl + 1) * 8);
  double* dens_arr = (double*)rd("dens_arr.f64", NL * 8);                   /* dens_arr(lnum, N) */
  double* ksca = (double*)rd("ksca_norm.f64", (size_t)L * 8);
  double* gsca = (double*)rd("gsca.f64", (size_t)L * 8);
  int32_t* kp_sca = (int32_t*)rd("kp_sca.i32", (size_t)L * 4);
  int32_t* npix = (int32_t*)rd("npix.i32", (size_t)L * 4);                  /* npix_arr */
  int32_t* npix_hp = (int32_t*)rd("npix_hp.i32", (size_t)L * 4);
  double* dir_i_out = (double*)rd("dir_i_out.f64", (size_t)ndir * 16);      /* dir_i_out(ndir, 2) */
  double* lumcell0 = (double*)rd("lumcell.f64", NL * 8);
  double* ccoord_p_src = (double*)rd("ccoord_p_src.f64", 0);
  double* lum_p_src = (double*)rd("lum_p_src.f64", 0);
  int32_t* src_cell = (int32_t*)rd("src_cell.i32", (size_t)N * 4);
  int32_t* cell_src = (int32_t*)rd("cell_src.i32", 0);
  dre_params* par_dir = (dre_params*)rd("params_direct.bin", sizeof(dre_params));
  dre_params* par_sca = (dre_params*)rd("params_scatt.bin", sizeof(dre_params));

  /* the Fortran host's arrays */
  double* u_fest = (double*)calloc(NL, 8);
  double* u_final = (double*)calloc(NL, 8);
  double* lumcell = (double*)calloc(NL, 8);
  double* lum_lost = (double*)calloc(L, 8);
  double* psel_av = (double*)calloc((size_t)N + S, 8);
  double* psel_av_direct = (double*)calloc((size_t)N + S, 8);
  double** scaspe = (double**)calloc(L, sizeof(double*));        /* array of c_ptr: scaspe_arr(k)%a(npix_arr(k), N) */
  double** scaspe_prev = (double**)calloc(L, sizeof(double*));
  for (int l = 0; l < L; l++) {
    scaspe[l] = (double*)calloc((size_t)npix[l] * N, 8);
    scaspe_prev[l] = (double*)calloc((size_t)npix[l] * N, 8);
  }

  dre_engine* e = NULL;
  if (dre_create(&e, 0, 0, 1, NULL) != 0) { fprintf(stderr, "abi_host: dre_create failed (no CUDA device?)\n"); return 3; }
  CHECK(dre_set_grid(e, N, max_lvl, base, cchild, lvl, cindex, ccoord, csize, csize_arr, carea_arr, cvol_arr));
  CHECK(dre_set_opacity(e, L, dens_arr, ksca, gsca, kp_sca, ndir, dir_i_out));
  CHECK(dre_set_sources(e, lumcell0, S, ccoord_p_src, lum_p_src, src_cell, cell_src));
  CHECK(dre_set_params(e, par_dir));
  dre_counters cnt;
  long long isect = 0;
  /* rtt_precalc_cell / rtt_precalc_src (dartray_hub.f90:21-25) */
  if (has_cell) { CHECK(dre_run_pass(e, DRE_RTT_PRECALC_CELL, 0, NULL, u_fest, NULL, NULL, NULL, NULL, &cnt)); isect += cnt.intersections; }
  if (S > 0) { CHECK(dre_run_pass(e, DRE_RTT_PRECALC_SRC, 0, NULL, u_fest, NULL, NULL, NULL, NULL, &cnt)); isect += cnt.intersections; }
  /* rtt_dir_cell / rtt_dir_src (:30-34) */
  if (has_cell) {
    CHECK(dre_run_pass(e, DRE_RTT_DIR_CELL, 0, u_fest, u_final, scaspe, (const double* const*)scaspe_prev, lum_lost, psel_av, &cnt));
    isect += cnt.intersections;
  }
  if (S > 0) {
    CHECK(dre_run_pass(e, DRE_RTT_DIR_SRC, 0, u_fest, u_final, scaspe, (const double* const*)scaspe_prev, lum_lost, psel_av, &cnt));
    isect += cnt.intersections;
  }
  memcpy(psel_av_direct, psel_av, ((size_t)N + S) * 8);
  wr("u_fest.out", u_fest, NL * 8);
  wr("u_final_direct.out", u_final, NL * 8);
  wr("psel_av_direct.out", psel_av_direct, ((size_t)N + S) * 8);
  /* rt_prepare, case (rtt_scatt), on the host arrays (dartray_hub.f90:432-457, calc_total_luminosity_sca rt_routines.f90:663-758) */
  memcpy(u_fest, u_final, NL * 8);
  int all_zero = 1;
  for (int l = 0; l < L; l++) {
    memcpy(scaspe_prev[l], scaspe[l], (size_t)npix[l] * N * 8);
    for (int c = 0; c < N; c++) {
      double s = 0;
      if (cchild[c] == -1) for (int p = 0; p < npix_hp[l]; p++) s += scaspe[l][(size_t)npix[l] * c + p];
      lumcell[(size_t)L * c + l] = s;
      if (s != 0) all_zero = 0;
    }
    memset(scaspe[l], 0, (size_t)npix[l] * N * 8);               /* sequential scattering: scaspe_arr = 0 */
  }
  memset(psel_av, 0, ((size_t)N + S) * 8);                      /* expand_psel_av_arr: a fresh slice */
  if (!all_zero) {
    CHECK(dre_set_params(e, par_sca));                          /* bm_par = bm_par_sca (:442) */
    CHECK(dre_set_sources(e, lumcell, S, ccoord_p_src, lum_p_src, src_cell, cell_src));
    CHECK(dre_run_pass(e, DRE_RTT_SCATT, 1, u_fest, u_final, scaspe, (const double* const*)scaspe_prev, lum_lost, psel_av, &cnt));
    isect += cnt.intersections;
  }
  wr("u_final.out", u_final, NL * 8);
  wr("lum_lost.out", lum_lost, (size_t)L * 8);
  wr("psel_av.out", psel_av, ((size_t)N + S) * 8);
  wr("lumcell.out", lumcell, NL * 8);
  for (int l = 0; l < L; l++) {
    char nm[64];
    snprintf(nm, sizeof nm, "scaspe_%d.out", l);
    wr(nm, scaspe[l], (size_t)npix[l] * N * 8);
  }
  printf("abi_host ok: %lld intersections, conv_en_lim %g\n", isect, conv[0]);
  dre_destroy(e);
  return 0;
}
