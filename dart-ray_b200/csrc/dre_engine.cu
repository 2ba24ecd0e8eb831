// dre_engine.cu — host side of the B200 ray-tracing engine and its C ABI (include/dartray_engine.h).
//
// Replaces rt_loop / rt_loop_iobs / rt_mpi of DART-Ray (rt_routines.f90:985-1058, :1470-1603,
// io_routines.f90:3667-3881).  No CPU fallback: every entry point that computes runs CUDA kernels.
#include <algorithm>
#include <chrono>
#include <dlfcn.h>
#include "dre_host.h"
#include "dre_lists.cuh"
#include "dre_kernels.cuh"
#include "dre_split.cuh"
#include "dre_maps.cuh"
#include "dre_dust.cuh"
#include "dre_sym2d.cuh"

namespace {

// ---- minimal NCCL binding (resolved with dlopen so the library loads without NCCL) --------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[DRE_NCCL_ID_BYTES]; } ncclUniqueId;
struct Nccl {
  void* h = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool load(std::string& err) {
    if (h) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
    if (!h) { err = std::string("cannot load NCCL: ") + dlerror(); return false; }
    GetUniqueId = (int (*)(ncclUniqueId*))dlsym(h, "ncclGetUniqueId");
    CommInitRank = (int (*)(ncclComm_t*, int, ncclUniqueId, int))dlsym(h, "ncclCommInitRank");
    AllReduce = (int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t))dlsym(h, "ncclAllReduce");
    CommDestroy = (int (*)(ncclComm_t))dlsym(h, "ncclCommDestroy");
    GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
    GroupStart = (int (*)())dlsym(h, "ncclGroupStart");
    GroupEnd = (int (*)())dlsym(h, "ncclGroupEnd");
    if (!GetUniqueId || !CommInitRank || !AllReduce || !CommDestroy) { err = "NCCL symbols missing"; return false; }
    return true;
  }
};
Nccl g_nccl;
const int NCCL_FLOAT64 = 8, NCCL_INT64 = 4, NCCL_SUM = 0;   // ncclDataType_t / ncclRedOp_t values (nccl.h)

inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

}  // namespace

struct dre_engine {
  int device = 0, rank = 0, nranks = 1;
  cudaStream_t st = nullptr;
  ncclComm_t comm = nullptr;
  std::string err;
  int sm_count = 148;
  // static inputs
  DreDev g;
  bool have_grid = false, have_opacity = false, have_sources = false, have_params = false;
  DBuf<int> cchild, src_cell, ik_sca, npix_hp, npix, off; DBuf<unsigned char> lvl; DBuf<long long> cindex;
  DBuf<DreCell> cells;
  DBuf<double> ccoord, csize, dens, ksca, gsca, lumcell, ccoord_p_src, lum_p_src;
  DBuf<unsigned short> lamof;
  DBuf<DreDir> class_dir[DRE_MAXCLASS], dirtab[DRE_MAXK + 1], obs_dir;
  DBuf<double> ftab[DRE_MAXK + 1];
  std::vector<int> h_npix, h_npix_hp, h_off, h_ik, h_cell_src;
  std::vector<double> h_dir_i_out, h_gsca;
  std::vector<double> det_peak;   // deterministic mode: largest phase-function value an observer slot can take, per wavelength
  double h_csize0 = 0;      // csize(0): the root cell
  std::vector<double> h_ccoord, h_csize, h_ccoord_p_src; std::vector<int> h_cchild;   // host copies for the map geometry
  int S = 0;
  // device-resident accumulators
  DBuf<double> acc, prev, tot, u_fest, lum_lost, lum_lost_prev, psel_sum, psel_av, i_obs, i_obs_in, tmp_d, red_d;
  DBuf<long long> ipsel;
  int i_obs_ndir_in = 0;
  DBuf<unsigned long long> counters;
  DBuf<int> errflag, srcflag, srcpos, srclist, counter, active;
  int nsrc_list = 0;
  // record
  long long rec_cap = 0; DBuf<dre_isect> rec; DBuf<unsigned long long> rec_n;
  // wavefront scratch (device-driven resolution step, dre_lists.cuh)
  ListStore H[2], W[2], OUT;
  DBuf<int> u_nc, u_gid, u_nk, u_step, u_mode, u_task_seg;
  DBuf<int> selH, posH, selW, posW, extcnt, ext_seg, ext, ipx, gone, cand, ncand, nxt, hasprev, unit_of_ext, ntask, nreco,
      task_off, reco_off, reco_cnt, fH, fW, posOH, posOW, keptH, keptW, newHcnt, newWcnt, hhead, hlink;
  DBuf<unsigned long long> hkey;   // sibling table of a step (dre_lists.cuh ph_fill_ext / ph_cand)
  DBuf<DreTask> tasks;
  DBuf<double> pkts;     // ray packets of the current step
  DBuf<DreCtl> ctl; DreCtl* h_ctl = nullptr; int* h_err = nullptr;   // control block of the step and its pinned host mirror
  DBuf<unsigned long long> scan_desc; int desc_stride = 0;          // tile descriptors of the chained scans (10 job slots)
  int capH = 0, capW = 0, capT = 0, capU = 0;
  DBuf<DreEvt> evt; DBuf<int> nev, evoff; int capE = 0;
  DBuf<int> sym_perm[DRE_MAXCLASS]; bool have_sym = false;   // 2D mode: mirrored-pixel tables of the scattering classes     // events of the rays of a step (dre_split.cuh)
  // replayable launch sequences of one resolution step (before / after the ray-march kernel), per list parity
  struct StepGraph { cudaGraphExec_t pre = nullptr, post = nullptr; std::vector<unsigned char> key; };
  StepGraph sgraph[2][2];   // [precalc / deposit][list parity]
  // options
  int opt_hash = 0, opt_use_bulk = 1, opt_batch_units = 1 << 20, opt_warps = 16, opt_stage_split = 2, opt_ctas_per_sm = 1, opt_dir_kmax = 9, opt_ftab_kmax = 8, opt_chunk = 8, opt_ring_slots = 1, opt_graphs = 1, opt_cap0 = 0, opt_split = 1, opt_coop = 0, opt_coop_work = 131072, opt_det = 0, opt_sector_rot = 1, opt_blocking_sync = 0;
  // deterministic mode: exponents of the per-wavelength bounds (running maxima; scaspe | u | lum_lost | psel) and the constants
  std::vector<int> det_e; std::vector<double> h_qc; DBuf<double> qc;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_blk = nullptr;      // blocking-sync event of the per-step wait (option blocking_sync)
  double last_kernel_ms = 0;
  long long peak_entries = 0;
  double est_H[16] = {0}, est_W[16] = {0}, est_T[16] = {0};   // list entries / tasks per unit seen so far, by rt_type
  size_t ftab_budget = (size_t)24 << 30, ftab_bytes = 0; unsigned ftab_skipped = 0;   // phase-function tables: at most 24 GiB or 15 % of the free memory
  size_t list_budget = 0;          // bytes the ray lists of a batch may take (half of the memory free after configuration)
  double per_unit_est[16] = {0};   // list entries per unit of the last pass, by rt_type
  double per_unit_evt[16] = {0};   // march events per unit (largest step), by rt_type
  long long last_launches = 0, n_launch = 0, launch_mark = 0;
  cudaEvent_t marks[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

  size_t rows_elems() const { return (size_t)g.N * g.ROW; }
};

namespace {

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of (function, device), shared by every engine of the process
int g_smem_set[64][3][3][2];

__global__ void k_bump_epoch(DreCtl* c) { c->scan_epoch += 16u; }
// descriptor regions of the chained scans: 12 job slots x (tiles of the largest input + 1) words; never cleared after
// the first allocation (epoch-tagged), re-created (zeroed) when they grow
void ensure_scan_desc(dre_engine* e, int tiles) {
  if (tiles + 1 <= e->desc_stride) return;
  const int stride = std::max(64, (int)((tiles + 1) * 1.5));
  e->scan_desc.release();
  e->scan_desc.zero((size_t)12 * stride, e->st);
  e->desc_stride = stride;
}

// exclusive scan of n ints (n known to the host) into out[0..n] (out[n] = total): job slot 9 of the chained scan
void scan_host_n(dre_engine* e, const int* in, int* out, int n) {
  const int nb = std::max(1, cdiv(std::max(n, 1), SCAN_TILE));
  ensure_scan_desc(e, nb);
  ScanJobs J; memset(&J, 0, sizeof(J));
  J.j[0].in = in; J.j[0].out = out; J.j[0].n_ptr = nullptr; J.j[0].n_const = n; J.j[0].slot = 9;
  CK(cudaMemsetAsync(&e->ctl.p->tickets[9], 0, sizeof(int), e->st));
  (e->n_launch++, k_scan_jobs)<<<dim3(std::min(nb, 2 * e->sm_count), 1), SCAN_THREADS, 0, e->st>>>(J, e->ctl.p, e->scan_desc.p, e->desc_stride, 9u, 0);
  (e->n_launch++, k_bump_epoch)<<<1, 1, 0, e->st>>>(e->ctl.p);
}

// publish the static inputs to the constant bank the kernels read
void sync_g(dre_engine* e) {
  CK(cudaMemcpyToSymbolAsync(c_g, &e->g, sizeof(DreDev), 0, cudaMemcpyHostToDevice, e->st));
}

int read_int(dre_engine* e, const int* d) {
  int v = 0; CK(cudaMemcpyAsync(&v, d, sizeof(int), cudaMemcpyDeviceToHost, e->st)); CK(cudaStreamSynchronize(e->st)); return v;
}

void build_dir_table(dre_engine* e, int k) {
  if (e->dirtab[k].p) return;
  const int nside = 1 << k; const size_t np = (size_t)12 * nside * nside;
  std::vector<DreDir> t(np);
  for (size_t p = 0; p < np; p++) { double th, ph; hp::pix2ang_nest(nside, (int)p, th, ph); t[p] = hp::make_dir(th, ph); }
  e->dirtab[k].upload(t.data(), np, e->st);
  CK(cudaStreamSynchronize(e->st));
  e->g.dirtab[k] = e->dirtab[k].p;
  if (k > e->g.dir_kmax) e->g.dir_kmax = k;
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
#define DRE_TRY(e) try {
#define DRE_CATCH(e) } catch (const DreError& x) { (e)->err = x.msg; return 1; } catch (const std::exception& x) { (e)->err = x.what(); return 1; } return 0;

extern "C" {

int dre_get_unique_id(void* id128) {
  std::string err;
  if (!g_nccl.load(err)) return 1;
  ncclUniqueId id; memset(&id, 0, sizeof(id));
  if (g_nccl.GetUniqueId(&id) != 0) return 2;
  memcpy(id128, &id, DRE_NCCL_ID_BYTES);
  return 0;
}

int dre_create(dre_engine** out, int device, int rank, int nranks, const void* nccl_id128) {
  if (!out) return 1;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    fprintf(stderr, "dre_create: no CUDA device (the engine has no CPU path)\n");
    return 2;
  }
  dre_engine* e = new dre_engine();
  memset(&e->g, 0, sizeof(e->g));
  e->g.dir_kmax = -1;
  e->device = device; e->rank = rank; e->nranks = nranks < 1 ? 1 : nranks;
  try {
    CK(cudaSetDevice(device));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device));
    e->sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking));
    for (int i = 0; i < 4; i++) CK(cudaEventCreate(&e->ev[i]));
    CK(cudaEventCreateWithFlags(&e->ev_blk, cudaEventBlockingSync | cudaEventDisableTiming));
    e->counters.zero(16, e->st); e->errflag.zero(1, e->st); e->counter.zero(4, e->st); e->active.zero(2, e->st);
    e->ctl.zero(1, e->st);
    CK(cudaMallocHost((void**)&e->h_ctl, sizeof(DreCtl))); CK(cudaMallocHost((void**)&e->h_err, sizeof(int)));
    memset(e->h_ctl, 0, sizeof(DreCtl)); *e->h_err = 0;
    e->rec_n.zero(1, e->st);
    if (e->nranks > 1) {
      DRE_REQUIRE(nccl_id128 != nullptr, "dre_create: nranks > 1 needs an NCCL unique id");
      std::string err;
      DRE_REQUIRE(g_nccl.load(err), err.c_str());
      ncclUniqueId id; memcpy(&id, nccl_id128, DRE_NCCL_ID_BYTES);
      const int rc = g_nccl.CommInitRank(&e->comm, e->nranks, id, rank);
      DRE_REQUIRE(rc == 0, "ncclCommInitRank failed");
    }
    CK(cudaStreamSynchronize(e->st));
  } catch (const DreError& x) {
    fprintf(stderr, "dre_create: %s\n", x.msg.c_str());
    delete e; return 3;
  }
  *out = e;
  return 0;
}

void dre_destroy(dre_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  if (e->comm) g_nccl.CommDestroy(e->comm);
  for (auto& sm : e->sgraph) for (auto& sg : sm) { if (sg.pre) cudaGraphExecDestroy(sg.pre); if (sg.post) cudaGraphExecDestroy(sg.post); }
  if (e->h_ctl) cudaFreeHost(e->h_ctl);
  if (e->h_err) cudaFreeHost(e->h_err);
  for (int i = 0; i < 4; i++) if (e->ev[i]) cudaEventDestroy(e->ev[i]);
  if (e->ev_blk) cudaEventDestroy(e->ev_blk);
  for (int i = 0; i < 8; i++) if (e->marks[i]) cudaEventDestroy(e->marks[i]);
  if (e->st) cudaStreamDestroy(e->st);
  delete e;
}

const char* dre_last_error(const dre_engine* e) { return e ? e->err.c_str() : "null engine"; }

int dre_set_grid(dre_engine* e, int32_t N, int32_t max_lvl, const int32_t base[2], const int32_t* cchild,
                 const int32_t* lvl, const int64_t* cindex, const double* ccoord, const double* csize,
                 const double* csize_arr, const double* carea_arr, const double* cvol_arr) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(N > 1 && max_lvl >= 1, "dre_set_grid: empty grid");
  DRE_REQUIRE(max_lvl <= DRE_MAXLVL, "dre_set_grid: max_lvl > 8 not supported");
  DRE_REQUIRE(base[0] >= 2 && base[1] >= 2 && base[0] < 256 && base[1] < 256, "dre_set_grid: base out of range");
  DreDev& g = e->g;
  g.N = N; g.max_lvl = max_lvl; g.base[0] = base[0]; g.base[1] = base[1];
  for (int i = 0; i < 2; i++) {   // set_base smooth_grid_routines.f90:622-649
    int j = 0; const long long b3 = (long long)base[i] * base[i] * base[i];
    while (!(b3 < (1ll << j))) j++;
    g.basediv[i] = 1ll << j; g.basemask[i] = (int)(g.basediv[i] - 1); g.baseshift[i] = j;
  }
  std::vector<unsigned char> l8(N);
  for (int i = 0; i < N; i++) { DRE_REQUIRE(lvl[i] >= 0 && lvl[i] <= max_lvl, "dre_set_grid: lvl out of range"); l8[i] = (unsigned char)lvl[i]; }
  e->cchild.upload(cchild, N, e->st); e->lvl.upload(l8.data(), N, e->st);
  e->cindex.upload((const long long*)cindex, N, e->st);
  e->ccoord.upload(ccoord, 3 * (size_t)N, e->st); e->csize.upload(csize, N, e->st); e->h_csize0 = csize[0];
  e->h_ccoord.assign(ccoord, ccoord + 3 * (size_t)N); e->h_csize.assign(csize, csize + N); e->h_cchild.assign(cchild, cchild + N);
  CK(cudaStreamSynchronize(e->st));
  g.cchild = e->cchild.p; g.lvl = e->lvl.p; g.cindex = e->cindex.p; g.ccoord = e->ccoord.p; g.csize = e->csize.p;
  for (int i = 0; i <= max_lvl; i++) { g.csize_arr[i] = csize_arr[i]; g.carea_arr[i] = carea_arr[i]; g.cvol_arr[i] = cvol_arr[i]; }
  DRE_REQUIRE(N < (1 << 27), "dre_set_grid: more than 2^27 cells");
  e->cells.ensure(N);
  sync_g(e);
  (e->n_launch++, k_build_nbr)<<<cdiv((size_t)N * 6, 256), 256, 0, e->st>>>(e->cells.p);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(e->st));
  g.cell = e->cells.p;
  e->have_grid = true; e->have_opacity = false; e->have_sources = false;
  e->det_e.clear(); e->h_qc.clear();   // rounding grids of the deterministic mode belong to one model
  DRE_CATCH(e)
}

int dre_set_opacity(dre_engine* e, int32_t L, const double* dens_arr, const double* ksca_arr_norm, const double* gsca_arr,
                    const int32_t* kp_sca_arr, int32_t ndir, const double* dir_i_out) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_grid, "dre_set_opacity: call dre_set_grid first");
  DRE_REQUIRE(L >= 1 && L <= 32 * DRE_KL, "dre_set_opacity: lnum must be in [1,128]");
  DRE_REQUIRE(ndir >= 0, "dre_set_opacity: negative tot_ndir_scaspe");
  DreDev& g = e->g;
  const int N = g.N;
  g.L = L; g.KLn = (L + 31) / 32; g.ndir = ndir;
  for (int k = 0; k <= DRE_MAXK; k++) { e->ftab[k].release(); g.ftab[k] = nullptr; }
  e->ftab_bytes = 0; e->ftab_skipped = 0; e->ftab_budget = (size_t)24 << 30;
  e->dens.upload(dens_arr, (size_t)L * N, e->st); e->ksca.upload(ksca_arr_norm, L, e->st); e->gsca.upload(gsca_arr, L, e->st);
  // classes: set_npix_arr rt_routines.f90:5289-5311
  e->h_npix.assign(L, 0); e->h_npix_hp.assign(L, 0); e->h_off.assign(L, 0); e->h_ik.assign(L, -1);
  int kmax = -1;
  for (int l = 0; l < L; l++) {
    DRE_REQUIRE(kp_sca_arr[l] >= -1 && kp_sca_arr[l] <= DRE_NSIDE_MIN_K, "dre_set_opacity: kp_sca must be in [-1,2]");
    e->h_npix_hp[l] = kp_sca_arr[l] < 0 ? 1 : 12 * (1 << (2 * kp_sca_arr[l]));
    e->h_npix[l] = e->h_npix_hp[l] + ndir;
    kmax = std::max(kmax, kp_sca_arr[l]);
  }
  g.nclass = 0;
  for (int k = 0; k <= kmax; k++) {
    bool used = false;
    for (int l = 0; l < L; l++) if (kp_sca_arr[l] == k) used = true;
    if (!used) continue;
    const int ic = g.nclass++;
    for (int l = 0; l < L; l++) if (kp_sca_arr[l] == k) e->h_ik[l] = ic;
    const int nhp = 12 * (1 << (2 * k)), np = nhp + ndir;
    g.class_kp[ic] = k; g.class_nhp[ic] = nhp; g.class_np[ic] = np;
    std::vector<DreDir> d(np);   // create_scaspe rt_routines.f90:2827-2848
    for (int j = 0; j < nhp; j++) { double th, ph; hp::pix2ang_nest(1 << k, j, th, ph); d[j] = hp::make_dir(th, ph); }
    for (int j = 0; j < ndir; j++) d[nhp + j] = hp::make_dir(dir_i_out[j], dir_i_out[ndir + j]);
    e->class_dir[ic].upload(d.data(), np, e->st);
    CK(cudaStreamSynchronize(e->st));
    g.class_dir[ic] = e->class_dir[ic].p;
  }
  // accumulator row layout: [0,L) u_final, then one even-padded block per wavelength
  int off = L + (L & 1);
  for (int l = 0; l < L; l++) { e->h_off[l] = off; off += e->h_npix[l] + (e->h_npix[l] & 1); }
  g.ROW = off;
  DRE_REQUIRE(g.ROW < 65536, "dre_set_opacity: accumulator row too long");
  std::vector<unsigned short> lam(g.ROW, 0);
  for (int l = 0; l < L; l++) for (int p = 0; p < e->h_npix[l] + (e->h_npix[l] & 1); p++) lam[e->h_off[l] + p] = (unsigned short)(8 * l);   // byte offset of en_sca(l) in the per-warp array
  e->lamof.upload(lam.data(), g.ROW, e->st);
  e->ik_sca.upload(e->h_ik.data(), L, e->st); e->npix_hp.upload(e->h_npix_hp.data(), L, e->st);
  e->npix.upload(e->h_npix.data(), L, e->st); e->off.upload(e->h_off.data(), L, e->st);
  e->h_dir_i_out.assign(dir_i_out, dir_i_out + 2 * (size_t)ndir);
  e->h_gsca.assign(gsca_arr, gsca_arr + L); e->det_peak.clear();
  std::vector<DreDir> od(std::max(ndir, 1));
  for (int j = 0; j < ndir; j++) od[j] = hp::make_dir(dir_i_out[j], dir_i_out[ndir + j]);
  e->obs_dir.upload(od.data(), od.size(), e->st);
  CK(cudaStreamSynchronize(e->st));
  g.dens = e->dens.p; g.ksca = e->ksca.p; g.gsca = e->gsca.p; g.ik_sca = e->ik_sca.p; g.npix_hp = e->npix_hp.p;
  g.npix = e->npix.p; g.off = e->off.p; g.lamof = e->lamof.p;
  for (int l = 0; l < 32 * DRE_KL; l++) { g.off_c[l] = l < L ? e->h_off[l] : 0; g.npix_c[l] = l < L ? e->h_npix[l] : 0; }
  // device-resident accumulators (zero-initialised)
  const size_t NL = (size_t)N * L;
  e->acc.zero(e->rows_elems(), e->st); e->prev.zero(e->rows_elems(), e->st); e->tot.zero(e->rows_elems(), e->st);
  e->u_fest.zero(NL, e->st); e->lum_lost.zero(L, e->st); e->lum_lost_prev.zero(L, e->st);
  e->lumcell.zero(NL, e->st); g.lumcell = e->lumcell.p;
  CK(cudaStreamSynchronize(e->st));
  for (double& x : e->per_unit_est) x = 0;
  for (double& x : e->per_unit_evt) x = 0;
  e->list_budget = 0;
  e->capH = e->capW = e->capT = 0;     // the luminosity rows of the lists are lnum wide: re-sized by the next pass
  e->have_sym = false;
  e->have_opacity = true; e->have_sources = false;
  e->det_e.clear(); e->h_qc.clear();
  DRE_CATCH(e)
}

int dre_set_sources(dre_engine* e, const double* lumcell, int32_t S, const double* ccoord_p_src, const double* lum_p_src_arr,
                    const int32_t* src_cell, const int32_t* cell_src) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_opacity, "dre_set_sources: call dre_set_opacity first");
  DreDev& g = e->g;
  const size_t NL = (size_t)g.N * g.L;
  if (lumcell) e->lumcell.upload(lumcell, NL, e->st); else e->lumcell.zero(NL, e->st);
  g.lumcell = e->lumcell.p;
  g.S = S; e->S = S;
  if (S > 0) {
    DRE_REQUIRE(ccoord_p_src && lum_p_src_arr && src_cell && cell_src, "dre_set_sources: point-source arrays missing");
    e->ccoord_p_src.upload(ccoord_p_src, 3 * (size_t)S, e->st); e->lum_p_src.upload(lum_p_src_arr, (size_t)S * g.L, e->st);
    e->h_ccoord_p_src.assign(ccoord_p_src, ccoord_p_src + 3 * (size_t)S);
    e->src_cell.upload(src_cell, g.N, e->st);
    e->h_cell_src.assign(cell_src, cell_src + S);
  } else {
    std::vector<int> none(g.N, -1);
    e->src_cell.upload(none.data(), g.N, e->st);
    e->ccoord_p_src.zero(3, e->st); e->lum_p_src.zero(g.L, e->st);
  }
  CK(cudaStreamSynchronize(e->st));
  g.ccoord_p_src = e->ccoord_p_src.p; g.lum_p_src = e->lum_p_src.p; g.src_cell = e->src_cell.p;
  const size_t NS = (size_t)g.N + S;
  e->psel_sum.zero(NS, e->st); e->ipsel.zero(NS, e->st); e->psel_av.zero(NS, e->st);
  e->have_sources = true;
  DRE_CATCH(e)
}

int dre_set_params(dre_engine* e, const dre_params* p) {
  DRE_TRY(e)
  DRE_REQUIRE(p != nullptr, "dre_set_params: null");
  DRE_REQUIRE(p->bm_par > 0 && p->bm_par_max > 0, "dre_set_params: bm_par / bm_par_max must be positive");
  e->g.par = *p;
  e->have_params = true;
  DRE_CATCH(e)
}

}  // extern "C"

// =========================================================================================
// device-resident state: upload / download / zero
// =========================================================================================
namespace {
double* rows_of(dre_engine* e, int which) {
  switch (which) {
    case DRE_ARR_SCASPE: return e->acc.p;
    case DRE_ARR_SCASPE_PREV: return e->prev.p;
    case DRE_ARR_SCASPE_TOT: return e->tot.p;
  }
  return nullptr;
}
void planes_io(dre_engine* e, double* rows, double* const* planes, int dir) {
  // The Fortran side holds one (P(l), N) plane per wavelength (ragged scaspe_arr(l)%a).  All planes go through
  // one device staging buffer: the copies are queued back to back on the stream, the transposes to / from the
  // accumulator-row layout follow, and the host waits once.
  const DreDev& g = e->g;
  size_t tot = 0;
  std::vector<size_t> offs(g.L);
  for (int l = 0; l < g.L; l++) { offs[l] = tot; tot += (size_t)g.N * e->h_npix[l]; }
  e->tmp_d.ensure(tot);
  for (int l = 0; l < g.L; l++) {
    const int P = e->h_npix[l];
    const size_t n = (size_t)g.N * P;
    double* t = e->tmp_d.p + offs[l];
    if (dir == 0) CK(cudaMemcpyAsync(t, planes[l], n * 8, cudaMemcpyHostToDevice, e->st));
    (e->n_launch++, k_plane_rows)<<<cdiv(n, 256), 256, 0, e->st>>>(rows, t, g.N, P, g.ROW, e->h_off[l], dir);
    if (dir == 1) CK(cudaMemcpyAsync(planes[l], t, n * 8, cudaMemcpyDeviceToHost, e->st));
  }
  CK(cudaStreamSynchronize(e->st));
}
void u_final_io(dre_engine* e, double* host, int dir) {
  const DreDev& g = e->g;
  const size_t n = (size_t)g.N * g.L;
  e->tmp_d.ensure(n);
  if (dir == 0) CK(cudaMemcpyAsync(e->tmp_d.p, host, n * 8, cudaMemcpyHostToDevice, e->st));
  (e->n_launch++, k_u_rows)<<<cdiv(n, 256), 256, 0, e->st>>>(e->acc.p, e->tmp_d.p, g.N, g.L, g.ROW, dir);
  if (dir == 1) CK(cudaMemcpyAsync(host, e->tmp_d.p, n * 8, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
}
}  // namespace

extern "C" {

int dre_upload(dre_engine* e, int which, const void* planes) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_opacity, "dre_upload: call dre_set_opacity first");
  DRE_REQUIRE(planes != nullptr, "dre_upload: null pointer");
  const DreDev& g = e->g;
  const size_t NL = (size_t)g.N * g.L;
  switch (which) {
    case DRE_ARR_U_FEST: e->u_fest.upload((const double*)planes, NL, e->st); break;
    case DRE_ARR_U_FINAL: u_final_io(e, (double*)planes, 0); break;
    case DRE_ARR_SCASPE: case DRE_ARR_SCASPE_PREV: case DRE_ARR_SCASPE_TOT:
      planes_io(e, rows_of(e, which), (double* const*)planes, 0); break;
    case DRE_ARR_LUM_LOST: e->lum_lost.upload((const double*)planes, g.L, e->st); break;
    case DRE_ARR_LUMCELL: e->lumcell.upload((const double*)planes, NL, e->st); break;
    default: DRE_REQUIRE(false, "dre_upload: array cannot be uploaded");
  }
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

int dre_download(dre_engine* e, int which, void* planes) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_opacity, "dre_download: call dre_set_opacity first");
  DRE_REQUIRE(planes != nullptr, "dre_download: null pointer");
  const DreDev& g = e->g;
  const size_t NL = (size_t)g.N * g.L, NS = (size_t)g.N + e->S;
  switch (which) {
    case DRE_ARR_U_FEST: CK(cudaMemcpyAsync(planes, e->u_fest.p, NL * 8, cudaMemcpyDeviceToHost, e->st)); break;
    case DRE_ARR_U_FINAL: u_final_io(e, (double*)planes, 1); break;
    case DRE_ARR_SCASPE: case DRE_ARR_SCASPE_PREV: case DRE_ARR_SCASPE_TOT:
      planes_io(e, rows_of(e, which), (double* const*)planes, 1); break;
    case DRE_ARR_LUM_LOST: CK(cudaMemcpyAsync(planes, e->lum_lost.p, g.L * 8, cudaMemcpyDeviceToHost, e->st)); break;
    case DRE_ARR_LUMCELL: CK(cudaMemcpyAsync(planes, e->lumcell.p, NL * 8, cudaMemcpyDeviceToHost, e->st)); break;
    case DRE_ARR_PSEL_AV:
      DRE_REQUIRE(e->have_sources, "dre_download: call dre_set_sources first");
      (e->n_launch++, k_psel_ratio)<<<cdiv(NS, 256), 256, 0, e->st>>>(e->psel_sum.p, e->ipsel.p, (int)NS, e->psel_av.p);
      CK(cudaMemcpyAsync(planes, e->psel_av.p, NS * 8, cudaMemcpyDeviceToHost, e->st)); break;
    case DRE_ARR_I_OBS:
      DRE_REQUIRE(e->i_obs.p, "dre_download: no i_obs pass was run");
      CK(cudaMemcpyAsync(planes, e->i_obs.p, (size_t)g.ndir * NS * g.L * 8, cudaMemcpyDeviceToHost, e->st)); break;
    case DRE_ARR_I_OBS_IN:
      DRE_REQUIRE(e->i_obs_in.p && e->i_obs_ndir_in > 0, "dre_download: no internal-observer pass was run");
      CK(cudaMemcpyAsync(planes, e->i_obs_in.p, (size_t)e->i_obs_ndir_in * NS * g.L * 8, cudaMemcpyDeviceToHost, e->st)); break;
    default: DRE_REQUIRE(false, "dre_download: unknown array");
  }
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

int dre_zero(dre_engine* e, int which) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_opacity, "dre_zero: call dre_set_opacity first");
  const DreDev& g = e->g;
  const size_t NL = (size_t)g.N * g.L, NS = (size_t)g.N + e->S;
  const size_t nsca = (size_t)g.N * (g.ROW - g.L);
  switch (which) {
    case DRE_ARR_U_FEST: e->u_fest.zero(NL, e->st); break;
    case DRE_ARR_U_FINAL: e->tmp_d.zero(NL, e->st); (e->n_launch++, k_u_rows)<<<cdiv(NL, 256), 256, 0, e->st>>>(e->acc.p, e->tmp_d.p, g.N, g.L, g.ROW, 0); break;
    case DRE_ARR_SCASPE: case DRE_ARR_SCASPE_PREV: case DRE_ARR_SCASPE_TOT:
      (e->n_launch++, k_zero_sca)<<<cdiv(nsca, 256), 256, 0, e->st>>>(rows_of(e, which), g.N, g.ROW, g.L); break;
    case DRE_ARR_LUM_LOST: e->lum_lost.zero(g.L, e->st); break;
    case DRE_ARR_PSEL_AV: e->psel_sum.zero(NS, e->st); e->ipsel.zero(NS, e->st); break;
    case DRE_ARR_LUMCELL: e->lumcell.zero(NL, e->st); break;
    case DRE_ARR_I_OBS: if (e->i_obs.p) e->i_obs.zero((size_t)g.ndir * NS * g.L, e->st); break;
    case DRE_ARR_I_OBS_IN: if (e->i_obs_in.p) e->i_obs_in.zero((size_t)e->i_obs_ndir_in * NS * g.L, e->st); break;
    default: DRE_REQUIRE(false, "dre_zero: unknown array");
  }
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

}  // extern "C"

// =========================================================================================
// rt_loop on the device: batches of (source, sector) units advanced as one wavefront
// =========================================================================================
namespace {

DreAccum make_accum(dre_engine* e, const double* prev_rows) {
  DreAccum A; memset(&A, 0, sizeof(A));
  A.acc = e->acc.p; A.u_fest = e->u_fest.p; A.prev = prev_rows; A.lum_lost = e->lum_lost.p;
  A.psel_sum = e->psel_sum.p; A.ipsel = e->ipsel.p; A.counters = e->counters.p;
  A.rec = e->rec.p; A.rec_cap = e->rec_cap; A.rec_n = e->rec_n.p; A.err = e->errflag.p; A.use_bulk = e->opt_use_bulk; A.want_hash = e->opt_hash; A.stage_split = e->opt_stage_split; A.ring_slots = e->opt_ring_slots;
  A.det = e->opt_det; A.qc = e->opt_det ? e->qc.p : nullptr;
  return A;
}

void check_ready(dre_engine* e) {
  DRE_REQUIRE(e->have_grid && e->have_opacity && e->have_sources && e->have_params,
              "engine not configured: dre_set_grid/opacity/sources/params must all be called");
}

const char* err_text(int code) {
  switch (code) {
    case 1: return "ray_tracing: negative path length (rt_routines.f90:1714)";
    case 2: return "ray_tracing: next cell not found (rt_routines.f90:1849)";
    case 3: return "find_cc_new2: host cell not found (rt_routines.f90:2037)";
    case 4: return "angular resolution exceeds nside 8192, the limit of the HEALPix routines (healpix_routines.f90:9)";
    case 5: return "define_next_level: nside fell below 2";
  }
  return "unknown device error";
}

// build the luminous-leaf list for the current rt_type (rt_routines.f90:1016)
int build_source_list(dre_engine* e) {
  const DreDev& g = e->g;
  e->srcflag.ensure(g.N); e->srcpos.ensure((size_t)g.N + 1); e->srclist.ensure(g.N);
  sync_g(e);
  (e->n_launch++, k_src_flags)<<<cdiv(g.N, 256), 256, 0, e->st>>>(e->srcflag.p);
  scan_host_n(e, e->srcflag.p, e->srcpos.p, g.N);
  (e->n_launch++, k_src_compact)<<<cdiv(g.N, 256), 256, 0, e->st>>>(e->srcflag.p, e->srcpos.p, g.N, e->srclist.p);
  e->nsrc_list = read_int(e, e->srcpos.p + g.N);
  return e->nsrc_list;
}

int trace_warps(dre_engine* e, size_t* smem) {
  const DreDev& g = e->g;
  int w = std::max(1, std::min(e->opt_warps, 32));
  w = std::min(w, DRE_TRACE_THREADS / 32);
  const size_t extra = e->opt_det ? dre_det_smem_bytes(g.L) : 0;
  while (w > 1 && (dre_trace_smem_bytes(g.ROW, g.L, w, e->opt_stage_split) + extra) * e->opt_ctas_per_sm > (size_t)224 * 1024) w--;
  *smem = dre_trace_smem_bytes(g.ROW, g.L, w, e->opt_stage_split) + extra;
  DRE_REQUIRE(*smem <= (size_t)227 * 1024, "accumulator row does not fit in shared memory");
  return w;
}

// phase-function rows for every pixel of nside 2^k (used by the deposit passes up to ftab_kmax)
void ensure_ftab(dre_engine* e, unsigned nkmask) {
  DreDev& g = e->g;
  for (int k = DRE_NSIDE_MIN_K; k <= std::min(e->opt_ftab_kmax, 8); k++) {
    if (!((nkmask >> k) & 1u) || g.ftab[k] || ((e->ftab_skipped >> k) & 1u)) continue;
    build_dir_table(e, k);
    const size_t npix = (size_t)12 << (2 * k);
    // the tables are a cache of calc_ffn_arr: beyond the byte budget a level stays untabulated and its rays evaluate the
    // phase function themselves (dre_phase), as the rays above ftab_kmax do
    if (e->ftab_bytes + npix * g.ROW * 8 > e->ftab_budget) { e->ftab_skipped |= 1u << k; continue; }
    e->ftab_bytes += npix * g.ROW * 8;
    e->ftab[k].ensure(npix * g.ROW, 1.0);
    sync_g(e);
    e->n_launch++;
    k_build_ftab<<<cdiv(npix * 32, 256), 256, 0, e->st>>>(k, e->ftab[k].p);
    CK(cudaGetLastError());
    g.ftab[k] = e->ftab[k].p;
    sync_g(e);
  }
}

// ---- one resolution step of the wavefront = [pre] list extraction, merge, tasks, ray packets -> [trace] the ray-march
// kernel -> [post] outcome flags, rebuilt lists, define_next_level.  Every kernel reads its counts from device memory
// (dre_lists.cuh), so the launch sequence is a constant: pre and post are replayed from CUDA graphs, the ray-march kernel
// is launched between them with CUDA events around it (its device time is the roofline denominator).
template <int MODE, int KL, int DET>
void launch_trace_kl(dre_engine* e, const DreAccum& A, const DreStep& S) {
  size_t smem = 0; int warps = 8;
  if (MODE == MODE_DEP) {
    warps = trace_warps(e, &smem);
    int& set = g_smem_set[e->device & 63][MODE][KL == 4 ? 2 : KL - 1][DET];
    if (set != (int)smem) { CK(cudaFuncSetAttribute(k_trace<MODE, KL, DET>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set = (int)smem; }
  } else smem = (size_t)warps * (8 + 32 * e->g.KLn + 2 + dre_pkt_stride(e->g.L)) * 8 + (DET ? dre_det_smem_bytes(e->g.L) : 0);
  const int ctas = e->sm_count * (MODE == MODE_DEP ? e->opt_ctas_per_sm : 4);
  CK(cudaEventRecord(e->ev[2], e->st));
  e->n_launch++;
  k_trace<MODE, KL, DET><<<ctas, warps * 32, smem, e->st>>>(A, S, e->opt_chunk);
  CK(cudaEventRecord(e->ev[3], e->st));
  CK(cudaGetLastError());
}

// deposit passes as geometry (thread per ray, two passes around a prefix sum) + deposit (warp per ray): dre_split.cuh.
// with_geom: the event offsets and pass 1 of the march are launched here (they are part of k_step_pre_coop otherwise).
template <int KL, int DET>
void launch_split_kl(dre_engine* e, const DreAccum& A, const DreStep& S, bool with_geom) {
  size_t smem = 0;
  const int warps = trace_warps(e, &smem);
  static_assert(KL == 1 || KL == 2 || KL == 4, "wavelengths per lane");
  int& set = g_smem_set[e->device & 63][2][KL == 4 ? 2 : KL - 1][DET];   // slot [2]: k_deposit
  if (set != (int)smem) { CK(cudaFuncSetAttribute(k_deposit<KL, DET>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set = (int)smem; }
  if (with_geom) {
    ScanJobs J; memset(&J, 0, sizeof(J));
    J.j[0].in = e->nev.p; J.j[0].out = e->evoff.p; J.j[0].n_ptr = &S.ctl->nT; J.j[0].slot = 10;
    k_scan_jobs<<<dim3(e->sm_count, 1), SCAN_THREADS, 0, e->st>>>(J, S.ctl, S.scan_desc, S.desc_stride, 10u, 1);
    k_geom_write<<<e->sm_count * 2 * DRE_GEOM_MINB, 128, 0, e->st>>>(A, S, e->evt.p, e->evoff.p, e->capE);
    e->n_launch += 2;
  }
  CK(cudaEventRecord(e->ev[2], e->st));
  k_deposit<KL, DET><<<e->sm_count * e->opt_ctas_per_sm, warps * 32, smem, e->st>>>(A, S, e->evt.p, e->opt_chunk);
  CK(cudaEventRecord(e->ev[3], e->st));
  CK(cudaGetLastError());
  e->n_launch += 1;
}
void launch_split(dre_engine* e, const DreAccum& A, const DreStep& S, bool with_geom) {
  const int det = A.det ? 1 : 0;
  switch (e->g.KLn * 2 + det) {
    case 2: launch_split_kl<1, 0>(e, A, S, with_geom); break;
    case 3: launch_split_kl<1, 1>(e, A, S, with_geom); break;
    case 4: launch_split_kl<2, 0>(e, A, S, with_geom); break;
    case 5: launch_split_kl<2, 1>(e, A, S, with_geom); break;
    default: if (det) launch_split_kl<4, 1>(e, A, S, with_geom); else launch_split_kl<4, 0>(e, A, S, with_geom); break;
  }
}

template <int MODE>
void launch_trace(dre_engine* e, const DreAccum& A, const DreStep& S) {
  if (MODE == MODE_DEP && e->opt_split) { launch_split(e, A, S, true); return; }
  if (A.det) {
    switch (e->g.KLn) {
      case 1: launch_trace_kl<MODE, 1, 1>(e, A, S); break;
      case 2: launch_trace_kl<MODE, 2, 1>(e, A, S); break;
      default: launch_trace_kl<MODE, 4, 1>(e, A, S); break;
    }
    return;
  }
  switch (e->g.KLn) {
    case 1: launch_trace_kl<MODE, 1, 0>(e, A, S); break;
    case 2: launch_trace_kl<MODE, 2, 0>(e, A, S); break;
    default: launch_trace_kl<MODE, 4, 0>(e, A, S); break;
  }
}

// Deterministic mode (option "deterministic"), start of a list pass.  Bounds of the sums, from the total luminosity
// Lt(l) of the sources of the pass (every unit of luminosity crosses a cell at most once, over at most sqrt(3) csize):
//   scaspe element <= Lt, u <= sqrt(3) Lt / (cs csize_min^2), lum_lost <= Lt   (x2 for slack; running maxima, so that
// what earlier passes left in the accumulators stays on the grid).  Quantum q = 2^(e-52) for a bound <= 2^e: sums stay
// exact up to 2x the bound; beyond that only bit-reproducibility degrades, not the value.  The price is dynamic range:
// an element keeps an absolute accuracy of ~sqrt(n) q / 2 after n contributions, i.e. elements below ~1e-10 of the bound
// of their wavelength no longer meet 1e-5 relative (deeply shadowed cells of an optically thick model).
// max(1, largest F of an observer slot): Henyey-Greenstein peak (1+|g|)/((1-|g|)^2 npix_hp) over the smallest normaliser
// sum_hp f, the latter sampled for ray directions at the centres of nside 8 (calc_ffn_arr rt_routines.f90:2186-2281)
double det_peak_factor(dre_engine* e, int l) {
  const DreDev& g = e->g;
  if (g.ndir == 0 || e->h_ik[l] < 0) return 1.0;
  if ((int)e->det_peak.size() != g.L) {
    e->det_peak.assign(g.L, 1.0);
    for (int w = 0; w < g.L; w++) {
      const int ic = e->h_ik[w];
      if (ic < 0) continue;
      const double gg = std::fabs(e->h_gsca[w]);
      const int kp = g.class_kp[ic], nhp = g.class_nhp[ic];
      std::vector<DreDir> cd(nhp);
      for (int j = 0; j < nhp; j++) { double th, ph; hp::pix2ang_nest(1 << kp, j, th, ph); cd[j] = hp::make_dir(th, ph); }
      double tot_min = 1e300;
      for (int r = 0; r < 768; r++) {
        double th, ph; hp::pix2ang_nest(8, r, th, ph);
        const DreDir rd = hp::make_dir(th, ph);
        double tot = 0;
        for (int j = 0; j < nhp; j++) {
          const double ads = cd[j].ct * rd.ct + cd[j].st * rd.st * (cd[j].cp * rd.cp + cd[j].sp * rd.sp);
          const double x = 1.0 + gg * gg - 2.0 * gg * ads;
          tot += (1.0 - gg * gg) / nhp / (x * std::sqrt(x));
        }
        tot_min = std::min(tot_min, tot);
      }
      const double fpeak = (1.0 + gg) / ((1.0 - gg) * (1.0 - gg) * nhp);
      e->det_peak[w] = std::max(1.0, fpeak / tot_min);
    }
  }
  return e->det_peak[l];
}

void det_prepare(dre_engine* e, int rt_type, int mode) {
  const DreDev& g = e->g;
  const int L = g.L, Lq = dre_lq(L), nq = 3 * Lq + 2;
  const bool src_pass = (rt_type == DRE_RTT_PRECALC_SRC || rt_type == DRE_RTT_DIR_SRC);
  if ((int)e->det_e.size() != nq) e->det_e.assign(nq, INT_MIN);
  e->red_d.ensure((size_t)2 * L);
  std::vector<double> lt(L, 0.0);
  if (!src_pass || e->S > 0) {
    e->n_launch++;
    if (src_pass) k_colsum<<<L, 256, 0, e->st>>>(e->lum_p_src.p, e->S, L, e->red_d.p);
    else k_colsum<<<L, 256, 0, e->st>>>(e->lumcell.p, g.N, L, e->red_d.p);
    CK(cudaMemcpyAsync(lt.data(), e->red_d.p, (size_t)L * 8, cudaMemcpyDeviceToHost, e->st));
    CK(cudaStreamSynchronize(e->st));
  }
  const double csmin = g.csize_arr[g.max_lvl];
  auto bump = [&](int idx, double bound) {
    if (!(bound > 0) || !std::isfinite(bound)) return;
    int ex; std::frexp(bound, &ex);          // bound = m 2^ex, m in [0.5, 1)
    e->det_e[idx] = std::max(e->det_e[idx], ex);
  };
  for (int l = 0; l < L; l++) {
    // scaspe is zero at the start of a scattering iteration (rt_routines.f90:747-756): its grid follows the
    // luminosity of the iteration instead of the running maximum
    if (rt_type == DRE_RTT_SCATT && lt[l] > 0) e->det_e[l] = INT_MIN;
    // an element of a HEALPix slot is at most Lt (the row is normalised over those slots); an observer-direction slot
    // shares the normaliser but not the sum: its phase-function value reaches f_peak / min(normaliser) for forward-peaked g
    bump(l, 2.0 * lt[l] * det_peak_factor(e, l));
    bump(Lq + l, 2.0 * std::sqrt(3.0) * lt[l] / (g.par.cs * csmin * csmin));
    bump(2 * Lq + l, 2.0 * lt[l]);
  }
  bump(3 * Lq, g.par.modelsize * 17179869184.0);   // psel_sum of a source: up to 2^34 rays of at most ~modelsize each
  e->h_qc.assign(nq, 0.0);
  for (int i = 0; i < nq; i++) if (e->det_e[i] != INT_MIN) e->h_qc[i] = std::ldexp(1.5, e->det_e[i]);
  e->qc.upload(e->h_qc.data(), nq, e->st);
  const size_t NL = (size_t)g.N * L;
  if (mode == MODE_PRECALC) {
    (e->n_launch++, k_quantize_cols)<<<cdiv(NL, 256), 256, 0, e->st>>>(e->u_fest.p, NL, L, e->qc.p + Lq, nullptr);
  } else {
    const size_t n = e->rows_elems();
    (e->n_launch++, k_quantize_rows)<<<cdiv(n, 256), 256, 0, e->st>>>(e->acc.p, n, g.ROW, L, g.lamof, e->qc.p);
    (e->n_launch++, k_quantize_cols)<<<1, 256, 0, e->st>>>(e->lum_lost.p, (size_t)L, L, e->qc.p + 2 * Lq, nullptr);
  }
  CK(cudaGetLastError());
}

// Buffers of the wavefront for U units and the given list / task capacities.  `keep_cur` >= 0: the contents of the
// lists H[keep_cur] / W[keep_cur] (nH / nW entries) survive the growth (a refused step is repeated after it).
void ensure_step(dre_engine* e, int U, int capH, int capW, int capT, int keep_cur, int nH, int nW) {
  const int L = e->g.L;
  if (U > e->capU) {
    e->u_nc.ensure(U); e->u_gid.ensure(U); e->u_nk.ensure(U); e->u_step.ensure(U); e->u_mode.ensure(U); e->u_task_seg.ensure((size_t)U + 1);
    e->extcnt.ensure(U); e->ext_seg.ensure((size_t)U + 1); e->reco_cnt.ensure(U);
    e->keptH.ensure(U); e->keptW.ensure(U); e->newHcnt.ensure(U); e->newWcnt.ensure(U);
    for (int i = 0; i < 2; i++) { e->H[i].seg.ensure((size_t)U + 1); e->W[i].seg.ensure((size_t)U + 1); }
    e->capU = (int)std::min<size_t>(e->u_nc.cap, e->reco_cnt.cap);
  }
  capH = std::max(capH, e->capH); capW = std::max(capW, e->capW); capT = std::max(capT, e->capT);
  if (capH > e->capH || capW > e->capW) {
    for (int i = 0; i < 2; i++) {
      if (i == keep_cur) { e->H[i].grow_keep(capH, nH, L, e->st); e->W[i].grow_keep(capW, nW, L, e->st); }
      else { e->H[i].grow_keep(capH, 0, L, e->st); e->W[i].grow_keep(capW, 0, L, e->st); }
    }
    e->selH.ensure(capH); e->posH.ensure((size_t)capH + 1); e->selW.ensure(capW); e->posW.ensure((size_t)capW + 1);
    const size_t nx = (size_t)capH + capW;
    e->ext.ensure(nx); e->ipx.ensure(nx); e->gone.ensure(nx); e->cand.ensure(3 * nx); e->ncand.ensure(nx); e->nxt.ensure(nx); e->hasprev.ensure(nx);
    e->unit_of_ext.ensure(nx); e->ntask.ensure(nx); e->nreco.ensure(nx); e->task_off.ensure(nx + 1); e->reco_off.ensure(nx + 1);
    e->hlink.ensure(nx); e->hkey.ensure(2 * (size_t)capW + 64); e->hhead.ensure(2 * (size_t)capW + 64);
    e->capH = capH; e->capW = capW;
  }
  if (capT > e->capT) {
    e->tasks.ensure(capT); e->OUT.grow_keep(capT, 0, L, e->st);
    e->fH.ensure(capT); e->fW.ensure(capT); e->posOH.ensure((size_t)capT + 1); e->posOW.ensure((size_t)capT + 1);
    e->pkts.ensure((size_t)capT * dre_pkt_stride(L));
    e->nev.ensure(capT); e->evoff.ensure((size_t)capT + 1);
    e->capT = capT;
  }
  if (e->capE < 1024) { e->evt.ensure(1 << 20); e->capE = 1 << 20; }
  ensure_scan_desc(e, cdiv((size_t)std::max(std::max(e->capH + e->capW, e->capT), std::max(U, e->g.N)) + 1, SCAN_TILE));
}

DreStep make_step(dre_engine* e, int U) {
  DreStep S; memset(&S, 0, sizeof(S));
  for (int i = 0; i < 2; i++) { S.H[i] = e->H[i].view(); S.W[i] = e->W[i].view(); S.H[i].n = S.W[i].n = 0; S.H[i].cap = e->capH; S.W[i].cap = e->capW; }
  S.OUT = e->OUT.view(); S.OUT.n = 0; S.OUT.cap = e->capT; S.OUT.seg = nullptr;
  S.u.nc = e->u_nc.p; S.u.gid = e->u_gid.p; S.u.nk = e->u_nk.p; S.u.step = e->u_step.p; S.u.mode = e->u_mode.p; S.u.task_seg = e->u_task_seg.p; S.u.U = 0;   // the unit count lives in DreCtl::U (set by k_seed): a graph serves every batch size
  (void)U;
  S.selH = e->selH.p; S.posH = e->posH.p; S.selW = e->selW.p; S.posW = e->posW.p; S.extcnt = e->extcnt.p; S.ext_seg = e->ext_seg.p;
  S.ext = e->ext.p; S.ipx = e->ipx.p; S.gone = e->gone.p; S.cand = e->cand.p; S.ncand = e->ncand.p; S.nxt = e->nxt.p; S.hasprev = e->hasprev.p;
  S.unit_of_ext = e->unit_of_ext.p; S.ntask = e->ntask.p; S.nreco = e->nreco.p; S.task_off = e->task_off.p; S.reco_off = e->reco_off.p;
  S.reco_cnt = e->reco_cnt.p; S.keptH = e->keptH.p; S.keptW = e->keptW.p; S.newHcnt = e->newHcnt.p; S.newWcnt = e->newWcnt.p;
  S.fH = e->fH.p; S.fW = e->fW.p; S.posOH = e->posOH.p; S.posOW = e->posOW.p; S.tasks = e->tasks.p; S.pkts = e->pkts.p;
  S.hkey = e->hkey.p; S.hhead = e->hhead.p; S.hlink = e->hlink.p;
  S.hcap = (int)std::min<size_t>(std::min(e->hkey.cap, e->hhead.cap), 2 * (size_t)e->capW + 64 > 2147483000ull ? 2147483000ull : 2 * (size_t)e->capW + 64);
  S.pstride = dre_pkt_stride(e->g.L); S.capH = e->capH; S.capW = e->capW; S.capT = e->capT; S.ctl = e->ctl.p;
  S.scan_desc = e->scan_desc.p; S.desc_stride = e->desc_stride; S.err = e->errflag.p; S.steps = e->counters.p + CNT_N; S.kmax = DRE_MAXK;
  return S;
}

ScanJobs scan_jobs(const int* in0, int* out0, const int* n0, int nc0, int slot0, const int* in1, int* out1, const int* n1, int nc1, int slot1,
                   int* total0 = nullptr) {
  ScanJobs J; memset(&J, 0, sizeof(J));
  J.j[0].total_out = total0;
  J.j[0].in = in0; J.j[0].out = out0; J.j[0].n_ptr = n0; J.j[0].n_const = nc0; J.j[0].slot = slot0;
  J.j[1].in = in1; J.j[1].out = out1; J.j[1].n_ptr = n1; J.j[1].n_const = nc1; J.j[1].slot = slot1;
  return J;
}

// launches of a step in front of the ray-march kernel; returns their number
template <int MODE>
int enqueue_pre(dre_engine* e, const DreAccum& A, const DreStep& S, int cur, cudaStream_t st) {
  const int G = e->sm_count * 2, Gs = e->sm_count;
  const int* Un = &S.ctl->U;
  k_step_select<<<G, STEP_THREADS, 0, st>>>(S, cur);
  k_scan_jobs<<<dim3(Gs, 2), SCAN_THREADS, 0, st>>>(scan_jobs(S.selH, S.posH, &S.ctl->nH, 0, 0, S.selW, S.posW, &S.ctl->nW, 0, 1), S.ctl, S.scan_desc, S.desc_stride, 0u, 0);
  k_step_unit_ext<<<G, STEP_THREADS, 0, st>>>(S, cur);
  k_scan_jobs<<<dim3(Gs, 1), SCAN_THREADS, 0, st>>>(scan_jobs(S.extcnt, S.ext_seg, Un, 0, 2, nullptr, nullptr, nullptr, 0, 0, &S.ctl->nExt), S.ctl, S.scan_desc, S.desc_stride, 2u, 0);
  k_step_fill_ext<<<G, STEP_THREADS, 0, st>>>(S, cur);
  k_step_cand<<<G, STEP_THREADS, 0, st>>>(S, cur);
  k_step_resolve<<<G, STEP_THREADS, 0, st>>>(S, cur);
  k_scan_jobs<<<dim3(Gs, 2), SCAN_THREADS, 0, st>>>(scan_jobs(S.ntask, S.task_off, &S.ctl->nExt, 0, 3, S.nreco, S.reco_off, &S.ctl->nExt, 0, 4), S.ctl, S.scan_desc, S.desc_stride, 3u, 0);
  k_step_tasks<<<G, STEP_THREADS, 0, st>>>(S, cur);
  if (MODE == MODE_DEP && e->opt_split) k_prep_geom<<<e->sm_count * 2 * DRE_GEOM_MINB, 128, 0, st>>>(A, S, cur, e->nev.p);   // packets + pass 0 of the march
  else k_prep_rays<MODE><<<e->sm_count * 8, 128, 0, st>>>(A, S, cur);
  return 10;
}
// launches of a step behind the ray-march kernel, ending with the copy of the control block to the host
int enqueue_post(dre_engine* e, const DreStep& S, int cur, cudaStream_t st, bool split) {
  const int G = e->sm_count * 2, Gs = e->sm_count;
  const int* Un = &S.ctl->U;
  if (!split) k_step_out_flags<<<G, STEP_THREADS, 0, st>>>(S);   // k_geom / k_deposit leave the outcome flags themselves
  k_scan_jobs<<<dim3(Gs, 2), SCAN_THREADS, 0, st>>>(scan_jobs(S.fH, S.posOH, &S.ctl->nT, 0, 5, S.fW, S.posOW, &S.ctl->nT, 0, 6), S.ctl, S.scan_desc, S.desc_stride, 5u, 1);
  k_step_unit_new<<<G, STEP_THREADS, 0, st>>>(S, cur);
  k_scan_jobs<<<dim3(Gs, 2), SCAN_THREADS, 0, st>>>(scan_jobs(S.newHcnt, S.H[1 - cur].seg, Un, 0, 7, S.newWcnt, S.W[1 - cur].seg, Un, 0, 8), S.ctl, S.scan_desc, S.desc_stride, 7u, 1);
  k_step_build<<<e->sm_count * 6, STEP_THREADS, 0, st>>>(S, cur, e->g.L);
  k_step_next<<<G, STEP_THREADS, 0, st>>>(S, cur);
  CK(cudaMemcpyAsync(e->h_ctl, S.ctl, sizeof(DreCtl), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(e->h_err, S.err, sizeof(int), cudaMemcpyDeviceToHost, st));
  return split ? 5 : 6;
}

// the two replayable halves of a step for list parity `cur`; re-captured when a pointer, a capacity or an option changed
template <int MODE>
void step_graphs(dre_engine* e, const DreAccum& A, const DreStep& S, int cur) {
  dre_engine::StepGraph& sg = e->sgraph[MODE == MODE_DEP ? 1 : 0][cur];
  // the key lists every value a captured launch freezes (field by field: struct padding is not comparable)
  std::vector<unsigned char> key;
  auto kp = [&key](const auto& v) { const unsigned char* q = reinterpret_cast<const unsigned char*>(&v); key.insert(key.end(), q, q + sizeof(v)); };
  auto kl = [&kp](const DreList& l) { kp(l.unit); kp(l.dir_ipix); kp(l.meta); kp(l.cc_old); kp(l.dplane); kp(l.prev); kp(l.lum); kp(l.seg); kp(l.cap); };
  kp(A.acc); kp(A.u_fest); kp(A.prev); kp(A.lum_lost); kp(A.psel_sum); kp(A.ipsel); kp(A.counters); kp(A.rec); kp(A.rec_cap); kp(A.rec_n);
  kp(A.err); kp(A.stage_split); kp(A.want_hash); kp(A.ring_slots); kp(A.use_bulk); kp(A.det); kp(A.qc);
  for (int i = 0; i < 2; i++) { kl(S.H[i]); kl(S.W[i]); }
  kl(S.OUT);
  kp(S.u.nc); kp(S.u.gid); kp(S.u.nk); kp(S.u.step); kp(S.u.mode); kp(S.u.task_seg);
  kp(S.selH); kp(S.posH); kp(S.selW); kp(S.posW); kp(S.extcnt); kp(S.ext_seg); kp(S.ext); kp(S.ipx); kp(S.gone); kp(S.cand); kp(S.ncand);
  kp(S.nxt); kp(S.hasprev); kp(S.unit_of_ext); kp(S.ntask); kp(S.nreco); kp(S.task_off); kp(S.reco_off); kp(S.reco_cnt); kp(S.keptH);
  kp(S.keptW); kp(S.newHcnt); kp(S.newWcnt); kp(S.fH); kp(S.fW); kp(S.posOH); kp(S.posOW); kp(S.tasks); kp(S.pkts); kp(S.pstride);
  kp(S.hkey); kp(S.hhead); kp(S.hlink); kp(S.hcap);
  kp(S.capH); kp(S.capW); kp(S.capT); kp(S.ctl); kp(S.scan_desc); kp(S.desc_stride); kp(S.err); kp(S.steps); kp(S.kmax);
  const int mode = MODE, L = e->g.L, smc = e->sm_count, split = e->opt_split;
  kp(mode); kp(L); kp(smc); kp(split); kp(e->nev.p);
  if (sg.pre && sg.key == key) return;
  if (sg.pre) { cudaGraphExecDestroy(sg.pre); sg.pre = nullptr; }
  if (sg.post) { cudaGraphExecDestroy(sg.post); sg.post = nullptr; }
  cudaGraph_t gr = nullptr;
  CK(cudaStreamBeginCapture(e->st, cudaStreamCaptureModeThreadLocal));
  enqueue_pre<MODE>(e, A, S, cur, e->st);
  CK(cudaStreamEndCapture(e->st, &gr));
  CK(cudaGraphInstantiate(&sg.pre, gr, 0)); CK(cudaGraphDestroy(gr));
  CK(cudaStreamBeginCapture(e->st, cudaStreamCaptureModeThreadLocal));
  enqueue_post(e, S, cur, e->st, MODE == MODE_DEP && e->opt_split);
  CK(cudaStreamEndCapture(e->st, &gr));
  CK(cudaGraphInstantiate(&sg.post, gr, 0)); CK(cudaGraphDestroy(gr));
  sg.key = key;
}

struct BatchStats { long long peak = 0; int maxH = 0, maxW = 0, maxT = 0, maxE = 0; };

// one batch of units through all its resolution steps (main_dir_loop rt_routines.f90:1366-1442 for every unit at once)
template <int MODE>
BatchStats run_batch_mode(dre_engine* e, DreAccum A, int unit0, int U, int capH0, int capW0, int capT0, double max_cap, int pmul) {
  BatchStats bs; bs.peak = U;
  const DreDev& g = e->g;
  unsigned nkmask = 1u << DRE_NSIDE_MIN_K;
  ensure_step(e, U, std::max(capH0, U), std::max(capW0, 64), std::max(capT0, 64), -1, 0, 0);
  DreStep S = make_step(e, U);
  int cur = 0, nH = U, nW = 0;
  (e->n_launch++, k_seed)<<<cdiv((size_t)U + 1, 256), 256, 0, e->st>>>(S, U, e->srclist.p, unit0, e->nranks, e->rank, e->nsrc_list, pmul, e->opt_sector_rot);
  static const int verbose = getenv("DRE_VERBOSE") ? atoi(getenv("DRE_VERBOSE")) : 0;
  for (;;) {
    const auto t_0 = std::chrono::steady_clock::now();
    // direction tables are built up to nside 512 at the start of a pass; a unit that refines further (the reference allows
    // nside <= 8192, healpix_routines.f90:9) gets its table here, before the step that uses it
    for (int k = e->g.dir_kmax + 1; k <= DRE_MAXK && (nkmask >> k); k++) { build_dir_table(e, k); sync_g(e); }
    if (MODE == MODE_DEP) ensure_ftab(e, nkmask);
    // Option coop (off by default; A/B switch): the list side of a step as two cooperative launches with grid-wide barriers
    // between the phases -- 1: on one CTA per SM, 2: on 4-32 CTAs for steps with little work.  Both measured SLOWER than the
    // graph-replayed kernels on B200 (TRUST-I rtt_dir_src, 37 small steps: 7.1 ms graphs, 7.7 ms coop=1, 9.6 ms coop=2):
    // a grid barrier costs more than a graph-node boundary here.
    const long long step_work = (long long)U / 8 + 5ll * ((long long)nH + nW);      // rays <= 4 x extracted entries
    const bool coop_step = MODE == MODE_DEP && e->opt_split && (e->opt_coop == 1 || (e->opt_coop == 2 && step_work <= e->opt_coop_work));
    if (coop_step) {
      const int cgrid = (e->opt_coop == 1) ? e->sm_count : (int)std::max<long long>(4, std::min<long long>(32, (step_work + 4095) / 4096));
      DreCoop cq; cq.evt = e->evt.p; cq.nev = e->nev.p; cq.evoff = e->evoff.p; cq.capE = e->capE; cq.L = e->g.L;
      DreAccum Aa = A; DreStep Ss = S; int cc = cur;
      void* a_pre[] = {&Aa, &Ss, &cc, &cq};
      CK(cudaLaunchCooperativeKernel((const void*)k_step_pre_coop, dim3(cgrid), dim3(SCAN_THREADS), a_pre, 0, e->st));
      launch_split(e, A, S, false);
      void* a_post[] = {&Ss, &cc, &cq};
      CK(cudaLaunchCooperativeKernel((const void*)k_step_post_coop, dim3(cgrid), dim3(SCAN_THREADS), a_post, 0, e->st));
      CK(cudaMemcpyAsync(e->h_ctl, S.ctl, sizeof(DreCtl), cudaMemcpyDeviceToHost, e->st));
      CK(cudaMemcpyAsync(e->h_err, S.err, sizeof(int), cudaMemcpyDeviceToHost, e->st));
      e->n_launch += 2;
    } else if (e->opt_graphs) {
      step_graphs<MODE>(e, A, S, cur);
      CK(cudaGraphLaunch(e->sgraph[MODE == MODE_DEP ? 1 : 0][cur].pre, e->st));
      launch_trace<MODE>(e, A, S);
      CK(cudaGraphLaunch(e->sgraph[MODE == MODE_DEP ? 1 : 0][cur].post, e->st));
      e->n_launch += (MODE == MODE_DEP && e->opt_split) ? 15 : 16;
    } else {
      e->n_launch += enqueue_pre<MODE>(e, A, S, cur, e->st);
      launch_trace<MODE>(e, A, S);
      e->n_launch += enqueue_post(e, S, cur, e->st, MODE == MODE_DEP && e->opt_split);
    }
    const auto t_1 = std::chrono::steady_clock::now();
    // the one wait of a step.  Option blocking_sync: sleep on an event instead of spinning in cudaStreamSynchronize -- a thread
    // that sleeps is not the one a busy host preempts (A/B in profiles/r02_sweeps.md)
    if (e->opt_blocking_sync) { CK(cudaEventRecord(e->ev_blk, e->st)); CK(cudaEventSynchronize(e->ev_blk)); }
    else CK(cudaStreamSynchronize(e->st));
    CK(cudaGetLastError());
    const DreCtl c = *e->h_ctl;
    if (verbose >= 2) {
      const auto t_2 = std::chrono::steady_clock::now();
      float kms = 0; if (c.nT > 0) cudaEventElapsedTime(&kms, e->ev[2], e->ev[3]);
      fprintf(stderr, "[dre step] rt %d U %d nH %d nW %d nExt %d nT %d -> nH2 %d nW2 %d active %d abort %d | enqueue %.0f us, wait %.0f us, march %.0f us\n",
              e->g.rt_type, U, c.nH, c.nW, c.nExt, c.nT, c.nH2, c.nW2, c.active, c.abort,
              std::chrono::duration<double, std::micro>(t_1 - t_0).count(), std::chrono::duration<double, std::micro>(t_2 - t_1).count(), kms * 1e3);
    }
    if (*e->h_err != 0) throw DreError{err_text(*e->h_err)};
    if (c.abort) {          // a buffer was too small: the step changed nothing; grow and repeat it
      const int nh = std::max(c.needH, c.nH), nw = std::max(c.needW, c.nW);
      // the lists of this batch outgrow the memory budget: the units seen so far under-estimated the later ones by more
      // than the head-room.  Nothing of this step was deposited, but earlier steps of the batch were: stop with a message.
      if (std::max(std::max(nh, nw), c.needT) > 1.5 * max_cap)
        throw DreError{"the ray lists of a unit batch exceed the memory budget: lower option batch_units"};
      if (getenv("DRE_VERBOSE")) fprintf(stderr, "[dre] step refused: need H %d W %d T %d E %d (have %d %d %d %d)\n", c.needH, c.needW, c.needT, c.needE, e->capH, e->capW, e->capT, e->capE);
      ensure_step(e, U, nh > e->capH ? (int)std::min<long long>(2147483000ll, (long long)nh * 3 / 2) : e->capH,
                  nw > e->capW ? (int)std::min<long long>(2147483000ll, (long long)nw * 3 / 2) : e->capW,
                  c.needT > e->capT ? (int)std::min<long long>(2147483000ll, (long long)c.needT * 3 / 2) : e->capT, cur, c.nH, c.nW);
      if (c.needE > e->capE) {
        const size_t want = (size_t)std::min<long long>(2147483000ll, (long long)c.needE * 3 / 2);
        e->evt.release(); e->evt.ensure(want, 1.0); e->capE = (int)want;
      }
      S = make_step(e, U);
      continue;
    }
    if (c.nT > 0) { float ms = 0; CK(cudaEventElapsedTime(&ms, e->ev[2], e->ev[3])); e->last_kernel_ms += ms; e->last_launches++; }
    bs.peak = std::max(bs.peak, (long long)nH + nW + c.nT + c.nH2 + c.nW2);
    bs.maxH = std::max(bs.maxH, std::max(c.needH, c.nH2)); bs.maxW = std::max(bs.maxW, std::max(c.needW, c.nW2)); bs.maxT = std::max(bs.maxT, c.nT); bs.maxE = std::max(bs.maxE, c.needE);
    nH = c.nH2; nW = c.nW2; cur ^= 1;
    if (c.active == 0) break;
    nkmask = (unsigned)c.nkmask;
  }
  (void)g;
  return bs;
}

void fill_counters(dre_engine* e, dre_counters* out, float pass_ms) {
  unsigned long long c[16];
  CK(cudaMemcpyAsync(c, e->counters.p, sizeof(c), cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
#ifdef DRE_TIMING
  { unsigned long long t[16]; CK(cudaMemcpyFromSymbol(t, g_dre_tim, sizeof(t)));
    double tot = 0; for (int i = 0; i < 16; i++) tot += (double)t[i];
    const char* nm[16] = {"setup:runs+Fload", "psel", "math", "ringwait", "u+en", "next", "product-loop", "fence+issue", "setup:end+fetch",
                          "setup:record", "setup:dir+div", "setup:lum+mask", "setup:fwait", "end:return", "end:outputs", ""};
    fprintf(stderr, "[dre timing] rt_type %d (per intersection %.0f cycles):", e->g.rt_type, c[CNT_ISECT] ? tot / (double)c[CNT_ISECT] : 0.0);
    for (int i = 0; i < 15; i++) fprintf(stderr, " %s %.1f%%", nm[i], 100 * t[i] / tot);
    fprintf(stderr, "\n");
    memset(t, 0, sizeof(t)); CK(cudaMemcpyToSymbol(g_dre_tim, t, sizeof(t))); }
#endif
  if (!out) return;
  memset(out, 0, sizeof(*out));
  out->sources = e->nsrc_list; out->rays = c[CNT_RAYS]; out->intersections = c[CNT_ISECT]; out->deposits = c[CNT_DEP];
  out->beam_tests = c[CNT_BEAM]; out->splits = c[CNT_SPLIT]; out->merges = c[CNT_MERGE]; out->sca_elems = c[CNT_SCA];
  out->active_lambda = c[CNT_ACT]; out->steps = c[CNT_N]; out->isect_hash = c[CNT_HASH];
  out->kernel_ms = e->last_kernel_ms; out->pass_ms = pass_ms;
  out->active_lambda_dep = c[CNT_ACTDEP]; out->descent_loads = c[CNT_DESC];
  out->launches = e->n_launch - e->launch_mark; out->trace_launches = e->last_launches;
}

void ensure_dir_tables(dre_engine* e) {
  const int kmax = std::max(2, std::min(e->opt_dir_kmax, DRE_MAXK));
  for (int k = 0; k <= kmax; k++) build_dir_table(e, k);
}

}  // namespace

extern "C" {

int dre_run_pass_resident(dre_engine* e, int rt_type, int iterations, dre_counters* out) {
  (void)iterations;
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  int mode;
  switch (rt_type) {
    case DRE_RTT_PRECALC_CELL: case DRE_RTT_PRECALC_SRC: mode = MODE_PRECALC; break;
    case DRE_RTT_DIR_CELL: case DRE_RTT_DIR_SRC: case DRE_RTT_SCATT: mode = MODE_DEP; break;
    default: DRE_REQUIRE(false, "dre_run_pass: rt_type is not a list pass (use dre_run_iobs for i_obs types)"); mode = 0;
  }
  e->g.rt_type = rt_type;
  ensure_dir_tables(e);
  e->counters.zero(16, e->st); e->errflag.zero(1, e->st); e->rec_n.zero(1, e->st);
  // psel_av_arr(i0, iterations, :) is ONE slice for the direct-light phase -- rtt_dir_cell writes sid = cell, rtt_dir_src writes
  // sid = N + src, rt_mpi reduces once after both (rt_routines.f90:1456-1463, dartray_hub.f90:31-34) -- and a fresh slice in
  // every scattering iteration (expand_psel_av_arr, dartray_hub.f90:453).  A pass therefore clears only the sid range it writes.
  if (mode == MODE_DEP && e->g.i2d != 2) {      // (the second loop of rt_loop_2D continues the slice of the first)
    const size_t N0 = (size_t)e->g.N, NS = N0 + e->S;
    const size_t lo = (rt_type == DRE_RTT_DIR_SRC) ? N0 : 0, hi = (rt_type == DRE_RTT_DIR_CELL) ? N0 : NS;
    if (hi > lo) {
      CK(cudaMemsetAsync(e->psel_sum.p + lo, 0, (hi - lo) * sizeof(double), e->st));
      CK(cudaMemsetAsync(e->ipsel.p + lo, 0, (hi - lo) * sizeof(long long), e->st));
    }
  }
  e->last_kernel_ms = 0; e->last_launches = 0; e->launch_mark = e->n_launch;
  CK(cudaEventRecord(e->ev[0], e->st));
  const int nsrc = build_source_list(e);
  const long long total_units = (long long)nsrc * 48;
  const long long mine = total_units > e->rank ? (total_units - e->rank + e->nranks - 1) / e->nranks : 0;
  if (e->opt_det) det_prepare(e, rt_type, mode);
  DreAccum A = make_accum(e, e->prev.p);
  // Batch size: the ray lists of a batch grow with the angular refinement of its units, which is not known in advance.
  // The wavefront buffers hold `cap` entries per list (HIGH and LOW double-buffered, outcomes, tasks, packets, index arrays);
  // cap comes from the entries per unit seen so far (high-water mark per rt_type, kept across passes) and is bounded by
  // half of the memory left after configuration, 32 GiB at most.  A step that would overflow is refused on the device and repeated after
  // the buffers have grown (run_batch_mode), so an estimate that is too low costs time, not correctness.
  if (e->list_budget == 0) {     // once per configuration: cudaMemGetInfo costs ~10 ms, and the list buffers are kept afterwards
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    e->list_budget = std::min((size_t)(0.5 * (double)free_b), (size_t)32 << 30);   // larger batches gain nothing: 32 GiB at most
    e->ftab_budget = std::min(e->ftab_budget, e->ftab_bytes + (size_t)(0.15 * (double)free_b));
  }
  const int L = e->g.L;
  const double bytes_per_cap = 1.25 * (5.0 * (28.0 + 8.0 * L) + 8.0 * dre_pkt_stride(L) + 32.0 + 16.0 + 16.0 + 2.0 * 15.0 * 4.0 + 2.0 * 12.0 + 8.0) + 12.0 * 24.0 * 1.5;   // + ~12 events of 24 B per ray
  const double max_cap = std::min(2.0e9, std::max(1.0e6, (double)e->list_budget / bytes_per_cap));
  double& est = e->per_unit_est[rt_type % 16];   // max over steps of the per-list need / units
  double& est_e = e->per_unit_evt[rt_type % 16];
  double per_unit = 0, per_unit_e = 0;
  // nothing known about this pass type yet: a probe batch of 16 384 units measures the entries per unit first
  long long U_next = est > 0 ? (long long)(max_cap / (1.25 * est)) : 16384;
  U_next = std::max<long long>(48, std::min<long long>(U_next, e->opt_batch_units));
  // multiplier of the scrambled source order (k_seed): near nsrc / golden ratio, coprime to nsrc
  int pmul = 1;
  if (nsrc > 2) {
    auto gcd = [](long long a, long long b) { while (b) { const long long t = a % b; a = b; b = t; } return a; };
    pmul = std::max(1, (int)(nsrc * 0.6180339887));
    while (gcd(pmul, nsrc) != 1) pmul++;
  }
  for (long long u0 = 0; u0 < mine;) {
    const int U = (int)std::min<long long>(U_next, mine - u0);
    int cap0 = (int)std::min(max_cap, std::max(4.0 * U, 1.25 * std::max(per_unit, est) * U) + 1024.0);
    if (e->opt_cap0 > 0) cap0 = e->opt_cap0;   // test hook: start small so that steps are refused and the buffers grow
    // the event buffer of the march follows the same rule: sized before the batch from the events per unit seen so far, so
    // that a larger batch than the last one does not have its largest step refused first
    if (mode == MODE_DEP && e->opt_split && e->opt_cap0 <= 0) {
      const double want = std::min(2147483000.0, 1.25 * std::max(per_unit_e, est_e) * U + 1024.0);
      if (want > (double)e->capE) { e->evt.release(); e->evt.ensure((size_t)want, 1.0); e->capE = (int)want; }
    }
    const BatchStats bs = (mode == MODE_PRECALC) ? run_batch_mode<MODE_PRECALC>(e, A, (int)u0, U, cap0, cap0, cap0, max_cap, pmul)
                                                 : run_batch_mode<MODE_DEP>(e, A, (int)u0, U, cap0, cap0, cap0, max_cap, pmul);
    e->peak_entries = std::max(e->peak_entries, bs.peak);
    per_unit = std::max(per_unit, (double)std::max(bs.maxH, std::max(bs.maxW, bs.maxT)) / U);
    per_unit_e = std::max(per_unit_e, (double)bs.maxE / U);
    if (getenv("DRE_VERBOSE")) fprintf(stderr, "[dre] rt_type %d batch at unit %lld: %d units, peak %lld list entries, per-list need %d / %d / %d (%.1f per unit), capacity %d\n",
                                       rt_type, u0, U, bs.peak, bs.maxH, bs.maxW, bs.maxT, per_unit, e->capH);
    u0 += U;
    U_next = (long long)(max_cap / (1.25 * std::max(per_unit, est)));   // x1.25: headroom for units richer than those seen so far
    U_next = std::max<long long>(48, std::min<long long>(U_next, e->opt_batch_units));
  }
  est_e = std::max(est_e, per_unit_e);
  est = std::max(est, per_unit);   // high-water mark (the scattering iterations share one rt_type and shrink)
  CK(cudaEventRecord(e->ev[1], e->st));
  CK(cudaStreamSynchronize(e->st));
  float ms = 0; CK(cudaEventElapsedTime(&ms, e->ev[0], e->ev[1]));
  fill_counters(e, out, ms);
  DRE_CATCH(e)
}

}  // extern "C"

// =========================================================================================
// drop-in calls with host arrays, rt_loop_iobs, rt_mpi, rt_prepare(rtt_scatt)
// =========================================================================================
extern "C" {

int dre_run_pass(dre_engine* e, int rt_type, int iterations, const double* u_fest, double* u_accum, double* const* scaspe,
                 const double* const* scaspe_prev, double* lum_lost, double* psel_av, dre_counters* out) {
  const bool precalc = (rt_type == DRE_RTT_PRECALC_CELL || rt_type == DRE_RTT_PRECALC_SRC);
  int rc;
  if (precalc) {
    if (u_accum && (rc = dre_upload(e, DRE_ARR_U_FEST, u_accum))) return rc;
  } else {
    if (u_fest && (rc = dre_upload(e, DRE_ARR_U_FEST, u_fest))) return rc;
    if (u_accum && (rc = dre_upload(e, DRE_ARR_U_FINAL, u_accum))) return rc;
    if (scaspe && (rc = dre_upload(e, DRE_ARR_SCASPE, scaspe))) return rc;
    if (scaspe_prev && rt_type == DRE_RTT_SCATT && (rc = dre_upload(e, DRE_ARR_SCASPE_PREV, scaspe_prev))) return rc;
    if (lum_lost && (rc = dre_upload(e, DRE_ARR_LUM_LOST, lum_lost))) return rc;
  }
  if ((rc = dre_run_pass_resident(e, rt_type, iterations, out))) return rc;
  if (precalc) {
    if (u_accum && (rc = dre_download(e, DRE_ARR_U_FEST, u_accum))) return rc;
  } else {
    if (u_accum && (rc = dre_download(e, DRE_ARR_U_FINAL, u_accum))) return rc;
    if (scaspe && (rc = dre_download(e, DRE_ARR_SCASPE, (void*)scaspe))) return rc;
    if (lum_lost && (rc = dre_download(e, DRE_ARR_LUM_LOST, lum_lost))) return rc;
    if (psel_av) {   // only the sid range of the pass just run: the other half of the slice belongs to the other direct pass
      DRE_TRY(e)
      const size_t N0 = (size_t)e->g.N, NS = N0 + e->S;
      const size_t lo = (rt_type == DRE_RTT_DIR_SRC) ? N0 : 0, hi = (rt_type == DRE_RTT_DIR_CELL) ? N0 : NS;
      if (hi > lo) {
        (e->n_launch++, k_psel_ratio)<<<cdiv(NS, 256), 256, 0, e->st>>>(e->psel_sum.p, e->ipsel.p, (int)NS, e->psel_av.p);
        CK(cudaMemcpyAsync(psel_av + lo, e->psel_av.p + lo, (hi - lo) * 8, cudaMemcpyDeviceToHost, e->st));
        CK(cudaStreamSynchronize(e->st));
      }
      DRE_CATCH(e)
    }
  }
  return 0;
}

int dre_run_iobs(dre_engine* e, int rt_type, const double* const* scaspe_tot, int32_t ndir_in, const double* ccoord_obs,
                 double* i_obs, double* i_obs_in, dre_counters* out) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  DRE_REQUIRE(rt_type == DRE_RTT_I_OBS_DIR_CELL || rt_type == DRE_RTT_I_OBS_DIR_SRC || rt_type == DRE_RTT_I_OBS,
              "dre_run_iobs: rt_type is not an i_obs pass");
  DRE_REQUIRE(ndir_in >= 0 && (ndir_in == 0 || ccoord_obs), "dre_run_iobs: internal observers need ccoord_obs");
  DreDev& g = e->g;
  g.rt_type = rt_type;
  const int L = g.L, ndir = g.ndir;
  const size_t NS = (size_t)g.N + e->S;
  sync_g(e);
  if (scaspe_tot && rt_type == DRE_RTT_I_OBS) planes_io(e, e->tot.p, (double* const*)scaspe_tot, 0);
  // device i_obs arrays: "+=" semantics on the caller's arrays when given, else resident
  const size_t n_out = (size_t)std::max(ndir, 1) * NS * L, n_in = (size_t)std::max(ndir_in, 1) * NS * L;
  if (!e->i_obs.p || e->i_obs.cap < n_out) { e->i_obs.release(); e->i_obs.zero(n_out, e->st); }
  if (!e->i_obs_in.p || e->i_obs_in.cap < n_in || e->i_obs_ndir_in != ndir_in) { e->i_obs_in.release(); e->i_obs_in.zero(n_in, e->st); e->i_obs_ndir_in = ndir_in; }
  if (i_obs && ndir > 0) CK(cudaMemcpyAsync(e->i_obs.p, i_obs, (size_t)ndir * NS * L * 8, cudaMemcpyHostToDevice, e->st));
  if (i_obs_in && ndir_in > 0) CK(cudaMemcpyAsync(e->i_obs_in.p, i_obs_in, (size_t)ndir_in * NS * L * 8, cudaMemcpyHostToDevice, e->st));
  e->counters.zero(16, e->st); e->errflag.zero(1, e->st); e->rec_n.zero(1, e->st);
  e->last_kernel_ms = 0; e->last_launches = 0; e->launch_mark = e->n_launch;
  CK(cudaEventRecord(e->ev[0], e->st));
  if (rt_type == DRE_RTT_I_OBS)   // calc_total_luminosity_sca case(rtt_i_obs) rt_routines.f90:760-771
    (e->n_launch++, k_lumcell_sca)<<<cdiv((size_t)g.N * L, 256), 256, 0, e->st>>>(e->tot.p, e->lumcell.p, 1);
  const int nsrc = build_source_list(e);
  DreIobs q; memset(&q, 0, sizeof(q));
  q.srclist = e->srclist.p; q.nsrc = nsrc; q.ndir = ndir; q.ndir_in = ndir_in; q.obs_dir = e->obs_dir.p;
  q.i_obs = e->i_obs.p; q.i_obs_in = e->i_obs_in.p;
  DBuf<DreDir> d_in_dir; DBuf<double> d_vm; DBuf<int> d_ip;
  if (ndir_in > 0 && nsrc > 0) {
    // find_theta_phi_obs_in rt_routines.f90:3164-3217, evaluated with the host libm
    std::vector<int> hl(nsrc); CK(cudaMemcpy(hl.data(), e->srclist.p, (size_t)nsrc * 4, cudaMemcpyDeviceToHost));
    std::vector<double> hc((size_t)3 * g.N), hps((size_t)3 * std::max(e->S, 1)); std::vector<int> hsc(g.N);
    CK(cudaMemcpy(hc.data(), e->ccoord.p, hc.size() * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hsc.data(), e->src_cell.p, (size_t)g.N * 4, cudaMemcpyDeviceToHost));
    if (e->S > 0) CK(cudaMemcpy(hps.data(), e->ccoord_p_src.p, (size_t)3 * e->S * 8, cudaMemcpyDeviceToHost));
    const size_t m = (size_t)nsrc * ndir_in;
    std::vector<DreDir> dd(m); std::vector<double> vm(m); std::vector<int> ip(m * std::max(g.nclass, 1), 0);
    const double E8 = (double)1E-8f;
    for (int si = 0; si < nsrc; si++) {
      const int nc = hl[si];
      const double* sc = (rt_type == DRE_RTT_I_OBS_DIR_SRC) ? &hps[3 * (size_t)hsc[nc]] : &hc[3 * (size_t)nc];
      for (int j = 0; j < ndir_in; j++) {
        double v[3]; for (int a = 0; a < 3; a++) v[a] = ccoord_obs[3 * (size_t)j + a] - sc[a];
        double vmod = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), theta, phi;
        if (vmod > E8) {
          double tr = (0 * v[0] + 0 * v[1] + 1 * v[2]) / vmod;
          if (fabs(tr - 1.0) < E8) tr = 1.0;
          if (fabs(tr + 1.0) < E8) tr = -1.0;
          theta = acos(tr);
          const double vxy = sin(theta);
          double pr = (vxy > E8) ? v[0] / (vmod * vxy) : 1.;
          if (fabs(pr - 1.0) < E8) pr = 1.0;
          if (fabs(pr + 1.0) < E8) pr = -1.0;
          phi = acos(pr);
          if (v[1] < 0) phi = 2 * hp::PI_ - phi;
        } else { theta = 0; phi = 0; vmod = 0; }
        const size_t o = (size_t)si * ndir_in + j;
        dd[o] = hp::make_dir(theta, phi); vm[o] = vmod;
        for (int c = 0; c < g.nclass; c++) ip[o * g.nclass + c] = hp::ang2pix_nest(1 << g.class_kp[c], theta, phi);
      }
    }
    d_in_dir.upload(dd.data(), m, e->st); d_vm.upload(vm.data(), m, e->st); d_ip.upload(ip.data(), ip.size(), e->st);
    CK(cudaStreamSynchronize(e->st));
    q.in_dir = d_in_dir.p; q.in_vecmod = d_vm.p; q.in_ip = d_ip.p;
  }
  if (nsrc > 0 && ndir + ndir_in > 0) {
    DreAccum A = make_accum(e, e->tot.p);
    const long long nt = (long long)nsrc * (ndir + ndir_in);
    DRE_REQUIRE(nt < (1ll << 31), "dre_run_iobs: too many rays for one launch");
    const int ctas = (int)std::max<long long>(1, std::min<long long>((long long)e->sm_count * 4, (nt + 7) / 8));
    CK(cudaMemsetAsync(e->counter.p, 0, sizeof(int), e->st));
    CK(cudaEventRecord(e->ev[2], e->st));
    e->n_launch++;
    if (g.KLn == 1) k_iobs<1><<<ctas, 256, 0, e->st>>>(A, q, e->counter.p);
    else if (g.KLn == 2) k_iobs<2><<<ctas, 256, 0, e->st>>>(A, q, e->counter.p);
    else k_iobs<4><<<ctas, 256, 0, e->st>>>(A, q, e->counter.p);
    CK(cudaEventRecord(e->ev[3], e->st));
    CK(cudaGetLastError());
    e->last_launches++;
  }
  CK(cudaEventRecord(e->ev[1], e->st));
  CK(cudaStreamSynchronize(e->st));
  if (nsrc > 0 && ndir + ndir_in > 0) { float ms = 0; CK(cudaEventElapsedTime(&ms, e->ev[2], e->ev[3])); e->last_kernel_ms = ms; }
  int herr = read_int(e, e->errflag.p);
  if (herr != 0) throw DreError{err_text(herr)};
  float ms = 0; CK(cudaEventElapsedTime(&ms, e->ev[0], e->ev[1]));
  fill_counters(e, out, ms);
  if (i_obs && ndir > 0) CK(cudaMemcpyAsync(i_obs, e->i_obs.p, (size_t)ndir * NS * L * 8, cudaMemcpyDeviceToHost, e->st));
  if (i_obs_in && ndir_in > 0) CK(cudaMemcpyAsync(i_obs_in, e->i_obs_in.p, (size_t)ndir_in * NS * L * 8, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

int dre_allreduce(dre_engine* e, int rt_type) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  if (e->nranks <= 1) return 0;
  DRE_REQUIRE(e->comm != nullptr, "dre_allreduce: no communicator");
  const DreDev& g = e->g;
  const size_t NL = (size_t)g.N * g.L, NS = (size_t)g.N + e->S;
  auto ar = [&](void* p, size_t n, int dt) {
    const int rc = g_nccl.AllReduce(p, p, n, dt, NCCL_SUM, e->comm, e->st);
    if (rc != 0) throw DreError{std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error")};
  };
  auto grp = [&](bool start) {
    if (!g_nccl.GroupStart || !g_nccl.GroupEnd) return;
    const int rc = start ? g_nccl.GroupStart() : g_nccl.GroupEnd();
    if (rc != 0) throw DreError{std::string("ncclGroup: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error")};
  };
  switch (rt_type) {
    case DRE_RTT_PRECALC_CELL: case DRE_RTT_PRECALC_SRC:       // reduce_u_fest_arr io_routines.f90:3708-3730
      ar(e->u_fest.p, NL, NCCL_FLOAT64); break;
    case DRE_RTT_DIR_CELL: case DRE_RTT_DIR_SRC:               // reduce_u_final_arr / reduce_scaspe_arr / reduce_lum_lost
      grp(true);
      ar(e->acc.p, e->rows_elems(), NCCL_FLOAT64); ar(e->lum_lost.p, g.L, NCCL_FLOAT64);
      ar(e->psel_sum.p, NS, NCCL_FLOAT64); ar(e->ipsel.p, NS, NCCL_INT64);
      grp(false); break;
    case DRE_RTT_SCATT: {
      // delta trick io_routines.f90:3743-3757 and lum_lost correction :3846-3848, all on the device: the four reductions
      // go out as one NCCL group and nothing waits for the host
      const bool det = e->opt_det && e->qc.p && (int)e->h_qc.size() == 3 * dre_lq(g.L) + 2;
      const double* qc_u = det ? e->qc.p + dre_lq(g.L) : nullptr;
      const double* qc_l = det ? e->qc.p + 2 * dre_lq(g.L) : nullptr;
      e->red_d.ensure((size_t)2 * g.L);
      (e->n_launch++, k_u_delta)<<<cdiv(NL, 256), 256, 0, e->st>>>(e->acc.p, e->u_fest.p, g.N, g.L, g.ROW, 0, qc_u);
      (e->n_launch++, k_lum_lost_corr)<<<cdiv(g.L, 128), 128, 0, e->st>>>(e->lum_lost.p, e->lum_lost_prev.p, e->red_d.p, g.L, e->nranks, qc_l, 0);
      grp(true);
      ar(e->acc.p, e->rows_elems(), NCCL_FLOAT64); ar(e->lum_lost.p, g.L, NCCL_FLOAT64);
      ar(e->psel_sum.p, NS, NCCL_FLOAT64); ar(e->ipsel.p, NS, NCCL_INT64);
      grp(false);
      (e->n_launch++, k_u_delta)<<<cdiv(NL, 256), 256, 0, e->st>>>(e->acc.p, e->u_fest.p, g.N, g.L, g.ROW, 1, qc_u);
      if (det) (e->n_launch++, k_lum_lost_corr)<<<cdiv(g.L, 128), 128, 0, e->st>>>(e->lum_lost.p, e->lum_lost_prev.p, e->red_d.p, g.L, e->nranks, qc_l, 1);
      break;
    }
    default: break;   // i_obs passes are not sharded (rt_loop_iobs has no MPI split, rt_routines.f90:1498)
  }
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

int dre_prepare_scatt(dre_engine* e, int iterations, double conv_en_lim, double* tot_rad_en, int32_t* cnflag) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  DRE_REQUIRE(tot_rad_en && cnflag, "dre_prepare_scatt: null output");
  const DreDev& g = e->g;
  const size_t NL = (size_t)g.N * g.L, nsca = (size_t)g.N * (g.ROW - g.L);
  sync_g(e);
  (e->n_launch++, k_u_rows)<<<cdiv(NL, 256), 256, 0, e->st>>>(e->acc.p, e->u_fest.p, g.N, g.L, g.ROW, 1);          // u_fest <- u_final
  CK(cudaMemcpyAsync(e->lum_lost_prev.p, e->lum_lost.p, g.L * 8, cudaMemcpyDeviceToDevice, e->st));
  if (iterations == 1)   // create_scaspe_tot rt_routines.f90:3012-3040: scaspe_tot_arr = scaspe_arr
    CK(cudaMemcpyAsync(e->tot.p, e->acc.p, e->rows_elems() * 8, cudaMemcpyDeviceToDevice, e->st));
  CK(cudaMemcpyAsync(e->prev.p, e->acc.p, e->rows_elems() * 8, cudaMemcpyDeviceToDevice, e->st));  // scaspe_prev <- scaspe
  e->red_d.ensure((size_t)2 * g.L);
  (e->n_launch++, k_lumcell_sca)<<<cdiv(NL, 256), 256, 0, e->st>>>(e->acc.p, e->lumcell.p, 0);
  (e->n_launch++, k_colsum)<<<g.L, 256, 0, e->st>>>(e->lumcell.p, g.N, g.L, e->red_d.p + g.L);
  std::vector<double> tl(g.L);
  CK(cudaMemcpyAsync(tl.data(), e->red_d.p + g.L, g.L * 8, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  // tot_rad_en(l) = sum over ALL cells of the HEALPix part of scaspe (rt_routines.f90:690-698): only leaf cells are ever
  // deposited into, so it is the column sum of lumcell just computed (k_totrad re-read every row for the same number)
  if (iterations == 1) for (int l = 0; l < g.L; l++) tot_rad_en[l] = tl[l];
  double sum_tot = 0; for (int l = 0; l < g.L; l++) sum_tot += tot_rad_en[l];
  int flag = 1;
  if (sum_tot > 0) for (int l = 0; l < g.L; l++) if (!(tl[l] / tot_rad_en[l] < conv_en_lim)) flag = 0;
  *cnflag = flag;
  if (iterations > 1) (e->n_launch++, k_add_sca)<<<cdiv(nsca, 256), 256, 0, e->st>>>(e->tot.p, e->acc.p, g.N, g.ROW, g.L);
  (e->n_launch++, k_zero_sca)<<<cdiv(nsca, 256), 256, 0, e->st>>>(e->acc.p, g.N, g.ROW, g.L);
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

// Mirrored HEALPix pixel of every pixel of nside 2^kp under the reflections X, Y, Z, XY, XZ, YZ, XYZ (what calc_scaspe_indices
// rt_routines.f90:4083-4217 finds by searching the ring for the nearest longitude): the centre direction is reflected and located
// with ang2pix_nest.  which = 0..6.
static void sym_index_host(int kp, int which, std::vector<int>& out) {
  const int nside = 1 << kp, n = 12 * nside * nside;
  static const int flip[7][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 1, 0}, {1, 0, 1}, {0, 1, 1}, {1, 1, 1}};
  out.resize(n);
  for (int p = 0; p < n; p++) {
    double th, ph; hp::pix2ang_nest(nside, p, th, ph);
    double t2 = th, p2 = ph;
    if (flip[which][2]) t2 = hp::PI_ - th;
    if (flip[which][0] && flip[which][1]) p2 = ph + hp::PI_;
    else if (flip[which][0]) p2 = hp::PI_ - ph;
    else if (flip[which][1]) p2 = hp::TWOPI - ph;
    out[p] = hp::ang2pix_nest(nside, t2, p2);
  }
}

int dre_scaspe_symmetry_index(dre_engine* e, int32_t kp, int32_t which, int32_t* out) {
  DRE_TRY(e)
  DRE_REQUIRE(kp >= 0 && kp <= DRE_NSIDE_MIN_K && which >= 0 && which < 7 && out, "dre_scaspe_symmetry_index: bad arguments");
  std::vector<int> v; sym_index_host(kp, which, v);
  for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
  DRE_CATCH(e)
}

static int sym2d_prepare(dre_engine* e, int rt_type) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  DRE_REQUIRE(rt_type == DRE_RTT_PRECALC_CELL || rt_type == DRE_RTT_DIR_CELL || rt_type == DRE_RTT_SCATT,
              "dre_run_pass_2d: rt_loop_2D runs rtt_precalc_cell, rtt_dir_cell and rtt_scatt (dartray_hub.f90:69-126)");
  DRE_REQUIRE(e->g.ndir == 0, "dre_run_pass_2d: the 2D mode needs tot_ndir_scaspe = 0 (fix_symmetry mirrors only the HEALPix part of scaspe)");
  if (!e->have_sym) {
    for (int ic = 0; ic < e->g.nclass; ic++) {
      std::vector<int> all, v;
      for (int w = 0; w < 7; w++) { sym_index_host(e->g.class_kp[ic], w, v); all.insert(all.end(), v.begin(), v.end()); }
      e->sym_perm[ic].upload(all.data(), all.size(), e->st);
    }
    CK(cudaStreamSynchronize(e->st));
    e->have_sym = true;
  }
  DRE_CATCH(e)
}
// fix_symmetry rt_routines.f90:3605-3635 on the device-resident arrays
static int sym2d_fix(dre_engine* e, int rt_type) {
  DRE_TRY(e)
  DreSym q; memset(&q, 0, sizeof(q));
  for (int ic = 0; ic < e->g.nclass; ic++) q.perm[ic] = e->sym_perm[ic].p;
  q.rows = e->acc.p; q.u_fest = e->u_fest.p;
  q.part = (rt_type == DRE_RTT_PRECALC_CELL) ? 1 : (rt_type == DRE_RTT_DIR_CELL ? 2 : 3);
  sync_g(e);
  e->errflag.zero(1, e->st);
  (e->n_launch++, k_fix_symmetry)<<<e->g.N, 128, 0, e->st>>>(q, e->errflag.p);
  (e->n_launch++, k_lum_lost_x8)<<<cdiv(e->g.L, 128), 128, 0, e->st>>>(e->lum_lost.p, e->lum_lost_prev.p, e->g.L);
  CK(cudaGetLastError());
  DRE_REQUIRE(read_int(e, e->errflag.p) == 0, "find_linked_cells: the grid is not symmetric (rt_routines.f90:3394)");
  DRE_CATCH(e)
}

int dre_run_pass_2d(dre_engine* e, int rt_type, int iterations, dre_counters* out) {
  if (!e) return 1;
  int rc = sym2d_prepare(e, rt_type);
  if (rc) return rc;
  dre_counters c0, c1; memset(&c0, 0, sizeof(c0)); memset(&c1, 0, sizeof(c1));
  e->g.i2d = 1;                                    // first loop: sources of the positive octant
  rc = dre_run_pass_resident(e, rt_type, iterations, &c0);
  e->g.i2d = 0;
  if (rc) return rc;
  if ((rc = sym2d_fix(e, rt_type))) return rc;
  e->g.i2d = 2;                                    // second loop: sources on the Cartesian planes
  rc = dre_run_pass_resident(e, rt_type, iterations, &c1);
  e->g.i2d = 0;
  if (rc) return rc;
  if (out) {
    *out = c0;
    out->sources += c1.sources; out->rays += c1.rays; out->intersections += c1.intersections; out->deposits += c1.deposits;
    out->beam_tests += c1.beam_tests; out->splits += c1.splits; out->merges += c1.merges; out->sca_elems += c1.sca_elems;
    out->active_lambda += c1.active_lambda; out->steps += c1.steps; out->kernel_ms += c1.kernel_ms; out->pass_ms += c1.pass_ms;
    out->isect_hash += c1.isect_hash; out->active_lambda_dep += c1.active_lambda_dep; out->descent_loads += c1.descent_loads;
    out->launches += c1.launches; out->trace_launches += c1.trace_launches;
  }
  return 0;
}

int dre_dust_emission(dre_engine* e, int32_t ls, int32_t ld, const double* kabs_stars, const double* kabs_dust, const double* lambda_si_dust,
                      const double* dl_stars, const double* dl_dust, double kext_ref, const double* dens_ref, const double* rf_stars,
                      const double* rf_dust, int32_t iterations_dustem, double* dens_stars, double* t_dust, int32_t to_lumcell) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_grid && e->have_opacity, "dre_dust_emission: grid / opacity not set");
  const DreDev& g = e->g;
  DRE_REQUIRE(ls >= 1 && ld >= 1 && ld <= 4096, "dre_dust_emission: bad wavelength counts");
  DRE_REQUIRE(kabs_stars && kabs_dust && lambda_si_dust && dl_stars && dl_dust && dens_ref && rf_stars && dens_stars, "dre_dust_emission: null input");
  DRE_REQUIRE(kext_ref != 0, "dre_dust_emission: kext_ref is zero");
  const bool resident = (rf_dust == nullptr) && iterations_dustem > 0;
  DRE_REQUIRE(!(resident || to_lumcell) || g.L == ld, "dre_dust_emission: the engine is not configured for the dust-emission wavelengths (lnum != lnum_dust)");
  const size_t N = g.N;
  DBuf<double> d_ks, d_kd, d_lam, d_dls, d_dld, d_ref, d_rfs, d_rfd, d_out, d_t;
  d_ks.upload(kabs_stars, ls, e->st); d_kd.upload(kabs_dust, ld, e->st); d_lam.upload(lambda_si_dust, ld, e->st);
  d_dls.upload(dl_stars, ls, e->st); d_dld.upload(dl_dust, ld, e->st); d_ref.upload(dens_ref, N, e->st);
  d_rfs.upload(rf_stars, N * ls, e->st);
  if (rf_dust && iterations_dustem > 0) d_rfd.upload(rf_dust, N * ld, e->st);
  d_out.ensure(N * ld); d_t.zero(N, e->st);
  e->errflag.zero(1, e->st);
  DreDust q; memset(&q, 0, sizeof(q));
  q.N = g.N; q.ls = ls; q.ld = ld; q.kabs_s = d_ks.p; q.kabs_d = d_kd.p; q.lam_d = d_lam.p; q.dl_s = d_dls.p; q.dl_d = d_dld.p;
  q.dens_ref = d_ref.p; q.cchild = e->cchild.p; q.csize = e->csize.p; q.rf_s = d_rfs.p; q.rf_d = d_rfd.p; q.rows = e->acc.p; q.ROW = g.ROW;
  q.kext_ref = kext_ref; q.iterations_dustem = iterations_dustem; q.dens_stars = d_out.p; q.t_dust = d_t.p;
  q.lumcell = to_lumcell ? e->lumcell.p : nullptr; q.err = e->errflag.p;
  (e->n_launch++, k_dust_emission)<<<cdiv(N, 128), 128, (size_t)3 * ld * sizeof(double), e->st>>>(q);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(dens_stars, d_out.p, N * ld * 8, cudaMemcpyDeviceToHost, e->st));
  if (t_dust) CK(cudaMemcpyAsync(t_dust, d_t.p, N * 8, cudaMemcpyDeviceToHost, e->st));
  const int herr = read_int(e, e->errflag.p);
  DRE_REQUIRE(herr == 0, herr == 6 ? "calc_t_dust_equil: negative absorbed energy (sed_routines.f90:865)"
                                   : (herr == 7 ? "zbrent_tdust: root not bracketed (sed_routines.f90:1005)" : "zbrent_tdust: too many iterations (sed_routines.f90:1071)"));
  if (to_lumcell) e->have_sources = e->have_sources;   // lumcell now holds the dust emission; point sources are untouched
  DRE_CATCH(e)
}

int dre_build_source_list(dre_engine* e, int rt_type, int32_t* n_sources, int32_t* list_out) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  e->g.rt_type = rt_type;
  const int n = build_source_list(e);
  if (n_sources) *n_sources = n;
  if (list_out && n > 0) CK(cudaMemcpy(list_out, e->srclist.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  DRE_CATCH(e)
}

int dre_locate_points(dre_engine* e, int32_t npts, const double* xyz, const int32_t* cc_old, int32_t* cell_out) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_grid && e->have_params, "dre_locate_points: grid/params not set");
  if (npts <= 0) return 0;
  DBuf<double> dx; DBuf<int> dc, dout;
  sync_g(e);
  dx.upload(xyz, (size_t)3 * npts, e->st); dc.upload(cc_old, npts, e->st); dout.ensure(npts);
  (e->n_launch++, k_locate)<<<cdiv(npts, 128), 128, 0, e->st>>>(npts, dx.p, dc.p, dout.p);
  CK(cudaMemcpyAsync(cell_out, dout.p, (size_t)npts * 4, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

// ---- dre_locate_points without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: graph capture, serving loops, harnesses that forbid it).  The staging buffers of
// k_locate -- the points, their cc_old and the located cells -- live in ONE file-scope device array that every GPU loading the
// library holds until exit: DRE_NOALLOC_MAX_POINTS * 32 B = 2 MiB.  Layout: xyz (3 doubles per point) | cc_old | cell_out.
// Host code must not use the symbol's name as a pointer (it is a host-side stand-in): the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.  All calls share the array and it keeps the last call's
// contents: calls that overlap on one GPU (two engines, two host threads) race.
#define DRE_NOALLOC_MAX_POINTS 65536
static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)DRE_NOALLOC_MAX_POINTS * 32];
static unsigned char* g_noalloc_ptr[64];     // device address of the array, per device

int dre_locate_points_noalloc(dre_engine* e, int32_t npts, const double* xyz, const int32_t* cc_old, int32_t* cell_out) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(e->have_grid && e->have_params, "dre_locate_points_noalloc: grid/params not set");
  DRE_REQUIRE(npts <= DRE_NOALLOC_MAX_POINTS, "dre_locate_points_noalloc: more than 65536 points (the static scratch array holds no more; use dre_locate_points)");
  DRE_REQUIRE(e->device >= 0 && e->device < 64, "dre_locate_points_noalloc: device index beyond 63");
  if (npts <= 0) return 0;
  DRE_REQUIRE(xyz && cc_old && cell_out, "dre_locate_points_noalloc: null argument");
  unsigned char*& base = g_noalloc_ptr[e->device];
  if (!base) {
    void* p = nullptr;
    const cudaError_t rc = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (rc != cudaSuccess || !p) throw DreError{std::string("dre_locate_points_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(rc)};
    base = static_cast<unsigned char*>(p);
  }
  double* dx = reinterpret_cast<double*>(base);
  int* dc = reinterpret_cast<int*>(base + (size_t)DRE_NOALLOC_MAX_POINTS * 24);
  int* dout = reinterpret_cast<int*>(base + (size_t)DRE_NOALLOC_MAX_POINTS * 28);
  sync_g(e);
  CK(cudaMemcpyAsync(dx, xyz, (size_t)3 * npts * 8, cudaMemcpyHostToDevice, e->st));
  CK(cudaMemcpyAsync(dc, cc_old, (size_t)npts * 4, cudaMemcpyHostToDevice, e->st));
  CK(cudaMemsetAsync(dout, 0xff, (size_t)npts * 4, e->st));      // the previous call's cells must not survive a failed launch
  (e->n_launch++, k_locate)<<<cdiv(npts, 128), 128, 0, e->st>>>(npts, dx, dc, dout);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(cell_out, dout, (size_t)npts * 4, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

int dre_pix2vec(dre_engine* e, int32_t nside, int32_t npix, const int32_t* ipix, double* theta, double* phi, double* vec) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  int k = 0; while ((1 << k) < nside) k++;
  DRE_REQUIRE((1 << k) == nside && k <= DRE_MAXK, "dre_pix2vec: nside must be a power of two <= 8192");
  DRE_REQUIRE(k <= 10, "dre_pix2vec: direction tables are limited to nside <= 1024");
  for (int i = 0; i < npix; i++) DRE_REQUIRE(ipix[i] >= 0 && ipix[i] < 12 * nside * nside, "dre_pix2vec: ipix out of range");
  build_dir_table(e, k);
  if (npix <= 0) return 0;
  DBuf<int> dp; DBuf<double> dv;
  dp.upload(ipix, npix, e->st); dv.ensure((size_t)3 * npix);
  (e->n_launch++, k_pix2vec)<<<cdiv(npix, 256), 256, 0, e->st>>>(e->dirtab[k].p, npix, dp.p, dv.p);
  CK(cudaMemcpyAsync(vec, dv.p, (size_t)3 * npix * 8, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  if (theta && phi) for (int i = 0; i < npix; i++) hp::pix2ang_nest(nside, ipix[i], theta[i], phi[i]);
  DRE_CATCH(e)
}

int dre_calc_sed(dre_engine* e, double dist_obs, int units, const double* lambda_si, double* sed) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  const DreDev& g = e->g;
  DRE_REQUIRE(sed != nullptr && dist_obs > 0, "dre_calc_sed: bad arguments");
  DRE_REQUIRE(units >= 0 && units <= 2 && (units != 2 || lambda_si), "dre_calc_sed: units must be 0, 1 or 2 (2 needs lambda_si)");
  DRE_REQUIRE(e->i_obs.p && g.ndir > 0, "dre_calc_sed: no i_obs pass was run");
  DBuf<double> dl, ds; DBuf<int> dc;
  std::vector<double> lam(g.L, 0.0);
  if (lambda_si) lam.assign(lambda_si, lambda_si + g.L);
  dl.upload(lam.data(), g.L, e->st);
  std::vector<int> cs(std::max(e->S, 1), 0);
  for (int i = 0; i < e->S; i++) cs[i] = e->h_cell_src[i];
  dc.upload(cs.data(), cs.size(), e->st);
  ds.ensure((size_t)g.ndir * g.L);
  sync_g(e);
  (e->n_launch++, k_sed)<<<g.ndir * g.L, 256, 0, e->st>>>(e->i_obs.p, dc.p, dist_obs, units, dl.p, ds.p);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(sed, ds.p, (size_t)g.ndir * g.L * 8, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

namespace {
// ---- host part of make_maps: the projected-cube textures of one line of sight -------------------------------
// calc_texture_array (visual_routines.f90:314-453) with find_ray_cell_intersections / calc_ray_inters (:456-566).
// Default-real literals of the reference (1E-3, 1E-6, 0.99, 2.5, 1+1E-7) are single precision and kept so.
struct HostTex { int nt = 0, im0 = 0, im1 = 0; std::vector<double> prof; };

double cube_chord(double half_tol, double half, const double* ro, const double* dir) {
  // cube centred on the origin; faces in the order x+, x-, y+, y-, z+, z- (j = 0..5 of calc_ray_inters);
  // the first two faces whose intersection point lies inside the face (tolerance (1+1E-7)*cellsize/2 on the two
  // other axes) give the chord
  double hit[2][3]; int nhit = 0;
  for (int j = 0; j < 6 && nhit < 2; j++) {
    const int ax = j >> 1;
    const double dx = (j & 1) ? -half : half;
    const double den = dir[ax];
    if (den == 0.) continue;
    const double dist = ((0.0 + dx) - ro[ax]) / den;
    double pc[3]; bool inside = true;
    for (int i = 0; i < 3; i++) { pc[i] = ro[i] + dist * dir[i]; if (i != ax && fabs(pc[i] - 0.0) > half_tol) inside = false; }
    if (!inside) continue;
    for (int i = 0; i < 3; i++) hit[nhit][i] = pc[i];
    nhit++;
  }
  if (nhit < 2) return 0.0;
  double s = 0;
  for (int i = 0; i < 3; i++) { const double d = hit[1][i] - hit[0][i]; s += d * d; }
  return pow(s, 0.5);      // (sum((p2-p1)**2.))**0.5
}

void host_textures(const dre_engine* e, int npix_hd, double size_map, double pixel_size_hd, double theta, double phi,
                   DreMapGeom& m, std::vector<HostTex>& tex) {
  const DreDev& g = e->g;
  double dtheta = (double)1E-3f;
  double* o = m.obs_vec; double* xi = m.xi_dir; double* yi = m.yi_dir;
  o[0] = sin(theta) * cos(phi); o[1] = sin(theta) * sin(phi); o[2] = cos(theta);
  double z_zi = 0.0 + o[0] * 0.0; z_zi = z_zi + o[1] * 0.0; z_zi = z_zi + o[2] * 1.0;      // scal_prod(obs_vec, z_dir)
  const double pi_ref = 2 * (double)1.57079637050628662109375f;                             // pi = 2*asin(1.)
  if (fabs(fabs(z_zi) - 1.) > (double)1E-6f) {
    yi[0] = 0.0 - z_zi * o[0]; yi[1] = 0.0 - z_zi * o[1]; yi[2] = 1.0 - z_zi * o[2];
  } else {                                                                                  // line of sight along Z :360-372
    if (fabs(theta - pi_ref) < (double)1E-6f) dtheta = -dtheta;
    o[0] = sin(theta + dtheta) * cos(phi); o[1] = sin(theta + dtheta) * sin(phi); o[2] = cos(theta + dtheta);
    z_zi = 0.0 + o[0] * 0.0; z_zi = z_zi + o[1] * 0.0; z_zi = z_zi + o[2] * 1.0;
    yi[0] = 0.0 - z_zi * o[0]; yi[1] = 0.0 - z_zi * o[1]; yi[2] = 1.0 - z_zi * o[2];
    z_zi = 0.0 + yi[0] * 0.0; z_zi = z_zi + yi[1] * 0.0; z_zi = z_zi + yi[2] * 1.0;
    yi[0] = yi[0] - z_zi * 0.0; yi[1] = yi[1] - z_zi * 0.0; yi[2] = yi[2] - z_zi * 1.0;
    o[0] = sin(theta - dtheta) * cos(phi); o[1] = sin(theta - dtheta) * sin(phi); o[2] = cos(theta - dtheta);
  }
  const double nrm = sqrt(yi[0] * yi[0] + yi[1] * yi[1] + yi[2] * yi[2]);
  for (int i = 0; i < 3; i++) yi[i] = yi[i] / nrm;
  xi[0] = yi[1] * o[2] - yi[2] * o[1];
  xi[1] = -(yi[0] * o[2] - yi[2] * o[0]);
  xi[2] = yi[0] * o[1] - yi[1] * o[0];
  tex.assign(g.max_lvl + 1, HostTex());
  for (int lev = 1; lev <= g.max_lvl; lev++) {
    int div = g.base[0]; for (int k = 1; k < lev; k++) div *= g.base[1];
    const double lside = e->h_csize0 / div;
    const int nt = (int)lround(npix_hd / size_map * lside * (double)2.5f);
    HostTex& t = tex[lev]; t.nt = nt; t.prof.assign((size_t)nt * nt, 0.0);
    std::vector<double> xpp(nt);
    for (int ip = 0; ip < nt; ip++) {
      double x = ip * pixel_size_hd - nt * pixel_size_hd / 2;
      if (nt & 1) x = x + pixel_size_hd / 2;
      xpp[ip] = x;
    }
    const double cellsize = g.csize_arr[lev] * (double)0.99f;
    const double half = cellsize / 2., half_tol = (double)(1.0f + 1E-7f) * cellsize / 2.;
    double tot = 0;
    for (int iy = 0; iy < nt; iy++)          // array-element order of prof2d (ix fastest), the order sum() adds in
      for (int ix = 0; ix < nt; ix++) {
        double ro[3];
        for (int a = 0; a < 3; a++) ro[a] = 0.0 + o[a] * g.par.modelsize + xpp[ix] * xi[a] + xpp[iy] * yi[a];
        const double len = cube_chord(half_tol, half, ro, o);
        if (len > 0) t.prof[ix + (size_t)nt * iy] = len;
        tot += t.prof[ix + (size_t)nt * iy];
      }
    for (double& v : t.prof) v = v / tot;
    std::vector<char> used(nt, 0);           // row ix or column ix holds a positive value
    for (int iy = 0; iy < nt; iy++) for (int ix = 0; ix < nt; ix++) if (t.prof[ix + (size_t)nt * iy] > 0) { used[ix] = 1; used[iy] = 1; }
    t.im0 = 0; t.im1 = nt - 1;
    for (int ix = 1; ix <= nt - 1; ix++) if (used[ix]) { t.im0 = ix - 1; break; }
    for (int ix = nt - 1; ix >= 1; ix--) if (used[ix]) { t.im1 = ix + 1; break; }
  }
}
}  // namespace

int dre_make_maps(dre_engine* e, int npixel_maps, double map_size_factor, int units, const double* lambda_si, double* maps) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  const DreDev& g = e->g;
  DRE_REQUIRE(maps != nullptr && npixel_maps > 0 && map_size_factor > 0, "dre_make_maps: bad arguments");
  DRE_REQUIRE(units >= 0 && units <= 2 && (units != 2 || lambda_si), "dre_make_maps: units must be 0, 1 or 2 (2 needs lambda_si)");
  DRE_REQUIRE(e->i_obs.p && g.ndir > 0, "dre_make_maps: no i_obs pass was run");
  DRE_REQUIRE(g.L <= 1024, "dre_make_maps: more than 1024 wavelengths");
  const int np = npixel_maps, nh = 4 * npixel_maps, L = g.L;
  const size_t NS = (size_t)g.N + g.S, plane = (size_t)np * np, plane_hd = (size_t)nh * nh;
  const double size_map = map_size_factor * g.par.modelsize;                   // create_map_arrays :256-269
  const double pixel_size = size_map / np, pixel_size_hd = size_map / nh;
  DreMapGeom m; m.npix = np; m.npix_hd = nh; m.area_pixel = pixel_size * pixel_size; m.area_pixel_hd = pixel_size_hd * pixel_size_hd;
  std::vector<double> xpp_hd(nh), xpp(np);
  for (int ip = 0; ip < nh; ip++) xpp_hd[ip] = ip * pixel_size_hd - nh * pixel_size_hd / 2;     // :593-596
  for (int ip = 0; ip < np; ip++) xpp[ip] = ip * pixel_size - np * pixel_size / 2;              // :747-750
  DBuf<double> d_xhd, d_x, d_hd, d_old, d_tmp, d_map, d_out, d_flux, d_lam, d_prof;
  DBuf<int> d_list, d_n, d_csrc;
  d_xhd.upload(xpp_hd.data(), nh, e->st); d_x.upload(xpp.data(), np, e->st);
  d_hd.ensure(plane_hd * L, 1.0); d_old.ensure(plane_hd * L, 1.0); d_tmp.ensure(plane * L, 1.0); d_map.ensure(plane * L, 1.0);
  d_out.ensure(plane * L, 1.0); d_flux.ensure(2 * (size_t)L, 1.0); d_list.ensure(g.N, 1.0); d_n.ensure(1, 1.0);
  std::vector<double> lam(L, 0.0);
  if (lambda_si) lam.assign(lambda_si, lambda_si + L);
  d_lam.upload(lam.data(), L, e->st);
  std::vector<int> cs(std::max(e->S, 1), 0);
  for (int i = 0; i < e->S; i++) cs[i] = e->h_cell_src[i];
  d_csrc.upload(cs.data(), cs.size(), e->st);
  sync_g(e);
  for (int idir = 0; idir < g.ndir; idir++) {
    const double theta = e->h_dir_i_out[idir], phi = e->h_dir_i_out[g.ndir + idir];
    std::vector<HostTex> tex;
    host_textures(e, nh, size_map, pixel_size_hd, theta, phi, m, tex);
    const double* io = e->i_obs.p + (size_t)idir * NS * L;
    CK(cudaMemsetAsync(d_map.p, 0, plane * L * 8, e->st));
    for (int lev = 1; lev <= g.max_lvl; lev++) {
      const HostTex& ht = tex[lev];
      int ncell = 0;
      CK(cudaMemsetAsync(d_n.p, 0, sizeof(int), e->st));
      (e->n_launch++, k_map_select)<<<cdiv(g.N, 256), 256, 0, e->st>>>(io, lev, d_list.p, d_n.p);
      CK(cudaMemcpyAsync(&ncell, d_n.p, sizeof(int), cudaMemcpyDeviceToHost, e->st));
      CK(cudaStreamSynchronize(e->st));
      if (ncell == 0) continue;              // an empty level adds nothing (tot_flux_new = 0, :732)
      if (ht.prof.empty()) continue;         // a level whose cells are narrower than 0.2 oversampled pixels
      d_prof.upload(ht.prof.data(), ht.prof.size(), e->st);
      DreMapTex t; t.prof = d_prof.p; t.nt = ht.nt; t.im0 = ht.im0; t.im1 = ht.im1;
      CK(cudaMemsetAsync(d_hd.p, 0, plane_hd * L * 8, e->st));
      (e->n_launch++, k_map_splat)<<<ncell, 256, 0, e->st>>>(m, t, io, d_xhd.p, d_list.p, d_hd.p);
      (e->n_launch++, k_map_flux)<<<L, 512, 0, e->st>>>(d_hd.p, plane_hd, m.area_pixel_hd, d_flux.p);
      (e->n_launch++, k_map_median)<<<dim3(cdiv(nh, 128), nh), 128, 0, e->st>>>(d_hd.p, d_old.p, nh, L);
      (e->n_launch++, k_map_rebin)<<<dim3(cdiv(np, 128), np, L), 128, 0, e->st>>>(d_old.p, d_tmp.p, np, L);
      (e->n_launch++, k_map_flux)<<<L, 512, 0, e->st>>>(d_tmp.p, plane, m.area_pixel, d_flux.p + L);
      (e->n_launch++, k_map_accum)<<<dim3(cdiv(plane, 256), L), 256, 0, e->st>>>(d_map.p, d_tmp.p, d_flux.p, d_flux.p + L, plane, L);
      CK(cudaStreamSynchronize(e->st));      // d_prof is re-uploaded for the next level
    }
    if (e->S > 0) (e->n_launch++, k_map_psrc)<<<e->S, std::max(32, (L + 31) / 32 * 32), 0, e->st>>>(m, io, d_x.p, d_csrc.p, d_map.p);
    (e->n_launch++, k_map_convert)<<<dim3(cdiv(plane, 256), L), 256, 0, e->st>>>(d_map.p, d_out.p, plane, units, d_lam.p);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(maps + (size_t)idir * L * plane, d_out.p, plane * L * 8, cudaMemcpyDeviceToHost, e->st));
    CK(cudaStreamSynchronize(e->st));
  }
  DRE_CATCH(e)
}

namespace {
// find_theta_phi_obs_in (rt_routines.f90:3164-3217) for the vector a - b, host libm
void host_theta_phi(const double* a, const double* b, double& theta, double& phi) {
  const double pi_ref = 2 * (double)1.57079637050628662109375f, e8 = (double)1E-8f;
  const double v[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
  const double vm = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  theta = 0; phi = 0;
  if (!(vm > e8)) return;
  double tr = (0 * v[0] + 0 * v[1] + 1 * v[2]) / vm;
  if (fabs(tr - 1.0) < e8) tr = 1.0;
  if (fabs(tr + 1.0) < e8) tr = -1.0;
  theta = acos(tr);
  const double vxy = sin(theta);
  double pr = (vxy > e8) ? v[0] / (vm * vxy) : 1.;
  if (fabs(pr - 1.0) < e8) pr = 1.0;
  if (fabs(pr + 1.0) < e8) pr = -1.0;
  phi = acos(pr);
  if (v[1] < 0) phi = 2 * pi_ref - phi;
}
}  // namespace

int dre_make_maps_in(dre_engine* e, int32_t ndir_in, const double* ccoord_obs, int kp_maps, int units, const double* lambda_si, double* maps_in) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  check_ready(e);
  const DreDev& g = e->g;
  DRE_REQUIRE(maps_in && ccoord_obs && ndir_in > 0, "dre_make_maps_in: bad arguments");
  DRE_REQUIRE(units >= 0 && units <= 2 && (units != 2 || lambda_si), "dre_make_maps_in: units must be 0, 1 or 2 (2 needs lambda_si)");
  DRE_REQUIRE(e->i_obs_in.p && e->i_obs_ndir_in == ndir_in, "dre_make_maps_in: run dre_run_iobs with the same internal observers first");
  DRE_REQUIRE(kp_maps >= 0 && kp_maps <= std::min(DRE_MAXK, 12), "dre_make_maps_in: kp_maps out of range");
  DRE_REQUIRE((long long)ndir_in * (g.N + 1) < 0x7fffffffLL, "dre_make_maps_in: too many observers x cells");
  build_dir_table(e, kp_maps);
  const int L = g.L, N = g.N, S = e->S;
  const size_t NS = (size_t)N + S;
  const int nside = 1 << kp_maps, npix = 12 * nside * nside;
  const double pi_ref = 2 * (double)1.57079637050628662109375f;
  const double omega_hp = 4 * pi_ref / npix;
  const int nblk = 2 * e->sm_count;
  DBuf<DreMapInCell> d_simple, d_flood;
  DBuf<double> d_map, d_unif, d_out, d_lam, d_len, d_obs;
  DBuf<int> d_mark, d_list, d_cnt;
  d_map.ensure((size_t)npix * L, 1.0); d_out.ensure((size_t)npix * L, 1.0); d_unif.ensure(L, 1.0); d_cnt.ensure(1, 1.0);
  d_mark.zero((size_t)nblk * npix, e->st); d_list.ensure((size_t)nblk * 3 * npix, 1.0); d_len.ensure((size_t)nblk * npix, 1.0);
  std::vector<double> lam(L, 0.0);
  if (lambda_si) lam.assign(lambda_si, lambda_si + L);
  d_lam.upload(lam.data(), L, e->st);
  sync_g(e);
  std::vector<DreMapInCell> simple, flood;
  for (int idir = 0; idir < ndir_in; idir++) {
    const double* ro = ccoord_obs + 3 * (size_t)idir;
    simple.clear(); flood.clear();
    for (int i = 0; i < N; i++) {                       // classification of :959-1014, geometry only
      if (e->h_cchild[i] != -1) continue;
      const double* rc = &e->h_ccoord[3 * (size_t)i];
      const double rel[3] = {rc[0] - ro[0], rc[1] - ro[1], rc[2] - ro[2]};
      double dist2 = 0; for (int a = 0; a < 3; a++) dist2 += rel[a] * rel[a];
      const double cellsize = e->h_csize[i];
      const double omega_cell = (dist2 != 0) ? cellsize * cellsize / dist2 : 4 * pi_ref;
      DreMapInCell c; c.cell = i; c.pad = 0; c.ip = 0; c.len0 = 0; c.ratio = omega_cell / omega_hp;
      if (c.ratio < 1) {
        double th, ph; host_theta_phi(rc, ro, th, ph);
        c.kind = 1; c.ip = hp::ang2pix_nest(nside, th, ph); simple.push_back(c); continue;
      }
      int outs = 0;                                      // calc_outcube rt_routines.f90:2072-2085
      for (int a = 0; a < 3; a++) if (fabs(rel[a]) > (double)(1.0f + 1E-7f) * cellsize / 2.0) outs++;
      if (outs == 0) { c.kind = 2; simple.push_back(c); continue; }
      double th, ph; host_theta_phi(rc, ro, th, ph);
      c.kind = 3; c.ip = hp::ang2pix_nest(nside, th, ph);
      const double dir[3] = {sin(th) * cos(ph), sin(th) * sin(ph), cos(th)};
      // chord of the centre line of sight (find_ray_cell_intersections, cube at rc): translate so that the cube sits at the origin
      { const double half = cellsize / 2., tol = (double)(1.0f + 1E-7f) * cellsize / 2.;
        double hit[2][3]; int nhit = 0;
        for (int j = 0; j < 6 && nhit < 2; j++) {
          const int ax = j >> 1; const double dx = (j & 1) ? -half : half, den = dir[ax];
          if (den == 0.) continue;
          const double dist = ((rc[ax] + dx) - ro[ax]) / den;
          double pc[3]; bool inside = true;
          for (int a = 0; a < 3; a++) { pc[a] = ro[a] + dist * dir[a]; if (a != ax && fabs(pc[a] - rc[a]) > tol) inside = false; }
          if (!inside) continue;
          for (int a = 0; a < 3; a++) hit[nhit][a] = pc[a];
          nhit++;
        }
        if (nhit == 2) { double q = 0; for (int a = 0; a < 3; a++) q += (hit[1][a] - hit[0][a]) * (hit[1][a] - hit[0][a]); c.len0 = pow(q, 0.5); }
      }
      flood.push_back(c);
    }
    for (int i = 0; i < S; i++) {                       // point sources :1038-1073 (kind 1, entry N + i)
      const double* rc = &e->h_ccoord_p_src[3 * (size_t)i];
      double dist2 = 0; for (int a = 0; a < 3; a++) dist2 += (rc[a] - ro[a]) * (rc[a] - ro[a]);
      if (dist2 == 0) continue;
      const double cellsize = e->h_csize[e->h_cell_src[i]];
      DreMapInCell c; c.cell = N + i; c.kind = 1; c.pad = 0; c.len0 = 0; c.ratio = cellsize * cellsize / dist2 / omega_hp;
      double th, ph; host_theta_phi(rc, ro, th, ph);
      c.ip = hp::ang2pix_nest(nside, th, ph);
      simple.push_back(c);
    }
    const double* io = e->i_obs_in.p + (size_t)idir * NS * L;
    CK(cudaMemsetAsync(d_map.p, 0, (size_t)npix * L * 8, e->st)); CK(cudaMemsetAsync(d_unif.p, 0, (size_t)L * 8, e->st));
    CK(cudaMemsetAsync(d_cnt.p, 0, sizeof(int), e->st));
    d_obs.upload(ro, 3, e->st);
    if (!simple.empty()) {
      d_simple.upload(simple.data(), simple.size(), e->st);
      (e->n_launch++, k_mapin_simple)<<<cdiv(simple.size() * L, 256), 256, 0, e->st>>>(d_simple.p, (int)simple.size(), io, npix, L, d_map.p, d_unif.p);
    }
    if (!flood.empty()) {
      d_flood.upload(flood.data(), flood.size(), e->st);
      (e->n_launch++, k_mapin_flood)<<<nblk, 128, 0, e->st>>>(d_flood.p, (int)flood.size(), io, d_obs.p, kp_maps, npix, L, e->dirtab[kp_maps].p,
                                                             d_mark.p, d_list.p, d_len.p, d_cnt.p, d_map.p, idir * (N + 1));
    }
    (e->n_launch++, k_mapin_finish)<<<dim3(cdiv(npix, 256), L), 256, 0, e->st>>>(d_map.p, d_unif.p, d_out.p, npix, units, d_lam.p);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(maps_in + (size_t)idir * L * npix, d_out.p, (size_t)npix * L * 8, cudaMemcpyDeviceToHost, e->st));
    CK(cudaStreamSynchronize(e->st));       // host vectors are reused for the next observer
  }
  DRE_CATCH(e)
}

int dre_eval_exp(dre_engine* e, int32_t n, const double* x, double* y) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  if (n <= 0) return 0;
  DRE_REQUIRE(x && y, "dre_eval_exp: null pointer");
  DBuf<double> dx, dy;
  dx.upload(x, n, e->st); dy.ensure(n);
  (e->n_launch++, k_eval_exp)<<<cdiv(n, 256), 256, 0, e->st>>>(n, dx.p, dy.p);
  CK(cudaMemcpyAsync(y, dy.p, (size_t)n * 8, cudaMemcpyDeviceToHost, e->st));
  CK(cudaStreamSynchronize(e->st));
  DRE_CATCH(e)
}

int dre_set_record(dre_engine* e, int64_t max_records) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  e->rec_cap = max_records < 0 ? 0 : max_records;
  if (e->rec_cap > 0) e->rec.ensure((size_t)e->rec_cap, 1.0);
  DRE_CATCH(e)
}
int dre_get_records(dre_engine* e, dre_isect* out, int64_t cap, int64_t* n) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  unsigned long long m = 0;
  CK(cudaMemcpy(&m, e->rec_n.p, 8, cudaMemcpyDeviceToHost));
  if (n) *n = (int64_t)m;
  const int64_t k = std::min<int64_t>(std::min<int64_t>(cap, (int64_t)m), e->rec_cap);
  if (out && k > 0) CK(cudaMemcpy(out, e->rec.p, (size_t)k * sizeof(dre_isect), cudaMemcpyDeviceToHost));
  DRE_CATCH(e)
}
int dre_set_option(dre_engine* e, const char* name, double value) {
  DRE_TRY(e)
  const std::string s(name ? name : "");
  if (s == "use_bulk") e->opt_use_bulk = value != 0;
  else if (s == "hash") e->opt_hash = value != 0;
  else if (s == "stage_split") e->opt_stage_split = std::max(1, std::min(8, (int)value));
  else if (s == "ftab_kmax") e->opt_ftab_kmax = std::max(0, std::min(8, (int)value));
  else if (s == "batch_units") e->opt_batch_units = std::max(48, (int)value);
  else if (s == "sector_rotation") e->opt_sector_rot = value != 0 ? 1 : 0;
  else if (s == "blocking_sync") e->opt_blocking_sync = value != 0 ? 1 : 0;
  else if (s == "warps") e->opt_warps = std::max(1, std::min(32, (int)value));
  else if (s == "ctas_per_sm") e->opt_ctas_per_sm = std::max(1, std::min(8, (int)value));
  else if (s == "graphs") e->opt_graphs = value ? 1 : 0;
  else if (s == "cap0") e->opt_cap0 = std::max(0, (int)value);
  else if (s == "split") e->opt_split = value ? 1 : 0;
  else if (s == "coop") e->opt_coop = std::max(0, std::min(2, (int)value));
  else if (s == "coop_work") e->opt_coop_work = std::max(0, (int)value);
  else if (s == "scan_chained") { /* accepted for compatibility: every prefix sum is the chained single-launch scan */ }
  else if (s == "ring_slots") e->opt_ring_slots = (value >= 2) ? 2 : 1;
  else if (s == "chunk") e->opt_chunk = std::max(1, std::min(64, (int)value));
  else if (s == "deterministic") e->opt_det = value != 0;
  else if (s == "dir_kmax") e->opt_dir_kmax = std::max(2, std::min(10, (int)value));
  else DRE_REQUIRE(false, "dre_set_option: unknown option");
  DRE_CATCH(e)
}
int dre_get_quanta(dre_engine* e, double* q_out) {
  DRE_TRY(e)
  DRE_REQUIRE(q_out != nullptr, "dre_get_quanta: null output");
  const int L = e->g.L, Lq = dre_lq(L);
  const bool have = (int)e->h_qc.size() == 3 * Lq + 2;
  for (int k = 0; k < 3; k++)
    for (int l = 0; l < L; l++) q_out[k * L + l] = have ? e->h_qc[k * Lq + l] / 6755399441055744.0 /* 1.5 * 2^52 */ : 0.0;
  DRE_CATCH(e)
}

int dre_mark(dre_engine* e, int slot) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(slot >= 0 && slot < 8, "dre_mark: slot out of range");
  if (!e->marks[slot]) CK(cudaEventCreate(&e->marks[slot]));
  CK(cudaEventRecord(e->marks[slot], e->st));
  DRE_CATCH(e)
}
int dre_elapsed_ms(dre_engine* e, int a, int b, double* ms) {
  DRE_TRY(e)
  CK(cudaSetDevice(e->device));
  DRE_REQUIRE(a >= 0 && a < 8 && b >= 0 && b < 8 && e->marks[a] && e->marks[b] && ms, "dre_elapsed_ms: bad slots");
  CK(cudaEventSynchronize(e->marks[b]));
  float f = 0; CK(cudaEventElapsedTime(&f, e->marks[a], e->marks[b]));
  *ms = f;
  DRE_CATCH(e)
}
int64_t dre_launch_count(const dre_engine* e) { return e ? e->n_launch : 0; }

}  // extern "C"
