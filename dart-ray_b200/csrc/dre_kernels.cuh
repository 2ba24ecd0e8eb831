// dre_kernels.cuh — persistent ray-march kernels and the separate small kernels
// (cell list, point locator, pix2vec, layout transposes, scattering-iteration bookkeeping).
#pragma once
#include "dre_trace.cuh"
#include "dre_lists.cuh"

__device__ __forceinline__ void dre_src_coord(const DreDev& g, int nc, double src[3]) {   // calc_src_ccoord :1922-1939
  const int rt = g.rt_type;
  if (rt == DRE_RTT_PRECALC_SRC || rt == DRE_RTT_DIR_SRC || rt == DRE_RTT_I_OBS_DIR_SRC) {
    const int s = g.src_cell[nc];
    for (int a = 0; a < 3; a++) src[a] = g.ccoord_p_src[3 * (size_t)s + a];
  } else for (int a = 0; a < 3; a++) src[a] = g.ccoord[3 * (size_t)nc + a];
}
__device__ __forceinline__ double dre_src_lum(const DreDev& g, int nc, int l) {            // src_lum_value :1897-1919
  const int rt = g.rt_type;
  if (rt == DRE_RTT_PRECALC_SRC || rt == DRE_RTT_DIR_SRC || rt == DRE_RTT_I_OBS_DIR_SRC)
    return (g.S > 0 && g.src_cell[nc] != -1) ? g.lum_p_src[(size_t)g.src_cell[nc] * g.L + l] : 0.0;
  return g.lumcell[(size_t)nc * g.L + l];
}

#define DRE_PKT_HDR 20      // doubles of a ray packet in front of the luminosities (record 16 + direction 4)
// The stride is a multiple of 4 doubles: the direction inside a packet is written as one 32-B aligned DreDir.
__host__ __device__ inline int dre_pkt_stride(int L) { return DRE_PKT_HDR + ((L + 3) & ~3); }
__host__ __device__ inline int dre_stage_len(int ROW, int split) {
  int n = (ROW + split - 1) / split; n += n & 1; return n < 2 ? 2 : n;
}
// qc != nullptr (deterministic build): the rounding constants are copied behind the per-warp regions.
__device__ __forceinline__ void dre_warp_smem(const DreDev& g, unsigned char* smem, int warp, int lane, WarpSmem& ws, bool need, int split,
                                              const double* __restrict__ qc = nullptr) {
  ws.cq = nullptr;
  const int nwarps = blockDim.x >> 5, nq = 3 * dre_lq(g.L) + 2;
  ws.pending = 0; ws.stage_len = 0; ws.slot = 0; ws.spos = 0; ws.run_valid = 0; ws.run_n = 0; ws.run_total = 0;
  ws.mbar = nullptr; ws.fphase = 0; ws.fwait = 0; ws.runs = nullptr;
  if (!need) {   // precalc: only the look-ahead buffer lives in (dynamic) shared memory
    ws.F = ws.stage = ws.en2 = nullptr; ws.lamof = nullptr;
    ws.pre = reinterpret_cast<double*>(smem) + (size_t)warp * (8 + 32 * g.KLn + 2 + dre_pkt_stride(g.L));
    ws.pmbar = (unsigned long long*)(ws.pre + 8 + 32 * g.KLn);
    ws.pkt = ws.pre + 8 + 32 * g.KLn + 2;
    if (lane == 0) mbar_init(ws.pmbar);
    if (qc) {
      double* cq = reinterpret_cast<double*>(smem) + (size_t)nwarps * (8 + 32 * g.KLn + 2 + dre_pkt_stride(g.L));
      for (int j = threadIdx.x; j < nq; j += blockDim.x) cq[j] = qc[j];
      ws.cq = cq;
    }
    __syncthreads();
    return;
  }
  const size_t lam_bytes = ((size_t)g.ROW * 2 + 15) & ~(size_t)15;
  unsigned short* lam = (unsigned short*)smem;
  for (int j = threadIdx.x; j < g.ROW; j += blockDim.x) lam[j] = g.lamof[j];
  const int sl = dre_stage_len(g.ROW, split);
  const size_t per = (size_t)(g.ROW + sl + 2 * g.L + (g.L & 1 ? 2 : 0) + 8 + 8 + 32 * g.KLn + 2 + dre_pkt_stride(g.L)) * 8;   // keeps 16-B alignment
  double* base = (double*)(smem + lam_bytes + per * warp);
  ws.lamof = lam; ws.stage = base; ws.F = base + sl; ws.en2 = base + sl + g.ROW; ws.stage_len = sl;
  ws.mbar = (unsigned long long*)(base + sl + g.ROW + 2 * g.L + (g.L & 1 ? 2 : 0));
  ws.runs = (int*)(ws.mbar + 2);
  ws.pre = base + sl + g.ROW + 2 * g.L + (g.L & 1 ? 2 : 0) + 8;
  ws.pmbar = (unsigned long long*)(ws.pre + 8 + 32 * g.KLn);
  ws.pkt = ws.pre + 8 + 32 * g.KLn + 2;
  if (lane == 0) mbar_init(ws.pmbar);
  ws.fphase = 0; ws.fwait = 0;
  if (lane == 0) mbar_init(ws.mbar);
  if (qc) {
    double* cq = (double*)(smem + lam_bytes + per * nwarps);
    for (int j = threadIdx.x; j < nq; j += blockDim.x) cq[j] = qc[j];
    ws.cq = cq;
  }
  __syncthreads();
}
static inline size_t dre_det_smem_bytes(int L) { return (size_t)(3 * dre_lq(L) + 2) * 8; }
static inline size_t dre_trace_smem_bytes(int ROW, int L, int warps, int split) {
  const size_t lam_bytes = ((size_t)ROW * 2 + 15) & ~(size_t)15;
  const size_t per = (size_t)(ROW + dre_stage_len(ROW, split) + 2 * L + (L & 1 ? 2 : 0) + 8 + 8 + 32 * ((L + 31) / 32) + 2 + dre_pkt_stride(L)) * 8;
  return lam_bytes + per * warps;
}

__device__ __forceinline__ void dre_flush_stats(const DreAccum& A, const WarpStats& st, int lane) {
  if (lane != 0) return;
  atomicAdd(&A.counters[CNT_RAYS], st.rays); atomicAdd(&A.counters[CNT_ISECT], st.isect);
  atomicAdd(&A.counters[CNT_DEP], st.dep); atomicAdd(&A.counters[CNT_BEAM], st.beam);
  atomicAdd(&A.counters[CNT_SPLIT], st.split); atomicAdd(&A.counters[CNT_MERGE], st.merge);
  atomicAdd(&A.counters[CNT_SCA], st.sca); atomicAdd(&A.counters[CNT_ACT], st.act);
  atomicAdd(&A.counters[CNT_HASH], st.hash);
  atomicAdd(&A.counters[CNT_ACTDEP], st.actdep); atomicAdd(&A.counters[CNT_DESC], st.desc);
#ifdef DRE_TIMING
  for (int i = 0; i < 16; i++) atomicAdd(&g_dre_tim[i], st.tim[i]);
#endif
}

// ---- ray records: everything one ray_tracing call needs, resolved by one thread per task before the march.
// The march kernel then reads ONE 128-B record per ray instead of chasing task -> unit arrays -> list entry ->
// direction table -> find_cc_new2 descent through dependent loads (measured: a third of its warp time).
struct __align__(16) DreRay {
  int unit, nc, gid, step;
  int ray_ord, nk, dk, dip;
  int type, cc_old, isel_ray, fpix;
  int cc0, lvl0, loc_rc, lvl_nc;
  int e0, e1, e2, e3;          // list entries carrying the luminosity (e1..e3 = -1 unless merged)
  double prev_ray, cs_nc, src[3];
  int kind, pad;
};
// A ray packet = [DreRay record (128 B) | direction (32 B) | launch luminosity of every wavelength (8 L B)], `pstride`
// doubles apart: the march kernel fetches the packet of ray t+1 with ONE TMA bulk copy while it traces ray t.
template <int MODE, int NT>
__device__ __forceinline__ void dre_prep_record(const DreDev& g, const DreUnits& u, const DreList& H, const DreList& W,
                                                const DreTask* __restrict__ tasks, int t, double* __restrict__ pkts, int pstride,
                                                int (*s_e)[NT], int* s_mode, int* s_w, int* s_nc, int* s_dk, int* s_dip) {
  const DreTask tk = tasks[t];
  const int un = tk.unit;
  DreRay q;
  q.unit = un; q.nc = u.nc[un]; q.nk = u.nk[un]; q.gid = u.gid[un]; q.step = u.step[un]; q.ray_ord = t - u.task_seg[un];
  double src[3];
  dre_src_coord(g, q.nc, src);
  q.src[0] = src[0]; q.src[1] = src[1]; q.src[2] = src[2];
  const DreList& S = (tk.kind == 2) ? W : H;
  const unsigned m = S.meta[tk.e0];
  if (tk.kind == 0) { q.dk = q.nk; q.dip = 4 * (S.dir_ipix[tk.e0] >> (2 * (META_DIRK(m) - (q.nk - 1)))) + tk.sub; q.type = RT_HIGH; }
  else if (tk.kind == 1) { q.dk = META_DIRK(m); q.dip = S.dir_ipix[tk.e0]; q.type = RT_RECO; }
  else { q.dk = q.nk; q.dip = S.dir_ipix[tk.e0] >> (2 * (META_DIRK(m) - q.nk)); q.type = RT_LOW; }
  q.fpix = -1;
  if (MODE == MODE_DEP && g.ftab[q.nk] != nullptr) q.fpix = (q.type == RT_RECO) ? 1 : q.dip;
  q.cc_old = S.cc_old[tk.e0]; q.isel_ray = META_ISEL(m);
  q.prev_ray = S.prev[tk.e0];
  if (tk.kind == 2) q.prev_ray = (((S.prev[tk.e0] + S.prev[tk.e1]) + S.prev[tk.e2]) + S.prev[tk.e3]) / 4;   // ray_list.f90:266
  q.e0 = tk.e0; q.e1 = tk.e1; q.e2 = tk.e2; q.e3 = tk.e3; q.kind = tk.kind; q.pad = 0;
  const DreCell* cn = &g.cell[q.nc];
  q.lvl_nc = cn->lvl; q.cs_nc = g.csize[q.nc];
  q.cc0 = -1; q.lvl0 = 0; q.loc_rc = 0;
  const DreDir dir = g.dirtab[q.dk][q.dip];
  if (q.cc_old != q.nc) {                                   // relaunched ray: find_cc_new2 :1962-2066
    const double d[3] = {dir.st * dir.cp, dir.st * dir.sp, dir.ct};   // calc_cproj :1945-1953
    q.loc_rc = dre_locate(g, src, d, q.prev_ray, q.cc_old, q.cc0, q.lvl0);
  }
  double* pk = pkts + (size_t)t * pstride;
  *reinterpret_cast<DreRay*>(pk) = q;
  *reinterpret_cast<DreDir*>(pk + 16) = dir;
  s_e[0][threadIdx.x] = tk.e0; s_e[1][threadIdx.x] = tk.e1; s_e[2][threadIdx.x] = tk.e2; s_e[3][threadIdx.x] = tk.e3;
  // mode of the luminosity gather: 0 list entry, 1 merged siblings, 2 scattering source (assign_src_lum), -1 no ray
  s_mode[threadIdx.x] = (g.rt_type == DRE_RTT_SCATT && q.cc_old == q.nc) ? 2 : (tk.kind == 2 ? 1 : 0);
  s_w[threadIdx.x] = (tk.kind == 2) ? 1 : 0;
  s_nc[threadIdx.x] = q.nc; s_dk[threadIdx.x] = q.dk; s_dip[threadIdx.x] = q.dip;
}

// luminosities of the packets of one block of rays, L consecutive threads per ray (coalesced rows)
template <int NT>
__device__ __forceinline__ void dre_prep_lum(const DreDev& g, const DreAccum& A, const DreList& H, const DreList& W, int t0, int nrays,
                                             double* __restrict__ pkts, int pstride, const int (*s_e)[NT], const int* s_mode,
                                             const int* s_w, const int* s_nc, const int* s_dk, const int* s_dip) {
  const int L = g.L, n = nrays * L;
  // element idx = ray i, wavelength l; four elements per round (loads first, then stores: four gathers in flight per
  // thread), (i, l) advanced without divisions
  const int di = NT / L, dl = NT - di * L;
  int i = (int)threadIdx.x / L, l = (int)threadIdx.x - i * L;
  for (int idx = threadIdx.x; idx < n;) {
    double v[4]; size_t dst[4]; int cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (idx < n) {
        const int mode = s_mode[i];
        const double* lumS = s_w[i] ? W.lum : H.lum;
        double x;
        if (mode == 2) {                                            // assign_src_lum :3043-3071
          const int ik = g.ik_sca[l];
          const int ip = (ik >= 0) ? (s_dip[i] >> (2 * (s_dk[i] - g.class_kp[ik]))) : 0;
          x = A.prev[(size_t)s_nc[i] * g.ROW + g.off[l] + ip] * g.npix_hp[l];
        } else if (mode == 1) {                                     // merged siblings: ray_list.f90:261-264
          const double a0 = lumS[(size_t)s_e[0][i] * L + l], a1 = lumS[(size_t)s_e[1][i] * L + l];
          const double a2 = lumS[(size_t)s_e[2][i] * L + l], a3 = lumS[(size_t)s_e[3][i] * L + l];
          x = a0; x = x + a1; x = x + a2; x = x + a3;
          x = x / 4;
        } else x = lumS[(size_t)s_e[0][i] * L + l];
        v[k] = x; dst[k] = (size_t)(t0 + i) * pstride + DRE_PKT_HDR + l; cnt = k + 1;
      }
      idx += NT; i += di; l += dl;
      if (l >= L) { l -= L; i++; }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) if (k < cnt) pkts[dst[k]] = v[k];
  }
}

template <int MODE>
__global__ void __launch_bounds__(128) k_prep_rays(DreAccum A, DreStep S, int cur) {
  const DreDev& g = c_g;
  __shared__ int s_e[4][128], s_mode[128], s_w[128], s_nc[128], s_dk[128], s_dip[128];
  const int nt = S.ctl->nT;          // 0 when the step was refused (DreCtl::abort)
  const DreList H = S.H[cur], W = S.W[cur];
  for (int t0 = blockIdx.x * 128; t0 < nt; t0 += gridDim.x * 128) {
    const int t = t0 + threadIdx.x;
    if (t < nt) dre_prep_record<MODE, 128>(g, S.u, H, W, S.tasks, t, S.pkts, S.pstride, s_e, s_mode, s_w, s_nc, s_dk, s_dip);
    __syncthreads();
    dre_prep_lum<128>(g, A, H, W, t0, min(128, nt - t0), S.pkts, S.pstride, s_e, s_mode, s_w, s_nc, s_dk, s_dip);
    __syncthreads();
  }
}

// ---- ray-march kernel of the list passes (precalc / direct / scattering) ----------------
template <int MODE, int KL, int DET>
__global__ void __launch_bounds__(DRE_TRACE_THREADS, 1) k_trace(DreAccum A, DreStep S, int chunk_max) {
  extern __shared__ __align__(128) unsigned char smem[];
  const DreDev& g = c_g;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // the number of rays of the step lives on the device (0 when the step was refused, DreCtl::abort)
  const int nt = S.ctl->nT;
  if (nt <= 0) return;
  const double* __restrict__ pkts = S.pkts;
  const int pstride = S.pstride;
  const DreList out = S.OUT;
  int* __restrict__ counter = &S.ctl->trace_counter;
  // rays are handed out in chunks of consecutive tasks: one atomic per chunk, >= 8 chunks per warp for balance
  const int chunk = max(1, min(chunk_max, nt / ((int)gridDim.x * (int)(blockDim.x >> 5) * 8)));
  WarpSmem ws;
  dre_warp_smem(g, smem, warp, lane, ws, MODE == MODE_DEP, A.stage_split, DET ? A.qc : nullptr);
  WarpStats st; memset(&st, 0, sizeof(st));
#ifdef DRE_TIMING
  st.tprev = clock64();
#endif
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
    DRE_TICK(8);    // setup: ray end (outputs) + chunk fetch + packet wait
    RayIn r;
    const int* qi = reinterpret_cast<const int*>(ws.pkt);
    const int un = qi[0];
    r.nc = qi[1]; r.gid = qi[2]; r.step = qi[3]; r.ray_ord = qi[4]; r.nk = qi[5];
    const int dk = qi[6], dip = qi[7];
    r.type = qi[8]; r.cc_old = qi[9]; r.isel_ray = qi[10]; r.fpix = qi[11];
    r.cc0 = qi[12]; r.lvl0 = qi[13]; r.loc_rc = qi[14]; r.lvl_nc = qi[15];
    r.prev_ray = ws.pkt[10]; r.cs_nc = ws.pkt[11]; r.src[0] = ws.pkt[12]; r.src[1] = ws.pkt[13]; r.src[2] = ws.pkt[14];
    r.vec_mod = 0;
    r.dir.st = ws.pkt[16]; r.dir.ct = ws.pkt[17]; r.dir.sp = ws.pkt[18]; r.dir.cp = ws.pkt[19];
    r.fdir = r.dir;
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
    DRE_TICK(9);    // setup: packet fields to registers, next packet requested
    RayOut o;
    st.psel = 0; st.npsel = 0;
    trace_ray<MODE, KL, DET>(g, A, ws, lane, r, I, lost, st, o, lk);
    if (MODE == MODE_DEP && lane == 0 && st.npsel > 0) {
      const int sid = (g.rt_type == DRE_RTT_DIR_SRC) ? g.N + g.src_cell[r.nc] : r.nc;
      atomicAdd(&A.psel_sum[sid], st.psel);
      atomicAdd((unsigned long long*)&A.ipsel[sid], (unsigned long long)st.npsel);
    }
    if (lane == 0) {
      const int ty = (o.kind == OUT_HIGH) ? RT_HIGH : (o.kind == OUT_LOW ? RT_LOW : RT_GONE);
      out.unit[t] = un; out.dir_ipix[t] = dip; out.meta[t] = dre_meta(dk, r.nk, o.isel_old, ty);
      out.cc_old[t] = o.cc; out.dplane[t] = o.dplane; out.prev[t] = o.prev;
    }
    if (o.kind != OUT_NONE) {
#pragma unroll
      for (int k = 0; k < KL; k++) {
        const int l = lane + 32 * k;
        if (k < g.KLn && l < L) out.lum[(size_t)t * L + l] = I[k];
      }
    }
    DRE_TICK(14);   // ray end: outcome record + luminosity written
    t++;
  }
  if (MODE == MODE_DEP) {
    if (ws.fwait) { mbar_wait(ws.mbar, ws.fphase); ws.fwait = 0; }
    if (lane == 0) bulk_wait_all();
#pragma unroll
    for (int k = 0; k < KL; k++) {
      const int l = lane + 32 * k;
      if (k < g.KLn && l < L && lost[k] != 0) atomicAdd(&A.lum_lost[l], lost[k]);
    }
  }
  dre_flush_stats(A, st, lane);
}

// ---- rt_loop_iobs rt_routines.f90:1470-1603: one attenuation-only ray per (source, direction)
struct DreIobs {
  const int* srclist; int nsrc;        // luminous leaf cells
  int ndir, ndir_in;
  const DreDir* obs_dir;               // [ndir] external directions dir_i_out
  const DreDir* in_dir;                // [nsrc][ndir_in] find_theta_phi_obs_in :3164-3217 (host libm)
  const double* in_vecmod;             // [nsrc][ndir_in]
  const int* in_ip;                    // [nsrc][ndir_in][nclass] ang2pix_nest of the observer direction
  double* i_obs;                       // [ndir][N+S][L]
  double* i_obs_in;                    // [ndir_in][N+S][L]
};
template <int KL>
__global__ void k_iobs(DreAccum A, DreIobs q, int* __restrict__ counter) {
  const DreDev& g = c_g;
  const int lane = threadIdx.x & 31;
  __shared__ double s_pre[8][8 + 32 * KL];      // look-ahead buffers of the 8 warps of a CTA
  WarpSmem ws; ws.cq = nullptr; ws.pending = 0; ws.slot = 0; ws.spos = 0; ws.run_valid = 0; ws.F = ws.stage = ws.en2 = nullptr; ws.lamof = nullptr; ws.mbar = nullptr; ws.fphase = 0; ws.fwait = 0; ws.runs = nullptr;
  ws.pre = s_pre[threadIdx.x >> 5];
  WarpStats st; memset(&st, 0, sizeof(st));
  double lost[KL];
#pragma unroll
  for (int k = 0; k < KL; k++) lost[k] = 0;
  const int L = g.L, nun = q.ndir + q.ndir_in;
  const long long nt = (long long)q.nsrc * nun;
  for (;;) {
    int t = 0;
    if (lane == 0) t = atomicAdd(counter, 1);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= nt) break;
    const int si = t / nun, j = t % nun;
    RayIn r;
    r.nc = q.srclist[si]; r.nk = DRE_NSIDE_MIN_K; r.gid = t; r.step = 0; r.ray_ord = 0;
    dre_src_coord(g, r.nc, r.src);
    r.cc_old = r.nc; r.isel_ray = 0; r.prev_ray = 0; r.vec_mod = 0;
    r.cc0 = -1; r.lvl0 = 0; r.loc_rc = 0; r.lvl_nc = g.cell[r.nc].lvl; r.cs_nc = g.csize[r.nc];
    const bool inner = (j >= q.ndir);
    const int jj = inner ? j - q.ndir : j;
    if (!inner) { r.dir = q.obs_dir[jj]; r.type = RT_IOBS; }
    else { r.dir = q.in_dir[(size_t)si * q.ndir_in + jj]; r.vec_mod = q.in_vecmod[(size_t)si * q.ndir_in + jj]; r.type = RT_IOBS_IN; }
    r.fdir = r.dir; r.fpix = -1;
    double I[KL];
#pragma unroll
    for (int k = 0; k < KL; k++) {
      const int l = lane + 32 * k;
      I[k] = 0;
      if (k < g.KLn && l < L) {
        if (g.rt_type == DRE_RTT_I_OBS) {
          const double* row = A.prev + (size_t)r.nc * g.ROW + g.off[l];
          if (!inner) I[k] = row[g.npix_hp[l] + jj] * g.npix_hp[l];                       // :1541
          else {
            const int ik = g.ik_sca[l];
            const int ip = (ik >= 0) ? q.in_ip[((size_t)si * q.ndir_in + jj) * g.nclass + ik] : 0;
            I[k] = row[ip] * g.npix_hp[l];
          }
        } else I[k] = dre_src_lum(g, r.nc, l);
      }
    }
    RayOut o;
    LamK lk0[KL];
#pragma unroll
    for (int k = 0; k < KL; k++) { lk0[k].npix = 0; lk0[k].alb = 0; }
    trace_ray<MODE_IOBS, KL>(g, A, ws, lane, r, I, lost, st, o, lk0);
    const int sid = (g.rt_type == DRE_RTT_I_OBS_DIR_SRC) ? g.N + g.src_cell[r.nc] : r.nc;
    double* dst = (inner ? q.i_obs_in : q.i_obs) + ((size_t)jj * (g.N + g.S) + sid) * L;
    if (o.iobs_valid) {
#pragma unroll
      for (int k = 0; k < KL; k++) {
        const int l = lane + 32 * k;
        if (k < g.KLn && l < L) dst[l] += I[k];
      }
    }
  }
  dre_flush_stats(A, st, lane);
}

// ---- phase-function table: one warp per pixel of nside 2^k writes calc_ffn_arr for all wavelengths
__global__ void k_build_ftab(int k, double* __restrict__ rows) {
  const DreDev& g = c_g;
  const int lane = threadIdx.x & 31;
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long npix = 12ll << (2 * k);
  if (w >= npix) return;
  WarpSmem ws; ws.cq = nullptr; ws.pending = 0; ws.slot = 0; ws.spos = 0; ws.run_valid = 0; ws.stage = ws.en2 = nullptr; ws.lamof = nullptr; ws.mbar = nullptr; ws.fphase = 0; ws.fwait = 0; ws.runs = nullptr;
  ws.F = rows + (size_t)w * g.ROW;
  unsigned act[DRE_KL];
  for (int q = 0; q < DRE_KL; q++) { const int rem = g.L - 32 * q; act[q] = rem >= 32 ? 0xffffffffu : (rem > 0 ? ((1u << rem) - 1u) : 0u); }
  dre_phase<DRE_KL>(g, ws, lane, g.dirtab[k][w], act);
}

// ---- cell-list kernel: leaf && sum(lum) /= 0 gate of rt_loop rt_routines.f90:1016 -------
__global__ void k_src_flags(int* __restrict__ flag) {
  const DreDev& g = c_g;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  double s = 0;
  for (int l = 0; l < g.L; l++) s += dre_src_lum(g, i, l);
  bool on = (g.cchild[i] == -1 && s != 0);
  if (on && g.i2d) {                       // skip_cell_2D_rt_loop :1174-1187
    const double* c = g.ccoord + 3 * (size_t)i;
    if (g.i2d == 1) on = !((c[0] <= 0.) || (c[1] <= 0.) || (c[2] <= 0.));      // only positive cells in the first loop
    else on = !((c[0] != 0.) && (c[1] != 0.) && (c[2] != 0.));                 // only cells on the Cartesian planes in the second
  }
  flag[i] = on ? 1 : 0;
}
__global__ void k_src_compact(const int* __restrict__ flag, const int* __restrict__ pos, int N, int* __restrict__ list) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N && flag[i]) list[pos[i]] = i;
}

// seeds of a batch of units: create_high_ray_list ray_list.f90:71-106 (nside 2 pixel = sector)
__global__ void k_seed(DreStep S, int U, const int* __restrict__ srclist, int unit0, int stride, int rank_off, int nsrc, int pmul, int rot) {
  const DreDev& g = c_g;
  const DreUnits u = S.u;
  const DreList H = S.H[0], W = S.W[0];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) S.ctl->U = U;
  if (i > U) return;
  H.seg[i] = i; W.seg[i] = 0;
  if (i == U) return;
  // Units are handed out in a scrambled source order -- source ordinal (s * pmul) mod nsrc, pmul coprime to nsrc -- so that every
  // batch is a sample of the whole model (neighbouring cells have similar ray trees: in cell order the first batch says little
  // about the list sizes of the later ones).  The global unit id (record / checksum key) stays source_ordinal*48 + sector.
  const long long gq = (long long)rank_off + (long long)(unit0 + i) * stride;
  // the sector is rotated by the source position: with 48 % nranks == 0 a rank would otherwise trace the same sectors of every
  // source (the up-looking ones of a disc, say) and the ranks would finish a pass at different times
  const int sector = rot ? (int)((gq % 48 + gq / 48) % 48) : (int)(gq % 48);
  const int sord = (int)(((gq / 48) * (long long)pmul) % nsrc);
  const int gid = sord * 48 + sector;
  const int nc = srclist[sord];
  u.nc[i] = nc; u.gid[i] = gid; u.nk[i] = DRE_NSIDE_MIN_K; u.step[i] = 0; u.mode[i] = 1;
  H.unit[i] = i; H.dir_ipix[i] = sector; H.meta[i] = dre_meta(1, 1, 0, RT_HIGH); H.cc_old[i] = nc;
  H.dplane[i] = 0; H.prev[i] = 0;
  for (int l = 0; l < g.L; l++) H.lum[(size_t)i * g.L + l] = dre_src_lum(g, nc, l);
}

// ---- traversal records: centre, level, size and the six tabulated neighbour steps of every cell.
// One thread per (cell, face): cindex_to_ccindd (smooth_grid_routines.f90:793-813), `increment`
// (:751-769) along the face axis, then ccindd_to_cc (:820-852) restricted to the levels <= lvl(cell).
__global__ void k_build_nbr(DreCell* __restrict__ cells) {
  const DreDev& g = c_g;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= g.N * 6) return;
  const int cc = t / 6, face = t % 6, axis = face >> 1;
  const int clvl = g.lvl[cc];
  if (face == 0) {
    DreCell* o = &cells[cc];
    for (int a = 0; a < 3; a++) o->c[a] = g.ccoord[3 * (size_t)cc + a];
    o->lvl = clvl; o->pad = 0; o->csize = g.csize[cc];
  }
  unsigned e = DRE_NBR_OUT;
  if (clvl >= 1) {
    unsigned long long c[3];
    dre_decode(g, cc, clvl, c);
    int k = (face & 1) ? 1 : -1;
    unsigned long long ca = c[axis];
    for (int j = clvl; j >= 1; j--) {
      const int b = dre_base(g, j);
      int v = cx_get(ca, j) + k;
      if (v >= b) v = 0; else if (v < 0) v = b - 1; else k = 0;
      ca = cx_set(ca, j, v);
      if (k == 0) break;
    }
    c[axis] = ca;
    if (k == 0) {
      int l = 1;
      for (int i = 1; i <= clvl; i++) {
        const int b = dre_base(g, i);
        l += (cx_get(c[2], i) * b + cx_get(c[1], i)) * b + cx_get(c[0], i);
        const int ch = g.cchild[l];
        if (ch == -1) { e = (unsigned)l | ((unsigned)i << 27); break; }
        if (i == clvl) { e = 0x80000000u | (unsigned)ch; break; }
        l = ch;
      }
    }
  }
  cells[cc].nbr[face] = e;
}

// ---- bounding-box / locator kernel: find_cc_new2 for arbitrary points ----------------------
__global__ void k_locate(int n, const double* __restrict__ xyz, const int* __restrict__ cc_old, int* __restrict__ out) {
  const DreDev& g = c_g;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double src[3] = {0, 0, 0};
  const double d[3] = {xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]};
  int cc_new = -1, clvl = 0;
  const int rc = dre_locate(g, src, d, 1.0, cc_old[i], cc_new, clvl);
  out[i] = (rc == 0) ? cc_new : -1;
}
__global__ void k_eval_exp(int n, const double* __restrict__ x, double* __restrict__ y) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = dre_exp(x[i]);
}
// pix2vec = pix2ang_nest + calc_cproj from the direction table
__global__ void k_pix2vec(const DreDir* __restrict__ tab, int n, const int* __restrict__ ipix, double* __restrict__ vec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const DreDir d = tab[ipix[i]];
  vec[3 * (size_t)i] = d.st * d.cp; vec[3 * (size_t)i + 1] = d.st * d.sp; vec[3 * (size_t)i + 2] = d.ct;
}

// ---- layout transposes between the Fortran arrays and the accumulator rows -----------------
// dir 0: rows[c][off+p] = plane[c][p] ; dir 1: plane[c][p] = rows[c][off+p]
__global__ void k_plane_rows(double* __restrict__ rows, double* __restrict__ plane, int N, int P, int ROW, int off, int dir) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)N * P) return;
  const size_t c = i / P; const int p = (int)(i % P);
  if (dir == 0) rows[c * ROW + off + p] = plane[i]; else plane[i] = rows[c * ROW + off + p];
}
__global__ void k_zero_sca(double* __restrict__ rows, int N, int ROW, int L) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int w = ROW - L;
  if (i >= (size_t)N * w) return;
  rows[(i / w) * ROW + L + (i % w)] = 0;
}
__global__ void k_add_sca(double* __restrict__ dst, const double* __restrict__ src, int N, int ROW, int L) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int w = ROW - L;
  if (i >= (size_t)N * w) return;
  const size_t j = (i / w) * ROW + L + (i % w);
  dst[j] += src[j];
}
// lumcell(l,c) = sum_{p<npix_hp} scaspe(l)(p,c) for leaves (calc_total_luminosity_sca rt_routines.f90:722-740);
// all = 1 sums the observer slots too (case rtt_i_obs :760-771)
__global__ void k_lumcell_sca(const double* __restrict__ rows, double* __restrict__ lumcell, int all) {
  const DreDev& g = c_g;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)g.N * g.L) return;
  const int c = (int)(i / g.L), l = (int)(i % g.L);
  double s = 0;
  if (g.cchild[c] == -1) {
    const double* r = rows + (size_t)c * g.ROW + g.off[l];
    const int n = all ? g.npix[l] : g.npix_hp[l];
    for (int p = 0; p < n; p++) s += r[p];
  }
  lumcell[i] = s;
}
// out[l] = sum_c a[c][l*stride...] in a fixed order: one block per wavelength, strided partials,
// tree reduction (deterministic for a fixed block size)
__global__ void k_colsum(const double* __restrict__ a, int N, int L, double* __restrict__ out) {
  __shared__ double sh[256];
  const int l = blockIdx.x;
  double s = 0;
  for (int c = threadIdx.x; c < N; c += blockDim.x) s += a[(size_t)c * L + l];
  sh[threadIdx.x] = s; __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) out[l] = sh[0];
}
// tot_rad_en(l) = sum over ALL cells of the HEALPix part of scaspe (rt_routines.f90:690-698)
__global__ void k_totrad(const double* __restrict__ rows, double* __restrict__ out) {
  const DreDev& g = c_g;
  __shared__ double sh[256];
  const int l = blockIdx.x;
  const int off = g.off[l], n = g.npix_hp[l];
  double s = 0;
  for (int c = threadIdx.x; c < g.N; c += blockDim.x) { const double* r = rows + (size_t)c * g.ROW + off; for (int p = 0; p < n; p++) s += r[p]; }
  sh[threadIdx.x] = s; __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) out[l] = sh[0];
}
// calc_sed / sum_i_obs sed_routines.f90:90-146, 212-246 on the device-resident i_obs array [ndir][N+S][L]:
// sed[j][l] = sum_c i_obs * conv(l) * (csize(c)/dist_obs)^2 + the point sources (size of their host cell).
// One block per (direction, wavelength); strided partial sums + a tree: fixed order for a fixed block size.
__global__ void k_sed(const double* __restrict__ i_obs, const int* __restrict__ cell_src, double dist_obs, int units,
                      const double* __restrict__ lambda_si, double* __restrict__ sed) {
  const DreDev& g = c_g;
  __shared__ double sh[256];
  const int l = blockIdx.x % g.L, j = blockIdx.x / g.L;
  const size_t NS = (size_t)g.N + g.S;
  const double parsec = (double)3.08567758e16f, cspeed = (double)2.99792458e8f;
  double s = 0;
  for (size_t q = threadIdx.x; q < NS; q += blockDim.x) {
    double v = i_obs[((size_t)j * NS + q) * g.L + l];
    if (units == 0) { v = v * 1e-7; v = v * 1e26 / (parsec * parsec); }
    else if (units == 1) { v = v * 1e26 / (parsec * parsec); }
    else { v = v * (lambda_si[l] * lambda_si[l]) / cspeed; v = v * 1e26 / (parsec * parsec); }
    const double cs = (q < (size_t)g.N) ? g.csize[q] : g.csize[cell_src[q - g.N]];
    s += v * ((cs / dist_obs) * (cs / dist_obs));
  }
  sh[threadIdx.x] = s; __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) sed[(size_t)j * g.L + l] = sh[0];
}
// u part of the rows <-> (N,L) array.  dir 0: rows <- u ; 1: u <- rows
__global__ void k_u_rows(double* __restrict__ rows, double* __restrict__ u, int N, int L, int ROW, int dir) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)N * L) return;
  const size_t j = (i / L) * ROW + (i % L);
  if (dir == 0) rows[j] = u[i]; else u[i] = rows[j];
}
// rt_mpi delta trick for the scattering u_final (io_routines.f90:3743-3757): rows.u = max(rows.u - u_fest, 0) / + u_fest
// qc_u != nullptr (deterministic mode): u_fest is taken on the rounding grid of the pass, as the accumulator rows were at
// its start, so the difference, its sum over the ranks and the final sum are all exact.
__global__ void k_u_delta(double* __restrict__ rows, const double* __restrict__ u_fest, int N, int L, int ROW, int dir,
                          const double* __restrict__ qc_u) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)N * L) return;
  const size_t j = (i / L) * ROW + (i % L);
  const double uf = qc_u ? dre_rq(u_fest[i], qc_u[i % L]) : u_fest[i];
  if (dir == 0) { const double v = rows[j] - uf; rows[j] = v < 0 ? 0 : v; } else rows[j] = rows[j] + uf;
}
// rt_mpi, lum_lost correction of the scattering iterations (io_routines.f90:3846-3848): every rank starts the pass from the
// reduced lum_lost of the previous one, so before the sum  lum_lost_r -= (R-1)/R lum_lost_prev.  Deterministic mode (qc):
// the exact form of the same correction -- subtract the grid value of lum_lost_prev (kept in `keep`) before the sum (phase 0)
// and add it back once after it (phase 1).
__global__ void k_lum_lost_corr(double* __restrict__ ll, const double* __restrict__ lp, double* __restrict__ keep, int L, int R,
                                const double* __restrict__ qc, int phase) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  if (qc) {
    if (phase == 0) { const double q = dre_rq(lp[l], qc[l]); keep[l] = q; ll[l] = ll[l] - q; }
    else ll[l] = ll[l] + keep[l];
  } else if (phase == 0) ll[l] = ll[l] - (R - 1) / (double)R * lp[l];
}
// Deterministic mode, start of a pass: what the accumulators already hold is put on the rounding grid of the pass.
__global__ void k_quantize_rows(double* __restrict__ rows, size_t n, int ROW, int L, const unsigned short* __restrict__ lamof,
                                const double* __restrict__ qc) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int j = (int)(i % ROW), Lq = dre_lq(L);
  const double c = (j < Lq) ? qc[Lq + min(j, L - 1)] : qc[lamof[j] >> 3];
  rows[i] = dre_rq(rows[i], c);
}
__global__ void k_quantize_cols(double* __restrict__ a, size_t n, int L, const double* __restrict__ qc_cols, const double* __restrict__ qc_one) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = dre_rq(a[i], qc_cols ? qc_cols[i % L] : qc_one[0]);
}
__global__ void k_psel_ratio(const double* __restrict__ s, const long long* __restrict__ n, int m, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) out[i] = n[i] > 0 ? s[i] / (double)n[i] : 0.0;
}
