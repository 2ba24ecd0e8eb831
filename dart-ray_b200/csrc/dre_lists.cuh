// dre_lists.cuh — wavefront form of the per-sector ray lists (ray_list.f90, main_dir_loop
// rt_routines.f90:1343-1467, define_next_level :2745-2770).
//
// Every work unit (source, sector) owns a HIGH and a LOW list, stored unit-grouped in two
// global SoA arrays.  One resolution step of ALL units of a batch is a fixed sequence of
// data-parallel kernels: select -> scan -> gather (extract_ray_list), sibling search + leader
// resolution (the 4-ray merge), task emission, the ray-march kernel, and a stable rebuild of
// both lists.  Prefix sums keep the reference's list order (append order == processing order),
// which the merge and therefore the intersection lists depend on.
//
// The step is DEVICE-DRIVEN: every count (list lengths, extracted entries, tasks) stays in device
// memory -- the length of a list is the last element of its segment array -- and every kernel is
// launched with a fixed grid that strides over a count it reads itself.  The host enqueues the whole
// step without a single read-back and looks at one small control block (DreCtl) per step: units still
// active, error flag, and "a buffer is too small" (the step then changes nothing and is repeated after
// the buffers have grown).  The launch sequence of a step is therefore a constant and is replayed
// from a CUDA graph.
#pragma once
#include "dre_common.cuh"

#define STEP_THREADS 256
#define GRID_LOOP(i, n) for (int i = blockIdx.x * blockDim.x + threadIdx.x, _n = (n); i < _n; i += gridDim.x * blockDim.x)

// Device-written control block of the wavefront; copied to the host once per step.
struct DreCtl {
  int U;               // units of the batch (set by k_seed)
  int active;          // units that still have rays after this step
  int nkmask;          // bit mask of their log2(nside)
  int abort;           // a buffer is too small for this step: nothing was changed, grow to need* and repeat
  int nH, nW;          // list lengths at the start of the step
  int nExt, nT;        // extracted entries, ray_tracing calls of the step
  int nH2, nW2;        // list lengths after the step
  int needH, needW, needT, needE;   // entries the step needs in the HIGH / LOW lists, the task arrays and the event buffer
  int trace_counter;   // task fetch counter of the ray-march kernel
  unsigned scan_epoch; // epoch base of the chained scans of the step
  int tickets[16];     // tile tickets, one per scan job of the step
};

// Pointers and capacities of the wavefront scratch (host-owned; passed to the kernels by value).
struct DreStep {
  DreList H[2], W[2], OUT;           // n unused: the length of H / W is seg[U]
  DreUnits u;
  int *selH, *posH, *selW, *posW;    // [capH (+1)], [capW (+1)]
  int *extcnt, *ext_seg;             // [U], [U+1]
  int *ext, *ipx, *gone, *cand, *ncand, *nxt, *hasprev, *unit_of_ext, *ntask, *nreco, *task_off, *reco_off;   // [capH + capW (+1)]
  int *reco_cnt, *keptH, *keptW, *newHcnt, *newWcnt;   // [U]
  int *fH, *fW, *posOH, *posOW;      // [capT (+1)]
  unsigned long long* hkey; int* hhead; int hcap;   // [hcap] open-addressing table of the step: (unit, parent pixel) -> chain of extracted LOW entries
  int* hlink;                        // [capH + capW] next entry of the chain
  DreTask* tasks;                    // [capT]
  double* pkts; int pstride;         // [capT] ray packets
  int capH, capW, capT;
  DreCtl* ctl;
  unsigned long long* scan_desc; int desc_stride;   // tile descriptors of the chained scans, one region per job slot
  int* err;                          // first reference STOP condition hit
  unsigned long long* steps;         // counter of (unit, step) pairs executed
  int kmax;                          // largest log2(nside) with a direction table
};

// ---------------------------------------------------------------------------------------
// exclusive scans: out[0..n] (n+1 values, out[n] = total), up to 2 independent jobs per launch
// ---------------------------------------------------------------------------------------
#define SCAN_THREADS 512
#define SCAN_ITEMS 4
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)
struct ScanJob { const int* in; int* out; const int* n_ptr; int n_const; int slot; int* total_out; };   // n = n_ptr ? *n_ptr : n_const; total also to *total_out
struct ScanJobs { ScanJob j[2]; };

// Chained scan with decoupled look-back, persistent form.  A block takes tile tickets from an atomic counter, so a
// tile's predecessors are always held by blocks that are already running (no assumption on the dispatch order of the
// grid).  Tile t publishes {epoch, state, value} as one 64-bit word (state 1 = tile aggregate, 2 = inclusive prefix) and
// adds up the aggregates of its predecessors until it meets an inclusive prefix.  The epoch makes the words of earlier
// scans on the same descriptor region read as "not published", so the region is never cleared.  The ticket counters of a
// step are cleared by its first kernel.  `skip` (DreCtl::abort) turns the scan into n = 0.
#define SCAN_DESC(epoch, state, val) (((unsigned long long)(epoch) << 34) | ((unsigned long long)(state) << 32) | (unsigned long long)(unsigned)(val))
// One job of the chained scan, run by every block of the calling kernel (blockDim.x == SCAN_THREADS, 1-D blocks).
__device__ __forceinline__ void dre_scan_job(const ScanJob& jb, DreCtl* ctl, unsigned long long* desc_all, int desc_stride,
                                             unsigned epoch_off, int skip_on_abort) {
  __shared__ int wsum[SCAN_THREADS / 32];
  __shared__ int s_prefix, s_tile;
  if (jb.out == nullptr) return;
  int n = jb.n_ptr ? *jb.n_ptr : jb.n_const;
  if (skip_on_abort && ctl->abort) return;
  if (n < 0) n = 0;
  const int ntiles = n > 0 ? (n + SCAN_TILE - 1) / SCAN_TILE : 1;
  unsigned epoch = (ctl->scan_epoch + epoch_off) % 0x3fffffffu + 1u;   // never 0: a cleared word reads as unpublished
  volatile unsigned long long* d = desc_all + (size_t)jb.slot * desc_stride;
  int* ticket = &ctl->tickets[jb.slot];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (;;) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    if (tile >= ntiles) break;
    const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS], tot = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = (base + i < n) ? jb.in[base + i] : 0; tot += v[i]; }
    int x = tot;
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    if (w == 0) {
      int s = (lane < SCAN_THREADS / 32) ? wsum[lane] : 0;
      for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
      if (lane < SCAN_THREADS / 32) wsum[lane] = s;
    }
    __syncthreads();
    if (w == 0) {          // look-back by one warp: 32 predecessors per round
      const int agg = wsum[SCAN_THREADS / 32 - 1];
      int prefix = 0;
      if (tile > 0) {
        if (lane == 0) { d[tile] = SCAN_DESC(epoch, 1u, agg); __threadfence(); }
        for (int j0 = tile - 1; ; ) {
          const int j = j0 - lane;
          // tile -1 stands for "inclusive prefix 0"
          const unsigned long long q = (j >= 0) ? d[j] : SCAN_DESC(epoch, 2u, 0);
          const unsigned state = (unsigned)(q >> 32) & 3u;
          const bool ready = ((unsigned)(q >> 34) == epoch) && state != 0;
          const unsigned rm = __ballot_sync(0xffffffffu, ready), im = __ballot_sync(0xffffffffu, ready && state == 2u);
          int upto;                                          // lanes 0..upto take part in this round
          if (im) {
            upto = __ffs(im) - 1;                            // nearest predecessor that knows its inclusive prefix
            const unsigned need = (upto == 31) ? 0xffffffffu : ((2u << upto) - 1u);
            if ((rm & need) != need) continue;               // an aggregate in between is not published yet
          } else {
            if (rm != 0xffffffffu) continue;
            upto = 31;
          }
          int val = (lane <= upto) ? (int)(unsigned)(q & 0xffffffffull) : 0;
          for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
          prefix += val;
          if (im) break;
          j0 -= 32;
        }
      }
      if (lane == 0) { d[tile] = SCAN_DESC(epoch, 2u, prefix + agg); s_prefix = prefix; }
    }
    __syncthreads();
    int excl = s_prefix + x - tot + (w > 0 ? wsum[w - 1] : 0);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) jb.out[base + i] = excl; excl += v[i]; }
    if (tile == ntiles - 1 && threadIdx.x == SCAN_THREADS - 1) { jb.out[n] = excl; if (jb.total_out) *jb.total_out = excl; }
    __syncthreads();       // wsum / s_tile are reused by the next tile
  }
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_jobs(ScanJobs J, DreCtl* ctl, unsigned long long* desc_all, int desc_stride,
                                                            unsigned epoch_off, int skip_on_abort) {
  dre_scan_job(J.j[blockIdx.y], ctl, desc_all, desc_stride, epoch_off + blockIdx.y, skip_on_abort);
}

// ---------------------------------------------------------------------------------------
// step kernels
// ---------------------------------------------------------------------------------------
// Sibling table of a step: slots in use for a LOW list of nW entries (at most nW entries are extracted: load <= 1/2).
// 2 * nW + 64 stays far inside an int: the list capacity is bounded by the 32-GiB list budget (< 3e7 entries, dre_run_pass_resident).
#define DRE_HEMPTY 0xffffffffffffffffull
__device__ __forceinline__ int dre_hsize(const DreStep& S, int nW) { return min(S.hcap, 2 * nW + 64); }
__device__ __forceinline__ unsigned dre_hslot(unsigned long long key, int hsize) {
  key ^= key >> 30; key *= 0xbf58476d1ce4e5b9ull; key ^= key >> 27; key *= 0x94d049bb133111ebull; key ^= key >> 31;
  return __umulhi((unsigned)key, (unsigned)hsize);
}

// Start of a step: mode of each unit -- extract from HIGH if it has entries, else from LOW (ray_list.f90:120,167) -- and the
// selection flags of extract_ray_list: HIGH entries with nside == nside/2, LOW entries with nside == 2*nside.
__device__ __forceinline__ void ph_select(const DreStep& S, int cur) {
  const DreList H = S.H[cur], W = S.W[cur];
  const DreUnits u = S.u;
  const int U = S.ctl->U, nH = H.seg[U], nW = W.seg[U];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    DreCtl* c = S.ctl;
    c->abort = 0; c->active = 0; c->nkmask = 0; c->trace_counter = 0; c->nH = nH; c->nW = nW;
    for (int i = 0; i < 16; i++) c->tickets[i] = 0;
  }
  GRID_LOOP(i, U) {
    const int nh = H.seg[i + 1] - H.seg[i], nw = W.seg[i + 1] - W.seg[i];
    u.mode[i] = nh > 0 ? 1 : (nw > 0 ? 2 : 0);
  }
  GRID_LOOP(q, dre_hsize(S, nW)) { S.hkey[q] = DRE_HEMPTY; S.hhead[q] = -1; }   // sibling table of this step (ph_fill_ext, ph_cand)
  GRID_LOOP(i, nH) {                     // a unit with HIGH entries is in mode 1
    const int un = H.unit[i];
    S.selH[i] = (META_NSK(H.meta[i]) == u.nk[un] - 1) ? 1 : 0;
  }
  GRID_LOOP(i, nW) {
    const int un = W.unit[i];
    const int nh = H.seg[un + 1] - H.seg[un];
    S.selW[i] = (nh == 0 && META_NSK(W.meta[i]) == u.nk[un] + 1) ? 1 : 0;
  }
}

__device__ __forceinline__ void ph_unit_ext(const DreStep& S, int cur) {
  const DreList H = S.H[cur], W = S.W[cur];
  const DreUnits u = S.u;
  GRID_LOOP(i, S.ctl->U) {
    int c = 0;
    if (u.mode[i] == 1) c = S.posH[H.seg[i + 1]] - S.posH[H.seg[i]];
    else if (u.mode[i] == 2) c = S.posW[W.seg[i + 1]] - S.posW[W.seg[i]];
    S.extcnt[i] = c;
  }
}

// EXT[x] = index of the extracted entry in its old list; for LOW-mode units also the parent
// pixel at the current nside (ang2pix_nest(nside, theta, phi), ray_list.f90:218) and gone = 0.
__device__ __forceinline__ void ph_fill_ext(const DreStep& S, int cur) {
  const DreUnits u = S.u;
  for (int which = 1; which <= 2; which++) {
    const DreList l = (which == 1) ? S.H[cur] : S.W[cur];
    const int* sel = (which == 1) ? S.selH : S.selW;
    const int* pos = (which == 1) ? S.posH : S.posW;
    GRID_LOOP(i, l.seg[S.ctl->U]) {
      if (!sel[i]) continue;
      const int un = l.unit[i];
      const int x = S.ext_seg[un] + (pos[i] - pos[l.seg[un]]);
      S.ext[x] = i; S.unit_of_ext[x] = un;
      S.gone[x] = 0; S.hasprev[x] = 0;
      if (which == 2) {
        const int ip = l.dir_ipix[i] >> (2 * (META_DIRK(l.meta[i]) - u.nk[un]));
        S.ipx[x] = ip;
        // chain of the extracted entries of this (unit, parent pixel): ph_cand finds the siblings there instead of
        // scanning the rest of the unit's segment
        const unsigned long long key = ((unsigned long long)(unsigned)un << 32) | (unsigned)ip;
        const int hs = dre_hsize(S, l.seg[S.ctl->U]);
        unsigned q = dre_hslot(key, hs);
        for (;;) {
          const unsigned long long o = atomicCAS(&S.hkey[q], DRE_HEMPTY, key);
          if (o == DRE_HEMPTY || o == key) break;
          q = (q + 1 == (unsigned)hs) ? 0u : q + 1;
        }
        S.hlink[x] = atomicExch(&S.hhead[q], x);
        S.ncand[x] = (int)q;            // slot, read back by ph_cand before it writes the candidate count
      } else S.ipx[x] = -1;
    }
  }
}

// sibling search of extract_ray_list (ray_list.f90:237-253): the first <= 3 later entries of the
// same unit with the same parent pixel, the same isel and dplane within 0.001 (relative to the
// leader).  `gone` is NOT consulted for the members, as in the reference (SURVEY A.3 note).
__device__ __forceinline__ void ph_cand(const DreStep& S, int cur) {
  const DreList w = S.W[cur];
  const DreUnits u = S.u;
  GRID_LOOP(x, S.ext_seg[S.ctl->U]) {
    const int un = S.unit_of_ext[x];
    if (u.mode[un] != 2) { S.ncand[x] = 0; S.nxt[x] = -1; continue; }
    const int i = S.ext[x];
    const int ip = S.ipx[x], isel = META_ISEL(w.meta[i]);
    const double dpl = w.dplane[i];
    // the later entries of the chain, in list order (a chain holds the <= 4 rays of one parent pixel; longer chains take
    // the scan below)
    int later[8], nl = 0;
    bool scan = false;
    for (int y = S.hhead[S.ncand[x]]; y >= 0; y = S.hlink[y]) {
      if (y <= x) continue;
      if (nl == 8) { scan = true; break; }
      int k = nl++;
      while (k > 0 && later[k - 1] > y) { later[k] = later[k - 1]; k--; }
      later[k] = y;
    }
    int n = 0, first = -1;
    if (!scan) {
      for (int q = 0; q < nl && (n < 3 || first < 0); q++) {
        const int y = later[q];
        if (first < 0) first = y;
        if (n == 3) break;
        const int j = S.ext[y];
        if (META_ISEL(w.meta[j]) != isel) continue;
        if (fabs(dpl - w.dplane[j]) / fabs(dpl) < (double)0.001f) S.cand[3 * (size_t)x + n++] = y;
      }
    } else {
      const int end = S.ext_seg[un + 1];
      for (int y = x + 1; y < end && (n < 3 || first < 0); y++) {
        if (S.ipx[y] != ip) continue;
        if (first < 0) first = y;
        if (n == 3) break;
        const int j = S.ext[y];
        if (META_ISEL(w.meta[j]) != isel) continue;
        if (fabs(dpl - w.dplane[j]) / fabs(dpl) < (double)0.001f) S.cand[3 * (size_t)x + n++] = y;
      }
    }
    S.ncand[x] = n;
    S.nxt[x] = first;                 // chain of the entries of this unit with the same parent pixel
    if (first >= 0) S.hasprev[first] = 1;
  }
}

// Leader resolution in list order (ray_list.f90:222-226, 268, 290) and the tasks / re-queued (ray_type_reco) entries each
// extracted entry produces.  An entry is a leader unless an earlier leader already consumed it.  Members always share the
// leader's parent pixel, so the entries of one (unit, parent pixel) chain are independent of every other chain: one thread
// walks each chain in list order (the head is the entry no earlier entry links to).
//   HIGH-mode entry: 4 child rays, or 1 for a ray_type_reco entry (rt_routines.f90:1389-1403)
//   LOW-mode leader: 1 merged ray when it has 3 siblings, else the group is re-queued on HIGH (ray_list.f90:255-291)
__device__ __forceinline__ void ph_resolve(const DreStep& S, int cur) {
  const DreList h = S.H[cur];
  const DreUnits u = S.u;
  GRID_LOOP(x0, S.ext_seg[S.ctl->U]) {
    const int md = u.mode[S.unit_of_ext[x0]];
    if (md == 1) { S.ntask[x0] = (META_TYPE(h.meta[S.ext[x0]]) == RT_RECO) ? 1 : 4; S.nreco[x0] = 0; continue; }
    if (md != 2) { S.ntask[x0] = 0; S.nreco[x0] = 0; continue; }
    if (S.hasprev[x0]) continue;
    for (int x = x0; x >= 0; x = S.nxt[x]) {
      if (S.gone[x]) { S.ntask[x] = 0; S.nreco[x] = 0; continue; }
      const int nc = S.ncand[x];
      for (int m = 0; m < nc; m++) S.gone[S.cand[3 * (size_t)x + m]] = 1;
      if (nc == 3) { S.ntask[x] = 1; S.nreco[x] = 0; } else { S.ntask[x] = 0; S.nreco[x] = 1 + nc; }
    }
  }
}

// Per-unit task segments (ray ordinal = task index - task_seg[unit]) and reco counts, the capacity check of the step, and
// the tasks themselves: one per ray_tracing call, in the reference's processing order.
__device__ __forceinline__ void ph_tasks(const DreStep& S, int cur) {
  const DreUnits u = S.u;
  const int U = S.ctl->U, nExt = S.ext_seg[U], nT = S.task_off[nExt], nR = S.reco_off[nExt];
  const int nH = S.H[cur].seg[U], nW = S.W[cur].seg[U];
  // worst case of the rebuilt lists: HIGH' = kept ++ reco ++ outcomes, LOW' = kept ++ outcomes, outcomes <= tasks
  const int needH = (nH - S.posH[nH]) + nR + nT, needW = (nW - S.posW[nW]) + nT;
  const bool ab = nT > S.capT || needH > S.capH || needW > S.capW;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    DreCtl* c = S.ctl;
    c->nExt = nExt; c->nT = ab ? 0 : nT; c->needH = needH; c->needW = needW; c->needT = nT; c->needE = 0; c->abort = ab ? 1 : 0;
  }
  if (ab) return;
  GRID_LOOP(un, U + 1) {
    u.task_seg[un] = S.task_off[S.ext_seg[un]];
    if (un < U) S.reco_cnt[un] = S.reco_off[S.ext_seg[un + 1]] - S.reco_off[S.ext_seg[un]];
  }
  GRID_LOOP(x, nExt) {
    const int nt = S.ntask[x];
    if (nt == 0) continue;
    const int un = S.unit_of_ext[x];
    for (int s = 0; s < nt; s++) {
      DreTask t; t.unit = un; t.e0 = S.ext[x]; t.e1 = t.e2 = t.e3 = -1; t.sub = s; t.pad = 0;
      if (u.mode[un] == 1) t.kind = (nt == 4) ? 0 : 1;
      else {
        t.kind = 2;
        t.e1 = S.ext[S.cand[3 * (size_t)x + 0]]; t.e2 = S.ext[S.cand[3 * (size_t)x + 1]]; t.e3 = S.ext[S.cand[3 * (size_t)x + 2]];
      }
      S.tasks[S.task_off[x] + s] = t;
    }
  }
}

__device__ __forceinline__ void ph_out_flags(const DreStep& S) {
  if (S.ctl->abort) return;
  GRID_LOOP(t, S.ctl->nT) {
    const int ty = META_TYPE(S.OUT.meta[t]);
    S.fH[t] = (ty == RT_HIGH) ? 1 : 0;
    S.fW[t] = (ty == RT_LOW) ? 1 : 0;
  }
}

// sizes of the rebuilt lists: HIGH' = kept ++ reco ++ go_high outcomes ; LOW' = kept ++ go_low outcomes
__device__ __forceinline__ void ph_unit_new(const DreStep& S, int cur) {
  if (S.ctl->abort) return;
  const DreList H = S.H[cur], W = S.W[cur];
  const DreUnits u = S.u;
  GRID_LOOP(i, S.ctl->U) {
    const int kh = (H.seg[i + 1] - H.seg[i]) - (S.posH[H.seg[i + 1]] - S.posH[H.seg[i]]);
    const int kw = (W.seg[i + 1] - W.seg[i]) - (S.posW[W.seg[i + 1]] - S.posW[W.seg[i]]);
    S.keptH[i] = kh; S.keptW[i] = kw;
    const int oh = S.posOH[u.task_seg[i + 1]] - S.posOH[u.task_seg[i]];
    const int ow = S.posOW[u.task_seg[i + 1]] - S.posOW[u.task_seg[i]];
    S.newHcnt[i] = kh + S.reco_cnt[i] + oh; S.newWcnt[i] = kw + ow;
  }
}

__device__ __forceinline__ void dre_copy_entry_hdr(const DreList& d, int di, const DreList& s, int si, int un) {
  d.unit[di] = un; d.dir_ipix[di] = s.dir_ipix[si]; d.meta[di] = s.meta[si]; d.cc_old[di] = s.cc_old[si];
  d.dplane[di] = s.dplane[si]; d.prev[di] = s.prev[si];
}

// destination slots of the four kinds of entries of the rebuilt lists (-1: the entry is not copied)
__device__ __forceinline__ int dre_dst_kept(const DreList& dst, const DreList& src, const int* sel, const int* pos, int i) {
  if (sel[i]) return -1;
  const int un = src.unit[i], s0 = src.seg[un];
  return dst.seg[un] + (i - s0) - (pos[i] - pos[s0]);          // kept entries of an old list (stable)
}
__device__ __forceinline__ int dre_dst_out(const DreStep& S, const DreList& Hn, const DreList& Wn, int t, int& toH) {
  const int ty = META_TYPE(S.OUT.meta[t]);
  if (ty != RT_HIGH && ty != RT_LOW) return -1;
  const int un = S.tasks[t].unit, t0 = S.u.task_seg[un];
  toH = (ty == RT_HIGH);
  return toH ? Hn.seg[un] + S.keptH[un] + S.reco_cnt[un] + (S.posOH[t] - S.posOH[t0])
             : Wn.seg[un] + S.keptW[un] + (S.posOW[t] - S.posOW[t0]);
}

// Rebuild of both lists in one launch: kept entries (stable), failed merges re-queued on HIGH as ray_type_reco with the
// current nside (ray_list.f90:276-291, add_to_high_ray_list :463-521), and the go_high / go_low outcomes of the ray-march
// kernel in task (= processing) order.  A group of G consecutive lanes moves one entry: its destination slot is looked up
// once, lane 0 copies the header and the lanes copy the luminosity row as 16-byte pieces (two entries per round, so that
// the two index chains overlap).
__device__ __forceinline__ void dre_copy_lum(double* __restrict__ d, const double* __restrict__ s, int L, int sub, int G) {
  if ((L & 1) == 0) {
    const double2* s2 = reinterpret_cast<const double2*>(s); double2* d2 = reinterpret_cast<double2*>(d);
    for (int l = sub; l < (L >> 1); l += G) d2[l] = s2[l];
  } else for (int l = sub; l < L; l += G) d[l] = s[l];
}
__device__ __forceinline__ void ph_build(const DreStep& S, int cur, int L) {
  if (S.ctl->abort) return;
  const DreList Hc = S.H[cur], Wc = S.W[cur], Hn = S.H[1 - cur], Wn = S.W[1 - cur];
  const DreUnits u = S.u;
  const int U = S.ctl->U, nH = Hc.seg[U], nW = Wc.seg[U], nExt = S.ext_seg[U], nT = S.ctl->nT;
  const int Lv = (L & 1) ? L : (L >> 1);
  const int G = Lv <= 4 ? 4 : (Lv <= 8 ? 8 : (Lv <= 16 ? 16 : 32));
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int sub = tid & (G - 1), gid = tid / G, ng = (gridDim.x * blockDim.x) / G;
  // ---- kept entries of the two old lists
  for (int w = 0; w < 2; w++) {
    const DreList& src = w ? Wc : Hc; const DreList& dst = w ? Wn : Hn;
    const int* sel = w ? S.selW : S.selH; const int* pos = w ? S.posW : S.posH;
    const int n = w ? nW : nH;
    for (int i = gid; i < n; i += 2 * ng) {
      const int i1 = i + ng;
      const int d0 = dre_dst_kept(dst, src, sel, pos, i);
      const int d1 = (i1 < n) ? dre_dst_kept(dst, src, sel, pos, i1) : -1;
      if (d0 >= 0) { if (sub == 0) dre_copy_entry_hdr(dst, d0, src, i, src.unit[i]); dre_copy_lum(dst.lum + (size_t)d0 * L, src.lum + (size_t)i * L, L, sub, G); }
      if (d1 >= 0) { if (sub == 0) dre_copy_entry_hdr(dst, d1, src, i1, src.unit[i1]); dre_copy_lum(dst.lum + (size_t)d1 * L, src.lum + (size_t)i1 * L, L, sub, G); }
    }
  }
  // ---- groups whose merge failed: leader and members go to HIGH as ray_type_reco
  for (int x = gid; x < nExt; x += ng) {
    const int nr = S.nreco[x];
    if (nr == 0) continue;
    const int un = S.unit_of_ext[x];
    const int base = Hn.seg[un] + S.keptH[un] + (S.reco_off[x] - S.reco_off[S.ext_seg[un]]);
    for (int m = 0; m < nr; m++) {
      const int si = (m == 0) ? S.ext[x] : S.ext[S.cand[3 * (size_t)x + m - 1]];
      const int di = base + m;
      if (sub == 0) {
        dre_copy_entry_hdr(Hn, di, Wc, si, un);
        const unsigned mt = Wc.meta[si];
        const int isel = META_ISEL(mt);
        Hn.meta[di] = dre_meta(META_DIRK(mt), u.nk[un], isel, RT_RECO);
        if (isel == 0) { Hn.dplane[di] = 0.; Hn.prev[di] = 0.; }
      }
      dre_copy_lum(Hn.lum + (size_t)di * L, Wc.lum + (size_t)si * L, L, sub, G);
    }
  }
  // ---- outcomes of the ray-march kernel
  for (int t = gid; t < nT; t += 2 * ng) {
    const int t1 = t + ng;
    int h0 = 0, h1 = 0;
    const int d0 = dre_dst_out(S, Hn, Wn, t, h0);
    const int d1 = (t1 < nT) ? dre_dst_out(S, Hn, Wn, t1, h1) : -1;
    if (d0 >= 0) { const DreList& dst = h0 ? Hn : Wn; if (sub == 0) dre_copy_entry_hdr(dst, d0, S.OUT, t, S.tasks[t].unit); dre_copy_lum(dst.lum + (size_t)d0 * L, S.OUT.lum + (size_t)t * L, L, sub, G); }
    if (d1 >= 0) { const DreList& dst = h1 ? Hn : Wn; if (sub == 0) dre_copy_entry_hdr(dst, d1, S.OUT, t1, S.tasks[t1].unit); dre_copy_lum(dst.lum + (size_t)d1 * L, S.OUT.lum + (size_t)t1 * L, L, sub, G); }
  }
}

// define_next_level rt_routines.f90:2745-2770 on the rebuilt lists; the last kernel of a step
__device__ __forceinline__ void ph_next(const DreStep& S, int cur) {
  DreCtl* c = S.ctl;
  if (blockIdx.x == 0 && threadIdx.x == 0) c->scan_epoch += 16u;
  if (c->abort) return;
  const DreList Hn = S.H[1 - cur], Wn = S.W[1 - cur];
  const DreUnits u = S.u;
  const int U = c->U;
  if (blockIdx.x == 0 && threadIdx.x == 0) { c->nH2 = Hn.seg[U]; c->nW2 = Wn.seg[U]; }
  GRID_LOOP(i, U) {
    if (u.mode[i] == 0) continue;
    u.step[i]++;
    atomicAdd(S.steps, 1ull);
    const int nh = Hn.seg[i + 1] - Hn.seg[i], nw = Wn.seg[i + 1] - Wn.seg[i];
    if (nh > 0) u.nk[i]++; else if (nw > 0) u.nk[i]--; else continue;
    if (u.nk[i] > S.kmax) { atomicCAS(S.err, 0, 4); continue; }
    if (u.nk[i] < 1) { atomicCAS(S.err, 0, 5); continue; }
    atomicAdd(&c->active, 1);
    atomicOr(&c->nkmask, 1 << u.nk[i]);
  }
}

// ---- one kernel per phase (the launch sequence of enqueue_pre / enqueue_post, replayed from CUDA graphs)
__global__ void __launch_bounds__(STEP_THREADS) k_step_select(DreStep S, int cur) { ph_select(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_unit_ext(DreStep S, int cur) { ph_unit_ext(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_fill_ext(DreStep S, int cur) { ph_fill_ext(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_cand(DreStep S, int cur) { ph_cand(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_resolve(DreStep S, int cur) { ph_resolve(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_tasks(DreStep S, int cur) { ph_tasks(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_out_flags(DreStep S) { ph_out_flags(S); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_unit_new(DreStep S, int cur) { ph_unit_new(S, cur); }
__global__ void __launch_bounds__(STEP_THREADS, 6) k_step_build(DreStep S, int cur, int L) { ph_build(S, cur, L); }
__global__ void __launch_bounds__(STEP_THREADS) k_step_next(DreStep S, int cur) { ph_next(S, cur); }
