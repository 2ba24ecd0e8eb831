// dre_lists.cuh — wavefront form of the per-sector ray lists (ray_list.f90, main_dir_loop
// rt_routines.f90:1343-1467, define_next_level :2745-2770).
//
// Every work unit (source, sector) owns a HIGH and a LOW list, stored unit-grouped in two
// global SoA arrays.  One resolution step of ALL units of a batch is a fixed sequence of
// data-parallel kernels: select -> scan -> gather (extract_ray_list), sibling search + leader
// resolution (the 4-ray merge), task emission, the ray-march kernel, and a stable rebuild of
// both lists.  Prefix sums keep the reference's list order (append order == processing order),
// which the merge and therefore the intersection lists depend on.
#pragma once
#include "dre_common.cuh"

// ---------------------------------------------------------------------------------------
// exclusive scan: out[0..n] (n+1 values, out[n] = total)
// ---------------------------------------------------------------------------------------
#define SCAN_THREADS 512
#define SCAN_ITEMS 4
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__global__ void k_scan_tile(const int* __restrict__ in, int* __restrict__ out, int* __restrict__ sums, int n) {
  __shared__ int wsum[SCAN_THREADS / 32];
  const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS], tot = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = (base + i < n) ? in[base + i] : 0; tot += v[i]; }
  int x = tot;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) wsum[w] = x;
  __syncthreads();
  if (w == 0) {
    int s = (lane < SCAN_THREADS / 32) ? wsum[lane] : 0;
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
    if (lane < SCAN_THREADS / 32) wsum[lane] = s;
  }
  __syncthreads();
  int excl = x - tot + (w > 0 ? wsum[w - 1] : 0);
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) out[base + i] = excl; excl += v[i]; }
  if (threadIdx.x == SCAN_THREADS - 1) sums[blockIdx.x] = excl;
}
// n <= SCAN_TILE: the whole scan (and the total out[n]) in one launch
__global__ void k_scan_small(const int* __restrict__ in, int* __restrict__ out, int n) {
  __shared__ int wsum[SCAN_THREADS / 32];
  const int base = threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS], tot = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = (base + i < n) ? in[base + i] : 0; tot += v[i]; }
  int x = tot;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) wsum[w] = x;
  __syncthreads();
  if (w == 0) {
    int s = (lane < SCAN_THREADS / 32) ? wsum[lane] : 0;
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
    if (lane < SCAN_THREADS / 32) wsum[lane] = s;
  }
  __syncthreads();
  int excl = x - tot + (w > 0 ? wsum[w - 1] : 0);
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) out[base + i] = excl; excl += v[i]; }
  if (threadIdx.x == SCAN_THREADS - 1) out[n] = excl;
}
// Any n in ONE launch: chained scan with decoupled look-back.  Tile t publishes {epoch, state, value} as one 64-bit word
// (state 1 = tile aggregate, 2 = inclusive prefix); a tile adds up the aggregates of its predecessors until it meets an
// inclusive prefix.  The epoch (one per call, 30 bits) makes the words of earlier scans on the same descriptor array read
// as "not published", so the array is never cleared.  Tiles are taken in blockIdx order: CUDA dispatches the blocks of
// a grid in index order, so every predecessor of a resident block has been dispatched and the wait terminates.
#define SCAN_DESC(epoch, state, val) (((unsigned long long)(epoch) << 34) | ((unsigned long long)(state) << 32) | (unsigned long long)(unsigned)(val))
__global__ void k_scan_chained(const int* __restrict__ in, int* __restrict__ out, int n, unsigned long long* desc, unsigned epoch) {
  __shared__ int wsum[SCAN_THREADS / 32];
  __shared__ int s_prefix;
  const int tile = blockIdx.x;
  const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS], tot = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = (base + i < n) ? in[base + i] : 0; tot += v[i]; }
  int x = tot;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
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
    volatile unsigned long long* d = desc;
    int prefix = 0;
    if (tile > 0) {
      if (lane == 0) d[tile] = SCAN_DESC(epoch, 1u, agg);
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
  for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) out[base + i] = excl; excl += v[i]; }
  if (tile == gridDim.x - 1 && threadIdx.x == SCAN_THREADS - 1) out[n] = excl;
}
__global__ void k_scan_add(int* __restrict__ out, const int* __restrict__ sums_scanned, int n, int nblocks) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] += sums_scanned[i / SCAN_TILE];
  if (i == 0) out[n] = sums_scanned[nblocks];
}
__global__ void k_set_int(int* p, int v) { *p = v; }

// ---------------------------------------------------------------------------------------
// step kernels
// ---------------------------------------------------------------------------------------
// mode of each unit: extract from HIGH if it has entries, else from LOW (ray_list.f90:120,167)
__global__ void k_plan(DreUnits u, const int* __restrict__ segH, const int* __restrict__ segW) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= u.U) return;
  const int nh = segH[i + 1] - segH[i], nw = segW[i + 1] - segW[i];
  u.mode[i] = nh > 0 ? 1 : (nw > 0 ? 2 : 0);
}

// which == 1: HIGH entries with nside == nside/2 ; which == 2: LOW entries with nside == 2*nside
__global__ void k_select(DreList l, DreUnits u, int which, int* __restrict__ sel) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= l.n) return;
  const int un = l.unit[i];
  const int want = (which == 1) ? u.nk[un] - 1 : u.nk[un] + 1;
  sel[i] = (u.mode[un] == which && META_NSK(l.meta[i]) == want) ? 1 : 0;
}

__global__ void k_unit_ext(DreUnits u, const int* __restrict__ segH, const int* __restrict__ segW,
                           const int* __restrict__ posH, const int* __restrict__ posW, int nH, int nW,
                           int* __restrict__ extcnt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= u.U) return;
  int c = 0;
  if (u.mode[i] == 1 && nH > 0) c = posH[segH[i + 1]] - posH[segH[i]];
  else if (u.mode[i] == 2 && nW > 0) c = posW[segW[i + 1]] - posW[segW[i]];
  extcnt[i] = c;
}

// EXT[x] = index of the extracted entry in its old list; for LOW-mode units also the parent
// pixel at the current nside (ang2pix_nest(nside, theta, phi), ray_list.f90:218) and gone = 0.
__global__ void k_fill_ext(DreList l, DreUnits u, int which, const int* __restrict__ sel, const int* __restrict__ pos,
                           const int* __restrict__ ext_seg, int* __restrict__ ext, int* __restrict__ ipx,
                           int* __restrict__ gone, int* __restrict__ hasprev) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= l.n || !sel[i]) return;
  const int un = l.unit[i];
  const int x = ext_seg[un] + (pos[i] - pos[l.seg[un]]);
  ext[x] = i;
  gone[x] = 0; hasprev[x] = 0;
  if (which == 2) {
    const unsigned m = l.meta[i];
    ipx[x] = l.dir_ipix[i] >> (2 * (META_DIRK(m) - u.nk[un]));
  } else ipx[x] = -1;
}

// sibling search of extract_ray_list (ray_list.f90:237-253): the first <= 3 later entries of the
// same unit with the same parent pixel, the same isel and dplane within 0.001 (relative to the
// leader).  `gone` is NOT consulted for the members, as in the reference (SURVEY A.3 note).
__global__ void k_cand(DreList w, DreUnits u, const int* __restrict__ ext_seg, const int* __restrict__ ext,
                       const int* __restrict__ unit_of_ext, const int* __restrict__ ipx, int next,
                       int* __restrict__ cand, int* __restrict__ ncand, int* __restrict__ nxt, int* __restrict__ hasprev) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= next) return;
  const int un = unit_of_ext[x];
  if (u.mode[un] != 2) { ncand[x] = 0; nxt[x] = -1; return; }
  const int i = ext[x];
  const int end = ext_seg[un + 1];
  const int ip = ipx[x], isel = META_ISEL(w.meta[i]);
  const double dpl = w.dplane[i];
  int n = 0, first = -1;
  for (int y = x + 1; y < end && (n < 3 || first < 0); y++) {
    if (ipx[y] != ip) continue;
    if (first < 0) first = y;
    if (n == 3) break;
    const int j = ext[y];
    if (META_ISEL(w.meta[j]) != isel) continue;
    if (fabs(dpl - w.dplane[j]) / fabs(dpl) < (double)0.001f) cand[3 * (size_t)x + n++] = y;
  }
  ncand[x] = n;
  nxt[x] = first;                 // chain of the entries of this unit with the same parent pixel
  if (first >= 0) hasprev[first] = 1;
}

// leader resolution in list order (ray_list.f90:222-226, 268, 290): an entry is a leader unless an
// earlier leader already consumed it.  Members always share the leader's parent pixel, so the entries
// of one (unit, parent pixel) chain are independent of every other chain: one thread walks each chain
// in list order (the head is the entry no earlier entry links to).
__global__ void k_resolve(DreUnits u, const int* __restrict__ unit_of_ext, int next, const int* __restrict__ cand,
                          const int* __restrict__ ncand, const int* __restrict__ nxt, const int* __restrict__ hasprev,
                          int* __restrict__ gone, int* __restrict__ leader) {
  const int x0 = blockIdx.x * blockDim.x + threadIdx.x;
  if (x0 >= next || u.mode[unit_of_ext[x0]] != 2 || hasprev[x0]) return;
  for (int x = x0; x >= 0; x = nxt[x]) {
    if (gone[x]) { leader[x] = 0; continue; }
    leader[x] = 1;
    for (int m = 0; m < ncand[x]; m++) gone[cand[3 * (size_t)x + m]] = 1;
  }
}

// tasks / re-queued (ray_type_reco) entries produced by each extracted entry
__global__ void k_count(DreList h, DreUnits u, const int* __restrict__ ext_seg, const int* __restrict__ ext, int next,
                        const int* __restrict__ unit_of_ext, const int* __restrict__ ncand, const int* __restrict__ leader,
                        int* __restrict__ ntask, int* __restrict__ nreco) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= next) return;
  const int un = unit_of_ext[x];
  int nt = 0, nr = 0;
  if (u.mode[un] == 1) nt = (META_TYPE(h.meta[ext[x]]) == RT_RECO) ? 1 : 4;     // rt_routines.f90:1389-1403
  else if (leader[x]) { if (ncand[x] == 3) nt = 1; else nr = 1 + ncand[x]; }      // ray_list.f90:255-291
  ntask[x] = nt; nreco[x] = nr;
}

__global__ void k_unit_of_ext(DreUnits u, const int* __restrict__ ext_seg, int* __restrict__ unit_of_ext) {
  const int un = blockIdx.x * blockDim.x + threadIdx.x;
  if (un >= u.U) return;
  for (int x = ext_seg[un]; x < ext_seg[un + 1]; x++) unit_of_ext[x] = un;
}

__global__ void k_emit_tasks(DreUnits u, const int* __restrict__ ext, int next, const int* __restrict__ unit_of_ext,
                             const int* __restrict__ cand, const int* __restrict__ ntask, const int* __restrict__ task_off,
                             DreTask* __restrict__ tasks) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= next) return;
  const int nt = ntask[x];
  if (nt == 0) return;
  const int un = unit_of_ext[x];
  for (int s = 0; s < nt; s++) {
    DreTask t; t.unit = un; t.e0 = ext[x]; t.e1 = t.e2 = t.e3 = -1; t.sub = s; t.pad = 0;
    if (u.mode[un] == 1) t.kind = (nt == 4) ? 0 : 1;
    else {
      t.kind = 2;
      t.e1 = ext[cand[3 * (size_t)x + 0]]; t.e2 = ext[cand[3 * (size_t)x + 1]]; t.e3 = ext[cand[3 * (size_t)x + 2]];
    }
    tasks[task_off[x] + s] = t;
  }
}

// per-unit task segment (ray ordinal = task index - task_seg[unit]) and reco count
__global__ void k_unit_tasks(DreUnits u, const int* __restrict__ ext_seg, const int* __restrict__ task_off,
                             const int* __restrict__ reco_off, int* __restrict__ reco_cnt) {
  const int un = blockIdx.x * blockDim.x + threadIdx.x;
  if (un > u.U) return;
  if (un == u.U) { u.task_seg[un] = task_off[ext_seg[un]]; return; }
  u.task_seg[un] = task_off[ext_seg[un]];
  reco_cnt[un] = reco_off[ext_seg[un + 1]] - reco_off[ext_seg[un]];
}

__global__ void k_out_flags(DreList out, int nt, int* __restrict__ fH, int* __restrict__ fW) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt) return;
  const int ty = META_TYPE(out.meta[t]);
  fH[t] = (ty == RT_HIGH) ? 1 : 0;
  fW[t] = (ty == RT_LOW) ? 1 : 0;
}

// sizes of the rebuilt lists: HIGH' = kept ++ reco ++ go_high outcomes ; LOW' = kept ++ go_low outcomes
__global__ void k_unit_new(DreUnits u, const int* __restrict__ segH, const int* __restrict__ segW,
                           const int* __restrict__ posH, const int* __restrict__ posW, int nH, int nW,
                           const int* __restrict__ reco_cnt, const int* __restrict__ posOH, const int* __restrict__ posOW,
                           int have_ext, int have_tasks, int* __restrict__ keptH, int* __restrict__ keptW,
                           int* __restrict__ newHcnt, int* __restrict__ newWcnt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= u.U) return;
  int kh = segH[i + 1] - segH[i], kw = segW[i + 1] - segW[i];
  if (nH > 0) kh -= posH[segH[i + 1]] - posH[segH[i]];
  if (nW > 0) kw -= posW[segW[i + 1]] - posW[segW[i]];
  keptH[i] = kh; keptW[i] = kw;
  int oh = 0, ow = 0, rc = 0;
  if (have_tasks) {
    oh = posOH[u.task_seg[i + 1]] - posOH[u.task_seg[i]];
    ow = posOW[u.task_seg[i + 1]] - posOW[u.task_seg[i]];
  }
  if (have_ext) rc = reco_cnt[i];
  newHcnt[i] = kh + rc + oh; newWcnt[i] = kw + ow;
}

__device__ __forceinline__ void dre_copy_entry(const DreList& d, int di, const DreList& s, int si, int L, int un) {
  d.unit[di] = un; d.dir_ipix[di] = s.dir_ipix[si]; d.meta[di] = s.meta[si]; d.cc_old[di] = s.cc_old[si];
  d.dplane[di] = s.dplane[si]; d.prev[di] = s.prev[si];
  const double* a = s.lum + (size_t)si * L; double* b = d.lum + (size_t)di * L;
  for (int l = 0; l < L; l++) b[l] = a[l];
}

// kept entries of an old list (stable)
__global__ void k_build_kept(DreList dst, DreList src, const int* __restrict__ sel, const int* __restrict__ pos, int L) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= src.n || sel[i]) return;
  const int un = src.unit[i];
  const int s0 = src.seg[un];
  const int rank = (i - s0) - (pos[i] - pos[s0]);
  dre_copy_entry(dst, dst.seg[un] + rank, src, i, L, un);
}
// failed merges re-queued on HIGH as ray_type_reco with the current nside (ray_list.f90:276-291,
// add_to_high_ray_list :463-521)
__global__ void k_build_reco(DreList dstH, DreList srcW, DreUnits u, const int* __restrict__ ext, int next,
                             const int* __restrict__ unit_of_ext, const int* __restrict__ ext_seg,
                             const int* __restrict__ cand, const int* __restrict__ nreco, const int* __restrict__ reco_off,
                             const int* __restrict__ keptH, int L) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= next) return;
  const int nr = nreco[x];
  if (nr == 0) return;
  const int un = unit_of_ext[x];
  const int base = dstH.seg[un] + keptH[un] + (reco_off[x] - reco_off[ext_seg[un]]);
  for (int m = 0; m < nr; m++) {
    const int si = (m == 0) ? ext[x] : ext[cand[3 * (size_t)x + m - 1]];
    const int di = base + m;
    dre_copy_entry(dstH, di, srcW, si, L, un);
    const unsigned mt = srcW.meta[si];
    const int isel = META_ISEL(mt);
    dstH.meta[di] = dre_meta(META_DIRK(mt), u.nk_prev[un], isel, RT_RECO);
    if (isel == 0) { dstH.dplane[di] = 0.; dstH.prev[di] = 0.; }
  }
}
// go_high / go_low outcomes of the ray-march kernel, in task (= processing) order
__global__ void k_build_out(DreList dstH, DreList dstW, DreList out, int nt, DreUnits u, const DreTask* __restrict__ tasks,
                            const int* __restrict__ posOH, const int* __restrict__ posOW, const int* __restrict__ keptH,
                            const int* __restrict__ keptW, const int* __restrict__ reco_cnt, int L) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt) return;
  const int ty = META_TYPE(out.meta[t]);
  if (ty != RT_HIGH && ty != RT_LOW) return;
  const int un = tasks[t].unit;
  const int t0 = u.task_seg[un];
  if (ty == RT_HIGH) dre_copy_entry(dstH, dstH.seg[un] + keptH[un] + reco_cnt[un] + (posOH[t] - posOH[t0]), out, t, L, un);
  else dre_copy_entry(dstW, dstW.seg[un] + keptW[un] + (posOW[t] - posOW[t0]), out, t, L, un);
}

// define_next_level rt_routines.f90:2745-2770 ; *active counts units that still have rays
__global__ void k_next(DreUnits u, const int* __restrict__ segH, const int* __restrict__ segW, int* __restrict__ active,
                       int* __restrict__ err, int kmax, unsigned long long* __restrict__ steps) {
  // active[0] = units that still have rays, active[1] = bit mask of their log2(nside)
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= u.U) return;
  u.nk_prev[i] = u.nk[i];
  if (u.mode[i] == 0) return;
  u.step[i]++;
  atomicAdd(steps, 1ull);
  const int nh = segH[i + 1] - segH[i], nw = segW[i + 1] - segW[i];
  if (nh > 0) u.nk[i]++; else if (nw > 0) u.nk[i]--; else return;
  if (u.nk[i] > kmax) { atomicCAS(err, 0, 4); return; }
  if (u.nk[i] < 1) { atomicCAS(err, 0, 5); return; }
  atomicAdd(active, 1);
  atomicOr(active + 1, 1 << u.nk[i]);
}
