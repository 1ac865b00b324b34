// dp.cu — banded integer dovetail aligner: G lanes per (A,B) pair, anti-diagonal wavefront
// in registers, neighbours through __shfl_*_sync, DPX max/add-max instructions.
// The math is csrc/dp_core.cuh (= generateFastDovetailAlignmentSet, src/BioLibs.scala:596-822,
// Alignment.valid / Overlap, src/ObjectStore.scala:89-142); this file is the machinery:
//   * persistent grid, warps fetch 32/G pairs at a time from a global counter;
//   * each lane slides 32-bit windows over the 2-bit packed reads (one refill per 16 rows),
//     so the inner loops touch memory only for the 16-entry cost table in shared memory;
//   * pairs are binned by band width into (G,D) classes so that G*D >= width+1 with little
//     padding; a class is one template instantiation.
#include "engine.cuh"
#include <algorithm>
#include "dp_core.cuh"
#include "dp_tp2.cuh"
#include "radix_sort.cuh"

namespace hovl {

constexpr int DP_THREADS = 256;

struct DpArgs {
  const uint32_t *packed; const uint64_t *woff; const uint32_t *len; const uint32_t *bad;
  const uint32_t *pa, *pb, *pcnt;       // pair arrays (1-based ids)
  const uint32_t *index;                // pairs of this class (nullptr = all pairs 0..np)
  uint32_t np;                          // pairs in this class
  hovl_record *out;
  unsigned long long *cells; uint32_t *counter;
  int32_t cost[16]; int32_t gO, gE, k, min_overlap; uint32_t min_identity, max_ignore;
};

// 16 bases starting at `pos`, first base in the top bits; pos may run off either end
struct Window {
  const uint32_t *w; int32_t nw, pos; uint32_t w0, w1, w2;
  __device__ __forceinline__ uint32_t word(int32_t i) const { return (i >= 0 && i < nw) ? __ldg(w + i) : 0u; }
  __device__ __forceinline__ void init(const uint32_t *words, int32_t nwords, int32_t p) {
    w = words; nw = nwords; pos = p;
    const int32_t wi = p >> 4;
    w0 = word(wi); w1 = word(wi + 1); w2 = word(wi + 2);
  }
  __device__ __forceinline__ uint32_t bits() const { return __funnelshift_l(w1, w0, 2 * (pos & 15)); }
  __device__ __forceinline__ void advance() {
    pos++;
    if ((pos & 15) == 0) { w0 = w1; w1 = w2; w2 = word((pos >> 4) + 2); }
  }
};

template <int G>
__device__ __forceinline__ Edge shfl_up_edge(const Edge &e) {
  Edge r;
  r.h = __shfl_up_sync(0xffffffffu, e.h, 1, G);
  r.t1 = __shfl_up_sync(0xffffffffu, e.t1, 1, G);
  r.t2 = __shfl_up_sync(0xffffffffu, e.t2, 1, G);
  return r;
}
template <int G>
__device__ __forceinline__ Edge shfl_down_edge(const Edge &e) {
  Edge r;
  r.h = __shfl_down_sync(0xffffffffu, e.h, 1, G);
  r.t1 = __shfl_down_sync(0xffffffffu, e.t1, 1, G);
  r.t2 = __shfl_down_sync(0xffffffffu, e.t2, 1, G);
  return r;
}

template <int G, int D>
__global__ void __launch_bounds__(DP_THREADS) dp_kernel(const DpArgs P) {
  static_assert(D >= 2 && D <= 8 && (G == 8 || G == 16 || G == 32), "class");
  __shared__ int32_t s_cost[16];
  if (threadIdx.x < 16) s_cost[threadIdx.x] = P.cost[threadIdx.x];
  __syncthreads();
  DpConst K; K.cost = s_cost; K.gO = P.gO; K.gE = P.gE;
  constexpr int GPW = 32 / G;                       // pairs per warp
  const int lane32 = threadIdx.x & 31;
  const int lane = lane32 & (G - 1);                // lane inside the group
  const int grp = lane32 / G;
  unsigned long long my_cells = 0, my_p1 = 0;

  for (;;) {
    uint32_t base = 0;
    if (lane32 == 0) base = atomicAdd(P.counter, (uint32_t)GPW);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= P.np) break;
    const uint32_t slot = base + grp;
    const bool have = slot < P.np;
    const uint32_t pidx = have ? (P.index ? P.index[slot] : slot) : 0u;
    int32_t n = 0, m = 0, width = 1; const uint32_t *wa = nullptr, *wb = nullptr; uint32_t ida = 0, idb = 0;
    int extra = 0; bool run = false;
    if (have) {
      ida = P.pa[pidx]; idb = P.pb[pidx];
      n = (int32_t)P.len[ida - 1]; m = (int32_t)P.len[idb - 1];
      wa = P.packed + P.woff[ida - 1]; wb = P.packed + P.woff[idb - 1];
      width = band_width(n, P.k, P.min_identity);
      if (P.bad[ida - 1] | P.bad[idb - 1]) extra = HOVL_CORE_F_BAD_BASE;       // MatchError in costFunc
      else if (m < width) extra = HOVL_CORE_F_SHORT_B;                          // charAt past the end of B
      run = (extra == 0) && (width + 1 <= G * D);
    }
    const int32_t nwa = run ? (n + 15) >> 4 : 0, nwb = run ? (m + 15) >> 4 : 0;
    DpOut R; dp_set_dud(R, extra);

    // ------------------------------ phase 1 ------------------------------
    int32_t ds = 0; bool ok1 = false;
    {
      P1Lane<D> L; p1_init<D>(L, lane);
      uint32_t bcol[D];
#pragma unroll
      for (int c = 0; c < D; c++) { const int32_t j0 = lane * D + c; bcol[c] = (run && j0 < m) ? packed_base(wb, j0) : 0u; }
      int32_t o_hx = 0, o_hm = 0; uint32_t o_t = (uint32_t)(lane * D + D);
      Window aw; aw.init(wa, nwa, -lane);                       // step s reads A[s - lane - 1]
      const int32_t steps = __reduce_max_sync(0xffffffffu, run ? n + G - 1 : 0);
      const int32_t nn = run ? n : 0;
      for (int32_t s = 1; s <= steps; s++) {
        const int32_t i = s - lane;
        int32_t in_hx = __shfl_up_sync(0xffffffffu, o_hx, 1, G);
        int32_t in_hm = __shfl_up_sync(0xffffffffu, o_hm, 1, G);
        uint32_t in_t = __shfl_up_sync(0xffffffffu, o_t, 1, G);
        if (lane == 0) { in_hx = 0; in_hm = 0; in_t = (uint32_t)i << 8; }
        const uint32_t a = aw.bits() >> 30;
        p1_row<D>(L, K, lane, i, nn, width, a, bcol, in_hx, in_hm, in_t, o_hx, o_hm, o_t);
        aw.advance();
      }
      // arg-max over the group: largest value, then smallest (i, j)  [strict '>' in row-major order]
      int32_t best = L.best, bi = L.best_i, bj = L.best_j; uint32_t bt = L.best_t;
#pragma unroll
      for (int o = G / 2; o; o >>= 1) {
        const int32_t ob = __shfl_xor_sync(0xffffffffu, best, o, G), oi = __shfl_xor_sync(0xffffffffu, bi, o, G),
                      oj = __shfl_xor_sync(0xffffffffu, bj, o, G);
        const uint32_t ot = __shfl_xor_sync(0xffffffffu, bt, o, G);
        if (ob > best || (ob == best && (oi < bi || (oi == bi && oj < bj)))) { best = ob; bi = oi; bj = oj; bt = ot; }
      }
      ok1 = run && best > 0 && (bt & 255u) == 0;               // trace reached column 0 (BioLibs.scala:464)
      ds = (int32_t)(bt >> 8);
      if (run && lane == 0) { my_cells += (unsigned long long)n * (unsigned long long)width; my_p1 += (unsigned long long)n * (unsigned long long)width; }
    }

    // ------------------------------ phase 2 ------------------------------
    {
      P2Geom g; g.ds = ds; g.dl = ok1 ? n - ds : -1; g.z = width / 2; g.width = width; g.m = m;
      P2Lane<D> L; p2_init<D>(L, lane);
      Edge up; up.h = 0; up.t1 = 0; up.t2 = 0;
      Window aw, bw;
      aw.init(wa, ok1 ? nwa : 0, ds - lane);                    // A[ds + u - 1], u = s - lane
      bw.init(wb, ok1 ? nwb : 0, lane * D - g.z - lane);        // B[lD + c - z + u - 1]
      const int32_t steps = __reduce_max_sync(0xffffffffu, ok1 ? g.dl + G - 1 : 0);
      for (int32_t s = 1; s <= steps; s++) {
        const int32_t u = s - lane;
        Edge in = shfl_up_edge<G>(up);
        if (lane == 0) { in.h = 0; in.t1 = 0; in.t2 = 0; }
        const uint32_t a = aw.bits() >> 30;
        const uint32_t bb = bw.bits();
        uint32_t bd[D];
#pragma unroll
        for (int c = 0; c < D; c++) bd[c] = (bb >> (30 - 2 * c)) & 3u;
        p2_first<D>(L, K, g, lane, u, a, bd[0], in);
        Edge right = shfl_down_edge<G>(p2_down<D>(L));
        if (lane == G - 1) { right.h = 0; right.t1 = 0; right.t2 = 0; }
        up = p2_rest<D>(L, K, g, lane, u, a, bd, right);
        aw.advance(); bw.advance();
      }
      int32_t best = L.best, bu = L.best_u, bk = L.best_kk; uint32_t t1 = L.best_t1, t2 = L.best_t2;
#pragma unroll
      for (int o = G / 2; o; o >>= 1) {
        const int32_t ob = __shfl_xor_sync(0xffffffffu, best, o, G), ou = __shfl_xor_sync(0xffffffffu, bu, o, G),
                      ok = __shfl_xor_sync(0xffffffffu, bk, o, G);
        const uint32_t o1 = __shfl_xor_sync(0xffffffffu, t1, o, G), o2 = __shfl_xor_sync(0xffffffffu, t2, o, G);
        if (ob > best || (ob == best && (ou < bu || (ou == bu && ok < bk)))) { best = ob; bu = ou; bk = ok; t1 = o1; t2 = o2; }
      }
      if (ok1) {
        if (lane == 0) my_cells += (unsigned long long)(g.dl + 1) * (unsigned long long)(width + 1);
        if (best > 0) {
          const int32_t su = (int32_t)(t1 >> 8), sk = (int32_t)(t1 & 255u);
          R.start_i = su + ds; R.start_j = sk - g.z + su;
          R.end_i = bu + ds;   R.end_j = bk - g.z + bu;
          R.correct = (int32_t)(t2 >> 16); R.error = (int32_t)(t2 & 0xffffu); R.score = best;
          dp_finish(R, n, m, P.min_overlap, P.min_identity, P.max_ignore);
        }
      }
    }
    if (have && lane == 0) {
      hovl_record rec;
      rec.a = (int32_t)ida; rec.b = (int32_t)idb; rec.count = P.pcnt ? (int32_t)P.pcnt[pidx] : 0;
      rec.flags = (!run && extra == 0) ? -1 : R.flags;          // -1: band wider than this class (host bug)
      rec.start_i = R.start_i; rec.start_j = R.start_j; rec.end_i = R.end_i; rec.end_j = R.end_j;
      rec.correct = R.correct; rec.error = R.error; rec.score = R.score; rec.ahg = R.ahg; rec.bhg = R.bhg;
      P.out[pidx] = rec;
    }
  }
  // one atomic per warp for the cell counter
  for (int o = 16; o; o >>= 1) { my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o); my_p1 += __shfl_xor_sync(0xffffffffu, my_p1, o); }
  if (lane32 == 0 && my_cells) atomicAdd(P.cells, my_cells);
  if (lane32 == 0 && my_p1) atomicAdd(P.cells + 2, my_p1);
}

// ---------------------------------------------------------------- thread-per-pair kernels
// Narrow bands (12 <= width <= 32): one thread per pair (csrc/dp_tp.cuh).  Phase 1 and phase 2
// are separate launches so that phase 2 can run the pairs in order of dovetail length: the 32
// pairs of a warp then finish together.  A warp fetches 32 pairs at a time from a global
// counter; its direction-code scratch is a private slice indexed by the resident warp.
constexpr int TP_THREADS = 128;
#ifndef HOVL_TP2_P2_BLOCKS
#define HOVL_TP2_P2_BLOCKS 3          // CTAs per SM the packed phase-2 kernel is compiled for; 4 (128 registers, spills) measured 28.4 vs 26.5 ms on K3
#endif

struct TpArgs {
  const uint32_t *packed; const uint64_t *woff; const uint32_t *len; const uint32_t *bad;
  const uint32_t *pa, *pb, *pcnt;       // pair arrays (1-based ids)
  const uint32_t *order;                // pair indices to run, in this order (nullptr = 0..np)
  uint32_t np;
  int32_t *p1;                          // per pair: doveStart >= 0, -1 dud, -2 short B, -3 bad base, -4 wrong class
  hovl_record *out;
  unsigned long long *cells; uint32_t *counter;
  uint32_t *scratch; size_t scratch_words;      // per thread
  int32_t cost4[16]; int32_t ge4, gox, goy, one, neg1, four, k, min_overlap; uint32_t min_identity, max_ignore;
  int32_t cost[16], gO, gE, max_len;    // unscaled, for the two-pairs-per-thread kernels (dp_tp2.cuh)
  int32_t tgt, floor;                   // phase-2 frame constants of the class being run (tp2::frame_fits)
  uint32_t *todo; uint32_t *todo_count; // tp_probe_kernel: pairs whose phase 1 still needs the DP
  int32_t wmin, wmax;                   // band widths of the class being run
};
// p1[] of a pair whose doveStart came from the exact-prefix probe carries this bit (its phase-1 cells
// were not evaluated); doveStart itself is < 2^15
constexpr int32_t TP_PROBED = 1 << 20;
__device__ __forceinline__ int32_t tp_dove_start(int32_t raw) { return raw >= 0 ? (raw & ~TP_PROBED) : raw; }

struct TpPair {
  uint32_t pidx, ida, idb; int32_t n, m, width; const uint32_t *wa, *wb;
};
__device__ __forceinline__ TpPair tp_fetch(const TpArgs &P, uint32_t slot) {
  TpPair q;
  q.pidx = P.order ? P.order[slot] : slot;
  q.ida = P.pa[q.pidx]; q.idb = P.pb[q.pidx];
  q.n = (int32_t)P.len[q.ida - 1]; q.m = (int32_t)P.len[q.idb - 1];
  q.wa = P.packed + P.woff[q.ida - 1]; q.wb = P.packed + P.woff[q.idb - 1];
  q.width = band_width(q.n, P.k, P.min_identity);
  return q;
}

template <int W, int WMIN, bool FAST>
__global__ void __launch_bounds__(TP_THREADS, 3) tp_phase1_kernel(const TpArgs P) {
  __shared__ int32_t s_cost4[16];
  if (threadIdx.x < 16) s_cost4[threadIdx.x] = P.cost4[threadIdx.x];
  __syncthreads();
  tp::Consts K; K.cost4 = s_cost4; K.ge4 = P.ge4; K.gox = P.gox; K.goy = P.goy; K.one = P.one; K.neg1 = P.neg1; K.four = P.four;
  const int lane = threadIdx.x & 31;
  const size_t gwarp = ((size_t)blockIdx.x * TP_THREADS + threadIdx.x) >> 5;
  uint32_t *sc = P.scratch + gwarp * 32 * P.scratch_words + lane;
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(P.counter, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= P.np) break;
    const uint32_t slot = base + lane;
    if (slot < P.np) {
      const TpPair q = tp_fetch(P, slot);
      int32_t r;
      if (P.bad[q.ida - 1] | P.bad[q.idb - 1]) r = -3;                 // MatchError in costFunc
      else if (q.m < q.width) r = -2;                                    // charAt past the end of B
      else if (q.width > W || q.width < WMIN) r = -4;
      else r = tp::phase1<W, WMIN, 32, FAST>(K, q.wa, q.n, q.wb, q.m, q.width, sc);
      P.p1[q.pidx] = r;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------- phase 1 without the DP
// tp::probe_prefix (dp_tp.cuh): where B's first `width` bases occur in A exactly, doveStart is the
// first occurrence and no DP is needed; the other pairs are appended to `todo` for the phase-1 DP.
__global__ void __launch_bounds__(TP_THREADS) tp_probe_kernel(const TpArgs P) {
  const int lane = threadIdx.x & 31;
  const uint32_t slot = blockIdx.x * TP_THREADS + threadIdx.x;
  bool need = false; uint32_t pidx = 0;
  if (slot < P.np) {
    const TpPair q = tp_fetch(P, slot);
    pidx = q.pidx;
    int32_t r;
    if (P.bad[q.ida - 1] | P.bad[q.idb - 1]) r = -3;                 // MatchError in costFunc
    else if (q.m < q.width) r = -2;                                   // charAt past the end of B
    else if (q.width > P.wmax || q.width < P.wmin) r = -4;
    else {
      const int32_t ds = tp::probe_prefix(q.wa, q.n, q.wb, q.m, q.width);
      r = ds >= 0 ? (ds | TP_PROBED) : -5;                            // -5: the DP decides
      need = r == -5;
    }
    if (!need) P.p1[pidx] = r;
  }
  __syncwarp();
  const uint32_t ball = __ballot_sync(0xffffffffu, need);
  if (ball) {
    const int leader = __ffs(ball) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(P.todo_count, (uint32_t)__popc(ball));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (need) P.todo[base + __popc(ball & ((1u << lane) - 1u))] = pidx;
  }
}

// sort key for phase 2: longest dovetail first, pairs without a phase 2 last
__global__ void __launch_bounds__(256) tp_dl_key_kernel(const uint32_t *__restrict__ order, uint32_t np,
                                                        const uint32_t *__restrict__ pa, const uint32_t *__restrict__ len,
                                                        const int32_t *__restrict__ p1, uint32_t *__restrict__ key,
                                                        uint32_t *__restrict__ val) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= np) return;
  const uint32_t pidx = order ? order[i] : i;
  const int32_t ds = tp_dove_start(p1[pidx]);
  const int32_t n = (int32_t)len[pa[pidx] - 1];
  key[i] = ds < 0 ? 0xffffu : (uint32_t)(0xfffe - (n - ds));            // n <= 32767
  val[i] = pidx;
}

template <int W, int WMIN, bool FAST>
__global__ void __launch_bounds__(TP_THREADS, 4) tp_phase2_kernel(const TpArgs P) {
  __shared__ int32_t s_cost4[16];
  if (threadIdx.x < 16) s_cost4[threadIdx.x] = P.cost4[threadIdx.x];
  __syncthreads();
  tp::Consts K; K.cost4 = s_cost4; K.ge4 = P.ge4; K.gox = P.gox; K.goy = P.goy; K.one = P.one; K.neg1 = P.neg1; K.four = P.four;
  const int lane = threadIdx.x & 31;
  const size_t gwarp = ((size_t)blockIdx.x * TP_THREADS + threadIdx.x) >> 5;
  uint32_t *sc = P.scratch + gwarp * 32 * P.scratch_words + lane;
  unsigned long long my_cells = 0, my_skipped = 0, my_p1 = 0;
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(P.counter, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= P.np) break;
    const uint32_t slot = base + lane;
    if (slot < P.np) {
      const TpPair q = tp_fetch(P, slot);
      const int32_t raw = P.p1[q.pidx], ds = tp_dove_start(raw);
      DpOut R;
      dp_set_dud(R, ds == -3 ? HOVL_CORE_F_BAD_BASE : (ds == -2 ? HOVL_CORE_F_SHORT_B : 0));
      if (ds >= -1) { my_cells += (unsigned long long)q.n * (unsigned long long)q.width; my_p1 += (unsigned long long)q.n * (unsigned long long)q.width; }
      if (raw >= 0 && (raw & TP_PROBED)) my_skipped += (unsigned long long)q.n * (unsigned long long)q.width;
      if (ds >= 0) {
        my_cells += (unsigned long long)(q.n - ds + 1) * (unsigned long long)(q.width + 1);
        tp::phase2<W, WMIN, 32, FAST>(K, q.wa, q.n, q.wb, q.m, q.width, ds, sc, P.min_overlap, P.min_identity, P.max_ignore, R);
      }
      hovl_record rec;
      rec.a = (int32_t)q.ida; rec.b = (int32_t)q.idb; rec.count = P.pcnt ? (int32_t)P.pcnt[q.pidx] : 0;
      rec.flags = ds == -4 ? -1 : R.flags;                               // -1: pair reached the wrong class (host bug)
      rec.start_i = R.start_i; rec.start_j = R.start_j; rec.end_i = R.end_i; rec.end_j = R.end_j;
      rec.correct = R.correct; rec.error = R.error; rec.score = R.score; rec.ahg = R.ahg; rec.bhg = R.bhg;
      P.out[q.pidx] = rec;
    }
    __syncwarp();
  }
  for (int o = 16; o; o >>= 1) {
    my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o); my_skipped += __shfl_xor_sync(0xffffffffu, my_skipped, o);
    my_p1 += __shfl_xor_sync(0xffffffffu, my_p1, o);
  }
  if (lane == 0 && my_cells) atomicAdd(P.cells, my_cells);
  if (lane == 0 && my_skipped) atomicAdd(P.cells + 1, my_skipped);
  if (lane == 0 && my_p1) atomicAdd(P.cells + 2, my_p1);
}

// ---------------------------------------------------------------- two pairs per thread
// csrc/dp_tp2.cuh: 16-bit half-word cells.  A warp fetches 64 pairs; lane l runs pairs 2l and
// 2l + 1.  A pair that cannot run (bad base, short B, no doveStart) lends its half word to a
// copy of its neighbour.
__device__ __forceinline__ void tp2_consts(const TpArgs &P, uint32_t *s_cost2, tp2::Consts2 &K) {
  for (int i = threadIdx.x; i < 256; i += TP_THREADS) {
    const int a1 = i >> 6, a0 = (i >> 4) & 3, b1 = (i >> 2) & 3, b0 = i & 3;
    s_cost2[i] = tp2::pack16(P.cost[a1 * 4 + b1] * 4 + 2, P.cost[a0 * 4 + b0] * 4 + 2);
  }
  __syncthreads();
  K.cost2 = s_cost2; K.ge = tp2::pack16(P.gE * 4, P.gE * 4); K.gox = tp2::pack16(P.gO * 4 + 1, P.gO * 4 + 1);
  K.goy = tp2::pack16(P.gO * 4, P.gO * 4); K.one = P.one; K.neg1 = P.neg1; K.four = P.four; K.tgt = P.tgt; K.floor = P.floor;
}
__device__ __forceinline__ tp2::Pair2 tp2_pair(const TpPair &q, int32_t ds) {
  tp2::Pair2 r; r.wa = q.wa; r.wb = q.wb; r.n = q.n; r.m = q.m; r.width = q.width; r.ds = ds; return r;
}

template <int W, int WMIN>
__global__ void __launch_bounds__(TP_THREADS, 3) tp2_phase1_kernel(const TpArgs P) {
  __shared__ uint32_t s_cost2[256];
  __shared__ uint32_t s_state[(2 * (W + 1) + 2) * TP_THREADS];
  tp2::Consts2 K; tp2_consts(P, s_cost2, K);
  const int lane = threadIdx.x & 31;
  const size_t gwarp = ((size_t)blockIdx.x * TP_THREADS + threadIdx.x) >> 5;
  uint32_t *sc = P.scratch + gwarp * 32 * P.scratch_words + lane;
  uint32_t *st = s_state + threadIdx.x;
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(P.counter, 64u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= P.np) break;
    TpPair q[2]; int32_t r[2]; bool have[2], run[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const uint32_t slot = base + 2 * lane + h;
      have[h] = slot < P.np; run[h] = false; r[h] = 0;
      if (have[h]) {
        q[h] = tp_fetch(P, slot);
        if (P.bad[q[h].ida - 1] | P.bad[q[h].idb - 1]) r[h] = -3;        // MatchError in costFunc
        else if (q[h].m < q[h].width) r[h] = -2;                           // charAt past the end of B
        else if (q[h].width > W || q[h].width < WMIN) r[h] = -4;
        else run[h] = true;
      }
    }
    if (run[0] || run[1]) {
      int32_t out[2];
      tp2::phase1x2<W, WMIN, 32, TP_THREADS>(K, tp2_pair(run[0] ? q[0] : q[1], 0), tp2_pair(run[1] ? q[1] : q[0], 0),
                                              P.max_len, sc, st, out);
      if (run[0]) r[0] = out[0];
      if (run[1]) r[1] = out[1];
    }
    if (have[0]) P.p1[q[0].pidx] = r[0];
    if (have[1]) P.p1[q[1].pidx] = r[1];
    __syncwarp();
  }
}

template <int W, int WMIN>
__global__ void __launch_bounds__(TP_THREADS, HOVL_TP2_P2_BLOCKS) tp2_phase2_kernel(const TpArgs P) {
  __shared__ uint32_t s_cost2[256];
  __shared__ uint32_t s_state[(2 * (W + 1) + 2) * TP_THREADS];
  tp2::Consts2 K; tp2_consts(P, s_cost2, K);
  const int lane = threadIdx.x & 31;
  const size_t gwarp = ((size_t)blockIdx.x * TP_THREADS + threadIdx.x) >> 5;
  uint32_t *sc = P.scratch + gwarp * 32 * P.scratch_words + lane;
  uint32_t *st = s_state + threadIdx.x;
  unsigned long long my_cells = 0, my_skipped = 0, my_p1 = 0;
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(P.counter, 64u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= P.np) break;
    TpPair q[2]; int32_t ds[2]; bool have[2]; DpOut R[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const uint32_t slot = base + 2 * lane + h;
      have[h] = slot < P.np; ds[h] = -1;
      dp_set_dud(R[h], 0);
      if (have[h]) {
        q[h] = tp_fetch(P, slot);
        const int32_t raw = P.p1[q[h].pidx];
        ds[h] = tp_dove_start(raw);
        dp_set_dud(R[h], ds[h] == -3 ? HOVL_CORE_F_BAD_BASE : (ds[h] == -2 ? HOVL_CORE_F_SHORT_B : 0));
        if (ds[h] >= -1) { my_cells += (unsigned long long)q[h].n * (unsigned long long)q[h].width; my_p1 += (unsigned long long)q[h].n * (unsigned long long)q[h].width; }
        if (raw >= 0 && (raw & TP_PROBED)) my_skipped += (unsigned long long)q[h].n * (unsigned long long)q[h].width;
        if (ds[h] >= 0) my_cells += (unsigned long long)(q[h].n - ds[h] + 1) * (unsigned long long)(q[h].width + 1);
      }
    }
    const bool run0 = have[0] && ds[0] >= 0, run1 = have[1] && ds[1] >= 0;
    if (run0 || run1) {
      DpOut unused; dp_set_dud(unused, 0);
      const tp2::Pair2 p0 = tp2_pair(q[0], ds[0]), p1 = tp2_pair(q[1], ds[1]);
      tp2::phase2x2<W, WMIN, 32, TP_THREADS>(K, run0 ? p0 : p1, run1 ? p1 : p0, P.max_len, sc, st, P.min_overlap,
                                              P.min_identity, P.max_ignore, run0 ? R[0] : unused, run1 ? R[1] : unused);
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
      if (!have[h]) continue;
      hovl_record rec;
      rec.a = (int32_t)q[h].ida; rec.b = (int32_t)q[h].idb; rec.count = P.pcnt ? (int32_t)P.pcnt[q[h].pidx] : 0;
      rec.flags = ds[h] == -4 ? -1 : R[h].flags;
      rec.start_i = R[h].start_i; rec.start_j = R[h].start_j; rec.end_i = R[h].end_i; rec.end_j = R[h].end_j;
      rec.correct = R[h].correct; rec.error = R[h].error; rec.score = R[h].score; rec.ahg = R[h].ahg; rec.bhg = R[h].bhg;
      P.out[q[h].pidx] = rec;
    }
    __syncwarp();
  }
  for (int o = 16; o; o >>= 1) {
    my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o); my_skipped += __shfl_xor_sync(0xffffffffu, my_skipped, o);
    my_p1 += __shfl_xor_sync(0xffffffffu, my_p1, o);
  }
  if (lane == 0 && my_cells) atomicAdd(P.cells, my_cells);
  if (lane == 0 && my_skipped) atomicAdd(P.cells + 1, my_skipped);
  if (lane == 0 && my_p1) atomicAdd(P.cells + 2, my_p1);
}

// ---------------------------------------------------------------- classes
// A class is one kernel instantiation; classes are ordered by band width.
//   lane-group classes (dp_kernel<G,D>): width + 1 <= G*D
//   thread-per-pair classes (tp_*<W,WMIN>): WMIN <= width <= W
//   wide-band classes (wb_kernel<D>, dp_wb.cu): width + 1 <= 32*D; (a, b) = the lane-group class that takes
//   their pairs when the scores do not fit wb's 25 bits (wb::fits) or option dp_wide is 0
struct DpClass { int tp; int a, b; int cap; int d; };       // tp 1: (W, WMIN); tp 0: (G, D); tp 2: wide band, d = D; cap = largest width
static const DpClass kClasses[] = {
    {0, 8, 2, 11, 0},
    {1, 12, 12, 12, 0}, {1, 16, 13, 16, 0}, {1, 20, 17, 20, 0}, {1, 24, 21, 24, 0}, {1, 28, 25, 28, 0}, {1, 32, 29, 32, 0},
    {2, 16, 4, 63, 2}, {2, 32, 3, 95, 3}, {2, 32, 4, 127, 4}, {2, 32, 6, 159, 5}, {2, 32, 6, 191, 6}, {2, 32, 8, 223, 7},
    {2, 32, 8, 255, 8}};
constexpr int N_CLASSES = (int)(sizeof(kClasses) / sizeof(kClasses[0]));

__host__ __device__ static inline int class_of_width(int32_t width) {
  const int cap[N_CLASSES] = {11, 12, 16, 20, 24, 28, 32, 63, 95, 127, 159, 191, 223, 255};
  for (int i = 0; i < N_CLASSES; i++) if (width <= cap[i]) return i;
  return N_CLASSES;                                                        // band too wide
}

__global__ void __launch_bounds__(256) classify_kernel(const uint32_t *__restrict__ pa, const uint32_t *__restrict__ len,
                                                       size_t np, int32_t k, uint32_t min_identity,
                                                       uint8_t *__restrict__ cls, uint32_t *__restrict__ counts) {
  const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= np) return;
  const int cidx = class_of_width(band_width((int32_t)len[pa[p] - 1], k, min_identity));
  cls[p] = (uint8_t)cidx;
  atomicAdd(&counts[cidx], 1u);
}
__global__ void __launch_bounds__(256) class_scatter_kernel(const uint8_t *__restrict__ cls, size_t np,
                                                            const uint32_t *__restrict__ class_base, uint32_t *__restrict__ cursor,
                                                            uint32_t *__restrict__ index) {
  const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= np) return;
  const int cidx = cls[p];
  index[class_base[cidx] + atomicAdd(&cursor[cidx], 1u)] = (uint32_t)p;
}

template <int G, int D>
static void launch_class(Ctx &c, DpArgs &a) {
  int per_sm = 0;
  HOVL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dp_kernel<G, D>, DP_THREADS, 0));
  if (per_sm < 1) per_sm = 1;
  const uint64_t want = div_up((uint64_t)a.np * G, DP_THREADS);
  const uint64_t full = (uint64_t)c.sm_count * per_sm;                    // persistent: one resident wave
  HOVL_CUDA(cudaMemsetAsync(a.counter, 0, 4, c.stream));
  HOVL_LAUNCH(c, (dp_kernel<G, D>), (unsigned)(want < full ? want : full), DP_THREADS, 0, a);
}

// Both phases of one thread-per-pair class over the pairs listed in `index` (nullptr = all).
// pk1 / pk2: the phase's values fit 16 bits -> two pairs per thread (dp_tp2.cuh).
template <typename Kern>
static unsigned tp_grid(Ctx &c, Kern kern, uint64_t np, int pairs_per_thread) {
  int per_sm = 0;
  HOVL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TP_THREADS, 0));
  if (per_sm < 1) per_sm = 1;
  const uint64_t want = div_up(np, (uint64_t)TP_THREADS * pairs_per_thread);
  const uint64_t full = (uint64_t)c.sm_count * per_sm;                      // persistent: one resident wave
  return (unsigned)(want < full ? want : full);
}
template <int W, int WMIN, bool FAST>
static void launch_tp_class(Ctx &c, TpArgs &a, const uint32_t *index, uint32_t np, uint32_t max_len, bool pk1, bool pk2) {
  const unsigned grid1 = pk1 ? tp_grid(c, tp2_phase1_kernel<W, WMIN>, np, 2) : tp_grid(c, tp_phase1_kernel<W, WMIN, FAST>, np, 1);
  const unsigned grid2 = pk2 ? tp_grid(c, tp2_phase2_kernel<W, WMIN>, np, 2) : tp_grid(c, tp_phase2_kernel<W, WMIN, FAST>, np, 1);
  const size_t words1 = pk1 ? tp2::scratch_words((int32_t)max_len, W) : tp::scratch_words((int32_t)max_len, W);
  const size_t words2 = pk2 ? tp2::scratch_words((int32_t)max_len, W) : tp::scratch_words((int32_t)max_len, W);
  const size_t need1 = (size_t)grid1 * TP_THREADS * words1, need2 = (size_t)grid2 * TP_THREADS * words2;
  DevBuf<uint32_t> scratch(c, need1 > need2 ? need1 : need2);
  DevBuf<uint32_t> k0(c, np), k1(c, np), v0(c, np), v1(c, np);
  a.scratch = scratch.get(); a.max_len = (int32_t)max_len;
  a.order = index; a.np = np; a.scratch_words = words1; a.wmin = WMIN; a.wmax = W;
  // phase 1 as a substring search where that is exact (tp_probe_kernel); the DP takes the rest
  DevBuf<uint32_t> todo, todo_count;
  uint32_t np1 = np;
  if (c.dp_probe && tp::probe_valid(a.cost, a.gE)) {
    todo.alloc(c, np); todo_count.alloc(c, 1);
    HOVL_CUDA(cudaMemsetAsync(todo_count.get(), 0, 4, c.stream));
    a.todo = todo.get(); a.todo_count = todo_count.get();
    HOVL_LAUNCH(c, tp_probe_kernel, (unsigned)div_up(np, TP_THREADS), TP_THREADS, 0, a);
    d2h(c, &np1, todo_count.get(), 1);
    sync(c);
    a.order = todo.get(); a.np = np1;
  }
  if (np1) {
    const unsigned g1 = pk1 ? tp_grid(c, tp2_phase1_kernel<W, WMIN>, np1, 2) : tp_grid(c, tp_phase1_kernel<W, WMIN, FAST>, np1, 1);
    HOVL_CUDA(cudaMemsetAsync(a.counter, 0, 4, c.stream));
    if (pk1) HOVL_LAUNCH(c, (tp2_phase1_kernel<W, WMIN>), g1, TP_THREADS, 0, a);
    else HOVL_LAUNCH(c, (tp_phase1_kernel<W, WMIN, FAST>), g1, TP_THREADS, 0, a);
  }
  a.np = np;
  HOVL_LAUNCH(c, tp_dl_key_kernel, (unsigned)div_up(np, 256), 256, 0, index, np, a.pa, a.len, a.p1, k0.get(), v0.get());
  const int r = radix_sort<uint32_t, uint32_t, true>(c, k0.get(), k1.get(), v0.get(), v1.get(), np, 0, 16);
  a.order = r ? v1.get() : v0.get(); a.scratch_words = words2;
  HOVL_CUDA(cudaMemsetAsync(a.counter, 0, 4, c.stream));
  if (pk2) HOVL_LAUNCH(c, (tp2_phase2_kernel<W, WMIN>), grid2, TP_THREADS, 0, a);
  else HOVL_LAUNCH(c, (tp_phase2_kernel<W, WMIN, FAST>), grid2, TP_THREADS, 0, a);
}

void align_pairs_wide(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                      const uint32_t *d_cnt, const uint32_t *index, uint32_t np, int D, hovl_record *d_out,
                      unsigned long long *d_cells, uint32_t *d_counter);
bool wide_band_fits(const hovl_settings &s, uint32_t max_len);

static void launch_by_index(Ctx &c, const hovl_settings &s, const Reads &R, int cidx, DpArgs &a, TpArgs &t, uint32_t max_len, bool fast,
                            bool packable) {
  HOVL_REQUIRE(cidx < N_CLASSES, HOVL_E_UNSUPPORTED, "band wider than 255 diagonals (read too long for this identity)");
  const DpClass &k = kClasses[cidx];
  if (k.tp == 1 && !packable) {
    // costs too large for (score << 2) + priority in 32 bits: the reference takes any Int costs, so do we, with the plain
    // 32-bit cells of the lane-group kernel
    if (k.cap <= 15) launch_class<8, 2>(c, a);
    else if (k.cap <= 23) launch_class<8, 3>(c, a);
    else if (k.cap <= 31) launch_class<8, 4>(c, a);
    else launch_class<16, 3>(c, a);
    return;
  }
  if (k.tp == 2 && c.dp_wide && wide_band_fits(s, max_len)) {
    align_pairs_wide(c, s, R, a.pa, a.pb, a.pcnt, a.index, a.np, k.d, a.out, a.cells, a.counter);
    return;
  }
  if (k.tp == 1) {
    const uint32_t *index = a.index; const uint32_t np = a.np;
    switch (k.a) {
#define HOVL_TP_LAUNCH(W_, WMIN_) case W_: \
        { const bool pk1 = c.dp_pack16 && tp2::fits16(t.cost, t.gO, t.gE, W_); \
          const bool pk2 = c.dp_pack16 && tp2::frame_fits(t.cost, t.gO, t.gE, W_, &t.tgt, &t.floor); \
          if (fast) launch_tp_class<W_, WMIN_, true>(c, t, index, np, max_len, pk1, pk2); \
          else launch_tp_class<W_, WMIN_, false>(c, t, index, np, max_len, pk1, pk2); } \
        break;
      HOVL_TP_CLASSES(HOVL_TP_LAUNCH)
#undef HOVL_TP_LAUNCH
      default: throw Error(HOVL_E_CUDA, "internal: unknown thread-per-pair class");
    }
    return;
  }
  switch (k.a * 100 + k.b) {
    case 802: launch_class<8, 2>(c, a); break;
    case 1603: launch_class<16, 3>(c, a); break;
    case 1604: launch_class<16, 4>(c, a); break;
    case 3203: launch_class<32, 3>(c, a); break;
    case 3204: launch_class<32, 4>(c, a); break;
    case 3206: launch_class<32, 6>(c, a); break;
    case 3208: launch_class<32, 8>(c, a); break;
    default: throw Error(HOVL_E_CUDA, "internal: unknown lane-group class");
  }
}

void align_pairs(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                 const uint32_t *d_cnt, size_t np, hovl_record *d_out, hovl_stats_t &st) {
  if (s.aligner == 1) { align_pairs_quadratic(c, s, R, d_a, d_b, d_cnt, np, d_out, st); return; }   // --quadratic-align
  StageTimer tm(c, &st.ns_dp, "dp");
  if (np == 0) return;
  HOVL_REQUIRE(np < 0xfffffff0ull, HOVL_E_UNSUPPORTED, "more than 2^32 pairs in one align call");
  HOVL_REQUIRE(R.max_len <= 32767, HOVL_E_UNSUPPORTED, "reads longer than 32767 bases are not supported by the DP kernel");
  DevBuf<unsigned long long> cells(c, 3);                 // [0] cells of the reference's matrices, [1] of them not evaluated, [2] phase-1 part of [0]
  DevBuf<uint32_t> counter(c, 1);
  HOVL_CUDA(cudaMemsetAsync(cells.get(), 0, 24, c.stream));
  DpArgs a;
  a.packed = R.packed.get(); a.woff = R.woff.get(); a.len = R.len.get(); a.bad = R.bad.get();
  a.pa = d_a; a.pb = d_b; a.pcnt = d_cnt; a.index = nullptr; a.np = (uint32_t)np; a.out = d_out;
  a.cells = cells.get(); a.counter = counter.get();
  for (int i = 0; i < 16; i++) a.cost[i] = s.cost[i / 4][i % 4];
  a.gO = s.gap_open; a.gE = s.gap_extend; a.k = s.k; a.min_overlap = s.min_overlap;
  a.min_identity = s.min_identity_f32; a.max_ignore = s.max_ignore_f32;

  const int c_lo = class_of_width(band_width((int32_t)R.min_len, s.k, s.min_identity_f32));
  const int c_hi = class_of_width(band_width((int32_t)R.max_len, s.k, s.min_identity_f32));
  HOVL_REQUIRE(c_hi < N_CLASSES, HOVL_E_UNSUPPORTED, "band wider than 255 diagonals (read too long for this identity)");

  TpArgs t;
  DevBuf<int32_t> p1;
  bool any_tp = false, packable = true;
  for (int i = c_lo; i <= c_hi; i++) any_tp = any_tp || kClasses[i].tp == 1;
  // the thread-per-pair cells carry 4*score + priority in an int32
  int64_t big = 1;
  for (int i = 0; i < 16; i++) { const int64_t v = s.cost[i / 4][i % 4]; big = std::max<int64_t>(big, v < 0 ? -v : v); }
  big = std::max<int64_t>(big, -(int64_t)s.gap_open - (int64_t)s.gap_extend);
  if (big * ((int64_t)R.max_len + 64) >= ((int64_t)1 << 28)) packable = false;
  // packed arg-max keys: score << 12 | (1023 - row) << 2 | priority in an int32
  const bool fast = R.max_len <= (uint32_t)tp::TP_FAST_MAX_LEN && big * ((int64_t)R.max_len + 64) < ((int64_t)1 << 18);
  if (any_tp && packable) {
    p1.alloc(c, np);
    t.packed = a.packed; t.woff = a.woff; t.len = a.len; t.bad = a.bad; t.pa = d_a; t.pb = d_b; t.pcnt = d_cnt;
    t.order = nullptr; t.np = (uint32_t)np; t.p1 = p1.get(); t.out = d_out; t.cells = cells.get(); t.counter = counter.get();
    t.scratch = nullptr; t.scratch_words = 0;
    const int32_t unit = fast ? 4096 : 4;                // values are (score << SH) + priority, dp_tp.cuh
    for (int i = 0; i < 16; i++) t.cost4[i] = unit * s.cost[i / 4][i % 4] + 2;
    t.ge4 = unit * s.gap_extend; t.gox = unit * s.gap_open + 1; t.goy = unit * s.gap_open; t.one = 1; t.neg1 = -1; t.four = 4;
    t.k = s.k; t.min_overlap = s.min_overlap; t.min_identity = s.min_identity_f32; t.max_ignore = s.max_ignore_f32;
    for (int i = 0; i < 16; i++) t.cost[i] = s.cost[i / 4][i % 4];
    t.gO = s.gap_open; t.gE = s.gap_extend; t.max_len = (int32_t)R.max_len;
    t.tgt = 0; t.floor = 0; t.todo = nullptr; t.todo_count = nullptr; t.wmin = 0; t.wmax = 0;
  }

  if (c_lo == c_hi) {
    launch_by_index(c, s, R, c_lo, a, t, R.max_len, fast, packable);   // the common case: one class, no index list
  } else {
    DevBuf<uint8_t> cls(c, np);
    DevBuf<uint32_t> counts(c, 2 * (N_CLASSES + 1)), index(c, np), base(c, N_CLASSES + 1);
    HOVL_CUDA(cudaMemsetAsync(counts.get(), 0, counts.bytes(), c.stream));
    HOVL_LAUNCH(c, classify_kernel, (unsigned)div_up(np, 256), 256, 0, d_a, R.len.get(), np, s.k, s.min_identity_f32,
                cls.get(), counts.get());
    uint32_t h[N_CLASSES + 1], hb[N_CLASSES + 1];
    d2h(c, h, counts.get(), (size_t)N_CLASSES + 1);
    sync(c);
    HOVL_REQUIRE(h[N_CLASSES] == 0, HOVL_E_UNSUPPORTED, "band wider than 255 diagonals (read too long for this identity)");
    uint32_t run = 0;
    for (int i = 0; i <= N_CLASSES; i++) { hb[i] = run; run += h[i]; }
    h2d(c, base.get(), hb, (size_t)N_CLASSES + 1);
    HOVL_LAUNCH(c, class_scatter_kernel, (unsigned)div_up(np, 256), 256, 0, cls.get(), np, base.get(),
                counts.get() + N_CLASSES + 1, index.get());
    for (int i = 0; i < N_CLASSES; i++) {
      if (!h[i]) continue;
      a.index = index.get() + hb[i]; a.np = h[i];
      launch_by_index(c, s, R, i, a, t, R.max_len, fast, packable);
    }
    sync(c);                                           // hb / index live until the launches are done
  }
  unsigned long long hc[3] = {0, 0, 0};
  HOVL_CUDA(cudaMemcpyAsync(hc, cells.get(), 24, cudaMemcpyDeviceToHost, c.stream));
  sync(c);
  st.dp_cells += hc[0]; st.dp_cells_skipped += hc[1]; st.dp_cells_phase1 += hc[2];
}

}  // namespace hovl
