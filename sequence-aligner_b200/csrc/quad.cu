// quad.cu — the quadratic local aligner (--quadratic-align; generateLocalAlignmentSet,
// src/BioLibs.scala:267-368, selected at src/Project4.scala:596-601): one warp per (A,B) pair,
// B in column tiles of 32*D columns, anti-diagonal wavefront in registers, neighbours through
// __shfl_up_sync, the tile's right-most column through a per-warp boundary buffer (16 B per row,
// double-buffered, L2-resident).  The math is csrc/quad_core.cuh; this file is the machinery:
// persistent grid, warps fetch pairs from a global counter, A slides through a 32-bit window of the
// 2-bit packed read, the boundary row for the NEXT step is loaded one step ahead.
// Bound: INT32 ALU pipe (lenA*lenB cells per pair, no reuse of anything in memory).
#include "engine.cuh"
#include "quad_core.cuh"

namespace hovl {

constexpr int QUAD_THREADS = 256;
constexpr int QUAD_D = 4;                 // columns per lane: a tile is 128 columns

struct QuadArgs {
  const uint32_t *packed; const uint64_t *woff; const uint32_t *len; const uint32_t *bad;
  const uint32_t *pa, *pb, *pcnt;       // pair arrays (1-based ids)
  uint32_t np;
  hovl_record *out;
  unsigned long long *cells; uint32_t *counter;
  uint4 *bnd; uint32_t bnd_rows;        // per resident warp: 2 * bnd_rows entries (row i at [i])
  int32_t cost[16]; int32_t gO, gE, min_overlap; uint32_t min_identity, max_ignore;
};

struct AWindow {                         // A[pos], first base in the top bits; pos may be negative or past the end
  const uint32_t *w; int32_t nw, pos; uint32_t w0, w1;
  __device__ __forceinline__ uint32_t word(int32_t i) const { return (i >= 0 && i < nw) ? __ldg(w + i) : 0u; }
  __device__ __forceinline__ void init(const uint32_t *words, int32_t nwords, int32_t p) {
    w = words; nw = nwords; pos = p;
    const int32_t wi = p >> 4;
    w0 = word(wi); w1 = word(wi + 1);
  }
  __device__ __forceinline__ uint32_t base() const { return (w0 >> (30 - 2 * (pos & 15))) & 3u; }
  __device__ __forceinline__ void advance() {
    pos++;
    if ((pos & 15) == 0) { w0 = w1; w1 = word((pos >> 4) + 1); }
  }
};

template <int D>
__global__ void __launch_bounds__(QUAD_THREADS) quad_kernel(const QuadArgs P) {
  __shared__ int32_t s_cost[16];
  if (threadIdx.x < 16) s_cost[threadIdx.x] = P.cost[threadIdx.x];
  __syncthreads();
  DpConst K; K.cost = s_cost; K.gO = P.gO; K.gE = P.gE;
  const int lane = threadIdx.x & 31;
  const size_t warp_global = ((size_t)blockIdx.x * QUAD_THREADS + threadIdx.x) >> 5;
  uint4 *bnd0 = P.bnd + warp_global * 2 * (size_t)P.bnd_rows, *bnd1 = bnd0 + P.bnd_rows;
  unsigned long long my_cells = 0;

  for (;;) {
    uint32_t pidx = 0;
    if (lane == 0) pidx = atomicAdd(P.counter, 1u);
    pidx = __shfl_sync(0xffffffffu, pidx, 0);
    if (pidx >= P.np) break;
    const uint32_t ida = P.pa[pidx], idb = P.pb[pidx];
    const int32_t n = (int32_t)P.len[ida - 1], m = (int32_t)P.len[idb - 1];
    const uint32_t *wa = P.packed + P.woff[ida - 1], *wb = P.packed + P.woff[idb - 1];
    const int extra = (P.bad[ida - 1] | P.bad[idb - 1]) ? HOVL_CORE_F_BAD_BASE : 0;      // MatchError in costFunc
    DpOut R; dp_set_dud(R, extra);
    if (!extra) {
      QLane<D> L; q_pair_init<D>(L);
      const int32_t nwa = (n + 15) >> 4;
      const int32_t tiles = (m + 32 * D - 1) / (32 * D);
      uint4 *bin = bnd0, *bout = bnd1;
      for (int32_t t = 0; t < tiles; t++) {
        const int32_t j0 = t * 32 * D + lane * D;
        q_tile_init<D>(L, j0);
        uint32_t bcol[D];
#pragma unroll
        for (int c = 0; c < D; c++) bcol[c] = (j0 + c < m) ? packed_base(wb, j0 + c) : 0u;
        QEdge o = q_border(0, j0 + D);
        AWindow aw; aw.init(wa, nwa, -lane);                      // step s reads A[s - lane - 1]
        const bool from_buf = (lane == 0) && (t > 0);
        const bool to_buf = (lane == 31) && (t + 1 < tiles);
        uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
        if (from_buf) nxt = bin[1];                               // row 1 (n >= 1 whenever tiles > 0 and a pair exists)
        const int32_t steps = n + 31;
        for (int32_t s = 1; s <= steps; s++) {
          const int32_t i = s - lane;
          QEdge in;
          in.hx = __shfl_up_sync(0xffffffffu, o.hx, 1);
          in.hm = __shfl_up_sync(0xffffffffu, o.hm, 1);
          in.t1 = __shfl_up_sync(0xffffffffu, o.t1, 1);
          in.t2 = __shfl_up_sync(0xffffffffu, o.t2, 1);
          if (lane == 0) {
            if (from_buf && i <= n) { in.hx = (int32_t)nxt.x; in.hm = (int32_t)nxt.y; in.t1 = nxt.z; in.t2 = nxt.w; }
            else in = q_border(i, j0);
            if (from_buf && i + 1 <= n) nxt = bin[i + 1];         // one step ahead of its use
          }
          const uint32_t a = aw.base();
          o = q_row<D>(L, K, j0, i, n, m, a, bcol, in);
          if (to_buf && i >= 1 && i <= n) bout[i] = make_uint4((uint32_t)o.hx, (uint32_t)o.hm, o.t1, o.t2);
          aw.advance();
        }
        __syncwarp();                                             // lane 31's rows are read by lane 0 in the next tile
        uint4 *tmp = bin; bin = bout; bout = tmp;
      }
      int32_t best = L.best, bi = L.best_i, bj = L.best_j; uint32_t t1 = L.best_t1, t2 = L.best_t2;
#pragma unroll
      for (int off = 16; off; off >>= 1) {
        const int32_t ob = __shfl_xor_sync(0xffffffffu, best, off), oi = __shfl_xor_sync(0xffffffffu, bi, off),
                      oj = __shfl_xor_sync(0xffffffffu, bj, off);
        const uint32_t o1 = __shfl_xor_sync(0xffffffffu, t1, off), o2 = __shfl_xor_sync(0xffffffffu, t2, off);
        if (ob > 0 && q_better(ob, oi, oj, best, bi, bj)) { best = ob; bi = oi; bj = oj; t1 = o1; t2 = o2; }
      }
      if (lane == 0) my_cells += (unsigned long long)n * (unsigned long long)m;
      if (best > 0) q_finish(R, best, bi, bj, t1, t2, n, m, P.min_overlap, P.min_identity, P.max_ignore);
    }
    if (lane == 0) {
      hovl_record rec;
      rec.a = (int32_t)ida; rec.b = (int32_t)idb; rec.count = P.pcnt ? (int32_t)P.pcnt[pidx] : 0;
      rec.flags = R.flags;
      rec.start_i = R.start_i; rec.start_j = R.start_j; rec.end_i = R.end_i; rec.end_j = R.end_j;
      rec.correct = R.correct; rec.error = R.error; rec.score = R.score; rec.ahg = R.ahg; rec.bhg = R.bhg;
      P.out[pidx] = rec;
    }
  }
  if (lane == 0 && my_cells) atomicAdd(P.cells, my_cells);
}

void align_pairs_quadratic(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                           const uint32_t *d_cnt, size_t np, hovl_record *d_out, hovl_stats_t &st) {
  StageTimer tm(c, &st.ns_dp, "dp");
  if (np == 0) return;
  HOVL_REQUIRE(np < 0xfffffff0ull, HOVL_E_UNSUPPORTED, "more than 2^32 pairs in one align call");
  HOVL_REQUIRE(R.max_len <= 32767, HOVL_E_UNSUPPORTED, "reads longer than 32767 bases are not supported by the DP kernel");
  int64_t big = 1;
  for (int i = 0; i < 16; i++) { const int64_t v = s.cost[i / 4][i % 4]; big = std::max<int64_t>(big, v < 0 ? -v : v); }
  big = std::max<int64_t>(big, -(int64_t)s.gap_open - (int64_t)s.gap_extend);
  HOVL_REQUIRE(big * ((int64_t)R.max_len + 2) < ((int64_t)1 << 31), HOVL_E_UNSUPPORTED, "cost matrix / gap penalties overflow int32 on these reads");
  DevBuf<unsigned long long> cells(c, 1);
  DevBuf<uint32_t> counter(c, 1);
  HOVL_CUDA(cudaMemsetAsync(cells.get(), 0, 8, c.stream));
  HOVL_CUDA(cudaMemsetAsync(counter.get(), 0, 4, c.stream));
  int per_sm = 0;
  HOVL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, quad_kernel<QUAD_D>, QUAD_THREADS, 0));
  if (per_sm < 1) per_sm = 1;
  const uint64_t want = div_up((uint64_t)np * 32, QUAD_THREADS);
  const uint64_t full = (uint64_t)c.sm_count * per_sm;                      // persistent: one resident wave
  const unsigned grid = (unsigned)(want < full ? want : full);
  QuadArgs a;
  a.bnd_rows = R.max_len + 2;
  DevBuf<uint4> bnd(c, (size_t)grid * (QUAD_THREADS / 32) * 2 * a.bnd_rows);
  a.packed = R.packed.get(); a.woff = R.woff.get(); a.len = R.len.get(); a.bad = R.bad.get();
  a.pa = d_a; a.pb = d_b; a.pcnt = d_cnt; a.np = (uint32_t)np; a.out = d_out;
  a.cells = cells.get(); a.counter = counter.get(); a.bnd = bnd.get();
  for (int i = 0; i < 16; i++) a.cost[i] = s.cost[i / 4][i % 4];
  a.gO = s.gap_open; a.gE = s.gap_extend; a.min_overlap = s.min_overlap;
  a.min_identity = s.min_identity_f32; a.max_ignore = s.max_ignore_f32;
  HOVL_LAUNCH(c, (quad_kernel<QUAD_D>), grid, QUAD_THREADS, 0, a);
  unsigned long long hc = 0;
  HOVL_CUDA(cudaMemcpyAsync(&hc, cells.get(), 8, cudaMemcpyDeviceToHost, c.stream));
  sync(c);
  st.dp_cells += hc;
}

}  // namespace hovl
