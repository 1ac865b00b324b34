// rebalance.cu — alignment re-balance by DP cells (north_star: "the deduplicated pairs are re-balanced across GPUs for
// alignment").  The reference has no counterpart (one JVM, futures over dispatch blocks,
// src/Project4.scala:748-760).
//
// In the sharded path a pair is counted by the rank that owns its fst read (a contiguous range of read ids), so the
// dispatched pairs — and with them the DP work — follow the reads: a rank whose reads are longer or overlap more has more
// cells to evaluate and everybody waits for it.  After the count merge every rank knows its dispatched pairs (sorted by
// (a, b)) and what each costs: lenA * width + (lenA + 1) * (width + 1) cells of the reference's matrices
// (src/BioLibs.scala:619-624, :703-724).  With the ranks' totals gathered, pair i of rank r sits at position
// (cost of lower ranks) + (cost of r's pairs before i) of the job's cost line, and goes to rank
// floor(position * parts / total).  Positions ascend with (rank, i), so every rank sends CONTIGUOUS segments and what a
// rank receives, in source-rank order, is again sorted by (a, b): rank order stays the canonical record order.
#include "handles.cuh"
#include "scan.cuh"
#include "dp_core.cuh"

namespace hovl {

struct CostIn {
  const uint32_t *a, *len; int32_t k; uint32_t mi;
  __device__ unsigned long long operator()(size_t i) const {
    const int32_t n = (int32_t)len[a[i] - 1u];
    const int32_t w = band_width(n, k, mi);
    return (unsigned long long)n * (unsigned long long)w + (unsigned long long)(n + 1) * (unsigned long long)(w + 1);
  }
};
struct CostOut {
  unsigned long long *excl; size_t n;
  __device__ void operator()(size_t i, const unsigned long long &v, const unsigned long long &p) const {
    excl[i] = p;
    if (i == n - 1) excl[n] = p + v;
  }
};
// cut[p] = first pair whose position on the cost line is >= p * total / parts   (p = 1 .. parts - 1)
__global__ void split_kernel(const unsigned long long *__restrict__ excl, size_t n, unsigned long long before,
                             unsigned long long total, uint32_t parts, unsigned long long *__restrict__ cut) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x + 1u;
  if (p >= parts) return;
  // smallest position that belongs to rank p: ceil(p * total / parts), in 128-bit-free arithmetic
  const unsigned long long q = total / parts, r = total % parts;
  const unsigned long long thr = q * p + (r * p + parts - 1) / parts;
  size_t lo = 0, hi = n;                               // first i with before + excl[i] >= thr
  while (lo < hi) { const size_t mid = (lo + hi) >> 1; if (before + excl[mid] >= thr) hi = mid; else lo = mid + 1; }
  cut[p] = lo;
}
__global__ void __launch_bounds__(256) interleave3_kernel(const uint32_t *__restrict__ a, const uint32_t *__restrict__ b,
                                                          const uint32_t *__restrict__ c, size_t n, uint32_t *__restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { out[3 * i] = a[i]; out[3 * i + 1] = b[i]; out[3 * i + 2] = c[i]; }
}
__global__ void __launch_bounds__(256) split3_kernel(const uint32_t *__restrict__ in, size_t n, uint32_t *__restrict__ a,
                                                     uint32_t *__restrict__ b, uint32_t *__restrict__ c) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { a[i] = in[3 * i]; b[i] = in[3 * i + 1]; c[i] = in[3 * i + 2]; }
}

// the rank's dispatched pairs and the exclusive prefix of their cost (n + 1 entries), kept in the context
void rebalance_select(hovl_ctx *h, const hovl_settings &s, uint64_t *d_keys, uint32_t *d_counts, size_t n_partial,
                      const uint32_t *d_len_all, uint32_t n_reads_total, uint64_t *n_pairs_out, uint64_t *cost_out) {
  Ctx &c = h->c;
  hovl_stats_t &st = h->sh_st;
  Pairs P;
  merge_partial_pairs(c, d_keys, d_counts, n_partial, n_reads_total, P, st);
  const size_t nd = select_pairs(c, s, P, h->sh_da, h->sh_db, h->sh_dc);
  P = Pairs();
  h->sh_npairs = nd;
  h->sh_cost.alloc(c, nd + 1);
  HOVL_CUDA(cudaMemsetAsync(h->sh_cost.get(), 0, 8, c.stream));
  DevBuf<unsigned long long> part;
  const unsigned long long *tot = scan_act<unsigned long long, CostIn, CostOut>(
      c, nd, CostIn{h->sh_da.get(), d_len_all, s.k, s.min_identity_f32}, CostOut{h->sh_cost.get(), nd}, part);
  *cost_out = read_total(c, tot);
  *n_pairs_out = nd;
}

void rebalance_split(hovl_ctx *h, uint64_t cost_before, uint64_t cost_total, uint32_t parts, uint64_t *counts_out) {
  Ctx &c = h->c;
  const size_t n = h->sh_npairs;
  std::vector<unsigned long long> cut((size_t)parts + 1, 0);
  cut[parts] = n;
  if (parts > 1 && n && cost_total) {
    DevBuf<unsigned long long> d_cut(c, (size_t)parts + 1);
    HOVL_LAUNCH(c, split_kernel, (parts + 63) / 64, 64, 0, h->sh_cost.get(), n, (unsigned long long)cost_before,
                (unsigned long long)cost_total, parts, d_cut.get());
    d2h(c, cut.data() + 1, d_cut.get() + 1, (size_t)parts - 1);
    sync(c);
  } else {
    for (uint32_t p = 1; p < parts; p++) cut[p] = n;        // nothing to cut: everything stays on the lowest rank of its range
    if (parts > 1 && n && !cost_total) cut[1] = n;
  }
  for (uint32_t p = 0; p < parts; p++) counts_out[p] = cut[p + 1] >= cut[p] ? cut[p + 1] - cut[p] : 0;
}

void rebalance_fetch(hovl_ctx *h, uint32_t *d_abc_out) {
  Ctx &c = h->c;
  const size_t n = h->sh_npairs;
  if (n) HOVL_LAUNCH(c, interleave3_kernel, (unsigned)div_up(n, 256), 256, 0, h->sh_da.get(), h->sh_db.get(), h->sh_dc.get(), n, d_abc_out);
  sync(c);
  h->sh_da.release(); h->sh_db.release(); h->sh_dc.release(); h->sh_cost.release(); h->sh_npairs = 0;
}

void rebalance_unpack(Ctx &c, const uint32_t *d_abc, size_t n, DevBuf<uint32_t> &a, DevBuf<uint32_t> &b, DevBuf<uint32_t> &cn) {
  a.alloc(c, n); b.alloc(c, n); cn.alloc(c, n);
  if (n) HOVL_LAUNCH(c, split3_kernel, (unsigned)div_up(n, 256), 256, 0, d_abc, n, a.get(), b.get(), cn.get());
}

}  // namespace hovl
