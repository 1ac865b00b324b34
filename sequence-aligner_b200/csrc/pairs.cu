// pairs.cu — the edge x middle match filter and PairData on the device.
//   count_pairs : per bucket, every head x middle then tail x middle instance
//                 (KmerTable.scala:118-128) is written as one packed (fst,snd) key by a
//                 load-balanced expansion kernel (one output slot per instance, however
//                 skewed the buckets are), the keys are radix-sorted and run-length
//                 encoded: key -> collision count (addKmerPair, KmerTable.scala:57-80).
//                 Instances are materialised in chunks of ctx.pair_chunk and the partial
//                 (key,count) lists merged, so PairData never has to fit at once.
//   select_pairs: minCollisions <= count <= maxCollisions (calcDispatchData, :155-187).
#include "engine.cuh"
#include "scan.cuh"
#include "radix_sort.cuh"

namespace hovl {

// ---------------------------------------------------------------- work list
struct WorkIn {
  const uint32_t *perm, *st_base, *md_base, *en_base;
  __device__ Duo operator()(size_t j) const {
    const uint32_t b = perm ? perm[j] : (uint32_t)j;
    const unsigned long long nst = st_base[b + 1] - st_base[b], nmd = md_base[b + 1] - md_base[b],
                             nen = en_base[b + 1] - en_base[b];
    const unsigned long long cnt = (nst + nen) * nmd;
    return Duo(cnt ? 1ull : 0ull, cnt);
  }
};
struct WorkOut {
  const uint32_t *perm; uint32_t *work_bucket; uint64_t *work_off; size_t n;
  __device__ void operator()(size_t j, const Duo &v, const Duo &p) const {
    if (v.a) { work_bucket[p.a] = perm ? perm[j] : (uint32_t)j; work_off[p.a] = p.b; }
    if (j == n - 1) work_off[p.a + v.a] = p.b + v.b;
  }
};

// ---------------------------------------------------------------- expansion
constexpr int EX_THREADS = 256;
constexpr int EX_ITEMS = 4;
constexpr int EX_TILE = EX_THREADS * EX_ITEMS;

template <typename K, bool SEQ>
__global__ void __launch_bounds__(EX_THREADS)
expand_kernel(uint64_t o_lo, uint64_t o_hi, const uint64_t *__restrict__ work_off, const uint32_t *__restrict__ work_bucket,
              uint32_t nw, const uint32_t *__restrict__ st_base, const uint32_t *__restrict__ md_base,
              const uint32_t *__restrict__ en_base, const uint64_t *__restrict__ st_list,
              const uint64_t *__restrict__ md_list, const uint64_t *__restrict__ en_list, int id_bits,
              K *__restrict__ keys, uint32_t *__restrict__ seq) {
  __shared__ uint64_t soff[EX_TILE + 1];
  __shared__ uint32_t s_wlo, s_m;
  const uint64_t tile_lo = o_lo + (uint64_t)blockIdx.x * EX_TILE;
  const uint64_t tile_hi = (tile_lo + EX_TILE < o_hi) ? tile_lo + EX_TILE : o_hi;
  if (threadIdx.x == 0) {
    uint32_t lo = 0, hi = nw;                       // work_off[lo] <= tile_lo < work_off[hi]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (work_off[mid] <= tile_lo) lo = mid; else hi = mid; }
    s_wlo = lo;
    s_m = (nw - lo) < (uint32_t)EX_TILE ? (nw - lo) : (uint32_t)EX_TILE;   // every work bucket holds >= 1 instance
  }
  __syncthreads();
  const uint32_t wlo = s_wlo, m = s_m;
  for (uint32_t i = threadIdx.x; i <= m; i += EX_THREADS) soff[i] = work_off[wlo + i];
  __syncthreads();
#pragma unroll
  for (int it = 0; it < EX_ITEMS; it++) {
    const uint64_t o = tile_lo + (uint64_t)it * EX_THREADS + threadIdx.x;
    if (o >= tile_hi) continue;
    uint32_t lo = 0, hi = m;                        // soff[lo] <= o < soff[hi]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (soff[mid] <= o) lo = mid; else hi = mid; }
    const uint32_t b = work_bucket[wlo + lo];
    uint64_t q = o - soff[lo];
    const uint32_t stb = st_base[b], mdb = md_base[b], enb = en_base[b];
    const uint32_t nst = st_base[b + 1] - stb, nmd = md_base[b + 1] - mdb;
    const uint64_t n1 = (uint64_t)nst * nmd;
    const uint64_t *edge = st_list + stb;           // phase 1: head x middle (KmerTable.scala:118-122)
    if (q >= n1) { q -= n1; edge = en_list + enb; } // phase 2: tail x middle (:124-128)
    uint64_t x, y;
    if (q < 0x100000000ull) { x = (uint32_t)q / nmd; y = (uint32_t)q % nmd; } else { x = q / nmd; y = q % nmd; }
    const uint64_t e = edge[x], md = md_list[mdb + y];
    const uint32_t erid = (uint32_t)(e >> 32), mrid = (uint32_t)(md >> 32);
    K key = 0;                                      // same read: no pair (KmerTable.scala:61)
    if (erid != mrid) {
      // the read whose k-mer sits further along is `fst` (:63-71); ties and NaN -> the middle one
      const bool edge_first = __uint_as_float((uint32_t)e) > __uint_as_float((uint32_t)md);
      const uint32_t fst = edge_first ? erid : mrid, snd = edge_first ? mrid : erid;
      key = ((K)fst << id_bits) | (K)snd;
    }
    keys[o - o_lo] = key;
    if (SEQ) seq[o - o_lo] = (uint32_t)o;
  }
}

// ---------------------------------------------------------------- run-length encode
template <typename K>
struct RleIn {
  const K *key;
  __device__ Quad operator()(size_t i) const {
    const K k = key[i];
    const bool head = (i == 0) || (k != key[i - 1]);
    return Quad((head && k != 0) ? 1u : 0u, k == 0 ? 1u : 0u, 0u, 0u);
  }
};
template <typename K>
struct RleOut {
  const K *key; const uint32_t *seq; size_t n;
  uint64_t *run_key; uint32_t *run_start; uint32_t *run_seq;
  __device__ void operator()(size_t i, const Quad &v, const Quad &p) const {
    if (v.a) { run_key[p.a] = (uint64_t)key[i]; run_start[p.a] = (uint32_t)i; if (seq) run_seq[p.a] = seq[i]; }
    if (i == n - 1) run_start[p.a + v.a] = (uint32_t)n;
  }
};
__global__ void __launch_bounds__(256) run_count_kernel(const uint32_t *__restrict__ run_start, size_t nr, uint32_t *__restrict__ count) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nr) count[j] = run_start[j + 1] - run_start[j];
}

// merge of per-chunk partial lists: keys sorted (stable), value = (count << 32) | seq
struct MergeIn {
  const uint64_t *key;
  __device__ uint32_t operator()(size_t i) const { return (i == 0 || key[i] != key[i - 1]) ? 1u : 0u; }
};
struct MergeOut {
  const uint64_t *key; size_t n; uint64_t *out_key; uint32_t *run_start;
  __device__ void operator()(size_t i, const uint32_t &v, const uint32_t &p) const {
    if (v) { out_key[p] = key[i]; run_start[p] = (uint32_t)i; }
    if (i == n - 1) run_start[p + v] = (uint32_t)n;
  }
};
__global__ void __launch_bounds__(256) merge_sum_kernel(const uint32_t *__restrict__ run_start, const uint64_t *__restrict__ val,
                                                        size_t nr, uint32_t *__restrict__ count, uint32_t *__restrict__ seq) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nr) return;
  uint32_t s = 0;
  for (uint32_t i = run_start[j]; i < run_start[j + 1]; i++) s += (uint32_t)(val[i] >> 32);
  count[j] = s;
  if (seq) seq[j] = (uint32_t)val[run_start[j]];     // stable sort: first = earliest chunk = smallest rank
}
__global__ void __launch_bounds__(256) zip_kernel(const uint32_t *__restrict__ count, const uint32_t *__restrict__ seq, size_t n,
                                                  uint64_t *__restrict__ out) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = ((uint64_t)count[j] << 32) | (seq ? seq[j] : 0u);
}

struct Partial { size_t n = 0; DevBuf<uint64_t> key; DevBuf<uint32_t> count, seq; };

template <typename K, bool SEQ>
static void pairs_chunk(Ctx &c, const Table &B, const uint64_t *work_off, const uint32_t *work_bucket, uint32_t nw,
                        uint64_t o_lo, uint64_t o_hi, int id_bits, Partial &out, uint64_t &n_self, hovl_stats_t &st) {
  const size_t n = (size_t)(o_hi - o_lo);
  DevBuf<K> k0(c, n), k1(c, n);
  DevBuf<uint32_t> s0, s1;
  if (SEQ) { s0.alloc(c, n); s1.alloc(c, n); }
  {
    StageTimer tm(c, &st.ns_expand, "expand");
    HOVL_LAUNCH(c, (expand_kernel<K, SEQ>), (unsigned)div_up(n, EX_TILE), EX_THREADS, 0, o_lo, o_hi, work_off, work_bucket,
                nw, B.st_base.get(), B.md_base.get(), B.en_base.get(), B.st_list.get(), B.md_list.get(), B.en_list.get(),
                id_bits, k0.get(), s0.get());
  }
  int r;
  {
    StageTimer tm(c, &st.ns_pair_sort, "pair_sort");
    r = radix_sort<K, uint32_t, SEQ>(c, k0.get(), k1.get(), s0.get(), s1.get(), n, 0, 2 * id_bits);
  }
  StageTimer tm(c, &st.ns_pair_reduce, "pair_reduce");
  const K *ks = r ? k1.get() : k0.get();
  const uint32_t *ss = SEQ ? (r ? s1.get() : s0.get()) : nullptr;
  RleIn<K> in{ks};
  DevBuf<Quad> part;
  const Quad t = read_total(c, scan_prepare<Quad, RleIn<K>>(c, n, in, part));
  out.n = t.a; n_self += t.b;
  out.key.alloc(c, out.n); out.count.alloc(c, out.n);
  if (SEQ) out.seq.alloc(c, out.n);
  DevBuf<uint32_t> run_start(c, out.n + 1);
  RleOut<K> ro{ks, ss, n, out.key.get(), run_start.get(), SEQ ? out.seq.get() : nullptr};
  scan_apply<Quad, RleIn<K>, RleOut<K>>(c, n, in, ro, part);
  if (out.n) HOVL_LAUNCH(c, run_count_kernel, (unsigned)div_up(out.n, 256), 256, 0, run_start.get(), out.n, out.count.get());
}

void count_pairs(Ctx &c, const hovl_settings &s, const Table &B, const uint32_t *d_perm, uint32_t n_reads,
                 bool want_seq, Pairs &P, hovl_stats_t &st) {
  (void)s;
  int id_bits = 1;
  while (((uint64_t)1 << id_bits) <= (uint64_t)n_reads) id_bits++;       // ids are 1..n_reads
  P.id_bits = id_bits; P.n = 0; P.n_instances = 0;
  const size_t nb = B.n_buckets;
  HOVL_REQUIRE(nb < 0xffffffffull, HOVL_E_UNSUPPORTED, "too many buckets");
  DevBuf<uint32_t> work_bucket(c, nb + 1);
  DevBuf<uint64_t> work_off(c, nb + 1);
  HOVL_CUDA(cudaMemsetAsync(work_off.get(), 0, 8, c.stream));
  Duo tot;
  {
    StageTimer tm(c, &st.ns_expand, "expand");
    WorkIn in{d_perm, B.st_base.get(), B.md_base.get(), B.en_base.get()};
    WorkOut out{d_perm, work_bucket.get(), work_off.get(), nb};
    DevBuf<Duo> part;
    tot = read_total(c, scan_act<Duo, WorkIn, WorkOut>(c, nb, in, out, part));
  }
  const uint32_t nw = (uint32_t)tot.a;
  const uint64_t total = tot.b;
  if (want_seq) HOVL_REQUIRE(total < 0xffffffffull, HOVL_E_UNSUPPORTED, "reference order needs < 2^32 pair instances");
  const uint64_t chunk = c.pair_chunk < 0x7fffffffull ? c.pair_chunk : 0x7fffffffull;
  std::vector<Partial> parts;
  uint64_t n_self = 0;
  for (uint64_t lo = 0; lo < total; lo += chunk) {
    const uint64_t hi = lo + chunk < total ? lo + chunk : total;
    parts.emplace_back();
    Partial &pp = parts.back();
    if (2 * id_bits <= 32) {
      if (want_seq) pairs_chunk<uint32_t, true>(c, B, work_off.get(), work_bucket.get(), nw, lo, hi, id_bits, pp, n_self, st);
      else pairs_chunk<uint32_t, false>(c, B, work_off.get(), work_bucket.get(), nw, lo, hi, id_bits, pp, n_self, st);
    } else {
      if (want_seq) pairs_chunk<uint64_t, true>(c, B, work_off.get(), work_bucket.get(), nw, lo, hi, id_bits, pp, n_self, st);
      else pairs_chunk<uint64_t, false>(c, B, work_off.get(), work_bucket.get(), nw, lo, hi, id_bits, pp, n_self, st);
    }
  }
  P.n_instances = total - n_self;
  st.n_pair_instances = P.n_instances;
  if (parts.empty()) { P.key.alloc(c, 0); P.count.alloc(c, 0); if (want_seq) P.seq.alloc(c, 0); st.n_distinct_pairs = 0; return; }
  if (parts.size() == 1) {
    P.n = parts[0].n; P.key = std::move(parts[0].key); P.count = std::move(parts[0].count);
    if (want_seq) P.seq = std::move(parts[0].seq);
    st.n_distinct_pairs = P.n;
    return;
  }
  // ---- merge the per-chunk lists: concatenate in chunk order, stable sort by key, sum runs ----
  StageTimer tm(c, &st.ns_pair_reduce, "pair_reduce");
  size_t m = 0;
  for (auto &p : parts) m += p.n;
  HOVL_REQUIRE(m < 0xffffffffull, HOVL_E_UNSUPPORTED, "too many partial pairs to merge");
  DevBuf<uint64_t> k0(c, m), k1(c, m), v0(c, m), v1(c, m);
  size_t at = 0;
  for (auto &p : parts) {
    if (!p.n) continue;
    HOVL_CUDA(cudaMemcpyAsync(k0.get() + at, p.key.get(), p.n * 8, cudaMemcpyDeviceToDevice, c.stream));
    HOVL_LAUNCH(c, zip_kernel, (unsigned)div_up(p.n, 256), 256, 0, p.count.get(), want_seq ? p.seq.get() : nullptr, p.n, v0.get() + at);
    at += p.n;
  }
  parts.clear();
  const int r = radix_sort<uint64_t, uint64_t, true>(c, k0.get(), k1.get(), v0.get(), v1.get(), m, 0, 2 * id_bits);
  const uint64_t *ks = r ? k1.get() : k0.get(); const uint64_t *vs = r ? v1.get() : v0.get();
  MergeIn in{ks};
  DevBuf<uint32_t> part;
  const uint32_t nr = read_total(c, scan_prepare<uint32_t, MergeIn>(c, m, in, part));
  P.n = nr; P.key.alloc(c, nr); P.count.alloc(c, nr);
  if (want_seq) P.seq.alloc(c, nr);
  DevBuf<uint32_t> run_start(c, (size_t)nr + 1);
  MergeOut mo{ks, m, P.key.get(), run_start.get()};
  scan_apply<uint32_t, MergeIn, MergeOut>(c, m, in, mo, part);
  if (nr) HOVL_LAUNCH(c, merge_sum_kernel, (unsigned)div_up(nr, 256), 256, 0, run_start.get(), vs, (size_t)nr, P.count.get(),
                      want_seq ? P.seq.get() : nullptr);
  st.n_distinct_pairs = P.n;
}

// ---------------------------------------------------------------- count filter
struct SelIn {
  const uint32_t *count; long long lo, hi;
  __device__ uint32_t operator()(size_t i) const { const long long x = count[i]; return (lo <= x && x <= hi) ? 1u : 0u; }
};
struct SelOut {
  const uint64_t *key; const uint32_t *count; int id_bits; uint32_t *a, *b, *cnt;
  __device__ void operator()(size_t i, const uint32_t &v, const uint32_t &p) const {
    if (v) { const uint64_t k = key[i]; a[p] = (uint32_t)(k >> id_bits); b[p] = (uint32_t)(k & ((1ull << id_bits) - 1)); cnt[p] = count[i]; }
  }
};
size_t select_pairs(Ctx &c, const hovl_settings &s, const Pairs &P, DevBuf<uint32_t> &a, DevBuf<uint32_t> &b,
                    DevBuf<uint32_t> &cnt) {
  SelIn in{P.count.get(), (long long)s.min_coll, (long long)s.max_coll};
  DevBuf<uint32_t> part;
  const uint32_t nd = read_total(c, scan_prepare<uint32_t, SelIn>(c, P.n, in, part));
  a.alloc(c, nd); b.alloc(c, nd); cnt.alloc(c, nd);
  SelOut out{P.key.get(), P.count.get(), P.id_bits, a.get(), b.get(), cnt.get()};
  scan_apply<uint32_t, SelIn, SelOut>(c, P.n, in, out, part);
  return nd;
}

}  // namespace hovl
