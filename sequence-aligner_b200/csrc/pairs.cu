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
#include <algorithm>
#include <memory>
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

static void count_pairs_sorted(Ctx &c, const Table &B, const uint32_t *d_perm, uint32_t n_reads,
                               bool want_seq, Pairs &P, hovl_stats_t &st) {
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


// ---------------------------------------------------------------- per-read gather (canonical order)
// The same PairData (KmerTable.scala:57-149) computed read-major instead of bucket-major: a
// CTA owns one read `a`, walks a's k-mers, and for each one scans the section lists of its
// bucket for the instances in which a is `fst`:
//   a's k-mer is a head/tail one  x  middle k-mer of y :  a is fst  iff  loc_a > loc_y      (:63-71)
//   a's k-mer is a middle one     x  head/tail k-mer of y :  a is fst  iff  !(loc_y > loc_a)
// (every instance is seen from both of its reads and counted by exactly one of them).  The
// counts of a's partners live in a shared-memory hash table, so PairData's 10^9 instances
// are never written to memory, let alone sorted; the distinct (key,count) entries are appended
// to one list which is sorted once at the end.  A read with more partners than the table holds
// is redone with a table in global memory.
#ifndef HOVL_PG_WINDOWS
#define HOVL_PG_WINDOWS 2
#endif
constexpr int PG_WINDOWS = HOVL_PG_WINDOWS;   // 32-entry windows whose loads are in flight together
constexpr int PG_THREADS = 256;     // most threads a CTA may have; Ctx::gather_threads picks the launch size

struct PgArgs {
  const uint32_t *list; uint32_t n_items;            // reads to do (0-based), nullptr = first_item .. first_item + n_items
  uint32_t first_item;
  uint32_t keep_lo, keep_hi;                          // only pairs with keep_lo <= count <= keep_hi are written out
                                                      // (calcDispatchData's filter, KmerTable.scala:165-167, fused)
  const uint64_t *koff; const uint32_t *len; int32_t k;
  const uint64_t *ent;                                // sharded table: tuple koff[r]+i of read r = (bucket << 32) | loc bits
  const uint32_t *bucket_of; const uint4 *base3;      // base3[b] = (st, md, en) list starts of bucket b
  const uint64_t *st_list, *md_list, *en_list;
  float head, tail, mlead, mtail;
  int id_bits, log_size;                              // table slots = 1 << log_size
  int threads;                                        // CTA size of this launch
  uint32_t *gtable;                                   // global tables (2 * slots words per CTA) or nullptr = shared memory
  uint64_t *out_key; uint32_t *out_count; unsigned long long cap;
  // ctl[0] next item, [1] entries reserved, [2] sum of counts, [3] first failed reservation,
  // [4] entries the failed reads need, [5] reads whose table overflowed, [6] reads that did not fit the output,
  // [7] distinct pairs seen (before the count filter), [8] section-list entries read by the reads that completed
  unsigned long long *ctl;
  uint32_t *tlist, *olist;
};

// Open addressing with DOUBLE hashing, at most half full.  The 32 lanes of a warp insert together, so an insert costs the
// LONGEST probe sequence among them: on K4 (2,600 partners per read) linear probing in a table filled to 64 % took
// 560 thread-instructions per counted instance (ncu, profiles/r02_pair_gather_full_K4.txt: 7.5e11 warp instructions).
struct PgTable {
  uint32_t *keys, *cnt; uint32_t mask, shift, limit; uint32_t *ndist; volatile uint32_t *over;
  __device__ __forceinline__ void add(uint32_t y, uint32_t inc) {
    uint32_t slot = (y * 0x9E3779B1u) >> shift;
    const uint32_t step = ((y * 0x85EBCA6Bu) >> shift) | 1u;       // odd: visits every slot of a power-of-two table
    for (;;) {
      uint32_t cur = keys[slot];
      if (cur == 0u) {
        if (*over) return;
        cur = atomicCAS(&keys[slot], 0u, y);
        if (cur == 0u) { if (atomicAdd(ndist, 1u) + 1u >= limit) *over = 1u; cur = y; }
      }
      if (cur == y) { atomicAdd(&cnt[slot], inc); return; }
      if (*over) return;                                // table half full: the read is redone with a bigger one
      slot = (slot + step) & mask;
    }
  }
};

// One warp takes 32 consecutive k-mers of the read at a time: lane l looks up the bucket and
// the section ranges of k-mer 32*chunk + l, the non-empty ranges are laid end to end (warp scan,
// descriptors compacted in shared memory) and the warp then strides over the concatenated partner
// entries 32 at a time, so every lane has work and the loads of one step are independent of each
// other.  Which descriptor owns slot t of a 32-entry window is two warp-wide bit operations: the
// descriptors that START inside the window set one bit each (redux.or), and the owner of slot s is
// (descriptors that start before the window) + popc(bits <= s) - 1.
struct PgDesc {                       // per lane, in shared memory (warp-private)
  uint32_t off[3];                    // entry x of the concatenation is list[sect][off[sect] + x], sect = md, st, en
  uint32_t n01, n012;                 // |md part|, |md| + |st| parts
  uint32_t inc; float loc; uint32_t start;   // start = exclusive scan of the lane totals
};

// The table pointer is a run-time choice (shared memory, or global memory for the overflow tier), so its loads and atomics
// are generic ones (LD.E / ATOM.E).  A variant templated on the table's address space (LDS / ATOMS.CAS) was measured
// SLOWER on B200: 16.4 vs 11.5 ms on K3 (profiles/r02_bench_K3_v4.json), no change on K4 — kept generic.
__global__ void __launch_bounds__(PG_THREADS) pair_gather_kernel(const PgArgs A) {
  extern __shared__ __align__(16) uint32_t pg_smem[];
  __shared__ uint32_t s_item, s_ndist, s_over, s_emit, s_keep;
  __shared__ unsigned long long s_off, s_sum, s_read;
  __shared__ const uint64_t *s_lists[3];                 // md, st, en
  if (threadIdx.x == 0) { s_lists[0] = A.md_list; s_lists[1] = A.st_list; s_lists[2] = A.en_list; }
  const uint32_t slots = 1u << A.log_size;
  PgTable T;
  // dynamic shared memory: one PgDesc per thread, then the table (unless it lives in global memory)
  PgDesc *s_desc = reinterpret_cast<PgDesc *>(pg_smem);
  T.keys = A.gtable ? A.gtable + (size_t)blockIdx.x * 2 * slots : pg_smem + (size_t)blockDim.x * (sizeof(PgDesc) / 4);
  T.cnt = T.keys + slots;
  T.mask = slots - 1; T.shift = 32 - A.log_size; T.limit = slots / 2;
  T.ndist = &s_ndist; T.over = &s_over;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  PgDesc *wd = s_desc + warp * 32;
  for (;;) {
    __syncthreads();
    if (tid == 0) { s_item = (uint32_t)atomicAdd(&A.ctl[0], 1ull); s_ndist = 0; s_over = 0; s_emit = 0; s_sum = 0; s_keep = 0; s_read = 0; }
    for (uint32_t i = tid; i < 2 * slots; i += blockDim.x) T.keys[i] = 0u;
    __syncthreads();
    const uint32_t item = s_item;
    if (item >= A.n_items) break;
    const uint32_t r = A.list ? A.list[item] : A.first_item + item;
    const uint32_t a_id = r + 1u;
    const uint64_t base = A.koff[r];
    uint32_t nk; float denom = 0.f;
    if (A.ent) nk = (uint32_t)(A.koff[r + 1] - base);
    else {
      const uint32_t L = A.len[r];
      nk = L >= (uint32_t)A.k ? L - (uint32_t)A.k + 1u : 0u;
      denom = (float)(L - (uint32_t)A.k);
    }
    // Warp w takes the k-mers p with p mod (warps per CTA) == w, 32 of them at a time.  What a k-mer costs depends on where
    // it sits in the read (a head k-mer with loc <= mlead reads nothing, a k-mer in middle AND tail reads three lists), so
    // consecutive blocks of 32 k-mers per warp left most warps waiting for one at the barrier below: on K4 (139 k-mers,
    // one block per warp) half of all warp time (ncu, profiles/r02_pair_gather_full_K4.txt: 48.6 % of the stall samples).
    const uint32_t nwarp = blockDim.x >> 5;
    for (uint32_t p0 = 0; p0 + warp < nk; p0 += blockDim.x) {
      if (__any_sync(0xffffffffu, *(volatile uint32_t *)&s_over)) break;
      const uint32_t p = p0 + lane * nwarp + warp;
      uint32_t n_md = 0, n_st = 0, n_en = 0;
      PgDesc d; d.inc = 0; d.loc = 0.f;
      uint32_t lo_md = 0, lo_st = 0, lo_en = 0;
      if (p < nk) {
        uint32_t b;
        if (A.ent) { const uint64_t e = A.ent[base + p]; b = (uint32_t)(e >> 32); d.loc = __uint_as_float((uint32_t)e); }
        else { b = A.bucket_of[base + p]; d.loc = __fdiv_rn((float)p, denom); }   // NaN when len == k: in no section
        d.inc = (d.loc <= A.head ? 1u : 0u) + (A.tail <= d.loc ? 1u : 0u);
        const uint4 q0 = A.base3[b], q1 = A.base3[b + 1];
        // Only the instances in which a is fst are counted here, and the section a list belongs to bounds
        // the loc of its entries: a middle entry has loc_y >= mlead, so an edge k-mer of a with
        // loc <= mlead is never fst (needs loc > loc_y); a tail entry has loc_y >= tail, so a middle
        // k-mer of a with loc < tail is never fst (needs loc >= loc_y).  Those lists are not read.
        if (d.inc && d.loc > A.mlead) { lo_md = q0.y; n_md = q1.y - q0.y; }
        if (A.mlead <= d.loc && d.loc <= A.mtail) {
          lo_st = q0.x; n_st = q1.x - q0.x;
          if (!(d.loc < A.tail)) { lo_en = q0.z; n_en = q1.z - q0.z; }
        }
      }
      d.n01 = n_md; d.n012 = n_md + n_st;
      d.off[0] = lo_md; d.off[1] = lo_st - d.n01; d.off[2] = lo_en - d.n012;     // wraps are undone by + x
      const uint32_t tot = n_md + n_st + n_en;
      uint32_t incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
      d.start = incl - tot;
      const uint32_t nonempty = __ballot_sync(0xffffffffu, tot != 0u);
      const uint32_t m = __popc(nonempty);
      if (tot) wd[__popc(nonempty & ((1u << lane) - 1u))] = d;
      __syncwarp();
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      if (lane == 0 && total) atomicAdd(&s_read, (unsigned long long)total);
      const uint32_t S = (uint32_t)lane < m ? wd[lane].start : 0xffffffffu;        // start of compacted descriptor `lane`
      for (uint32_t w0 = 0; w0 < total; w0 += 32 * PG_WINDOWS) {
        uint64_t v[PG_WINDOWS]; uint32_t inc[PG_WINDOWS]; float la[PG_WINDOWS]; bool edge[PG_WINDOWS], live[PG_WINDOWS];
#pragma unroll
        for (int j = 0; j < PG_WINDOWS; j++) {
          const uint32_t w = w0 + j * 32, t = w + lane;
          const uint32_t heads = __reduce_or_sync(0xffffffffu, (S - w) < 32u ? 1u << (S - w) : 0u);
          const uint32_t before = __popc(__ballot_sync(0xffffffffu, S < w));
          live[j] = t < total;
          v[j] = 0; inc[j] = 0; la[j] = 0.f; edge[j] = false;
          if (live[j]) {
            const PgDesc &q = wd[before + __popc(heads & (0xffffffffu >> (31 - lane))) - 1u];
            const uint32_t x = t - q.start;
            const uint32_t sect = (x >= q.n01 ? 1u : 0u) + (x >= q.n012 ? 1u : 0u);
            edge[j] = sect == 0u; inc[j] = sect ? 1u : q.inc; la[j] = q.loc;
            v[j] = s_lists[sect][q.off[sect] + x];
          }
        }
#pragma unroll
        for (int j = 0; j < PG_WINDOWS; j++) {
          if (!live[j]) continue;
          const uint32_t y = (uint32_t)(v[j] >> 32);
          const float loc_y = __uint_as_float((uint32_t)v[j]);
          const bool mine = edge[j] ? (la[j] > loc_y) : !(loc_y > la[j]);
          if (y != a_id && mine) T.add(y, inc[j]);
        }
      }
      __syncwarp();
    }
    __syncthreads();
    if (s_over) { if (tid == 0) A.tlist[atomicAdd(&A.ctl[5], 1ull)] = r; continue; }
    const uint32_t ndist = s_ndist;
    if (ndist == 0) { if (tid == 0 && s_read) atomicAdd(&A.ctl[8], s_read); continue; }
    // pass over the table: sum of counts, how many entries survive the count filter
    unsigned long long sum = 0; uint32_t keep = 0;
    for (uint32_t i = tid; i < slots; i += blockDim.x) {
      if (T.keys[i]) { const uint32_t cn = T.cnt[i]; sum += cn; keep += (cn >= A.keep_lo && cn <= A.keep_hi) ? 1u : 0u; }
    }
    for (int o = 16; o; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); keep += __shfl_xor_sync(0xffffffffu, keep, o); }
    if (lane == 0) { if (sum) atomicAdd(&s_sum, sum); if (keep) atomicAdd(&s_keep, keep); }
    __syncthreads();
    const uint32_t nd = s_keep;
    if (tid == 0) {
      unsigned long long off = 0;
      if (nd) off = atomicAdd(&A.ctl[1], (unsigned long long)nd);
      s_off = off;
      if (nd && off + nd > A.cap) {
        atomicMin(&A.ctl[3], off); atomicAdd(&A.ctl[4], (unsigned long long)nd);
        A.olist[atomicAdd(&A.ctl[6], 1ull)] = r;
      } else {
        atomicAdd(&A.ctl[2], s_sum); atomicAdd(&A.ctl[7], (unsigned long long)ndist); atomicAdd(&A.ctl[8], s_read);
      }
    }
    __syncthreads();
    const unsigned long long off = s_off;
    if (nd == 0 || off + nd > A.cap) continue;
    for (uint32_t i = tid; i < slots; i += blockDim.x) {
      const uint32_t y = T.keys[i];
      if (y) {
        const uint32_t cn = T.cnt[i];
        if (cn >= A.keep_lo && cn <= A.keep_hi) {
          const uint32_t at = atomicAdd(&s_emit, 1u);
          A.out_key[off + at] = ((uint64_t)a_id << A.id_bits) | y;
          A.out_count[off + at] = cn;
        }
      }
    }
  }
}

struct PgPiece { DevBuf<uint64_t> key; DevBuf<uint32_t> count; size_t n = 0; };

// Runs the gather over `list` (or all reads) until every read either has been emitted or is on
// the table-overflow list; returns the number of reads on that list (device array tlist).
static uint32_t gather_run(Ctx &c, PgArgs a, bool global_table, size_t cap_guess, std::vector<PgPiece> &pieces,
                           DevBuf<uint32_t> &tlist, uint64_t &sum_counts, uint64_t &n_distinct, uint64_t &n_entries) {
  DevBuf<unsigned long long> ctl(c, 9);
  DevBuf<uint32_t> olist[2];
  tlist.alloc(c, a.n_items ? a.n_items : 1);
  a.tlist = tlist.get();
  const uint32_t slots = 1u << a.log_size;
  const int threads = a.threads;
  const size_t smem = (global_table ? 0 : (size_t)slots * 8) + (size_t)threads * sizeof(PgDesc);
  HOVL_CUDA(cudaFuncSetAttribute(pair_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  HOVL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pair_gather_kernel, threads, smem));
  if (per_sm < 1) per_sm = 1;
  uint32_t n_tover = 0;
  int cur = 0;
  size_t cap = cap_guess;
  for (int round = 0; a.n_items; round++) {
    unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)c.sm_count * per_sm, a.n_items);
    DevBuf<uint32_t> gt;
    if (global_table) {
      const uint64_t budget = (uint64_t)1 << 31;                        // bytes of global tables
      grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(grid, budget / ((uint64_t)slots * 8)));
      gt.alloc(c, (size_t)grid * 2 * slots);
      a.gtable = gt.get();
    }
    pieces.emplace_back();
    PgPiece &pc = pieces.back();
    pc.key.alloc(c, cap); pc.count.alloc(c, cap);
    olist[cur].alloc(c, a.n_items);
    a.out_key = pc.key.get(); a.out_count = pc.count.get(); a.cap = cap; a.ctl = ctl.get(); a.olist = olist[cur].get();
    unsigned long long h[9] = {0, 0, 0, ~0ull, 0, n_tover, 0, 0, 0};
    h2d(c, ctl.get(), h, 9);
    HOVL_LAUNCH(c, pair_gather_kernel, grid, threads, smem, a);
    d2h(c, h, ctl.get(), 9);
    sync(c);
    pc.n = (size_t)std::min<unsigned long long>(h[1], h[3]);
    sum_counts += h[2]; n_distinct += h[7]; n_entries += h[8];
    n_tover = (uint32_t)h[5];
    if (h[6] == 0) break;
    HOVL_REQUIRE(round < 64, HOVL_E_CUDA, "internal: pair gather does not converge");
    a.list = olist[cur].get(); a.n_items = (uint32_t)h[6];
    cap = (size_t)h[4];
    cur ^= 1;
  }
  return n_tover;
}

void count_pairs_gather(Ctx &c, const hovl_settings &s, const GatherSrc &R, const Table &B, bool sort_result, bool filter,
                        Pairs &P, hovl_stats_t &st) {
  int id_bits = 1;
  while (((uint64_t)1 << id_bits) <= (uint64_t)R.n_reads) id_bits++;
  P.id_bits = id_bits; P.n = 0; P.n_instances = 0;
  std::vector<PgPiece> pieces;
  uint64_t sum_counts = 0, n_distinct = 0, n_entries = 0;
  {
    StageTimer tm(c, &st.ns_expand, "pair_gather");
    PgArgs a;
    a.list = nullptr; a.n_items = R.count ? R.count : R.n_reads; a.first_item = R.count ? R.first : 0;
    a.koff = R.koff; a.len = R.len; a.k = s.k; a.ent = R.ent;
    a.bucket_of = B.bucket_of.get(); a.base3 = B.base3.get();
    a.st_list = B.st_list.get(); a.md_list = B.md_list.get(); a.en_list = B.en_list.get();
    a.head = __builtin_bit_cast(float, s.head_f32); a.tail = __builtin_bit_cast(float, s.tail_f32);
    a.mlead = __builtin_bit_cast(float, s.mid_lead_f32); a.mtail = __builtin_bit_cast(float, s.mid_tail_f32);
    a.id_bits = id_bits; a.gtable = nullptr;
    // A warp takes 32 k-mers of the read at a time and the CTA meets at a barrier when the read is done: a read of 139
    // k-mers (K4) on 128 threads is one full round and one round in which a single warp works on 11 k-mers (ncu: 11.3
    // barrier stall cycles per issue).  Reads whose k-mers fit one round get a CTA of exactly that many warps.
    a.threads = c.gather_threads;
    if (c.gather_threads_auto && !R.ent && R.max_kmers && R.max_kmers <= (uint32_t)PG_THREADS) a.threads = std::max(64, (int)((R.max_kmers + 31) / 32) * 32);
    a.keep_lo = 0; a.keep_hi = 0xffffffffu;
    if (filter) {                                          // counts are >= 1 and < 2^32
      a.keep_lo = s.min_coll > 0 ? (uint32_t)s.min_coll : 0u;
      a.keep_hi = s.max_coll < 0 ? 0u : (uint32_t)s.max_coll;
      if (s.max_coll < 0) a.keep_lo = 1, a.keep_hi = 0;    // nothing passes
    }
    size_t guess = c.pair_cap_hint ? c.pair_cap_hint + c.pair_cap_hint / 16 + 1024
                                   : std::max<size_t>((size_t)1 << 20, (size_t)R.n_tuples / 8);
    // Shared-memory tables of 1024 .. 16384 slots, then a table in global memory (a 256-slot
    // tier was measured slower on K3: 14.6 vs 13.8 ms).  The table size sets how many CTAs fit an SM.  A pilot over
    // the first reads finds the smallest size that holds (nearly) every read's partners; a read
    // that overflows its table is redone one size up.
    std::vector<int> logs = {10, 11, 12, 13, 14};
    if (c.pair_table_log) logs = {c.pair_table_log, c.pair_table_log + 2};
    std::vector<std::unique_ptr<DevBuf<uint32_t>>> lists;
    std::vector<uint32_t> overflow_at(logs.size() + 1, 0);
    auto cascade = [&](uint32_t first, uint32_t count, size_t tier0, size_t cap0) {
      PgArgs q = a;
      q.list = nullptr; q.first_item = first; q.n_items = count;
      uint32_t left = count;
      for (size_t t = tier0; t < logs.size() && left; t++) {
        q.log_size = logs[t];
        lists.emplace_back(new DevBuf<uint32_t>());
        left = gather_run(c, q, false, t == tier0 ? cap0 : std::max<size_t>((size_t)1 << 16, (size_t)left * 1024), pieces,
                          *lists.back(), sum_counts, n_distinct, n_entries);
        overflow_at[t] = left;
        q.list = lists.back()->get(); q.n_items = left;
      }
      if (left) {
        // partners of one read <= reads - 1: a global table of more than twice that many slots always fits them
        q.log_size = 1;
        while (((uint64_t)1 << q.log_size) < 2 * ((uint64_t)R.n_reads + 2)) q.log_size++;
        HOVL_REQUIRE(q.log_size <= 28, HOVL_E_UNSUPPORTED, "too many reads for the pair-count overflow table");
        lists.emplace_back(new DevBuf<uint32_t>());
        left = gather_run(c, q, true, std::max<size_t>((size_t)1 << 16, (size_t)left * 4096), pieces, *lists.back(),
                          sum_counts, n_distinct, n_entries);
        HOVL_REQUIRE(left == 0, HOVL_E_CUDA, "internal: pair-count overflow table overflowed");
      }
    };
    const uint32_t first = R.count ? R.first : 0u, n_sub = R.count ? R.count : R.n_reads;   // the reads whose pairs are counted here
    const uint32_t n_pilot = std::min<uint32_t>(n_sub, 4096);
    const size_t pilot_cap = std::max<size_t>((size_t)1 << 16, n_sub ? guess / n_sub * n_pilot * 2 : 0);
    cascade(first, n_pilot, 0, pilot_cap);
    size_t tier = 0;
    while (tier + 1 < logs.size() && overflow_at[tier] > n_pilot / 64) tier++;
    if (n_sub > n_pilot) cascade(first + n_pilot, n_sub - n_pilot, tier, guess);
  }
  StageTimer tm(c, &st.ns_pair_sort, "pair_sort");
  size_t m = 0;
  for (auto &p : pieces) m += p.n;
  c.pair_cap_hint = m;
  P.n_instances = sum_counts; st.n_pair_instances = sum_counts; st.n_distinct_pairs = n_distinct;
  st.n_list_entries = n_entries;
  P.n = m;
  if (m == 0) { P.key.alloc(c, 0); P.count.alloc(c, 0); return; }
  DevBuf<uint64_t> k0(c, m), k1(c, m);
  DevBuf<uint32_t> v0(c, m), v1(c, m);
  size_t at = 0;
  for (auto &p : pieces) {
    if (!p.n) continue;
    HOVL_CUDA(cudaMemcpyAsync(k0.get() + at, p.key.get(), p.n * 8, cudaMemcpyDeviceToDevice, c.stream));
    HOVL_CUDA(cudaMemcpyAsync(v0.get() + at, p.count.get(), p.n * 4, cudaMemcpyDeviceToDevice, c.stream));
    at += p.n;
  }
  pieces.clear();
  const int r = sort_result ? radix_sort<uint64_t, uint32_t, true>(c, k0.get(), k1.get(), v0.get(), v1.get(), m, 0, 2 * id_bits) : 0;
  P.key = std::move(r ? k1 : k0); P.count = std::move(r ? v1 : v0);
}

void count_pairs(Ctx &c, const hovl_settings &s, const Reads &R, const Table &B, const uint32_t *d_perm,
                 bool want_seq, bool filter, Pairs &P, hovl_stats_t &st) {
  if (!want_seq && !d_perm && B.bucket_of.get() != nullptr && c.pair_mode == 0) {
    GatherSrc src;
    src.n_reads = R.n_reads; src.koff = R.koff.get(); src.len = R.len.get(); src.ent = nullptr; src.n_tuples = R.n_kmers;
    src.max_kmers = R.max_len >= (uint32_t)s.k ? R.max_len - (uint32_t)s.k + 1u : 0u;
    count_pairs_gather(c, s, src, B, true, filter, P, st);
  } else count_pairs_sorted(c, B, d_perm, R.n_reads, want_seq, P, st);
}

// Partial (key, count) lists from several ranks, concatenated: sort by key, add up equal keys
// (the second exchange of the sharded path).  Clobbers keys/counts.
__global__ void __launch_bounds__(256) merge_sum32_kernel(const uint32_t *__restrict__ run_start, const uint32_t *__restrict__ val,
                                                          size_t nr, uint32_t *__restrict__ count) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nr) return;
  uint32_t s = 0;
  for (uint32_t i = run_start[j]; i < run_start[j + 1]; i++) s += val[i];
  count[j] = s;
}
void merge_partial_pairs(Ctx &c, uint64_t *keys, uint32_t *counts, size_t m, uint32_t n_reads, Pairs &P, hovl_stats_t &st) {
  StageTimer tm(c, &st.ns_pair_reduce, "pair_reduce");
  int id_bits = 1;
  while (((uint64_t)1 << id_bits) <= (uint64_t)n_reads) id_bits++;
  P.id_bits = id_bits; P.n = 0;
  HOVL_REQUIRE(m < 0xffffffffull, HOVL_E_UNSUPPORTED, "too many partial pairs to merge");
  if (m == 0) { P.key.alloc(c, 0); P.count.alloc(c, 0); return; }
  DevBuf<uint64_t> k1(c, m); DevBuf<uint32_t> v1(c, m);
  const int r = radix_sort<uint64_t, uint32_t, true>(c, keys, k1.get(), counts, v1.get(), m, 0, 2 * id_bits);
  const uint64_t *ks = r ? k1.get() : keys; const uint32_t *vs = r ? v1.get() : counts;
  MergeIn in{ks};
  DevBuf<uint32_t> part;
  const uint32_t nr = read_total(c, scan_prepare<uint32_t, MergeIn>(c, m, in, part));
  P.n = nr; P.key.alloc(c, nr); P.count.alloc(c, nr);
  DevBuf<uint32_t> run_start(c, (size_t)nr + 1);
  MergeOut mo{ks, m, P.key.get(), run_start.get()};
  scan_apply<uint32_t, MergeIn, MergeOut>(c, m, in, mo, part);
  if (nr) HOVL_LAUNCH(c, merge_sum32_kernel, (unsigned)div_up(nr, 256), 256, 0, run_start.get(), vs, (size_t)nr, P.count.get());
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
