// kmer.cu — splitter and KmerTable stages on the device.
//   build_reads   : per-read metadata (one scan) + ASCII -> 2-bit packing
//   encode_kmers  : every k-window -> (code, read id, loc) tuple   [BioLibs.scala:54-61]
//   sort_kmers    : stable LSD radix sort by code                   [KmerTable.scala:41-53]
//   group_kmers   : bucket boundaries + head/middle/tail section lists [KmerTable.scala:105-116]
#include "engine.cuh"
#include "scan.cuh"
#include "radix_sort.cuh"
#include "dp_core.cuh"

namespace hovl {

// ---------------------------------------------------------------- read metadata
struct MetaIn {
  const uint64_t *off; int32_t k;
  __device__ Duo operator()(size_t r) const {
    const uint64_t L = off[r + 1] - off[r];
    return Duo((L + 15) >> 4, L >= (uint64_t)k ? L - (uint64_t)k + 1 : 0);
  }
};
struct MetaOut {
  const uint64_t *off; uint64_t *woff, *koff; uint32_t *len, *bad; uint32_t *max_len; size_t n;   // max_len[0] = max, max_len[1] = min
  __device__ void operator()(size_t r, const Duo &v, const Duo &p) const {
    woff[r] = p.a; koff[r] = p.b;
    const uint32_t L = (uint32_t)(off[r + 1] - off[r]);
    len[r] = L; bad[r] = 0;
    atomicMax(max_len, L); atomicMin(max_len + 1, L);
    if (r == n - 1) { woff[n] = p.a + v.a; koff[n] = p.b + v.b; }
  }
};

// upper_bound(prefix, x) - 1 : the segment that holds element x
__device__ __forceinline__ uint32_t find_segment(const uint64_t *__restrict__ prefix, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;                 // prefix[lo] <= x < prefix[hi]
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (prefix[mid] <= x) lo = mid; else hi = mid; }
  return lo;
}

// One thread per packed word: 16 ASCII bases -> 32 bits, first base in the top bits.
__global__ void __launch_bounds__(256) pack_kernel(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ off,
                                                   const uint64_t *__restrict__ woff, uint32_t n_reads, uint64_t n_words,
                                                   uint32_t *__restrict__ packed, uint32_t *__restrict__ bad) {
  const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_words) return;
  const uint32_t r = find_segment(woff, n_reads, g);
  const uint64_t L = off[r + 1] - off[r];
  const uint64_t p0 = (g - woff[r]) << 4;
  const uint8_t *src = bases + off[r] + p0;
  const int cnt = (int)((L - p0) < 16 ? (L - p0) : 16);
  uint32_t w = 0; bool any_bad = false;
#pragma unroll
  for (int t = 0; t < 16; t++) {
    uint32_t v = 0;
    if (t < cnt) {
      const uint8_t ch = src[t];
      // A0 C1 G2 T3 (costFunc order, BioLibs.scala:142-160); anything else counts as 0 in the
      // k-mer code (ObjectStore.scala:60-63) and marks the read as unalignable.
      if (ch == 'A') v = 0; else if (ch == 'C') v = 1; else if (ch == 'G') v = 2; else if (ch == 'T') v = 3;
      else any_bad = true;
    }
    w |= v << (30 - 2 * t);
  }
  packed[g] = w;
  if (any_bad) bad[r] = 1;
}

void build_reads(Ctx &c, const hovl_settings &s, Reads &R, hovl_stats_t &st) {
  StageTimer tm(c, &st.ns_pack, "pack");
  const size_t n = R.n_reads;
  R.woff.alloc(c, n + 1); R.koff.alloc(c, n + 1); R.len.alloc(c, n); R.bad.alloc(c, n);
  DevBuf<uint32_t> maxlen(c, 2);
  HOVL_CUDA(cudaMemsetAsync(maxlen.get(), 0, 4, c.stream));
  HOVL_CUDA(cudaMemsetAsync(maxlen.get() + 1, 0xff, 4, c.stream));
  HOVL_CUDA(cudaMemsetAsync(R.woff.get(), 0, 8, c.stream));
  HOVL_CUDA(cudaMemsetAsync(R.koff.get(), 0, 8, c.stream));
  DevBuf<Duo> part;
  MetaIn in{R.d_offsets, s.k};
  MetaOut out{R.d_offsets, R.woff.get(), R.koff.get(), R.len.get(), R.bad.get(), maxlen.get(), n};
  const Duo *tot = scan_act<Duo, MetaIn, MetaOut>(c, n, in, out, part);
  const Duo t = read_total(c, tot);
  uint32_t mm[2] = {0, 0};
  HOVL_CUDA(cudaMemcpyAsync(mm, maxlen.get(), 8, cudaMemcpyDeviceToHost, c.stream));
  sync(c);
  R.n_words = t.a; R.n_kmers = t.b; R.max_len = mm[0]; R.min_len = n ? mm[1] : 0;
  R.packed.alloc(c, R.n_words + 2);
  HOVL_CUDA(cudaMemsetAsync(R.packed.get() + R.n_words, 0, 8, c.stream));
  if (R.n_words)
    HOVL_LAUNCH(c, pack_kernel, (unsigned)div_up(R.n_words, 256), 256, 0, R.d_bases, R.d_offsets, R.woff.get(),
                (uint32_t)n, R.n_words, R.packed.get(), R.bad.get());
  st.n_reads = n; st.n_bases = R.n_bases; st.n_kmers = R.n_kmers;
}

// ---------------------------------------------------------------- splitter
// One thread per packed word = 16 consecutive window starts.  The next word of the read is
// taken from the neighbouring lane (warp-shuffle rolling window); a window of 16 bases
// starting t bases into the word is one funnel shift.  The tuples of a CTA's 256 words are
// consecutive in the output (koff counts exactly the windows that exist), so they are staged in
// shared memory (XOR-swizzled: a thread's 16 tuples are 16 elements apart from its neighbour's)
// and written out as full 128-byte lines.
constexpr int ENC_THREADS = 256, ENC_TILE = ENC_THREADS * 16;
__device__ __forceinline__ uint32_t enc_swz(uint32_t i) { return i ^ ((i >> 4) & 15u); }

__global__ void __launch_bounds__(ENC_THREADS) encode_kernel(const uint32_t *__restrict__ packed, const uint64_t *__restrict__ woff,
                                                             const uint64_t *__restrict__ koff, const uint32_t *__restrict__ len,
                                                             uint32_t n_reads, uint64_t n_words, int32_t k, uint32_t id_base,
                                                             uint32_t *__restrict__ keys, uint64_t *__restrict__ vals,
                                                             uint32_t *__restrict__ idx /* compact form: vals == nullptr */) {
  extern __shared__ __align__(16) unsigned char enc_smem[];
  uint64_t *s_val = reinterpret_cast<uint64_t *>(enc_smem);                   // ENC_TILE
  uint32_t *s_key = reinterpret_cast<uint32_t *>(enc_smem + (size_t)ENC_TILE * 8);   // ENC_TILE
  __shared__ unsigned long long s_begin;
  __shared__ uint32_t s_count;
  const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = g < n_words;
  const uint32_t w0 = live ? packed[g] : 0u;
  uint32_t w1 = __shfl_down_sync(0xffffffffu, w0, 1);
  if ((threadIdx.x & 31) == 31) w1 = live ? packed[g + 1] : 0u;      // slack words are zero
  uint32_t r = 0, L = 0, p0 = 0, nwin = 0;                             // nwin = windows of read r
  uint64_t o0 = 0;
  if (live) {
    r = find_segment(woff, n_reads, g);
    L = len[r];
    nwin = L >= (uint32_t)k ? L - (uint32_t)k + 1u : 0u;
    p0 = (uint32_t)(g - woff[r]) << 4;
    if (g + 1 >= woff[r + 1]) w1 = 0;                                 // neighbour belongs to the next read
    o0 = koff[r] + (p0 < nwin ? p0 : nwin);
  }
  if (threadIdx.x == 0) { s_begin = o0; s_count = 0; }                // thread 0 is always live
  __syncthreads();
  const uint32_t rel = (uint32_t)(o0 - s_begin);
  const uint32_t mine = (live && p0 < nwin) ? (nwin - p0 < 16u ? nwin - p0 : 16u) : 0u;
  if (mine) {
    const uint32_t last = L - (uint32_t)k;                            // last window start
    const int kk = k < 16 ? k : 16;                                   // only 16 bases are hashed
    const float denom = (float)last;                                  // (len - k).toFloat
    const uint64_t rid = (uint64_t)(id_base + r + 1) << 32;           // ids are 1-based (BioLibs.scala:27)
#pragma unroll
    for (int t = 0; t < 16; t++) {
      const uint32_t p = p0 + t;
      if (p <= last) {
        const uint32_t win = __funnelshift_l(w1, w0, 2 * t);
        const uint32_t idx = kk == 16 ? win : (win >> (32 - 2 * kk));
        // A0 C1 G2 T3 -> the reference's A0 C1 T2 G3 (ObjectStore.scala:55-59): swap codes 2,3
        const uint32_t code = idx ^ ((idx >> 1) & 0x55555555u);
        const float loc = __fdiv_rn((float)p, denom);                 // i.toFloat / (len-k).toFloat
        const uint32_t at = enc_swz(rel + t);
        s_key[at] = code;
        // L == k gives 0f/0f: keep Java's canonical NaN bits (Float.floatToIntBits), not the GPU's 0x7fffffff
        if (vals) s_val[at] = rid | (uint64_t)(loc != loc ? 0x7fc00000u : __float_as_uint(loc));
      }
    }
    atomicMax(&s_count, rel + mine);
  }
  __syncthreads();
  const uint32_t count = s_count;
  const uint64_t begin = s_begin;
  for (uint32_t i = threadIdx.x; i < count; i += ENC_THREADS) {
    const uint32_t at = enc_swz(i);
    keys[begin + i] = s_key[at];
    if (vals) vals[begin + i] = s_val[at];
    else idx[begin + i] = (uint32_t)(begin + i);
  }
}

void encode_kmers(Ctx &c, const hovl_settings &s, const Reads &R, Tuples &T, hovl_stats_t &st, bool allow_compact) {
  StageTimer tm(c, &st.ns_encode, "encode");
  HOVL_REQUIRE(R.n_kmers < 0xfffffff0ull, HOVL_E_UNSUPPORTED, "more than 2^32 k-mers on one device");
  T.n = R.n_kmers; T.cur = 0;
  T.compact = allow_compact && c.compact_tuples && R.n_reads && R.min_len == R.max_len && R.max_len >= (uint32_t)s.k;
  st.kmer_tuple_bytes = T.compact ? 8 : 12;
  for (int i = 0; i < 2; i++) {
    T.key[i].alloc(c, T.n);
    if (T.compact) T.idx[i].alloc(c, T.n); else T.val[i].alloc(c, T.n);
  }
  const size_t smem = (size_t)ENC_TILE * 12;
  HOVL_CUDA(cudaFuncSetAttribute(encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (R.n_words)
    HOVL_LAUNCH(c, encode_kernel, (unsigned)div_up(R.n_words, ENC_THREADS), ENC_THREADS, smem, R.packed.get(), R.woff.get(),
                R.koff.get(), R.len.get(), R.n_reads, R.n_words, s.k, R.id_base, T.k(), T.compact ? nullptr : T.v(),
                T.compact ? T.ix() : nullptr);
}

void sort_kmers(Ctx &c, const hovl_settings &s, Tuples &T, hovl_stats_t &st) {
  StageTimer tm(c, &st.ns_kmer_sort, "kmer_sort");
  const int bits = 2 * (s.k < 16 ? s.k : 16);
  const int r = T.compact
      ? radix_sort<uint32_t, uint32_t, true>(c, T.key[T.cur].get(), T.key[T.cur ^ 1].get(), T.idx[T.cur].get(),
                                             T.idx[T.cur ^ 1].get(), T.n, 0, bits)
      : radix_sort<uint32_t, uint64_t, true>(c, T.key[T.cur].get(), T.key[T.cur ^ 1].get(), T.val[T.cur].get(),
                                             T.val[T.cur ^ 1].get(), T.n, 0, bits);
  T.cur ^= r;
}

// ---------------------------------------------------------------- buckets + sections
// calcPairData's three filters (KmerTable.scala:105-116): head  loc <= e,
// middle  midLead <= loc <= midTail,  tail  1-e <= loc.  NaN (len == k) fails all of them.
// The value (read id << 32 | loc bits) of sorted tuple i: stored, or recomputed from the read-major index
// when every read has the same length (idx = read * nwin + window start; loc as the splitter computes it).
struct ValSrc {
  const uint64_t *val; const uint32_t *idx; uint32_t nwin, id_base; float denom;
  __device__ __forceinline__ uint64_t at(size_t i) const {
    if (!idx) return val[i];
    const uint32_t t = idx[i], r = t / nwin, pos = t - r * nwin;
    const float loc = __fdiv_rn((float)pos, denom);
    return ((uint64_t)(id_base + r + 1u) << 32) | (uint64_t)(loc != loc ? 0x7fc00000u : __float_as_uint(loc));
  }
};
struct GroupIn {
  const uint32_t *key; ValSrc val; float head, tail, mlead, mtail;
  __device__ Quad operator()(size_t i) const {
    const float loc = __uint_as_float((uint32_t)val.at(i));
    const uint32_t hd = (i == 0 || key[i] != key[i - 1]) ? 1u : 0u;
    return Quad(hd, loc <= head ? 1u : 0u, (mlead <= loc && loc <= mtail) ? 1u : 0u, tail <= loc ? 1u : 0u);
  }
};
struct GroupOut {
  ValSrc val; size_t n;
  uint64_t *bucket_start; uint32_t *st_base, *md_base, *en_base;
  uint64_t *st_list, *md_list, *en_list;
  const uint32_t *len; const uint64_t *koff; int32_t k; uint32_t *bucket_of;   // read-major tuple index -> bucket
  uint64_t *ent; const uint64_t *ekoff; uint32_t *ecursor;                     // k-mer shard: per-read (bucket, loc) entries
  uint4 *base3;                                                                // (st, md, en) bases of a bucket side by side
  uint32_t uni_len, id0;                                                       // common read length (0 = ragged); id of the first local read - 1
  // read-major index of the tuple: koff[read] + window start, where
  // window start = round(loc * (len-k)): loc = fl(i / (len-k)) is within 2^-23 * i of i / (len-k)
  __device__ uint64_t dest_of(uint64_t x) const {
    const uint32_t r = (uint32_t)(x >> 32) - 1u;
    if (uni_len) {                          // every read has the same length: no lookups
      const uint32_t d = uni_len - (uint32_t)k;
      const uint32_t pos = d ? __float2uint_rn(__fmul_rn(__uint_as_float((uint32_t)x), (float)d)) : 0u;
      return (uint64_t)(r - id0) * (d + 1u) + pos;
    }
    const uint32_t d = len[r - id0] - (uint32_t)k;
    const uint32_t pos = d ? __float2uint_rn(__fmul_rn(__uint_as_float((uint32_t)x), (float)d)) : 0u;
    return koff[r - id0] + pos;
  }
  __device__ void operator()(size_t i, const Quad &v, const Quad &p) const {
    const uint64_t x = val.at(i);
    if (v.a) {
      bucket_start[p.a] = i; st_base[p.a] = p.b; md_base[p.a] = p.c; en_base[p.a] = p.d;
      if (base3) base3[p.a] = make_uint4(p.b, p.c, p.d, 0u);
    }
    if (bucket_of) bucket_of[dest_of(x)] = p.a + v.a - 1u;
    if (ent) {                              // order inside a read does not matter to the pair counts
      const uint32_t r = (uint32_t)(x >> 32) - 1u;
      ent[ekoff[r] + atomicAdd(&ecursor[r], 1u)] = ((uint64_t)(p.a + v.a - 1u) << 32) | (uint32_t)x;
    }
    if (v.b) st_list[p.b] = x;
    if (v.c) md_list[p.c] = x;
    if (v.d) en_list[p.d] = x;
    if (i == n - 1) {                       // closing entry so that sizes are differences
      const uint32_t nb = p.a + v.a;
      bucket_start[nb] = n; st_base[nb] = p.b + v.b; md_base[nb] = p.c + v.c; en_base[nb] = p.d + v.d;
      if (base3) base3[nb] = make_uint4(p.b + v.b, p.c + v.c, p.d + v.d, 0u);
    }
  }
};
// bucket_of is the inverse of the sort's permutation: 4 random bytes per tuple.  Written straight
// from the scan it costs a 32 B sector fill and a 32 B write-back per tuple (ncu: 4.4 x the
// algorithmic bytes).  Every read-major index is hit exactly once, so the index space cut into
// windows of 2^shift entries gives partitions of exactly known size: the scan's third pass drops
// (index, bucket) into the partition of its window (one shared-memory counter per window, one global
// cursor bump per window per CTA), and bucket_of_apply_kernel then walks the partitions in order, so
// that its random writes stay inside one L2-resident window at a time.
constexpr int GROUP_MAX_PARTS = 1024;
struct PartOut { uint2 *buf; unsigned long long *cursor; int shift; uint32_t parts; };

__global__ void __launch_bounds__(SCAN_THREADS, 4) group_apply_kernel(size_t n, GroupIn in, GroupOut out, const Quad *partials,
                                                                   PartOut po) {
  __shared__ __align__(16) unsigned char raw_[sizeof(Quad) * SCAN_THREADS];
  __shared__ uint32_t s_cnt[GROUP_MAX_PARTS];
  __shared__ unsigned long long s_base[GROUP_MAX_PARTS];
  Quad *sh = reinterpret_cast<Quad *>(raw_);
  for (uint32_t q = threadIdx.x; q < po.parts; q += SCAN_THREADS) s_cnt[q] = 0;
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  Quad v[SCAN_ITEMS]; Quad acc = Quad();
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { const size_t i = base + k; v[k] = (i < n) ? in(i) : Quad(); acc = acc + v[k]; }
  Quad excl = block_exclusive_scan<Quad>(acc, sh, (Quad *)nullptr);        // its barriers also order the s_cnt reset
  Quad run = partials[blockIdx.x] + excl;
  uint32_t lo[SCAN_ITEMS], bucket[SCAN_ITEMS], pr[SCAN_ITEMS];       // index inside the window, bucket, window << 16 | rank in CTA
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    const size_t i = base + k;
    if (i < n) {
      out(i, v[k], run);
      const uint64_t dest = out.val.idx ? (uint64_t)out.val.idx[i] : out.dest_of(out.val.at(i));
      const uint32_t q = (uint32_t)(dest >> po.shift);
      lo[k] = (uint32_t)(dest & ((1ull << po.shift) - 1ull));
      bucket[k] = run.a + v[k].a - 1u;
      pr[k] = (q << 16) | atomicAdd(&s_cnt[q], 1u);                    // rank < SCAN_TILE = 2048, q < 1024
    }
    run = run + v[k];
  }
  __syncthreads();
  for (uint32_t q = threadIdx.x; q < po.parts; q += SCAN_THREADS)
    if (s_cnt[q]) s_base[q] = ((unsigned long long)q << po.shift) + atomicAdd(&po.cursor[q], (unsigned long long)s_cnt[q]);
  __syncthreads();
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    const size_t i = base + k;
    if (i < n) po.buf[s_base[pr[k] >> 16] + (pr[k] & 0xffffu)] = make_uint2(lo[k], bucket[k]);
  }
}
__global__ void __launch_bounds__(256) bucket_of_apply_kernel(const uint2 *__restrict__ buf, size_t n, int shift,
                                                              uint32_t *__restrict__ bucket_of) {
  const size_t i = (size_t)blockIdx.x * 1024 + threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const size_t j = i + (size_t)k * 256;
    if (j < n) { const uint2 e = buf[j]; bucket_of[((j >> shift) << shift) + e.x] = e.y; }
  }
}
__global__ void __launch_bounds__(256) max_bucket_kernel(const uint64_t *__restrict__ bucket_start, size_t nb,
                                                         unsigned long long *out) {
  __shared__ unsigned long long s_max[8];
  unsigned long long v = 0;
  for (size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += (size_t)gridDim.x * blockDim.x) {
    const unsigned long long y = bucket_start[b + 1] - bucket_start[b];
    v = y > v ? y : v;
  }
  for (int o = 16; o; o >>= 1) { const unsigned long long y = __shfl_xor_sync(0xffffffffu, v, o); v = y > v ? y : v; }
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; w++) v = s_max[w] > v ? s_max[w] : v;
    if (v) atomicMax(out, v);                                      // one atomic per CTA
  }
}

void group_kmers(Ctx &c, const hovl_settings &s, const Reads &R, Tuples &T, Table &B, bool want_bucket_of, hovl_stats_t &st,
                 const EntIndex *ei) {
  StageTimer tm(c, &st.ns_group, "group");
  const size_t n = T.n;
  const uint32_t nwin = T.compact ? R.max_len - (uint32_t)s.k + 1u : 1u;
  const ValSrc vs{T.compact ? nullptr : T.v(), T.compact ? T.ix() : nullptr, nwin, R.id_base, (float)(nwin - 1u)};
  GroupIn in{T.k(), vs, f32_from_bits(s.head_f32), f32_from_bits(s.tail_f32), f32_from_bits(s.mid_lead_f32),
             f32_from_bits(s.mid_tail_f32)};
  DevBuf<Quad> part;
  const Quad *tot = scan_prepare<Quad, GroupIn>(c, n, in, part);
  const Quad t = read_total(c, tot);
  B.n_buckets = t.a; B.n_st = t.b; B.n_md = t.c; B.n_en = t.d;
  B.bucket_start.alloc(c, B.n_buckets + 1);
  B.st_base.alloc(c, B.n_buckets + 1); B.md_base.alloc(c, B.n_buckets + 1); B.en_base.alloc(c, B.n_buckets + 1);
  B.st_list.alloc(c, B.n_st); B.md_list.alloc(c, B.n_md); B.en_list.alloc(c, B.n_en);
  if (n == 0) {
    HOVL_CUDA(cudaMemsetAsync(B.bucket_start.get(), 0, 8, c.stream));
    HOVL_CUDA(cudaMemsetAsync(B.st_base.get(), 0, 4, c.stream));
    HOVL_CUDA(cudaMemsetAsync(B.md_base.get(), 0, 4, c.stream));
    HOVL_CUDA(cudaMemsetAsync(B.en_base.get(), 0, 4, c.stream));
  }
  if (want_bucket_of) B.bucket_of.alloc(c, n);
  const bool want_base3 = want_bucket_of || ei != nullptr;
  if (want_base3) B.base3.alloc(c, B.n_buckets + 1);
  GroupOut out{vs, n, B.bucket_start.get(), B.st_base.get(), B.md_base.get(), B.en_base.get(),
               B.st_list.get(), B.md_list.get(), B.en_list.get(),
               R.len.get(), R.koff.get(), s.k, want_bucket_of ? B.bucket_of.get() : nullptr,
               ei ? ei->ent : nullptr, ei ? ei->ekoff : nullptr, ei ? ei->ecursor : nullptr,
               want_base3 ? B.base3.get() : nullptr,
               (want_bucket_of && R.min_len == R.max_len) ? R.max_len : 0u, R.id_base};
  if (want_bucket_of && n && c.group_partition) {
    // windows of 2^shift entries: 32 MB of bucket_of each, more only if that needs > GROUP_MAX_PARTS windows
    int shift = c.group_window_log2;
    while (div_up(n, (size_t)1 << shift) > (size_t)GROUP_MAX_PARTS) shift++;
    PartOut po; po.shift = shift; po.parts = (uint32_t)div_up(n, (size_t)1 << shift);
    DevBuf<uint2> pbuf(c, n);
    DevBuf<unsigned long long> cursor(c, po.parts);
    HOVL_CUDA(cudaMemsetAsync(cursor.get(), 0, cursor.bytes(), c.stream));
    po.buf = pbuf.get(); po.cursor = cursor.get();
    out.bucket_of = nullptr;
    HOVL_LAUNCH(c, group_apply_kernel, (unsigned)div_up(n, SCAN_TILE), SCAN_THREADS, 0, n, in, out, part.get(), po);
    HOVL_LAUNCH(c, bucket_of_apply_kernel, (unsigned)div_up(n, 1024), 256, 0, pbuf.get(), n, shift, B.bucket_of.get());
  } else {
    scan_apply<Quad, GroupIn, GroupOut>(c, n, in, out, part);
  }
  DevBuf<unsigned long long> mx(c, 1);
  HOVL_CUDA(cudaMemsetAsync(mx.get(), 0, 8, c.stream));
  if (B.n_buckets)
    HOVL_LAUNCH(c, max_bucket_kernel, (unsigned)std::min<size_t>(div_up(B.n_buckets, 256), (size_t)c.sm_count * 8), 256, 0,
                B.bucket_start.get(), B.n_buckets, mx.get());
  unsigned long long h = 0;
  HOVL_CUDA(cudaMemcpyAsync(&h, mx.get(), 8, cudaMemcpyDeviceToHost, c.stream));
  sync(c);
  B.max_bucket = h;
  st.n_buckets = B.n_buckets; st.max_bucket = B.max_bucket;
}

}  // namespace hovl
