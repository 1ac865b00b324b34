// shard.cu — device helpers of the sharded (multi-GPU) path, SURVEY §8(e):
//   reads by id -> encode locally -> [all-to-all #1: k-mer tuples by mixed hash of the code]
//   -> local KmerTable + per-read pair gather -> [all-to-all #2: partial (pair, count) by the
//   owner of the pair's fst read] -> sum, count filter, DP on the owner -> records to rank 0.
// The exchanges themselves are the caller's (NCCL through torch.distributed); these kernels
// lay the data out for them: one contiguous segment per destination rank.
#include "engine.cuh"
#include "scan.cuh"

namespace hovl {

constexpr int PT_THREADS = 256;
constexpr int PT_ITEMS = 8;
constexpr int PT_TILE = PT_THREADS * PT_ITEMS;
constexpr int PT_MAX_PARTS = 256;

struct CodeOwner {
  const uint32_t *codes; uint32_t parts; int range_bits;
  __device__ uint32_t operator()(size_t i) const { return shard_owner(codes[i], range_bits, parts); }
};
struct PairOwner {
  const uint64_t *keys; int id_bits; uint32_t n_reads, parts;
  __device__ uint32_t operator()(size_t i) const { return shard_of_read((uint32_t)(keys[i] >> id_bits), n_reads, parts); }
};

template <typename Owner>
__global__ void __launch_bounds__(PT_THREADS) part_count_kernel(size_t n, Owner owner, unsigned long long *__restrict__ hist) {
  __shared__ uint32_t sh[PT_MAX_PARTS];
  for (int i = threadIdx.x; i < PT_MAX_PARTS; i += PT_THREADS) sh[i] = 0;
  __syncthreads();
  const size_t stride = (size_t)gridDim.x * PT_THREADS;
  for (size_t i = (size_t)blockIdx.x * PT_THREADS + threadIdx.x; i < n; i += stride) atomicAdd(&sh[owner(i)], 1u);
  __syncthreads();
  for (int i = threadIdx.x; i < PT_MAX_PARTS; i += PT_THREADS) if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// hist[0..parts) counts -> base[p] exclusive prefix, cursor[p] = base[p]
__global__ void part_base_kernel(const unsigned long long *hist, uint32_t parts, unsigned long long *cursor) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    unsigned long long run = 0;
    for (uint32_t p = 0; p < parts; p++) { cursor[p] = run; run += hist[p]; }
  }
}

// every CTA reserves one run per destination for its tile, then writes its items into the runs
template <typename Owner, typename A, typename B>
__global__ void __launch_bounds__(PT_THREADS) part_scatter_kernel(size_t n, Owner owner, const A *__restrict__ a, const B *__restrict__ b,
                                                                  unsigned long long *__restrict__ cursor,
                                                                  A *__restrict__ a_out, B *__restrict__ b_out) {
  __shared__ uint32_t sh_cnt[PT_MAX_PARTS];
  __shared__ unsigned long long sh_base[PT_MAX_PARTS];
  for (int i = threadIdx.x; i < PT_MAX_PARTS; i += PT_THREADS) sh_cnt[i] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * PT_TILE;
  uint32_t dst[PT_ITEMS], rank[PT_ITEMS];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    const size_t i = base + (size_t)k * PT_THREADS + threadIdx.x;
    if (i < n) { dst[k] = owner(i); rank[k] = atomicAdd(&sh_cnt[dst[k]], 1u); }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < PT_MAX_PARTS; i += PT_THREADS)
    if (sh_cnt[i]) sh_base[i] = atomicAdd(&cursor[i], (unsigned long long)sh_cnt[i]);
  __syncthreads();
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    const size_t i = base + (size_t)k * PT_THREADS + threadIdx.x;
    if (i < n) { const unsigned long long at = sh_base[dst[k]] + rank[k]; a_out[at] = a[i]; b_out[at] = b[i]; }
  }
}

template <typename Owner, typename A, typename B>
static void partition2(Ctx &c, size_t n, uint32_t parts, Owner owner, const A *a, const B *b, A *a_out, B *b_out,
                       uint64_t *counts_host) {
  HOVL_REQUIRE(parts >= 1 && parts <= (uint32_t)PT_MAX_PARTS, HOVL_E_ARG, "number of shards must be 1..256");
  DevBuf<unsigned long long> hist(c, 2 * PT_MAX_PARTS);
  HOVL_CUDA(cudaMemsetAsync(hist.get(), 0, hist.bytes(), c.stream));
  unsigned long long h[PT_MAX_PARTS];
  if (n) {
    const unsigned grid = (unsigned)std::min<uint64_t>(div_up(n, PT_THREADS), (uint64_t)c.sm_count * 8);
    HOVL_LAUNCH(c, (part_count_kernel<Owner>), grid, PT_THREADS, 0, n, owner, hist.get());
    HOVL_LAUNCH(c, part_base_kernel, 1, 32, 0, hist.get(), parts, hist.get() + PT_MAX_PARTS);
    HOVL_LAUNCH(c, (part_scatter_kernel<Owner, A, B>), (unsigned)div_up(n, PT_TILE), PT_THREADS, 0, n, owner, a, b,
                hist.get() + PT_MAX_PARTS, a_out, b_out);
  }
  d2h(c, h, hist.get(), (size_t)parts);
  sync(c);
  for (uint32_t p = 0; p < parts; p++) counts_host[p] = h[p];
}

void partition_tuples(Ctx &c, const uint32_t *codes, const uint64_t *vals, size_t n, uint32_t parts, int range_bits,
                      uint32_t *codes_out, uint64_t *vals_out, uint64_t *counts_host) {
  partition2(c, n, parts, CodeOwner{codes, parts, range_bits}, codes, vals, codes_out, vals_out, counts_host);
}
void partition_pairs(Ctx &c, const uint64_t *keys, const uint32_t *counts, size_t n, int id_bits, uint32_t n_reads,
                     uint32_t parts, uint64_t *keys_out, uint32_t *counts_out, uint64_t *counts_host) {
  partition2(c, n, parts, PairOwner{keys, id_bits, n_reads, parts}, keys, counts, keys_out, counts_out, counts_host);
}

// ---------------------------------------------------------------- partition fused with its all-to-all
// The tuple exchange (all-to-all #1) is a partition by owner followed by a transfer of each part.  With
// the ranks' receive buffers mapped into this process (CUDA IPC, NVLink / NVSwitch peer memory) the
// partition kernel stores every tuple straight into its owner's buffer: no local partitioned copy, no
// separate collective.  A CTA stages its 2048 tuples by destination in shared memory, reserves one run
// per destination with a single cursor bump (cursor[p] starts where this rank's segment begins in rank
// p's buffer), and writes each run as consecutive 4 B / 8 B stores, i.e. full lines on the wire.
constexpr int PEER_MAX_PARTS = 32;
struct PeerDst { uint32_t *codes[PEER_MAX_PARTS]; uint64_t *vals[PEER_MAX_PARTS]; };

void count_owners(Ctx &c, const uint32_t *codes, size_t n, uint32_t parts, int range_bits, uint64_t *counts_host) {
  HOVL_REQUIRE(parts >= 1 && parts <= (uint32_t)PT_MAX_PARTS, HOVL_E_ARG, "number of shards must be 1..256");
  DevBuf<unsigned long long> hist(c, PT_MAX_PARTS);
  HOVL_CUDA(cudaMemsetAsync(hist.get(), 0, hist.bytes(), c.stream));
  unsigned long long h[PT_MAX_PARTS];
  if (n) {
    const unsigned grid = (unsigned)std::min<uint64_t>(div_up(n, PT_THREADS), (uint64_t)c.sm_count * 8);
    HOVL_LAUNCH(c, (part_count_kernel<CodeOwner>), grid, PT_THREADS, 0, n, CodeOwner{codes, parts, range_bits}, hist.get());
  }
  d2h(c, h, hist.get(), (size_t)parts);
  sync(c);
  for (uint32_t p = 0; p < parts; p++) counts_host[p] = h[p];
}

__global__ void __launch_bounds__(PT_THREADS) peer_scatter_kernel(size_t n, uint32_t parts, int range_bits, const uint32_t *__restrict__ codes,
                                                                  const uint64_t *__restrict__ vals,
                                                                  unsigned long long *__restrict__ cursor, const PeerDst D) {
  __shared__ uint32_t s_cnt[PEER_MAX_PARTS], s_start[PEER_MAX_PARTS];
  __shared__ unsigned long long s_base[PEER_MAX_PARTS];
  __shared__ uint32_t s_code[PT_TILE];
  __shared__ uint64_t s_val[PT_TILE];
  __shared__ uint8_t s_dst[PT_TILE];
  if (threadIdx.x < PEER_MAX_PARTS) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * PT_TILE;
  uint32_t dst[PT_ITEMS], rank[PT_ITEMS], code[PT_ITEMS];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    const size_t i = base + (size_t)k * PT_THREADS + threadIdx.x;
    if (i < n) { code[k] = codes[i]; dst[k] = shard_owner(code[k], range_bits, parts); rank[k] = atomicAdd(&s_cnt[dst[k]], 1u); }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (uint32_t p = 0; p < parts; p++) { s_start[p] = run; run += s_cnt[p]; }
  }
  if (threadIdx.x < parts && s_cnt[threadIdx.x])
    s_base[threadIdx.x] = atomicAdd(&cursor[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
  __syncthreads();
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    const size_t i = base + (size_t)k * PT_THREADS + threadIdx.x;
    if (i < n) {
      const uint32_t at = s_start[dst[k]] + rank[k];
      s_code[at] = code[k]; s_val[at] = vals[i]; s_dst[at] = (uint8_t)dst[k];
    }
  }
  __syncthreads();
  const uint32_t count = (uint32_t)((n - base) < (size_t)PT_TILE ? (n - base) : (size_t)PT_TILE);
  for (uint32_t j = threadIdx.x; j < count; j += PT_THREADS) {
    const uint32_t p = s_dst[j];
    const unsigned long long at = s_base[p] + (j - s_start[p]);
    D.codes[p][at] = s_code[j];                         // peer memory: the store goes over NVLink
    D.vals[p][at] = s_val[j];
  }
}

void scatter_tuples_peer(Ctx &c, const uint32_t *codes, const uint64_t *vals, size_t n, uint32_t parts, int range_bits,
                         const uint64_t *peer_codes, const uint64_t *peer_vals, const uint64_t *peer_base) {
  HOVL_REQUIRE(parts >= 1 && parts <= (uint32_t)PEER_MAX_PARTS, HOVL_E_UNSUPPORTED, "peer scatter supports at most 32 ranks");
  PeerDst D;
  unsigned long long hb[PEER_MAX_PARTS];
  for (uint32_t p = 0; p < (uint32_t)PEER_MAX_PARTS; p++) {
    D.codes[p] = p < parts ? reinterpret_cast<uint32_t *>(peer_codes[p]) : nullptr;
    D.vals[p] = p < parts ? reinterpret_cast<uint64_t *>(peer_vals[p]) : nullptr;
    hb[p] = p < parts ? peer_base[p] : 0;
  }
  DevBuf<unsigned long long> cursor(c, PEER_MAX_PARTS);
  h2d(c, cursor.get(), hb, (size_t)PEER_MAX_PARTS);
  if (n) HOVL_LAUNCH(c, peer_scatter_kernel, (unsigned)div_up(n, PT_TILE), PT_THREADS, 0, n, parts, range_bits, codes, vals, cursor.get(), D);
  sync(c);                                              // the stores have left this GPU; the caller's barrier orders the ranks
}

// ---------------------------------------------------------------- per-read offsets of a shard
__global__ void __launch_bounds__(256) read_hist_kernel(const uint64_t *__restrict__ vals, size_t n, uint32_t *__restrict__ cnt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) atomicAdd(&cnt[(uint32_t)(vals[i] >> 32) - 1u], 1u);
}
struct CntIn { const uint32_t *cnt; __device__ unsigned long long operator()(size_t r) const { return cnt[r]; } };
struct CntOut {
  uint64_t *off; size_t n;
  __device__ void operator()(size_t r, const unsigned long long &v, const unsigned long long &p) const {
    off[r] = p;
    if (r == n - 1) off[n] = p + v;
  }
};
void shard_read_offsets(Ctx &c, const uint64_t *vals, size_t n, uint32_t n_reads, DevBuf<uint64_t> &ekoff) {
  DevBuf<uint32_t> cnt(c, (size_t)n_reads + 1);
  HOVL_CUDA(cudaMemsetAsync(cnt.get(), 0, cnt.bytes(), c.stream));
  if (n) HOVL_LAUNCH(c, read_hist_kernel, (unsigned)div_up(n, 256), 256, 0, vals, n, cnt.get());
  ekoff.alloc(c, (size_t)n_reads + 1);
  HOVL_CUDA(cudaMemsetAsync(ekoff.get(), 0, 8, c.stream));
  DevBuf<unsigned long long> part;
  scan_act<unsigned long long, CntIn, CntOut>(c, n_reads, CntIn{cnt.get()}, CntOut{ekoff.get(), n_reads}, part);
  sync(c);                                              // cnt / part die here
}

// ---------------------------------------------------------------- replicated reads
struct LayIn { const uint32_t *len; __device__ unsigned long long operator()(size_t r) const { return (len[r] + 15u) >> 4; } };
struct LayOut {
  const uint32_t *len; uint64_t *woff; uint32_t *mm; size_t n;
  __device__ void operator()(size_t r, const unsigned long long &v, const unsigned long long &p) const {
    woff[r] = p;
    atomicMax(mm, len[r]); atomicMin(mm + 1, len[r]);
    if (r == n - 1) woff[n] = p + v;
  }
};
void shard_read_layout(Ctx &c, const uint32_t *len, uint32_t n_reads, Reads &R) {
  R.n_reads = n_reads;
  R.woff.alloc(c, (size_t)n_reads + 1);
  HOVL_CUDA(cudaMemsetAsync(R.woff.get(), 0, 8, c.stream));
  DevBuf<uint32_t> mm(c, 2);
  HOVL_CUDA(cudaMemsetAsync(mm.get(), 0, 4, c.stream));
  HOVL_CUDA(cudaMemsetAsync(mm.get() + 1, 0xff, 4, c.stream));
  DevBuf<unsigned long long> part;
  const unsigned long long *tot = scan_act<unsigned long long, LayIn, LayOut>(c, n_reads, LayIn{len},
                                                                              LayOut{len, R.woff.get(), mm.get(), n_reads}, part);
  uint32_t h[2] = {0, 0};
  d2h(c, h, mm.get(), 2);
  R.n_words = read_total(c, tot);
  R.max_len = h[0]; R.min_len = n_reads ? h[1] : 0;
}

struct KoffIn {
  const uint32_t *len; uint32_t k;
  __device__ unsigned long long operator()(size_t r) const { return len[r] >= k ? len[r] - k + 1u : 0u; }
};
void shard_read_koff(Ctx &c, const uint32_t *len, uint32_t n_reads, int32_t k, Reads &R) {
  R.koff.alloc(c, (size_t)n_reads + 1);
  HOVL_CUDA(cudaMemsetAsync(R.koff.get(), 0, 8, c.stream));
  DevBuf<unsigned long long> part;
  const unsigned long long *tot = scan_act<unsigned long long, KoffIn, CntOut>(c, n_reads, KoffIn{len, (uint32_t)k},
                                                                               CntOut{R.koff.get(), n_reads}, part);
  R.n_kmers = read_total(c, tot);
}

}  // namespace hovl
