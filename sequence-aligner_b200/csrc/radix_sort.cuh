// radix_sort.cuh — stable LSD radix sort (8-bit digits), one read + one write of the data
// per digit ("onesweep": every tile publishes its digit counts and resolves its global
// offsets with a decoupled look-back over earlier tiles, so there is no separate
// count pass per digit).
//
// This is the KmerTable of the reference (src/KmerTable.scala:41-53: group k-mers by code,
// bucket contents in arrival order) and its PairData (:57-80) re-expressed as sorts:
// stability gives the arrival order inside a bucket for free.
//
//   rs_hist_kernel   one pass over the keys, shared-memory histograms for ALL digits
//   rs_bases_kernel  exclusive scan of each 256-bin histogram -> digit base offsets
//   rs_pass_kernel   per digit: tile load, warp-private shared-memory histograms, ranking inside the warp from the
//                    mask of lanes that hold the same digit, look-back, shared-memory staged coalesced scatter.
//                    The mask comes from eight ballots (one per digit bit), not from match.any: on B200 match.any costs
//                    about 8 cycles per DISTINCT value in the warp (tools/warp_op_probe.cu, profiles/r02_warp_op_probe.txt:
//                    244 cycles per sub-partition for a random 8-bit digit, 95 for the eight ballots) and was 56 % of
//                    the pass; option "sort_rank_ballot" = 0 keeps it for comparison.
//                    The tile (4096 keys and their values, 16 + 16..32 KB) is staged by TMA: one thread arms an
//                    mbarrier with the byte count and issues two cp.async.bulk copies global -> shared (SASS UBLKCP);
//                    the threads then pick their keys out of shared memory instead of issuing 32 LDGs each.  The
//                    staging arrays are the ones the scatter reuses later.  A partial last tile and unaligned inputs
//                    take the plain-load path (option "sort_tma" = 0 forces it, for A/B timing).
#pragma once
#include "common.cuh"

namespace hovl {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;      // 4096 keys per tile
constexpr int RS_RADIX = 256;
constexpr int RS_MAX_PASSES = 8;

constexpr uint64_t RS_FLAG_LOCAL = 1ull << 62;
constexpr uint64_t RS_FLAG_INCL = 2ull << 62;
constexpr uint64_t RS_VALUE_MASK = (1ull << 62) - 1;

template <typename K>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const K *__restrict__ keys, size_t n, int begin_bit,
                                                            int npass, unsigned long long *__restrict__ ghist) {
  __shared__ uint32_t sh[RS_MAX_PASSES * RS_RADIX];
  for (int i = threadIdx.x; i < npass * RS_RADIX; i += RS_THREADS) sh[i] = 0;
  __syncthreads();
  const size_t stride = (size_t)gridDim.x * RS_THREADS;
  for (size_t i = (size_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
    const K k = keys[i];
#pragma unroll
    for (int p = 0; p < RS_MAX_PASSES; p++)
      if (p < npass) atomicAdd(&sh[p * RS_RADIX + (uint32_t)((k >> (begin_bit + 8 * p)) & 0xff)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < npass * RS_RADIX; i += RS_THREADS)
    if (sh[i]) atomicAdd(&ghist[i], (unsigned long long)sh[i]);
}

// one block per pass: ghist[p][*] -> exclusive prefix in place
static __global__ void __launch_bounds__(RS_RADIX) rs_bases_kernel(unsigned long long *ghist) {
  __shared__ unsigned long long sh[RS_RADIX];
  unsigned long long *h = ghist + (size_t)blockIdx.x * RS_RADIX;
  const int t = threadIdx.x;
  const unsigned long long v = h[t];
  sh[t] = v;
  __syncthreads();
  for (int d = 1; d < RS_RADIX; d <<= 1) {
    unsigned long long x = sh[t];
    if (t >= d) x += sh[t - d];
    __syncthreads();
    sh[t] = x;
    __syncthreads();
  }
  h[t] = sh[t] - v;
}

// ---- TMA bulk copy of a tile (sm_90+ PTX; 16-byte aligned source, destination and size)
__device__ __forceinline__ uint32_t rs_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void rs_mbar_init(uint64_t *bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(rs_smem_addr(bar)));
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");         // the async proxy sees the initialised barrier
}
__device__ __forceinline__ void rs_mbar_expect(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rs_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rs_bulk_load(void *dst_smem, const void *src_global, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(rs_smem_addr(dst_smem)), "l"(src_global), "r"(bytes), "r"(rs_smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void rs_mbar_wait(uint64_t *bar, uint32_t phase) {
  asm volatile("{\n\t.reg .pred p;\n\tRS_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra RS_DONE;\n\tbra RS_WAIT;\n\tRS_DONE:\n\t}"
               ::"r"(rs_smem_addr(bar)), "r"(phase) : "memory");
}

template <typename K, typename V, bool HAS_V>
struct RsSmem {
  alignas(16) uint64_t mbar;
  uint32_t warp_hist[RS_WARPS * RS_RADIX];
  uint32_t tile_start[RS_RADIX];
  uint32_t tile_cnt[RS_RADIX];           // early counts: the tile's digit histogram, known before the ranking
  unsigned long long gbase[RS_RADIX];
  uint32_t warp_tot[RS_WARPS];
  uint32_t tile_id;
  alignas(16) K keys[RS_TILE];
  alignas(16) V vals[HAS_V ? RS_TILE : 1];
};

template <typename K, typename V, bool HAS_V, bool BALLOT>
__global__ void __launch_bounds__(RS_THREADS, (sizeof(K) + (HAS_V ? sizeof(V) : 0) <= 8) ? 4 : 3)
rs_pass_kernel(const K *__restrict__ kin, K *__restrict__ kout, const V *__restrict__ vin, V *__restrict__ vout,
               size_t n, int shift, const unsigned long long *__restrict__ digit_base,
               unsigned long long *state, uint32_t *ticket, int use_tma, int early) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RsSmem<K, V, HAS_V> &S = *reinterpret_cast<RsSmem<K, V, HAS_V> *>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;

  if (tid == 0) { S.tile_id = atomicAdd(ticket, 1u); if (use_tma) rs_mbar_init(&S.mbar); }   // tiles are started in ticket order
  for (int i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) S.warp_hist[i] = 0;
  S.tile_cnt[tid] = 0;                                  // RS_THREADS == RS_RADIX
  __syncthreads();
  const size_t tile = S.tile_id;
  const size_t tile_base = tile * (size_t)RS_TILE;
  const uint32_t nvalid = (uint32_t)((n - tile_base) < (size_t)RS_TILE ? (n - tile_base) : (size_t)RS_TILE);

  // ---- load (warp-striped: item i of lane l = warp chunk [i*32 + l]) and rank ----
  K key[RS_ITEMS]; uint16_t rank[RS_ITEMS];
  V val[HAS_V ? RS_ITEMS : 1];
  const uint32_t wbase = warp * (32 * RS_ITEMS);
  const bool tma = use_tma && nvalid == (uint32_t)RS_TILE;            // uniform for the CTA
  if (tma) {
    // TMA-staged tile: one thread arms the barrier with the byte count and issues the bulk copies
    if (tid == 0) {
      rs_mbar_expect(&S.mbar, (uint32_t)(RS_TILE * sizeof(K) + (HAS_V ? RS_TILE * sizeof(V) : 0)));
      rs_bulk_load(S.keys, kin + tile_base, (uint32_t)(RS_TILE * sizeof(K)), &S.mbar);
      if (HAS_V) rs_bulk_load(S.vals, vin + tile_base, (uint32_t)(RS_TILE * sizeof(V)), &S.mbar);
    }
    rs_mbar_wait(&S.mbar, 0);
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) key[i] = S.keys[wbase + i * 32 + lane];
    // the values stay where TMA put them until the keys have been scattered (fewer live registers while ranking)
  } else {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
      const uint32_t loc = wbase + i * 32 + lane;
      key[i] = (loc < nvalid) ? kin[tile_base + loc] : (K)~(K)0;
    }
    // the values are not needed until the scatter; asking for them now keeps their loads in
    // flight under the ranking and the look-back
    if (HAS_V) {
#pragma unroll
      for (int i = 0; i < RS_ITEMS; i++) {
        const uint32_t loc = wbase + i * 32 + lane;
        if (loc < nvalid) val[i] = vin[tile_base + loc];
      }
    }
  }
  if (early) {
    // Early counts: the tile's digit histogram by shared-memory atomics, published as this tile's LOCAL state BEFORE the
    // ranking.  A successor's look-back then never waits for our ranking, and our own look-back (after the ranking)
    // finds its predecessors' counts long since published (ncu, profiles/r02_rs_pass_ballot_full_K3x0.25.txt: 26 % of
    // the stall samples sat on the look-back's spin when the counts were published after the ranking).
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) atomicAdd(&S.tile_cnt[(uint32_t)((key[i] >> shift) & 0xff)], 1u);
    __syncthreads();
    unsigned long long *st0 = state + tile * RS_RADIX + tid;
    *(volatile unsigned long long *)st0 = (tile == 0 ? RS_FLAG_INCL : RS_FLAG_LOCAL) | S.tile_cnt[tid];
  }
  uint32_t *wh = S.warp_hist + warp * RS_RADIX;
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const uint32_t d = (uint32_t)((key[i] >> shift) & 0xff);
    uint32_t mask;
    if (BALLOT) {
      // the lanes holding the same digit, from eight ballots (one per digit bit) instead of match.any.  Written so that a
      // bit costs four instructions: key & uniform bit mask -> predicate, VOTE, predicated NOT, OR into `diff`
      uint32_t diff = 0;                      // lanes whose digit differs from ours in some bit
#pragma unroll
      for (int b = 0; b < 8; b++) {
        const bool bit = (key[i] & ((K)1 << (shift + b))) != 0;
        const uint32_t v = __ballot_sync(0xffffffffu, bit);
        diff |= v ^ (bit ? 0xffffffffu : 0u);
      }
      mask = ~diff;
    } else {
      mask = __match_any_sync(0xffffffffu, d);
    }
    const uint32_t below = mask & lt_mask;
    uint32_t prev = 0;
    if (below == 0) { prev = wh[d]; wh[d] = prev + __popc(mask); }
    __syncwarp();
    prev = __shfl_sync(0xffffffffu, prev, __ffs(mask) - 1);
    rank[i] = (uint16_t)(prev + __popc(below));
  }
  __syncthreads();

  // ---- per-digit: exclusive scan over warps, tile count, publish, tile-local digit start ----
  uint32_t count = 0;
  {
    const int d = tid;                                  // RS_THREADS == RS_RADIX
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) { const uint32_t c = S.warp_hist[w * RS_RADIX + d]; S.warp_hist[w * RS_RADIX + d] = count; count += c; }
    // sentinel padding of a partial tile lands in digit 255; it is never written out and no
    // later tile exists to look back at it, so the published count may include it.
    unsigned long long *st = state + tile * RS_RADIX + d;
    if (!early) {
      if (tile == 0) *(volatile unsigned long long *)st = RS_FLAG_INCL | count;
      else *(volatile unsigned long long *)st = RS_FLAG_LOCAL | count;
    }
    // exclusive scan of count over the 256 digits
    uint32_t x = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) S.warp_tot[warp] = x;
    __syncthreads();
    uint32_t wofs = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) if (w < warp) wofs += S.warp_tot[w];
    const uint32_t start = wofs + x - count;
    S.tile_start[d] = start;
    // ---- decoupled look-back over earlier tiles ----
    unsigned long long excl = 0;
    if (tile > 0) {
      size_t t = tile - 1;
      for (;;) {
        unsigned long long s;
        do { s = *(volatile unsigned long long *)(state + t * RS_RADIX + d); } while ((s >> 62) == 0);
        excl += s & RS_VALUE_MASK;
        if ((s >> 62) == 2) break;
        t--;
      }
      *(volatile unsigned long long *)st = RS_FLAG_INCL | (excl + count);
    }
    S.gbase[d] = digit_base[d] + excl - start;
  }
  __syncthreads();

  // ---- stage in shared memory in sorted order, then write coalesced runs ----
  uint16_t pos[RS_ITEMS];
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const uint32_t d = (uint32_t)((key[i] >> shift) & 0xff);
    pos[i] = (uint16_t)(S.tile_start[d] + wh[d] + rank[i]);
    S.keys[pos[i]] = key[i];
  }
  if (HAS_V && tma) {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) val[i] = S.vals[wbase + i * 32 + lane];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) {
    const uint32_t p = j * RS_THREADS + tid;
    if (p < nvalid) { const K k = S.keys[p]; kout[S.gbase[(uint32_t)((k >> shift) & 0xff)] + p] = k; }
  }
  if (HAS_V) {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
      const uint32_t loc = wbase + i * 32 + lane;
      if (loc < nvalid) S.vals[pos[i]] = val[i];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RS_ITEMS; j++) {
      const uint32_t p = j * RS_THREADS + tid;
      if (p < nvalid) { const K k = S.keys[p]; vout[S.gbase[(uint32_t)((k >> shift) & 0xff)] + p] = S.vals[p]; }
    }
  }
}

// Sorts n (key[,value]) items by key bits [begin_bit, end_bit), stable.  k0/v0 hold the input;
// k1/v1 are same-sized scratch.  Returns 0 when the result is in k0/v0, 1 when in k1/v1.
template <typename K, typename V, bool HAS_V>
static int radix_sort(Ctx &c, K *k0, K *k1, V *v0, V *v1, size_t n, int begin_bit, int end_bit) {
  if (n == 0 || end_bit <= begin_bit) return 0;
  const int npass = (end_bit - begin_bit + 7) / 8;
  HOVL_REQUIRE(npass <= RS_MAX_PASSES, HOVL_E_ARG, "radix_sort: too many digits");
  const size_t ntiles = div_up(n, RS_TILE);
  DevBuf<unsigned long long> ghist(c, (size_t)npass * RS_RADIX);
  DevBuf<unsigned long long> state(c, ntiles * RS_RADIX);
  DevBuf<uint32_t> ticket(c, 1);
  HOVL_CUDA(cudaMemsetAsync(ghist.get(), 0, ghist.bytes(), c.stream));
  const unsigned hgrid = (unsigned)(ntiles < (size_t)c.sm_count * 8 ? ntiles : (size_t)c.sm_count * 8);
  HOVL_LAUNCH(c, (rs_hist_kernel<K>), hgrid, RS_THREADS, 0, k0, n, begin_bit, npass, ghist.get());
  HOVL_LAUNCH(c, rs_bases_kernel, (unsigned)npass, RS_RADIX, 0, ghist.get());
  const size_t smem = sizeof(RsSmem<K, V, HAS_V>);
  static bool attr_set[64] = {};                      // the attribute is per device (one context per device, several devices per process)
  if (c.device < 0 || c.device >= 64 || !attr_set[c.device]) {
    HOVL_CUDA(cudaFuncSetAttribute(rs_pass_kernel<K, V, HAS_V, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    HOVL_CUDA(cudaFuncSetAttribute(rs_pass_kernel<K, V, HAS_V, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (c.device >= 0 && c.device < 64) attr_set[c.device] = true;
  }
  K *kin = k0, *kout = k1; V *vin = v0, *vout = v1;
  // bulk copies need 16-byte aligned sources (both ping-pong buffers serve as sources)
  auto al16 = [](const void *q) { return ((uintptr_t)q & 15u) == 0; };
  const int use_tma = c.sort_tma && al16(k0) && al16(k1) && (!HAS_V || (al16(v0) && al16(v1)));
  for (int p = 0; p < npass; p++) {
    HOVL_CUDA(cudaMemsetAsync(state.get(), 0, state.bytes(), c.stream));
    HOVL_CUDA(cudaMemsetAsync(ticket.get(), 0, sizeof(uint32_t), c.stream));
    if (c.sort_rank_ballot)
      HOVL_LAUNCH(c, (rs_pass_kernel<K, V, HAS_V, true>), (unsigned)ntiles, RS_THREADS, smem, kin, kout, vin, vout, n,
                  begin_bit + 8 * p, ghist.get() + (size_t)p * RS_RADIX, state.get(), ticket.get(), use_tma, c.sort_early_counts);
    else
      HOVL_LAUNCH(c, (rs_pass_kernel<K, V, HAS_V, false>), (unsigned)ntiles, RS_THREADS, smem, kin, kout, vin, vout, n,
                  begin_bit + 8 * p, ghist.get() + (size_t)p * RS_RADIX, state.get(), ticket.get(), use_tma, c.sort_early_counts);
    K *tk = kin; kin = kout; kout = tk; V *tv = vin; vin = vout; vout = tv;
  }
  return npass & 1;
}

}  // namespace hovl
