// scan.cuh — device-wide exclusive scan fused with its consumer ("scan, then act").
//
// scan_act(n, in, out): for every i in [0,n) computes p(i) = sum_{j<i} in(j) and calls
// out(i, in(i), p(i)); the grand total is left in partials[n_tiles].  `in` and `out` are
// functors evaluated on the device, so flags are never materialised and compaction /
// bucket-table construction happen inside the scan's second pass.  T is any POD with
// operator+ and a default constructor that yields zero (uint32_t, uint64_t, Quad).
//
// Three launches: per-tile reduce, one-block scan of the tile sums, per-tile scan + act.
#pragma once
#include "common.cuh"

namespace hovl {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

struct Quad {                        // four 32-bit counters scanned together
  uint32_t a, b, c, d;
  __host__ __device__ Quad() : a(0), b(0), c(0), d(0) {}
  __host__ __device__ Quad(uint32_t a_, uint32_t b_, uint32_t c_, uint32_t d_) : a(a_), b(b_), c(c_), d(d_) {}
  __host__ __device__ Quad operator+(const Quad &o) const { return Quad(a + o.a, b + o.b, c + o.c, d + o.d); }
};

// value of lane (lane - delta) of the warp, any POD whose size is a multiple of 4 bytes
template <typename T>
__device__ __forceinline__ T shfl_up_pod(const T &v, int delta) {
  static_assert(sizeof(T) % 4 == 0, "POD of 32-bit words");
  T r = v;
  const uint32_t *src = reinterpret_cast<const uint32_t *>(&v);
  uint32_t *dst = reinterpret_cast<uint32_t *>(&r);
#pragma unroll
  for (int w = 0; w < (int)(sizeof(T) / 4); w++) dst[w] = __shfl_up_sync(0xffffffffu, src[w], delta);
  return r;
}

// exclusive scan of one value per thread across the block; returns the block total in *total.
// Warp scans by shuffle, one shared-memory exchange of the warp totals: two barriers in all.
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T *sh /* SCAN_THREADS */, T *total) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  constexpr int NW = SCAN_THREADS / 32;
  T incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const T y = shfl_up_pod(incl, d); if (lane >= d) incl = y + incl; }
  if (lane == 31) sh[warp] = incl;
  __syncthreads();
  T before = T();                                   // totals of the warps before this one
  T all = T();
#pragma unroll
  for (int w = 0; w < NW; w++) { const T x = sh[w]; if (w < warp) before = before + x; all = all + x; }
  T excl = shfl_up_pod(incl, 1);
  if (lane == 0) excl = T();
  excl = before + excl;
  if (total) *total = all;
  __syncthreads();                                  // sh may be reused by the caller's next call
  return excl;
}

template <typename T, typename In>
__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(size_t n, In in, T *partials) {
  __shared__ __align__(16) unsigned char raw_[sizeof(T) * SCAN_THREADS];
  T *sh = reinterpret_cast<T *>(raw_);
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  T acc = T();
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { const size_t i = base + k; if (i < n) acc = acc + in(i); }
  T total;
  block_exclusive_scan<T>(acc, sh, &total);
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// one block: partials[0..nt) -> exclusive prefixes in place, partials[nt] = grand total
template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) scan_partials_kernel(size_t nt, T *partials) {
  __shared__ __align__(16) unsigned char raw_[sizeof(T) * (SCAN_THREADS + 1)];
  T *sh = reinterpret_cast<T *>(raw_);
  T &carry_sh = sh[SCAN_THREADS];
  if (threadIdx.x == 0) carry_sh = T();
  __syncthreads();
  for (size_t base = 0; base < nt; base += SCAN_TILE) {
    T v[SCAN_ITEMS]; T acc = T();
    const size_t b = base + (size_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = (b + k < nt) ? partials[b + k] : T(); acc = acc + v[k]; }
    T total;
    T excl = block_exclusive_scan<T>(acc, sh, &total);
    T run = carry_sh + excl;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { if (b + k < nt) partials[b + k] = run; run = run + v[k]; }
    __syncthreads();
    if (threadIdx.x == 0) carry_sh = carry_sh + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) partials[nt] = carry_sh;
}

template <typename T, typename In, typename Out>
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(size_t n, In in, Out out, const T *partials) {
  __shared__ __align__(16) unsigned char raw_[sizeof(T) * SCAN_THREADS];
  T *sh = reinterpret_cast<T *>(raw_);
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  T v[SCAN_ITEMS]; T acc = T();
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { const size_t i = base + k; v[k] = (i < n) ? in(i) : T(); acc = acc + v[k]; }
  T excl = block_exclusive_scan<T>(acc, sh, (T *)nullptr);
  T run = partials[blockIdx.x] + excl;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { const size_t i = base + k; if (i < n) out(i, v[k], run); run = run + v[k]; }
}

// Phase 1+2: tile sums and their scan.  Returns a device pointer to the grand total (valid
// while `partials` lives) so the caller can size its outputs before phase 3.
template <typename T, typename In>
static const T *scan_prepare(Ctx &c, size_t n, In in, DevBuf<T> &partials) {
  const size_t nt = div_up(n, SCAN_TILE);
  partials.alloc(c, nt + 1);
  if (nt) HOVL_LAUNCH(c, (scan_reduce_kernel<T, In>), (unsigned)nt, SCAN_THREADS, 0, n, in, partials.get());
  HOVL_LAUNCH(c, (scan_partials_kernel<T>), 1, SCAN_THREADS, 0, nt, partials.get());
  return partials.get() + nt;
}
// Phase 3: per-tile scan + out(i, in(i), exclusive prefix).
template <typename T, typename In, typename Out>
static void scan_apply(Ctx &c, size_t n, In in, Out out, DevBuf<T> &partials) {
  const size_t nt = div_up(n, SCAN_TILE);
  if (nt) HOVL_LAUNCH(c, (scan_apply_kernel<T, In, Out>), (unsigned)nt, SCAN_THREADS, 0, n, in, out, partials.get());
}
template <typename T, typename In, typename Out>
static const T *scan_act(Ctx &c, size_t n, In in, Out out, DevBuf<T> &partials) {
  const T *tot = scan_prepare<T, In>(c, n, in, partials);
  scan_apply<T, In, Out>(c, n, in, out, partials);
  return tot;
}
template <typename T>
static T read_total(Ctx &c, const T *d_total) {
  T h;
  HOVL_CUDA(cudaMemcpyAsync(&h, d_total, sizeof(T), cudaMemcpyDeviceToHost, c.stream));
  sync(c);
  return h;
}

struct Duo {                         // two 64-bit counters scanned together
  unsigned long long a, b;
  __host__ __device__ Duo() : a(0), b(0) {}
  __host__ __device__ Duo(unsigned long long a_, unsigned long long b_) : a(a_), b(b_) {}
  __host__ __device__ Duo operator+(const Duo &o) const { return Duo(a + o.a, b + o.b); }
};

}  // namespace hovl
