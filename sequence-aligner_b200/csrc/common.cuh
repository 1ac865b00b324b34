// common.cuh — context, error plumbing and stream-ordered device buffers for libhovl.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/hovl.h"

namespace hovl {

struct Error {
  int code; std::string msg;
  Error(int c, std::string m) : code(c), msg(std::move(m)) {}
};

#define HOVL_CUDA(call)                                                                   \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      char b_[512];                                                                       \
      snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),     \
               __FILE__, __LINE__);                                                       \
      throw ::hovl::Error(e_ == cudaErrorMemoryAllocation ? HOVL_E_NOMEM : HOVL_E_CUDA, b_); \
    }                                                                                     \
  } while (0)

#define HOVL_REQUIRE(cond, code, text)                                                    \
  do { if (!(cond)) throw ::hovl::Error((code), (text)); } while (0)

static inline uint64_t div_up(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// Device memory for one context: a few big cudaMalloc'ed slabs carved up by a host-side
// first-fit free list.  Everything a context does is ordered on ONE stream, so a block can be
// handed out again as soon as the host has freed it; no allocation touches the driver once the
// slabs are warm (stream-ordered cudaMallocAsync re-mapped gigabytes on every run).  When a
// run needed more than one slab they are merged into a single one at the end of the run.
struct Arena {
  struct Slab { char *base = nullptr; size_t size = 0; std::map<size_t, size_t> free; };   // offset -> bytes
  struct Live { size_t slab, off, bytes; };
  std::vector<Slab> slabs;
  std::map<void *, Live> live;
  size_t in_use = 0, peak = 0;
  static size_t round_up(size_t b) { return (b + 511) & ~(size_t)511; }
  void *take(size_t si, size_t bytes) {
    Slab &S = slabs[si];
    for (auto it = S.free.begin(); it != S.free.end(); ++it) {
      if (it->second < bytes) continue;
      const size_t off = it->first, rest = it->second - bytes;
      S.free.erase(it);
      if (rest) S.free[off + bytes] = rest;
      void *p = S.base + off;
      live[p] = Live{si, off, bytes};
      in_use += bytes; if (in_use > peak) peak = in_use;
      return p;
    }
    return nullptr;
  }
  void *alloc(size_t bytes) {          // returns nullptr when the device is out of memory
    bytes = round_up(bytes ? bytes : 1);
    for (size_t i = 0; i < slabs.size(); i++) if (void *p = take(i, bytes)) return p;
    size_t total = 0;
    for (auto &S : slabs) total += S.size;
    size_t want = bytes > total ? bytes : total;                 // at least double what we have
    if (want < ((size_t)256 << 20)) want = (size_t)256 << 20;
    Slab S;
    if (cudaMalloc((void **)&S.base, want) != cudaSuccess) {
      cudaGetLastError();
      want = bytes;
      if (cudaMalloc((void **)&S.base, want) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    }
    S.size = want; S.free[0] = want;
    slabs.push_back(std::move(S));
    return take(slabs.size() - 1, bytes);
  }
  void release(void *p) {
    auto it = live.find(p);
    if (it == live.end()) return;
    const Live l = it->second;
    live.erase(it);
    in_use -= l.bytes;
    auto &F = slabs[l.slab].free;
    size_t off = l.off, bytes = l.bytes;
    auto nx = F.lower_bound(off);
    if (nx != F.end() && off + bytes == nx->first) { bytes += nx->second; nx = F.erase(nx); }
    if (nx != F.begin()) { auto pv = std::prev(nx); if (pv->first + pv->second == off) { off = pv->first; bytes += pv->second; F.erase(pv); } }
    F[off] = bytes;
  }
  // End of a run (stream idle, nothing live): one slab big enough for the run's peak.
  void consolidate() {
    if (!live.empty() || slabs.size() < 2) return;
    size_t total = 0;
    for (auto &S : slabs) { total += S.size; cudaFree(S.base); }
    slabs.clear();
    size_t want = round_up(peak + peak / 8);
    if (want > total) want = total;
    Slab S;
    if (cudaMalloc((void **)&S.base, want) != cudaSuccess) { cudaGetLastError(); return; }
    S.size = want; S.free[0] = want;
    slabs.push_back(std::move(S));
  }
  void destroy() { for (auto &S : slabs) cudaFree(S.base); slabs.clear(); live.clear(); in_use = 0; }
};

// Page-locked host buffers for results: records are copied device -> host straight into one of
// these (no pageable staging, no first-touch page faults per run) and the buffer goes back to
// the pool when the result is freed.  Shared between a context and its live results.
struct HostPool {
  std::mutex mu;
  std::vector<std::pair<void *, size_t>> free_list;
  void *take(size_t bytes, size_t *cap) {
    {
      std::lock_guard<std::mutex> g(mu);
      size_t best = free_list.size();
      for (size_t i = 0; i < free_list.size(); i++)
        if (free_list[i].second >= bytes && (best == free_list.size() || free_list[i].second < free_list[best].second)) best = i;
      if (best != free_list.size()) {
        void *p = free_list[best].first; *cap = free_list[best].second;
        free_list.erase(free_list.begin() + (long)best);
        return p;
      }
    }
    void *p = nullptr;
    const size_t want = bytes + bytes / 8 + 4096;
    if (cudaHostAlloc(&p, want, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }   // any device copies into it (multi.cu)
    *cap = want;
    return p;
  }
  void give(void *p, size_t cap) {
    std::lock_guard<std::mutex> g(mu);
    free_list.emplace_back(p, cap);
    while (free_list.size() > 4) {                      // keep the largest few
      size_t small = 0;
      for (size_t i = 1; i < free_list.size(); i++) if (free_list[i].second < free_list[small].second) small = i;
      cudaFreeHost(free_list[small].first);
      free_list.erase(free_list.begin() + (long)small);
    }
  }
  ~HostPool() { for (auto &f : free_list) cudaFreeHost(f.first); }
};

struct PendingTimer { cudaEvent_t a = nullptr, b = nullptr; uint64_t *dst = nullptr; };
// One timed kernel launch (option "kernel_timing"): events on the launching stream around it.
struct KernelTimer { cudaEvent_t a = nullptr, b = nullptr; const char *stage; const char *kname; };
struct KernelTime { std::string name; uint64_t launches = 0, ns = 0; };

struct Ctx {
  int device = 0;
  std::vector<PendingTimer> timers;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = true;           // false after hovl_set_option("stream", ...)
  bool kernel_timing = false;       // hovl_set_option("kernel_timing", 1)
  const char *stage = "";           // name of the open StageTimer (prefix of kernel-time rows)
  std::vector<KernelTimer> ktimers;
  std::vector<KernelTime> ktimes;   // per (stage:kernel) totals of the last run
  Arena arena;
  std::shared_ptr<HostPool> host_pool = std::make_shared<HostPool>();
  std::string err;
  uint64_t launches = 0;            // kernels launched by the current run
  size_t pair_chunk = (size_t)1 << 31;   // pair instances materialised per expansion chunk
  int dp_probe = 1;                 // 1 = phase 1 by exact-prefix search where that is exact (dp.cu), 0 = always the DP
  int sort_tma = 1;                 // 1 = the radix passes stage their tiles with TMA bulk copies (radix_sort.cuh); 0 = plain loads
  int sort_early_counts = 1;        // 1 = a radix tile publishes its digit counts before the ranking (radix_sort.cuh); 0 = after
  int sort_rank_ballot = 1;         // radix ranking: 1 = eight ballots per key (one per digit bit), 0 = match.any (3 x slower
                                    // on random 8-bit digits: its cost grows with the distinct values in the warp)
  int shard_by_range = 0;           // sharded path: 1 = k-mer codes are owned by contiguous code ranges (replicated-table variant), 0 = by mixed hash
  int dp_wide = 1;                  // 1 = band widths 33..255 run the direction-code warp-per-pair kernel (dp_wb.cu); 0 = dp_kernel<G,D>
  int dp_pack16 = 1;                // 1 = two pairs per thread where a phase's scores fit 16 bits (dp_tp2.cuh); 0 = never
  int compact_tuples = 1;           // 1 = 8-byte (code, index) tuples when every read has the same length (kmer.cu)
  int gather_threads = 128;         // CTA size of the pair gather (one read per CTA): 32..256
  int gather_threads_auto = 1;      // 1 = reads of at most 256 k-mers get a CTA of exactly their warps (pairs.cu); 0 after set_option
  int group_window_log2 = 23;       // entries of bucket_of per partition window (2^23 x 4 B = 32 MB, L2-resident)
  int group_partition = 1;          // 1 = bucket_of through window partitions (kmer.cu), 0 = scattered straight from the scan
  int pair_mode = 0;                // 0 = per-read gather + shared-memory counting (canonical order),
                                    // 1 = expand + sort + run-length encode everywhere
  int pair_table_log = 0;           // test knob: log2 slots of the first shared-memory pair table (0 = 12)
  size_t pair_cap_hint = 0;         // distinct pairs the last gather run produced (sizes the next output buffer)
};

// Device array carved from the context's arena (valid in the context's stream order).
template <typename T>
struct DevBuf {
  T *p = nullptr; size_t n = 0; Ctx *c = nullptr;
  DevBuf() {}
  DevBuf(Ctx &ctx, size_t count) { alloc(ctx, count); }
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n), c(o.c) { o.p = nullptr; o.n = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept {
    if (this != &o) { release(); p = o.p; n = o.n; c = o.c; o.p = nullptr; o.n = 0; }
    return *this;
  }
  void alloc(Ctx &ctx, size_t count) {
    release();
    c = &ctx; n = count;
    p = (T *)ctx.arena.alloc(count * sizeof(T));
    if (!p) {
      n = 0;
      size_t fr = 0, tot = 0, slab = 0;
      cudaMemGetInfo(&fr, &tot);
      for (auto &S : ctx.arena.slabs) slab += S.size;
      char b[256];
      snprintf(b, sizeof b, "device out of memory: wanted %zu B; arena holds %zu B in %zu slabs (%zu B in use), device free %zu of %zu B",
               count * sizeof(T), slab, ctx.arena.slabs.size(), ctx.arena.in_use, fr, tot);
      throw ::hovl::Error(HOVL_E_NOMEM, b);
    }
  }
  // point at memory someone else owns (a caller's device buffer); never released here
  void borrow(T *ptr, size_t count) { release(); p = ptr; n = count; c = nullptr; }
  void release() { if (p && c) c->arena.release(p); p = nullptr; n = 0; }
  ~DevBuf() { release(); }
  T *get() const { return p; }
  size_t bytes() const { return n * sizeof(T); }
};

template <typename T>
static inline void d2h(Ctx &c, T *dst, const T *src, size_t count) {
  if (count) HOVL_CUDA(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
}
template <typename T>
static inline void h2d(Ctx &c, T *dst, const T *src, size_t count) {
  if (count) HOVL_CUDA(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyHostToDevice, c.stream));
}
static inline void sync(Ctx &c) { HOVL_CUDA(cudaStreamSynchronize(c.stream)); }

// Every kernel launch goes through this so that gpu_launches is a count, not a guess.
#define HOVL_LAUNCH(ctx, kernel, grid, block, smem, ...)                                  \
  do {                                                                                    \
    ::hovl::KernelTimer kt_;                                                              \
    if ((ctx).kernel_timing) {                                                            \
      kt_.stage = (ctx).stage; kt_.kname = #kernel;                                      \
      cudaEventCreate(&kt_.a); cudaEventCreate(&kt_.b); cudaEventRecord(kt_.a, (ctx).stream); \
    }                                                                                     \
    kernel<<<(grid), (block), (smem), (ctx).stream>>>(__VA_ARGS__);                       \
    if ((ctx).kernel_timing) { cudaEventRecord(kt_.b, (ctx).stream); (ctx).ktimers.push_back(kt_); } \
    (ctx).launches++;                                                                     \
    HOVL_CUDA(cudaGetLastError());                                                        \
  } while (0)

// Stage timer on the context's stream: events are recorded now and read back by
// resolve_timers() after the run's final synchronisation (no sync per stage).
struct StageTimer {
  Ctx *c; size_t idx; bool open; const char *prev;
  StageTimer(Ctx &ctx, uint64_t *ns, const char *name = nullptr) : c(&ctx), open(true), prev(ctx.stage) {
    if (name) ctx.stage = name;
    PendingTimer t; t.dst = ns;
    cudaEventCreate(&t.a); cudaEventCreate(&t.b); cudaEventRecord(t.a, ctx.stream);
    idx = ctx.timers.size(); ctx.timers.push_back(t);
  }
  void stop() { if (open) { cudaEventRecord(c->timers[idx].b, c->stream); open = false; c->stage = prev; } }
  ~StageTimer() { stop(); }
};
static inline void resolve_timers(Ctx &c) {
  for (auto &t : c.timers) {
    float ms = 0;
    if (cudaEventSynchronize(t.b) == cudaSuccess && cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess && t.dst)
      *t.dst += (uint64_t)((double)ms * 1e6);
    cudaEventDestroy(t.a); cudaEventDestroy(t.b);
  }
  c.timers.clear();
  for (auto &k : c.ktimers) {
    float ms = 0;
    if (cudaEventSynchronize(k.b) == cudaSuccess && cudaEventElapsedTime(&ms, k.a, k.b) == cudaSuccess) {
      std::string name = std::string(k.stage) + ":" + k.kname;
      size_t i = 0;
      while (i < c.ktimes.size() && c.ktimes[i].name != name) i++;
      if (i == c.ktimes.size()) { c.ktimes.emplace_back(); c.ktimes[i].name = name; }
      c.ktimes[i].launches++; c.ktimes[i].ns += (uint64_t)((double)ms * 1e6);
    }
    cudaEventDestroy(k.a); cudaEventDestroy(k.b);
  }
  c.ktimers.clear();
}

}  // namespace hovl
