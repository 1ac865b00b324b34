// common.cuh — context, error plumbing and stream-ordered device buffers for libhovl.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
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
  std::string err;
  uint64_t launches = 0;            // kernels launched by the current run
  size_t pair_chunk = (size_t)1 << 31;   // pair instances materialised per expansion chunk
};

// Stream-ordered device array (cudaMallocAsync on the context's stream).
template <typename T>
struct DevBuf {
  T *p = nullptr; size_t n = 0; cudaStream_t s = nullptr;
  DevBuf() {}
  DevBuf(Ctx &c, size_t count) { alloc(c, count); }
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n), s(o.s) { o.p = nullptr; o.n = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept {
    if (this != &o) { release(); p = o.p; n = o.n; s = o.s; o.p = nullptr; o.n = 0; }
    return *this;
  }
  void alloc(Ctx &c, size_t count) {
    release();
    s = c.stream; n = count;
    HOVL_CUDA(cudaMallocAsync((void **)&p, (count ? count : 1) * sizeof(T), s));
  }
  void release() { if (p) { cudaFreeAsync(p, s); p = nullptr; n = 0; } }
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
