// handles.cuh — the opaque handles of include/hovl.h (shared by engine.cu and multi.cu).
#pragma once
#include "engine.cuh"

using namespace hovl;      // internal header: included by engine.cu and multi.cu only

struct hovl_ctx {
  Ctx c;
  // sharded path: partial (pair, count) list of hovl_shard_pairs, laid out by destination rank,
  // and the statistics of the stages this rank has run so far
  DevBuf<uint64_t> sh_keys; DevBuf<uint32_t> sh_counts; size_t sh_n = 0;
  DevBuf<hovl_record> sh_recs; size_t sh_nrec = 0;   // hovl_shard_pairs_local: the rank's records until hovl_shard_fetch_records
  hovl_stats_t sh_st;
  Tuples sh_T;                              // hovl_shard_encode(codes_out = NULL): tuples kept for hovl_shard_scatter
  std::vector<cudaStream_t> sh_streams;     // hovl_shard_gather_peer: one stream per peer
  // hovl_shard_select: the rank's dispatched pairs (sorted by (a, b)) and the exclusive prefix of their DP cost (rebalance.cu)
  DevBuf<uint32_t> sh_da, sh_db, sh_dc; DevBuf<unsigned long long> sh_cost; size_t sh_npairs = 0;
  int sh_range_bits = 0;                    // owner function of the tuples in flight: 0 = mixed hash, else code bits (range mode)
  // hovl_create with n_devices > 1: this handle only holds one context per device and the exchange buffers; hovl_run on it
  // drives all of them from one host thread each (multi.cu)
  std::vector<hovl_ctx *> subs;
  std::vector<void *> mx_codes, mx_vals; std::vector<size_t> mx_cap;     // per device: receive buffers of all-to-all #1
};
struct hovl_result {
  hovl_record *recs = nullptr; size_t n = 0, cap_bytes = 0;      // page-locked, from the context's pool
  std::shared_ptr<HostPool> pool;
  hovl_stats_t st;
  void reserve(Ctx &c, size_t count) {
    release();
    pool = c.host_pool;
    recs = (hovl_record *)pool->take((count ? count : 1) * sizeof(hovl_record), &cap_bytes);
    if (!recs) throw Error(HOVL_E_NOMEM, "host out of page-locked memory");
    n = count;
  }
  void release() { if (recs) { pool->give(recs, cap_bytes); recs = nullptr; n = 0; } }
  ~hovl_result() { release(); }
};


namespace hovl {
// multi.cu: the sharded path over the sub-contexts of `h` inside this process (peer memory, no NCCL, no IPC)
// rebalance.cu: alignment re-balance by DP cells
void rebalance_select(hovl_ctx *h, const hovl_settings &s, uint64_t *d_keys, uint32_t *d_counts, size_t n_partial,
                      const uint32_t *d_len_all, uint32_t n_reads_total, uint64_t *n_pairs_out, uint64_t *cost_out);
void rebalance_split(hovl_ctx *h, uint64_t cost_before, uint64_t cost_total, uint32_t parts, uint64_t *counts_out);
void rebalance_fetch(hovl_ctx *h, uint32_t *d_abc_out);
void rebalance_unpack(Ctx &c, const uint32_t *d_abc, size_t n, DevBuf<uint32_t> &a, DevBuf<uint32_t> &b, DevBuf<uint32_t> &cn);
int multi_create(hovl_ctx *h, const int *devices, int n_devices, std::string &err);
void multi_destroy(hovl_ctx *h);
int multi_run(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads, hovl_result **out);
}  // namespace hovl
