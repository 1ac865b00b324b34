// multi.cu — the sharded path over several GPUs INSIDE one process: hovl_create(ctx, devices, n > 1) + hovl_run.
//
// SURVEY 8(b) asks for a context that "owns streams and scratch memory for n_devices GPUs in one process", so that a
// JVM host reaches 8 GPUs through include/hovl.h alone (no Python, no one-process-per-GPU launcher).  The reference has
// no counterpart (one JVM heap, src/KmerTable.scala:26-37).  This driver is hovl/dist.py's "partials" mode restated in
// C++ on top of the same stage entry points, with every exchange done through peer memory (cudaDeviceEnablePeerAccess:
// NVLink / NVSwitch) instead of NCCL and CUDA IPC:
//   one host thread per device; reads by id -> hovl_shard_encode -> [count matrix in host memory, barrier]
//   -> hovl_shard_scatter stores the tuples into the owners' buffers (all-to-all #1) -> every device copies the peers'
//   2-bit packed reads (all-gather) -> hovl_shard_pairs -> every device copies its segments of the peers' partial
//   (pair, count) lists (all-to-all #2) -> hovl_shard_align -> every device copies its records into ONE page-locked
//   result buffer at the offset its rank gives it (rank order == canonical (a, b) order).
// Needs id_mode = 1 and order = 1 like the sharded path; the records equal hovl_run's on one device.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include "handles.cuh"

namespace hovl {

extern "C" {
int hovl_create(hovl_ctx **out, const int *devices, int n_devices);
void hovl_destroy(hovl_ctx *h);
}

namespace {

// reusable barrier for the device threads; a failed thread keeps arriving so that nobody waits for ever
struct Barrier {
  std::mutex mu; std::condition_variable cv; int n, waiting = 0; unsigned long long round = 0;
  explicit Barrier(int n_) : n(n_) {}
  void wait() {
    std::unique_lock<std::mutex> lk(mu);
    const unsigned long long r = round;
    if (++waiting == n) { waiting = 0; round++; cv.notify_all(); }
    else cv.wait(lk, [&] { return round != r; });
  }
};

struct Shared {                       // what the device threads tell each other (host memory)
  int G = 0; uint32_t n_reads = 0;
  std::vector<uint64_t> counts;       // [r * G + p]: tuples rank r has for rank p
  std::vector<uint64_t> pair_counts;  // [r * G + p]: partial pairs rank r has for rank p
  std::vector<uint64_t> n_words, n_local, n_rec;
  std::vector<const uint32_t *> packed, len, bad;       // per rank: device arrays of its own reads
  std::vector<const uint64_t *> pkeys; std::vector<const uint32_t *> pcnts;   // per rank: partial lists laid out by destination
  std::vector<const hovl_record *> recs;
  std::vector<hovl_stats_t> st;
  std::vector<int> rc; std::vector<std::string> err;
  hovl_result *res = nullptr;
  std::atomic<int> failed{0};
};

#define MX_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { (void)cudaGetLastError(); \
  throw Error(e_ == cudaErrorMemoryAllocation ? HOVL_E_NOMEM : HOVL_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while (0)

}  // namespace

int multi_create(hovl_ctx *h, const int *devices, int n_devices, std::string &err) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) { (void)cudaGetLastError(); err = "no CUDA device (libhovl has no CPU fallback)"; return HOVL_E_NO_DEVICE; }
  if (!devices) { err = "devices is NULL"; return HOVL_E_ARG; }
  for (int i = 0; i < n_devices; i++) {
    if (devices[i] < 0 || devices[i] >= count) { err = "device index out of range"; return HOVL_E_ARG; }
    for (int j = 0; j < i; j++) if (devices[j] == devices[i]) { err = "a device is listed twice"; return HOVL_E_ARG; }
  }
  for (int i = 0; i < n_devices; i++) {
    hovl_ctx *sub = nullptr;
    const int rc = hovl_create(&sub, &devices[i], 1);
    if (rc != HOVL_OK) { err = "cannot create the context of a device"; multi_destroy(h); return rc; }
    h->subs.push_back(sub);
  }
  for (int i = 0; i < n_devices; i++) {                 // every device reads and writes every other one's memory
    cudaSetDevice(devices[i]);
    for (int j = 0; j < n_devices; j++) {
      if (i == j) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
      if (!can) { err = "the devices cannot access each other's memory (no NVLink / PCIe peer access)"; multi_destroy(h); return HOVL_E_UNSUPPORTED; }
      const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); multi_destroy(h); return HOVL_E_CUDA; }
      (void)cudaGetLastError();
    }
  }
  h->mx_codes.assign((size_t)n_devices, nullptr); h->mx_vals.assign((size_t)n_devices, nullptr); h->mx_cap.assign((size_t)n_devices, 0);
  h->c.device = devices[0];
  return HOVL_OK;
}

void multi_destroy(hovl_ctx *h) {
  for (size_t r = 0; r < h->subs.size(); r++) {
    cudaSetDevice(h->subs[r]->c.device);
    if (r < h->mx_codes.size()) { if (h->mx_codes[r]) cudaFree(h->mx_codes[r]); if (h->mx_vals[r]) cudaFree(h->mx_vals[r]); }
    hovl_destroy(h->subs[r]);
  }
  h->subs.clear(); h->mx_codes.clear(); h->mx_vals.clear(); h->mx_cap.clear();
}

static void device_thread(hovl_ctx *h, int r, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                          Shared &S, Barrier &B) {
  hovl_ctx *sub = h->subs[(size_t)r];
  Ctx &c = sub->c;
  const int G = S.G; const uint32_t n = S.n_reads;
  const uint32_t lo = (uint32_t)((uint64_t)n * r / G), hi = (uint32_t)((uint64_t)n * (r + 1) / G), n_local = hi - lo;
  const uint64_t b0 = offsets[lo], nb = offsets[hi] - b0;
  hovl_result *mine = nullptr;
  // a stage that throws marks the run failed; the thread still walks through every barrier below
  auto stage = [&](auto &&f) {
    if (S.failed.load()) return;
    try { cudaSetDevice(c.device); f(); }
    catch (const Error &e) { S.rc[(size_t)r] = e.code; S.err[(size_t)r] = e.msg; S.failed.store(1); }
    catch (const std::bad_alloc &) { S.rc[(size_t)r] = HOVL_E_NOMEM; S.err[(size_t)r] = "host out of memory"; S.failed.store(1); }
  };
  auto check = [&](int rc) { if (rc != HOVL_OK) throw Error(rc, c.err.empty() ? "stage failed" : c.err); };
  DevBuf<uint8_t> d_bases; DevBuf<uint64_t> d_offs; DevBuf<uint32_t> d_packed, d_len, d_bad;
  DevBuf<uint32_t> all_packed, all_len, all_bad; DevBuf<uint64_t> d_keys, r_keys; DevBuf<uint32_t> d_cnts, r_cnts;
  DevBuf<hovl_record> d_recs;
  uint64_t n_words = 0, n_recv = 0, words_all = 0;

  stage([&] {                                           // reads of this rank -> device, splitter, tuples kept for the scatter
    std::vector<uint64_t> loc((size_t)n_local + 1);
    for (uint32_t i = 0; i <= n_local; i++) loc[i] = offsets[lo + i] - b0;
    d_bases.alloc(c, nb); d_offs.alloc(c, (size_t)n_local + 1);
    h2d(c, d_bases.get(), bases + b0, nb); h2d(c, d_offs.get(), loc.data(), (size_t)n_local + 1);
    const uint64_t pcap = nb / 16 + n_local + 2;
    d_packed.alloc(c, pcap); d_len.alloc(c, n_local ? n_local : 1); d_bad.alloc(c, n_local ? n_local : 1);
    sync(c);
    check(hovl_shard_encode(sub, s, d_bases.get(), d_offs.get(), n_local, nb, lo, n, (uint32_t)G, nullptr, nullptr, nb ? nb : 1,
                            d_packed.get(), pcap, d_len.get(), d_bad.get(), &S.counts[(size_t)r * G], &n_words));
    S.n_words[(size_t)r] = n_words; S.n_local[(size_t)r] = n_local;
    S.packed[(size_t)r] = d_packed.get(); S.len[(size_t)r] = d_len.get(); S.bad[(size_t)r] = d_bad.get();
  });
  B.wait();                                             // the count matrix is complete
  uint64_t total = 0;
  stage([&] {                                           // this rank's receive buffers (kept across runs, grown when needed)
    for (int q = 0; q < G; q++) total += S.counts[(size_t)q * G + r];
    if (total > h->mx_cap[(size_t)r]) {
      if (h->mx_codes[(size_t)r]) { cudaFree(h->mx_codes[(size_t)r]); cudaFree(h->mx_vals[(size_t)r]); h->mx_codes[(size_t)r] = h->mx_vals[(size_t)r] = nullptr; h->mx_cap[(size_t)r] = 0; }
      const size_t cap = (size_t)(total + total / 4 + 4096);
      MX_CUDA(cudaMalloc(&h->mx_codes[(size_t)r], cap * 4));
      MX_CUDA(cudaMalloc(&h->mx_vals[(size_t)r], cap * 8));
      h->mx_cap[(size_t)r] = cap;
    }
  });
  B.wait();                                             // every rank's receive buffers exist
  stage([&] {                                           // all-to-all #1: the partition kernel stores into the owners' buffers
    std::vector<uint64_t> pc((size_t)G), pv((size_t)G), base((size_t)G);
    for (int p = 0; p < G; p++) {
      pc[(size_t)p] = (uint64_t)(uintptr_t)h->mx_codes[(size_t)p]; pv[(size_t)p] = (uint64_t)(uintptr_t)h->mx_vals[(size_t)p];
      uint64_t before = 0;
      for (int q = 0; q < r; q++) before += S.counts[(size_t)q * G + p];
      base[(size_t)p] = before;
    }
    check(hovl_shard_scatter(sub, (uint32_t)G, pc.data(), pv.data(), base.data()));
  });
  B.wait();                                             // every store has left its GPU (hovl_shard_scatter synchronises)
  stage([&] {                                           // all-gather of the packed reads, lengths, bad-base flags
    for (int q = 0; q < G; q++) words_all += S.n_words[(size_t)q];
    all_packed.alloc(c, words_all + 2); all_len.alloc(c, (size_t)n + 1); all_bad.alloc(c, (size_t)n + 1);
    uint64_t wat = 0, rat = 0;
    for (int q = 0; q < G; q++) {
      const int qd = h->subs[(size_t)q]->c.device;
      if (S.n_words[(size_t)q]) MX_CUDA(cudaMemcpyPeerAsync(all_packed.get() + wat, c.device, S.packed[(size_t)q], qd, S.n_words[(size_t)q] * 4, c.stream));
      if (S.n_local[(size_t)q]) {
        MX_CUDA(cudaMemcpyPeerAsync(all_len.get() + rat, c.device, S.len[(size_t)q], qd, S.n_local[(size_t)q] * 4, c.stream));
        MX_CUDA(cudaMemcpyPeerAsync(all_bad.get() + rat, c.device, S.bad[(size_t)q], qd, S.n_local[(size_t)q] * 4, c.stream));
      }
      wat += S.n_words[(size_t)q]; rat += S.n_local[(size_t)q];
    }
    sync(c);
    uint64_t n_partial = 0;                             // KmerTable slice + partial PairData, laid out by destination
    check(hovl_shard_pairs(sub, s, (uint32_t *)h->mx_codes[(size_t)r], (uint64_t *)h->mx_vals[(size_t)r], total, n, (uint32_t)G,
                           &S.pair_counts[(size_t)r * G], &n_partial));
    d_keys.alloc(c, n_partial ? n_partial : 1); d_cnts.alloc(c, n_partial ? n_partial : 1);
    check(hovl_shard_fetch_pairs(sub, d_keys.get(), d_cnts.get(), n_partial));
    S.pkeys[(size_t)r] = d_keys.get(); S.pcnts[(size_t)r] = d_cnts.get();
  });
  B.wait();                                             // every rank's partial lists are ready (and its reads were copied)
  stage([&] {                                           // all-to-all #2: this rank's segment of every partial list
    for (int q = 0; q < G; q++) n_recv += S.pair_counts[(size_t)q * G + r];
    r_keys.alloc(c, n_recv ? n_recv : 1); r_cnts.alloc(c, n_recv ? n_recv : 1);
    uint64_t at = 0;
    for (int q = 0; q < G; q++) {
      uint64_t off = 0;
      for (int p = 0; p < r; p++) off += S.pair_counts[(size_t)q * G + p];
      const uint64_t cnt = S.pair_counts[(size_t)q * G + r];
      const int qd = h->subs[(size_t)q]->c.device;
      if (cnt) {
        MX_CUDA(cudaMemcpyPeerAsync(r_keys.get() + at, c.device, S.pkeys[(size_t)q] + off, qd, cnt * 8, c.stream));
        MX_CUDA(cudaMemcpyPeerAsync(r_cnts.get() + at, c.device, S.pcnts[(size_t)q] + off, qd, cnt * 4, c.stream));
      }
      at += cnt;
    }
    sync(c);
  });
  B.wait();                                             // every segment has been copied: the partial lists may go
  stage([&] {
    d_keys.release(); d_cnts.release(); d_packed.release(); d_len.release(); d_bad.release(); d_bases.release(); d_offs.release();
    d_recs.alloc(c, n_recv ? n_recv : 1);               // sum, count filter, dovetail DP on the owner of fst
    const int rc = hovl_shard_align(sub, s, r_keys.get(), r_cnts.get(), n_recv, all_packed.get(), words_all, all_len.get(), all_bad.get(),
                                    n, d_recs.get(), n_recv ? n_recv : 1, &mine);
    if (!mine) check(rc);
    S.rc[(size_t)r] = rc; if (rc != HOVL_OK) S.err[(size_t)r] = c.err;      // HOVL_E_SHORT_READ / _BAD_BASE still deliver records
    S.st[(size_t)r] = mine->st; S.n_rec[(size_t)r] = mine->st.n_records;
  });
  B.wait();                                             // record counts are known: rank 0 sizes the result
  if (r == 0) stage([&] {
    uint64_t tot = 0;
    for (int q = 0; q < G; q++) tot += S.n_rec[(size_t)q];
    S.res = new hovl_result();
    memset(&S.res->st, 0, sizeof S.res->st);
    S.res->reserve(h->subs[0]->c, (size_t)tot);
  });
  B.wait();
  stage([&] {                                           // every device lands its records at its rank-order offset
    uint64_t at = 0;
    for (int q = 0; q < r; q++) at += S.n_rec[(size_t)q];
    if (S.res && S.n_rec[(size_t)r])
      MX_CUDA(cudaMemcpyAsync(S.res->recs + at, d_recs.get(), S.n_rec[(size_t)r] * sizeof(hovl_record), cudaMemcpyDeviceToHost, c.stream));
    sync(c);
  });
  if (mine) hovl_result_free(mine);
  d_recs.release(); r_keys.release(); r_cnts.release(); all_packed.release(); all_len.release(); all_bad.release();
  cudaSetDevice(c.device);
  c.arena.consolidate();
  B.wait();
}

int multi_run(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads, hovl_result **out) {
  const int G = (int)h->subs.size();
  if (!s || !out || !offsets || (!bases && offsets[n_reads])) { h->c.err = "NULL pointer"; return HOVL_E_ARG; }
  if (s->id_mode != 1 || s->order != 1) { h->c.err = "several devices: id_mode = 1 and order = 1 (wide pair key, canonical order) are required"; return HOVL_E_UNSUPPORTED; }
  if (s->aligner != 0 && s->aligner != 1) { h->c.err = "aligner must be 0 or 1"; return HOVL_E_ARG; }
  for (uint32_t i = 0; i < n_reads; i++) if (offsets[i] > offsets[i + 1]) { h->c.err = "offsets must be non-decreasing"; return HOVL_E_ARG; }
  Shared S;
  S.G = G; S.n_reads = n_reads;
  S.counts.assign((size_t)G * G, 0); S.pair_counts.assign((size_t)G * G, 0);
  S.n_words.assign((size_t)G, 0); S.n_local.assign((size_t)G, 0); S.n_rec.assign((size_t)G, 0);
  S.packed.assign((size_t)G, nullptr); S.len.assign((size_t)G, nullptr); S.bad.assign((size_t)G, nullptr);
  S.pkeys.assign((size_t)G, nullptr); S.pcnts.assign((size_t)G, nullptr); S.recs.assign((size_t)G, nullptr);
  S.st.resize((size_t)G); S.rc.assign((size_t)G, HOVL_OK); S.err.resize((size_t)G);
  for (auto &t : S.st) memset(&t, 0, sizeof t);
  const auto t0 = std::chrono::steady_clock::now();
  Barrier B(G);
  std::vector<std::thread> th;
  for (int r = 0; r < G; r++) th.emplace_back(device_thread, h, r, s, bases, offsets, std::ref(S), std::ref(B));
  for (auto &t : th) t.join();
  int rc = HOVL_OK;
  for (int r = 0; r < G; r++) if (S.rc[(size_t)r] != HOVL_OK && rc == HOVL_OK) { rc = S.rc[(size_t)r]; h->c.err = S.err[(size_t)r]; }
  if (S.failed.load() || !S.res) {
    delete S.res;
    if (rc == HOVL_OK) { rc = HOVL_E_CUDA; h->c.err = "a device thread failed"; }
    return rc;
  }
  // whole-job statistics: counters summed over the devices, stage times as the maximum (the devices run side by side)
  hovl_stats_t &W = S.res->st;
  for (int r = 0; r < G; r++) {
    const hovl_stats_t &a = S.st[(size_t)r];
    W.n_reads += a.n_reads; W.n_bases += a.n_bases; W.n_kmers += a.n_kmers; W.n_buckets += a.n_buckets;
    W.n_pair_instances += a.n_pair_instances; W.n_distinct_pairs += a.n_distinct_pairs; W.n_dispatched += a.n_dispatched;
    W.n_leads += a.n_leads; W.n_duds += a.n_duds; W.n_align_valid += a.n_align_valid; W.n_ovl_valid += a.n_ovl_valid;
    W.dp_cells += a.dp_cells; W.n_records += a.n_records; W.gpu_launches += a.gpu_launches; W.n_list_entries += a.n_list_entries;
    W.dp_cells_skipped += a.dp_cells_skipped; W.dp_cells_phase1 += a.dp_cells_phase1;
    auto mx = [](uint64_t &x, uint64_t y) { if (y > x) x = y; };
    mx(W.max_bucket, a.max_bucket); mx(W.ns_h2d, a.ns_h2d); mx(W.ns_pack, a.ns_pack); mx(W.ns_encode, a.ns_encode);
    mx(W.ns_kmer_sort, a.ns_kmer_sort); mx(W.ns_group, a.ns_group); mx(W.ns_expand, a.ns_expand); mx(W.ns_pair_sort, a.ns_pair_sort);
    mx(W.ns_pair_reduce, a.ns_pair_reduce); mx(W.ns_order, a.ns_order); mx(W.ns_dp, a.ns_dp); mx(W.ns_d2h, a.ns_d2h);
    W.kmer_tuple_bytes = a.kmer_tuple_bytes;
  }
  W.ns_total = (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
  *out = S.res;
  return rc;
}

}  // namespace hovl
