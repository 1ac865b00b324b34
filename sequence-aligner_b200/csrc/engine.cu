// engine.cu — the C ABI of include/hovl.h on top of the device stages (engine.cuh).
// No CPU fallback: every compute entry point needs a CUDA device and fails otherwise.
#include <algorithm>
#include <chrono>
#include <cstring>
#include <cstdlib>
#include <numeric>
#include "engine.cuh"
#include "scan.cuh"
#include "radix_sort.cuh"
#include "dp_core.cuh"

using namespace hovl;

#include "handles.cuh"

static thread_local std::string g_create_error;

static uint64_t wall_ns() {
  return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
             std::chrono::steady_clock::now().time_since_epoch()).count();
}
static uint32_t bits_of(float f) { uint32_t b; memcpy(&b, &f, 4); return b; }
static float float_of(uint32_t b) { float f; memcpy(&f, &b, 4); return f; }

// ---------------------------------------------------------------- settings
extern "C" void hovl_settings_set_sections(hovl_settings *s, uint32_t kmer_edge_f32, uint32_t kmer_center_f32) {
  // AlignSettings, src/ObjectStore.scala:32-35; each operation rounded to binary32
  const float e = float_of(kmer_edge_f32), c = float_of(kmer_center_f32);
  volatile float half_c = c * 0.5f;
  volatile float tail = 1.0f - e, lead = 0.5f - half_c, mtail = 0.5f + half_c;
  s->head_f32 = bits_of(e); s->tail_f32 = bits_of(tail); s->mid_lead_f32 = bits_of(lead); s->mid_tail_f32 = bits_of(mtail);
}
extern "C" void hovl_settings_default(hovl_settings *s) {
  static const int32_t hoxd[4][4] = {   // defaultHOXD, src/BioLibs.scala:119-141 (A,C,G,T)
    {91, -114, -31, -123}, {-114, 100, -125, -31}, {-31, -125, 100, -114}, {-123, -31, -114, 91}};
  memset(s, 0, sizeof *s);
  s->k = 12;                                        // src/Project4.scala:41
  memcpy(s->cost, hoxd, sizeof hoxd);
  s->gap_open = -200; s->gap_extend = -20;          // :109-110
  s->min_overlap = 40;                              // :106
  s->min_identity_f32 = bits_of(0.98f);             // :107
  s->max_ignore_f32 = bits_of(90.0f);               // :108, Float at ObjectStore.scala:25
  s->min_coll = 7; s->max_coll = 222;               // :111-112
  hovl_settings_set_sections(s, bits_of(0.4f), bits_of(0.4f));   // :113-114 (edge, center), tuple order :257
  s->aligner = 0; s->order = 0; s->id_mode = 0; s->filter_valid = 1;
}

static void check_settings(const hovl_settings *s, uint32_t n_reads) {
  HOVL_REQUIRE(s != nullptr, HOVL_E_ARG, "settings is NULL");
  HOVL_REQUIRE(s->k >= 1, HOVL_E_ARG, "k must be >= 1");
  HOVL_REQUIRE(s->gap_open <= 0 && s->gap_extend <= 0, HOVL_E_ARG, "gap penalties must be <= 0 (Project4.scala:160-165)");
  HOVL_REQUIRE(s->aligner == 0 || s->aligner == 1, HOVL_E_ARG, "aligner must be 0 (fast dovetail) or 1 (quadratic)");
  HOVL_REQUIRE(!(s->id_mode == 0 && n_reads > 32767), HOVL_E_UNSUPPORTED,
               "id_mode 0 is the reference's 16-bit pair key (KmerTable.scala:73): at most 32767 reads");
  HOVL_REQUIRE(!(s->id_mode != 0 && s->order == 0), HOVL_E_UNSUPPORTED,
               "reference (Trove) order is defined on the 16-bit pair key only; use order=1 with id_mode=1");
}

// ---------------------------------------------------------------- context
extern "C" int hovl_create(hovl_ctx **out, const int *devices, int n_devices) {
  if (!out) return HOVL_E_ARG;
  *out = nullptr;
  if (n_devices > 1) {                                  // one handle over n devices: the sharded path inside this process (multi.cu)
    hovl_ctx *m = new hovl_ctx();
    std::string err;
    const int rc = multi_create(m, devices, n_devices, err);
    if (rc != HOVL_OK) { g_create_error = err; delete m; return rc; }
    *out = m;
    return HOVL_OK;
  }
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) {
    cudaGetLastError();
    g_create_error = "no CUDA device (libhovl has no CPU fallback)";
    return HOVL_E_NO_DEVICE;
  }
  const int dev = (devices && n_devices == 1) ? devices[0] : 0;
  if (dev < 0 || dev >= count) { g_create_error = "device index out of range"; return HOVL_E_ARG; }
  hovl_ctx *h = new hovl_ctx();
  try {
    HOVL_CUDA(cudaSetDevice(dev));
    h->c.device = dev;
    cudaDeviceProp prop;
    HOVL_CUDA(cudaGetDeviceProperties(&prop, dev));
    h->c.sm_count = prop.multiProcessorCount;
    HOVL_CUDA(cudaStreamCreateWithFlags(&h->c.stream, cudaStreamNonBlocking));
    if (const char *e = getenv("HOVL_PAIR_CHUNK")) { const long long v = atoll(e); if (v > 0) h->c.pair_chunk = (size_t)v; }
  } catch (const Error &e) {
    g_create_error = e.msg; const int code = e.code; delete h; return code;
  }
  *out = h;
  // HOVL_OPTS="name=value,name=value": options applied to every handle of the process (A/B runs of the test suite and of
  // hosts that do not expose hovl_set_option); unknown names are an error here as they are there
  if (const char *e = getenv("HOVL_OPTS")) {
    std::string spec(e);
    size_t at = 0;
    while (at < spec.size()) {
      size_t end = spec.find(',', at);
      if (end == std::string::npos) end = spec.size();
      const std::string kv = spec.substr(at, end - at);
      at = end + 1;
      if (kv.empty()) continue;
      const size_t eq = kv.find('=');
      const int rc = eq == std::string::npos ? HOVL_E_ARG : hovl_set_option(h, kv.substr(0, eq).c_str(), atoll(kv.c_str() + eq + 1));
      if (rc != HOVL_OK) { g_create_error = "HOVL_OPTS: bad option '" + kv + "'"; hovl_destroy(h); *out = nullptr; return HOVL_E_ARG; }
    }
  }
  return HOVL_OK;
}
extern "C" void hovl_destroy(hovl_ctx *h) {
  if (!h) return;
  if (!h->subs.empty()) { multi_destroy(h); delete h; return; }
  cudaSetDevice(h->c.device);
  if (h->c.stream) { cudaStreamSynchronize(h->c.stream); if (h->c.own_stream) cudaStreamDestroy(h->c.stream); }
  h->sh_keys.release(); h->sh_counts.release(); h->sh_recs.release();
  h->sh_da.release(); h->sh_db.release(); h->sh_dc.release(); h->sh_cost.release();
  for (cudaStream_t st : h->sh_streams) cudaStreamDestroy(st);
  h->c.arena.destroy();
  delete h;
}
extern "C" const char *hovl_last_error(const hovl_ctx *h) { return h ? h->c.err.c_str() : g_create_error.c_str(); }
extern "C" int hovl_set_option(hovl_ctx *h, const char *name, int64_t value) {
  if (!h || !name) return HOVL_E_ARG;
  if (!h->subs.empty()) {                                // several devices: the knob goes to every device's context
    if (!strcmp(name, "stream")) { h->c.err = "a context over several devices launches on its own streams"; return HOVL_E_UNSUPPORTED; }
    for (hovl_ctx *sub : h->subs) { const int rc = hovl_set_option(sub, name, value); if (rc != HOVL_OK) { h->c.err = sub->c.err; return rc; } }
    return HOVL_OK;
  }
  if (!strcmp(name, "pair_chunk") && value > 0) { h->c.pair_chunk = (size_t)value; return HOVL_OK; }
  if (!strcmp(name, "kernel_timing")) { h->c.kernel_timing = value != 0; return HOVL_OK; }
  if (!strcmp(name, "pair_table_log") && (value == 0 || (value >= 4 && value <= 12))) { h->c.pair_table_log = (int)value; return HOVL_OK; }
  if (!strcmp(name, "pair_cap_hint") && value >= 0) { h->c.pair_cap_hint = (size_t)value; return HOVL_OK; }
  if (!strcmp(name, "pair_mode") && (value == 0 || value == 1)) { h->c.pair_mode = (int)value; return HOVL_OK; }
  if (!strcmp(name, "group_partition") && (value == 0 || value == 1)) { h->c.group_partition = (int)value; return HOVL_OK; }
  if (!strcmp(name, "group_window_log2") && value >= 4 && value <= 30) { h->c.group_window_log2 = (int)value; return HOVL_OK; }
  if (!strcmp(name, "gather_threads") && value >= 32 && value <= 256 && value % 32 == 0) { h->c.gather_threads = (int)value; h->c.gather_threads_auto = 0; return HOVL_OK; }
  if (!strcmp(name, "dp_probe") && (value == 0 || value == 1)) { h->c.dp_probe = (int)value; return HOVL_OK; }
  if (!strcmp(name, "compact_tuples") && (value == 0 || value == 1)) { h->c.compact_tuples = (int)value; return HOVL_OK; }
  if (!strcmp(name, "dp_pack16") && (value == 0 || value == 1)) { h->c.dp_pack16 = (int)value; return HOVL_OK; }
  if (!strcmp(name, "dp_wide") && (value == 0 || value == 1)) { h->c.dp_wide = (int)value; return HOVL_OK; }
  if (!strcmp(name, "shard_by_range") && (value == 0 || value == 1)) { h->c.shard_by_range = (int)value; return HOVL_OK; }
  if (!strcmp(name, "sort_tma") && (value == 0 || value == 1)) { h->c.sort_tma = (int)value; return HOVL_OK; }
  if (!strcmp(name, "sort_early_counts") && (value == 0 || value == 1)) { h->c.sort_early_counts = (int)value; return HOVL_OK; }
  if (!strcmp(name, "sort_rank_ballot") && (value == 0 || value == 1)) { h->c.sort_rank_ballot = (int)value; return HOVL_OK; }
  if (!strcmp(name, "stream")) {                     // run on the caller's stream (borrowed cudaStream_t)
    cudaSetDevice(h->c.device);
    if (h->c.stream) { cudaStreamSynchronize(h->c.stream); if (h->c.own_stream) cudaStreamDestroy(h->c.stream); }
    h->c.stream = (cudaStream_t)(uintptr_t)value; h->c.own_stream = false;
    return HOVL_OK;
  }
  h->c.err = std::string("unknown option: ") + name;
  return HOVL_E_ARG;
}
extern "C" void hovl_free(void *p) { free(p); }
extern "C" int hovl_kernel_times(const hovl_ctx *h, uint32_t i, char *name, uint32_t cap, uint64_t *launches, uint64_t *ns) {
  if (!h) return HOVL_E_ARG;
  const auto &kt = h->c.ktimes;
  if (i < kt.size()) {
    if (name && cap) { strncpy(name, kt[i].name.c_str(), cap - 1); name[cap - 1] = 0; }
    if (launches) *launches = kt[i].launches;
    if (ns) *ns = kt[i].ns;
  }
  return (int)kt.size();
}

template <typename F>
static int guarded(hovl_ctx *h, F &&f) {
  if (!h) return HOVL_E_ARG;
  if (!h->subs.empty()) { h->c.err = "this entry point needs a one-device context (hovl_create with n_devices = 1)"; return HOVL_E_UNSUPPORTED; }
  try {
    HOVL_CUDA(cudaSetDevice(h->c.device));
    h->c.err.clear();
    return f();
  } catch (const Error &e) {
    h->c.err = e.msg;
    for (auto &t : h->c.timers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    h->c.timers.clear();
    cudaStreamSynchronize(h->c.stream); cudaGetLastError();
    return e.code;
  } catch (const std::bad_alloc &) {
    h->c.err = "host out of memory"; return HOVL_E_NOMEM;
  } catch (...) {
    h->c.err = "unexpected exception"; return HOVL_E_CUDA;
  }
}

static void upload_reads(Ctx &c, Reads &R, const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads, uint64_t *ns) {
  HOVL_REQUIRE(offsets != nullptr, HOVL_E_ARG, "offsets is NULL");
  const uint64_t nb = offsets[n_reads];
  HOVL_REQUIRE(bases != nullptr || nb == 0, HOVL_E_ARG, "bases is NULL");
  for (uint32_t i = 0; i < n_reads; i++) HOVL_REQUIRE(offsets[i] <= offsets[i + 1], HOVL_E_ARG, "offsets must be non-decreasing");
  StageTimer tm(c, ns, "h2d");
  R.n_reads = n_reads; R.n_bases = nb;
  R.own_bases.alloc(c, nb); R.own_offsets.alloc(c, (size_t)n_reads + 1);
  h2d(c, R.own_bases.get(), bases, nb);
  h2d(c, R.own_offsets.get(), offsets, (size_t)n_reads + 1);
  R.d_bases = R.own_bases.get(); R.d_offsets = R.own_offsets.get();
}

// ---------------------------------------------------------------- small kernels of the driver
__global__ void __launch_bounds__(256) bucket_first_kernel(const uint64_t *__restrict__ bucket_start, const uint32_t *__restrict__ key,
                                                           const uint64_t *__restrict__ val, size_t nb,
                                                           uint32_t *__restrict__ code, uint64_t *__restrict__ first) {
  const size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b < nb) { const uint64_t i = bucket_start[b]; code[b] = key[i]; first[b] = val[i]; }
}
__global__ void __launch_bounds__(256) pair_zip_kernel(const uint64_t *__restrict__ key, const uint32_t *__restrict__ count, size_t n,
                                                       uint64_t *__restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (key[i] << 32) | count[i];
}

struct RecIn {
  const hovl_record *r; int filter; uint32_t *status;
  __device__ Quad operator()(size_t i) const {
    const int32_t f = r[i].flags;
    if (f < 0) { atomicOr(status, 4u); return Quad(); }
    if (f & HOVL_F_SHORT_B) atomicOr(status, 1u);
    if (f & HOVL_F_BAD_BASE) atomicOr(status, 2u);
    const uint32_t keep = (!filter || (f & HOVL_F_ALIGN_VALID)) ? 1u : 0u;
    return Quad(keep, (f & HOVL_F_DUD) ? 1u : 0u, (f & HOVL_F_ALIGN_VALID) ? 1u : 0u, (f & HOVL_F_OVL_VALID) ? 1u : 0u);
  }
};
struct RecOut {
  const hovl_record *r; hovl_record *out;
  __device__ void operator()(size_t i, const Quad &v, const Quad &p) const { if (v.a) out[p.a] = r[i]; }
};

// ---------------------------------------------------------------- reference (Trove) order
// KmerData insertion order = first occurrence of each code over reads (id up) and positions
// (up) (KmerTable.scala:45-52); the first tuple of a bucket carries exactly that (id, loc).
static void kmerdata_iteration_order(Ctx &c, Tuples &T, const Table &B, DevBuf<uint32_t> &d_perm, hovl_stats_t &st) {
  const size_t nb = B.n_buckets;
  DevBuf<uint32_t> d_code(c, nb); DevBuf<uint64_t> d_first(c, nb);
  if (nb) HOVL_LAUNCH(c, bucket_first_kernel, (unsigned)div_up(nb, 256), 256, 0, B.bucket_start.get(), T.k(), T.v(), nb,
                      d_code.get(), d_first.get());
  std::vector<uint32_t> code(nb); std::vector<uint64_t> first(nb);
  d2h(c, code.data(), d_code.get(), nb); d2h(c, first.data(), d_first.get(), nb);
  sync(c);
  const uint64_t t0 = wall_ns();
  std::vector<uint32_t> ins(nb);
  std::iota(ins.begin(), ins.end(), 0u);
  std::sort(ins.begin(), ins.end(), [&](uint32_t x, uint32_t y) { return first[x] < first[y]; });
  std::vector<int32_t> keys(nb);
  for (size_t i = 0; i < nb; i++) keys[i] = (int32_t)code[ins[i]];
  std::vector<uint32_t> ord;
  trove_iteration_order(keys.data(), nb, ord);
  std::vector<uint32_t> perm(nb);
  for (size_t j = 0; j < nb; j++) perm[j] = ins[ord[j]];
  st.ns_order += wall_ns() - t0;
  d_perm.alloc(c, nb);
  h2d(c, d_perm.get(), perm.data(), nb);
  sync(c);
}

// PairData insertion order -> PairData iteration -> DispatchData insertion -> DispatchData
// iteration (KmerTable.scala:162-176, :251-263).  Returns the dispatch list in final order.
// PairData (27.4 M keys on c_ruddii) is replayed on the device (trove.cu) and filtered there, in iteration order;
// only the dispatched pairs come to the host, where the lead map (<= 32767 keys) is replayed by order.cpp.
__global__ void __launch_bounds__(256) pair_javakey_kernel(const uint64_t *__restrict__ v, size_t n, int id_bits,
                                                           int32_t *__restrict__ pk) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t key = (uint32_t)(v[i] >> 32), fst = key >> id_bits, snd = key & ((1u << id_bits) - 1u);
  pk[i] = (int32_t)((fst << 16) ^ snd);                             // KmerTable.scala:73
}
struct DispIn {                       // q-th entry of the PairData iteration: dispatched?  (:165-167)
  const uint64_t *v; const uint32_t *it; long long lo, hi;
  __device__ uint32_t operator()(size_t q) const { const long long cnt = (uint32_t)v[it[q]]; return (lo <= cnt && cnt <= hi) ? 1u : 0u; }
};
struct DispOut {
  const uint64_t *v; const uint32_t *it; const int32_t *pk; int32_t *a, *b; uint32_t *cnt;
  __device__ void operator()(size_t q, const uint32_t &keep, const uint32_t &p) const {
    if (!keep) return;
    const uint32_t i = it[q];
    const int32_t k = pk[i];
    a[p] = k >> 16; b[p] = (int32_t)((uint32_t)k << 16) >> 16;      // :169-170 (sign-extending unpack)
    cnt[p] = (uint32_t)v[i];
  }
};
static void dispatch_reference_order(Ctx &c, const hovl_settings &s, Pairs &P, std::vector<uint32_t> &ha,
                                     std::vector<uint32_t> &hb, std::vector<uint32_t> &hc, hovl_stats_t &st) {
  const size_t np = P.n;
  DevBuf<uint64_t> v0(c, np), v1(c, np);
  DevBuf<uint32_t> s1(c, np);
  if (np) HOVL_LAUNCH(c, pair_zip_kernel, (unsigned)div_up(np, 256), 256, 0, P.key.get(), P.count.get(), np, v0.get());
  const int r = radix_sort<uint32_t, uint64_t, true>(c, P.seq.get(), s1.get(), v0.get(), v1.get(), np, 0, 32);
  const uint64_t *dv = r ? v1.get() : v0.get();                     // (key << 32 | count) in PairData insertion order
  DevBuf<int32_t> pk(c, np);
  if (np) HOVL_LAUNCH(c, pair_javakey_kernel, (unsigned)div_up(np, 256), 256, 0, dv, np, P.id_bits, pk.get());
  DevBuf<uint32_t> it;
  trove_iteration_order_device(c, pk.get(), np, it);
  DispIn in{dv, it.get(), (long long)s.min_coll, (long long)s.max_coll};
  DevBuf<uint32_t> part;
  const uint32_t nsel = read_total(c, scan_prepare<uint32_t, DispIn>(c, np, in, part));
  DevBuf<int32_t> da(c, nsel), db(c, nsel);
  DevBuf<uint32_t> dc(c, nsel);
  DispOut out{dv, it.get(), pk.get(), da.get(), db.get(), dc.get()};
  scan_apply<uint32_t, DispIn, DispOut>(c, np, in, out, part);
  std::vector<int32_t> sa(nsel), sb(nsel);
  std::vector<uint32_t> sel_c(nsel);
  d2h(c, sa.data(), da.get(), nsel); d2h(c, sb.data(), db.get(), nsel); d2h(c, sel_c.data(), dc.get(), nsel);
  sync(c);
  const uint64_t t0 = wall_ns();
  std::vector<int32_t> lead_slot(65536, -1), leads;
  std::vector<uint32_t> sel_lead(nsel), lead_count;
  for (size_t q = 0; q < nsel; q++) {                               // DispatchData insertion order (:171-176)
    int32_t &slot = lead_slot[(size_t)(sa[q] & 0xffff)];
    if (slot < 0) { slot = (int32_t)leads.size(); leads.push_back(sa[q]); lead_count.push_back(0); }
    sel_lead[q] = (uint32_t)slot;
    lead_count[(size_t)slot]++;
  }
  std::vector<size_t> lead_start(leads.size() + 1, 0);
  for (size_t l = 0; l < leads.size(); l++) lead_start[l + 1] = lead_start[l] + lead_count[l];
  std::vector<size_t> fill(lead_start.begin(), lead_start.end() - 1);
  std::vector<uint32_t> gb(nsel), gc(nsel);
  for (size_t q = 0; q < nsel; q++) { const size_t at = fill[sel_lead[q]]++; gb[at] = (uint32_t)sb[q]; gc[at] = sel_c[q]; }
  std::vector<uint32_t> lord;
  trove_iteration_order(leads.data(), leads.size(), lord);
  ha.clear(); hb.clear(); hc.clear();
  ha.reserve(gb.size()); hb.reserve(gb.size()); hc.reserve(gb.size());
  for (size_t j = 0; j < lord.size(); j++) {
    const size_t l = lord[j];
    for (size_t at = lead_start[l]; at < lead_start[l + 1]; at++) { ha.push_back((uint32_t)leads[l]); hb.push_back(gb[at]); hc.push_back(gc[at]); }
  }
  st.n_leads = leads.size();
  st.ns_order += wall_ns() - t0;
}

__global__ void __launch_bounds__(256) lead_count_kernel(const uint32_t *__restrict__ a, size_t n, unsigned long long *out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool head = i < n && (i == 0 || a[i] != a[i - 1]);
  const uint32_t m = __ballot_sync(0xffffffffu, head);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(out, (unsigned long long)__popc(m));
}

// ---------------------------------------------------------------- the path
// Align `np` pairs that are already on the device, compact per filter_valid, bring them home.
// d_dst != nullptr: the kept records stay on the device (caller's buffer of dst_cap records)
// keep_dev != nullptr: the kept records stay on the device in *keep_dev (the context's, for hovl_shard_fetch_records)
static int align_and_collect(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                             const uint32_t *d_cnt, size_t np, hovl_result &res, hovl_record *d_dst = nullptr,
                             uint64_t dst_cap = 0, DevBuf<hovl_record> *keep_dev = nullptr) {
  hovl_stats_t &st = res.st;
  DevBuf<hovl_record> d_rec(c, np);
  align_pairs(c, s, R, d_a, d_b, d_cnt, np, d_rec.get(), st);
  DevBuf<uint32_t> status(c, 1);
  HOVL_CUDA(cudaMemsetAsync(status.get(), 0, 4, c.stream));
  RecIn in{d_rec.get(), s.filter_valid, status.get()};
  DevBuf<Quad> part;
  const Quad t = read_total(c, scan_prepare<Quad, RecIn>(c, np, in, part));
  DevBuf<hovl_record> d_keep;
  if (d_dst) HOVL_REQUIRE(t.a <= dst_cap, HOVL_E_ARG, "device record buffer too small");
  else d_keep.alloc(c, t.a);
  RecOut out{d_rec.get(), d_dst ? d_dst : d_keep.get()};
  scan_apply<Quad, RecIn, RecOut>(c, np, in, out, part);
  res.reserve(c, (d_dst || keep_dev) ? 0 : t.a);
  uint32_t hstatus = 0;
  {
    StageTimer tm(c, &st.ns_d2h, "d2h");
    if (!d_dst && !keep_dev) d2h(c, res.recs, d_keep.get(), (size_t)t.a);
    d2h(c, &hstatus, status.get(), 1);
  }
  sync(c);
  if (keep_dev) *keep_dev = std::move(d_keep);
  st.n_dispatched = np; st.n_duds = t.b; st.n_align_valid = t.c; st.n_ovl_valid = t.d; st.n_records = t.a;
  HOVL_REQUIRE(!(hstatus & 4u), HOVL_E_CUDA, "internal: a pair reached a DP class narrower than its band");
  if (hstatus & 2u) { c.err = "a dispatched read holds a non-ACGT base (MatchError in costFunc, BioLibs.scala:144)"; return HOVL_E_BAD_BASE; }
  if (hstatus & 1u) { c.err = "a dispatched read B is shorter than the band width (BioLibs.scala:648)"; return HOVL_E_SHORT_READ; }
  return HOVL_OK;
}

static int run_core(Ctx &c, const hovl_settings &s, Reads &R, hovl_result &res) {
  hovl_stats_t &st = res.st;
  build_reads(c, s, R, st);
  Tuples T; Table B; Pairs P;
  encode_kmers(c, s, R, T, st, s.order != 0 && c.pair_mode == 0);     // the reference-order replay reads stored values
  sort_kmers(c, s, T, st);
  group_kmers(c, s, R, T, B, s.order != 0 && c.pair_mode == 0, st);
  int rc;
  if (s.order == 0) {
    DevBuf<uint32_t> d_perm;
    kmerdata_iteration_order(c, T, B, d_perm, st);
    T = Tuples();                                     // the sorted tuples are no longer needed
    count_pairs(c, s, R, B, d_perm.get(), true, false, P, st);
    B = Table();
    std::vector<uint32_t> ha, hb, hc;
    dispatch_reference_order(c, s, P, ha, hb, hc, st);
    P = Pairs();
    const size_t nd = ha.size();
    DevBuf<uint32_t> da(c, nd), db(c, nd), dc(c, nd);
    h2d(c, da.get(), ha.data(), nd); h2d(c, db.get(), hb.data(), nd); h2d(c, dc.get(), hc.data(), nd);
    rc = align_and_collect(c, s, R, da.get(), db.get(), dc.get(), nd, res);
  } else {
    T = Tuples();
    count_pairs(c, s, R, B, nullptr, false, true, P, st);
    B = Table();
    DevBuf<uint32_t> da, db, dc;
    const size_t nd = select_pairs(c, s, P, da, db, dc);
    P = Pairs();
    DevBuf<unsigned long long> nl(c, 1);
    HOVL_CUDA(cudaMemsetAsync(nl.get(), 0, 8, c.stream));
    if (nd) HOVL_LAUNCH(c, lead_count_kernel, (unsigned)div_up(nd, 256), 256, 0, da.get(), nd, nl.get());
    unsigned long long hl = 0;
    HOVL_CUDA(cudaMemcpyAsync(&hl, nl.get(), 8, cudaMemcpyDeviceToHost, c.stream));
    rc = align_and_collect(c, s, R, da.get(), db.get(), dc.get(), nd, res);
    st.n_leads = hl;
  }
  return rc;
}

static int run_entry(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads,
                     uint64_t n_bases, bool on_device, hovl_result **out) {
  if (out) *out = nullptr;
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(out != nullptr, HOVL_E_ARG, "out is NULL");
    check_settings(s, n_reads);
    Ctx &c = h->c;
    c.launches = 0; c.ktimes.clear();
    const uint64_t t0 = wall_ns();
    hovl_result *res = new hovl_result();
    memset(&res->st, 0, sizeof res->st);
    int rc;
    try {
      Reads R;
      if (on_device) {
        HOVL_REQUIRE(offsets != nullptr && (bases != nullptr || n_bases == 0), HOVL_E_ARG, "NULL device pointer");
        R.n_reads = n_reads; R.n_bases = n_bases; R.d_bases = bases; R.d_offsets = offsets;
      } else {
        upload_reads(c, R, bases, offsets, n_reads, &res->st.ns_h2d);
      }
      rc = run_core(c, *s, R, *res);
      sync(c);
      resolve_timers(c);
    } catch (...) { delete res; throw; }
    c.arena.consolidate();                              // all device arrays of the run are released by now
    res->st.gpu_launches = c.launches;
    res->st.ns_total = wall_ns() - t0;
    *out = res;
    return rc;
  });
}

extern "C" int hovl_run(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                        uint32_t n_reads, hovl_result **out) {
  if (h && !h->subs.empty()) { if (out) *out = nullptr; h->c.err.clear(); return multi_run(h, s, bases, offsets, n_reads, out); }
  return run_entry(h, s, bases, offsets, n_reads, 0, false, out);
}
extern "C" int hovl_run_device(hovl_ctx *h, const hovl_settings *s, const uint8_t *d_bases, const uint64_t *d_offsets,
                               uint32_t n_reads, uint64_t n_bases, hovl_result **out) {
  return run_entry(h, s, d_bases, d_offsets, n_reads, n_bases, true, out);
}
extern "C" const hovl_record *hovl_records(const hovl_result *r, uint64_t *n) {
  if (!r) { if (n) *n = 0; return nullptr; }
  if (n) *n = r->n;
  return r->recs;
}
extern "C" int hovl_stats(const hovl_result *r, hovl_stats_t *st) {
  if (!r || !st) return HOVL_E_ARG;
  *st = r->st;
  return HOVL_OK;
}
extern "C" void hovl_result_free(hovl_result *r) { delete r; }

// Overlap.print (ObjectStore.scala:127-135) + calcOverlaps (Project4.scala:806-820)
extern "C" uint64_t hovl_format_ovl(const hovl_record *recs, uint64_t n, char *buf, uint64_t cap) {
  uint64_t need = 0; char tmp[160];
  for (uint64_t i = 0; i < n; i++) {
    if (!(recs[i].flags & HOVL_F_OVL_VALID)) continue;
    const int l = snprintf(tmp, sizeof tmp, "{OVL\nadj:N\nrds:%d,%d\nscr:0\nahg:%d\nbhg:%d\n}\n", recs[i].a, recs[i].b,
                           recs[i].ahg, recs[i].bhg);
    if (buf && need + (uint64_t)l <= cap) memcpy(buf + need, tmp, (size_t)l);
    need += (uint64_t)l;
  }
  return need;
}

// Iteration order of a gnu.trove 3.0.3 TIntObjectHashMap filled with `keys` (distinct) in the
// given insertion order: order_out[j] = index of the j-th key the iterator yields.  Host only.
extern "C" int hovl_trove_order(const int32_t *keys, uint64_t n, uint32_t *order_out) {
  if ((!keys || !order_out) && n) return HOVL_E_ARG;
  try {
    std::vector<uint32_t> ord;
    trove_iteration_order(keys, (size_t)n, ord);
    if (ord.size() != n) return HOVL_E_ARG;
    if (n) memcpy(order_out, ord.data(), n * sizeof(uint32_t));
    return HOVL_OK;
  } catch (...) { return HOVL_E_NOMEM; }
}

// Same order, computed on the context's device (trove.cu): what hovl_run(order = 0) uses for PairData.  Host arrays in / out.
extern "C" int hovl_trove_order_device(hovl_ctx *h, const int32_t *keys, uint64_t n, uint32_t *order_out) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE((keys && order_out) || n == 0, HOVL_E_ARG, "NULL pointer");
    Ctx &c = h->c;
    DevBuf<int32_t> dk(c, n);
    h2d(c, dk.get(), keys, (size_t)n);
    DevBuf<uint32_t> d_order;
    trove_iteration_order_device(c, dk.get(), (size_t)n, d_order);
    d2h(c, order_out, d_order.get(), (size_t)n);
    sync(c);
    return HOVL_OK;
  });
}

// ---------------------------------------------------------------- stage entry points
template <typename T>
static T *host_array(size_t n) {
  T *p = (T *)malloc((n ? n : 1) * sizeof(T));
  if (!p) throw std::bad_alloc();
  return p;
}

extern "C" int hovl_stage_table(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                uint32_t n_reads, uint32_t **codes, uint32_t **read_ids, uint32_t **loc_f32, uint64_t *n_kmers) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(codes && read_ids && loc_f32 && n_kmers, HOVL_E_ARG, "NULL output pointer");
    check_settings(s, n_reads);
    Ctx &c = h->c; hovl_stats_t st; memset(&st, 0, sizeof st);
    Reads R; upload_reads(c, R, bases, offsets, n_reads, &st.ns_h2d);
    build_reads(c, *s, R, st);
    Tuples T;
    encode_kmers(c, *s, R, T, st);
    sort_kmers(c, *s, T, st);
    const size_t n = T.n;
    std::vector<uint64_t> v(n);
    uint32_t *hc = host_array<uint32_t>(n), *hr = host_array<uint32_t>(n), *hl = host_array<uint32_t>(n);
    d2h(c, hc, T.k(), n); d2h(c, v.data(), T.v(), n);
    sync(c); resolve_timers(c);
    for (size_t i = 0; i < n; i++) { hr[i] = (uint32_t)(v[i] >> 32); hl[i] = (uint32_t)v[i]; }
    *codes = hc; *read_ids = hr; *loc_f32 = hl; *n_kmers = n;
    return HOVL_OK;
  });
}

extern "C" int hovl_stage_cover(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                uint32_t n_reads, uint64_t *n_buckets, uint64_t **sizes, uint64_t **counts, uint64_t *n_sizes) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(n_buckets && sizes && counts && n_sizes, HOVL_E_ARG, "NULL output pointer");
    check_settings(s, n_reads);
    Ctx &c = h->c; hovl_stats_t st; memset(&st, 0, sizeof st);
    Reads R; upload_reads(c, R, bases, offsets, n_reads, &st.ns_h2d);
    build_reads(c, *s, R, st);
    Tuples T; Table B;
    encode_kmers(c, *s, R, T, st);
    sort_kmers(c, *s, T, st);
    group_kmers(c, *s, R, T, B, false, st);
    const size_t nb = B.n_buckets;
    std::vector<uint64_t> start(nb + 1);
    d2h(c, start.data(), B.bucket_start.get(), nb + 1);
    sync(c); resolve_timers(c);
    std::map<uint64_t, uint64_t> hist;                          // bucket size -> distinct k-mers of that size
    for (size_t b = 0; b < nb; b++) hist[start[b + 1] - start[b]]++;
    uint64_t *hs = host_array<uint64_t>(hist.size()), *hc = host_array<uint64_t>(hist.size());
    size_t i = 0;
    for (auto &kv : hist) { hs[i] = kv.first; hc[i] = kv.second; i++; }
    *n_buckets = nb; *sizes = hs; *counts = hc; *n_sizes = hist.size();
    return HOVL_OK;
  });
}

extern "C" int hovl_stage_pairs(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                uint32_t n_reads, uint64_t **keys, uint32_t **counts, uint64_t *n_pairs, uint64_t *n_instances) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(keys && counts && n_pairs, HOVL_E_ARG, "NULL output pointer");
    check_settings(s, n_reads);
    Ctx &c = h->c; hovl_stats_t st; memset(&st, 0, sizeof st);
    Reads R; upload_reads(c, R, bases, offsets, n_reads, &st.ns_h2d);
    build_reads(c, *s, R, st);
    Tuples T; Table B; Pairs P;
    encode_kmers(c, *s, R, T, st, c.pair_mode == 0);
    sort_kmers(c, *s, T, st);
    group_kmers(c, *s, R, T, B, c.pair_mode == 0, st);
    T = Tuples();
    count_pairs(c, *s, R, B, nullptr, false, false, P, st);
    const size_t n = P.n;
    uint64_t *hk = host_array<uint64_t>(n); uint32_t *hc = host_array<uint32_t>(n);
    d2h(c, hk, P.key.get(), n); d2h(c, hc, P.count.get(), n);
    sync(c); resolve_timers(c);
    const uint64_t mask = (1ull << P.id_bits) - 1;
    for (size_t i = 0; i < n; i++) hk[i] = ((hk[i] >> P.id_bits) << 32) | (hk[i] & mask);
    *keys = hk; *counts = hc; *n_pairs = n;
    if (n_instances) *n_instances = P.n_instances;
    return HOVL_OK;
  });
}

extern "C" int hovl_stage_align(hovl_ctx *h, const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                uint32_t n_reads, const uint32_t *pair_a, const uint32_t *pair_b, uint64_t n_pairs,
                                hovl_record *records_out) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE((pair_a && pair_b && records_out) || n_pairs == 0, HOVL_E_ARG, "NULL pair/record pointer");
    check_settings(s, n_reads);
    for (uint64_t i = 0; i < n_pairs; i++)
      HOVL_REQUIRE(pair_a[i] >= 1 && pair_a[i] <= n_reads && pair_b[i] >= 1 && pair_b[i] <= n_reads, HOVL_E_ARG,
                   "pair id out of range (ids are 1-based)");
    Ctx &c = h->c;
    hovl_result res; memset(&res.st, 0, sizeof res.st);
    Reads R; upload_reads(c, R, bases, offsets, n_reads, &res.st.ns_h2d);
    build_reads(c, *s, R, res.st);
    DevBuf<uint32_t> da(c, n_pairs), db(c, n_pairs);
    h2d(c, da.get(), pair_a, n_pairs); h2d(c, db.get(), pair_b, n_pairs);
    hovl_settings s2 = *s; s2.filter_valid = 0;               // one record per input pair, in order
    const int rc = align_and_collect(c, s2, R, da.get(), db.get(), nullptr, n_pairs, res);
    resolve_timers(c);
    if (n_pairs) memcpy(records_out, res.recs, n_pairs * sizeof(hovl_record));
    return rc;
  });
}

// ---------------------------------------------------------------- sharded path
extern "C" uint32_t hovl_shard_of_code(uint32_t code, uint32_t n_parts) { return shard_of_code(code, n_parts ? n_parts : 1); }
extern "C" uint32_t hovl_shard_of_code_range(uint32_t code, int32_t k, uint32_t n_parts) {
  return shard_of_code_range(code, 2 * (k < 16 ? k : 16), n_parts ? n_parts : 1);
}
extern "C" uint32_t hovl_shard_of_read(uint32_t read_id, uint32_t n_reads_total, uint32_t n_parts) {
  return shard_of_read(read_id ? read_id : 1, n_reads_total, n_parts ? n_parts : 1);
}

extern "C" int hovl_shard_encode(hovl_ctx *h, const hovl_settings *s, const uint8_t *d_bases, const uint64_t *d_offsets,
                                 uint32_t n_local, uint64_t n_bases_local, uint32_t id_base, uint32_t n_reads_total,
                                 uint32_t n_parts, uint32_t *d_codes_out, uint64_t *d_vals_out, uint64_t tuple_cap,
                                 uint32_t *d_packed_out, uint64_t packed_cap, uint32_t *d_len_out, uint32_t *d_bad_out,
                                 uint64_t *counts_out, uint64_t *n_words_out) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(d_offsets && counts_out && n_words_out && d_len_out && d_bad_out, HOVL_E_ARG, "NULL pointer");
    HOVL_REQUIRE(s != nullptr && s->k >= 1, HOVL_E_ARG, "bad settings");
    HOVL_REQUIRE(s->id_mode == 1 && s->order == 1, HOVL_E_UNSUPPORTED, "the sharded path needs id_mode=1, order=1");
    HOVL_REQUIRE((uint64_t)id_base + n_local <= n_reads_total, HOVL_E_ARG, "read slice outside 1..n_reads_total");
    Ctx &c = h->c;
    c.launches = 0; c.ktimes.clear();
    memset(&h->sh_st, 0, sizeof h->sh_st);
    h->sh_keys.release(); h->sh_counts.release(); h->sh_n = 0;
    const uint64_t t0 = wall_ns();
    Reads R;
    R.n_reads = n_local; R.n_bases = n_bases_local; R.d_bases = d_bases; R.d_offsets = d_offsets; R.id_base = id_base;
    build_reads(c, *s, R, h->sh_st);
    HOVL_REQUIRE(R.n_words <= packed_cap && R.n_kmers <= tuple_cap, HOVL_E_ARG, "output buffer too small");
    h->sh_T = Tuples();
    h->sh_range_bits = c.shard_by_range ? 2 * (s->k < 16 ? s->k : 16) : 0;
    Tuples T;
    encode_kmers(c, *s, R, T, h->sh_st);
    if (d_codes_out && d_vals_out) partition_tuples(c, T.k(), T.v(), T.n, n_parts, h->sh_range_bits, d_codes_out, d_vals_out, counts_out);
    else {                                  // peer-memory exchange: count now, hovl_shard_scatter stores them remotely
      count_owners(c, T.k(), T.n, n_parts, h->sh_range_bits, counts_out);
      T.key[T.cur ^ 1].release(); T.val[T.cur ^ 1].release();
      h->sh_T = std::move(T);
    }
    if (R.n_words) HOVL_CUDA(cudaMemcpyAsync(d_packed_out, R.packed.get(), R.n_words * 4, cudaMemcpyDeviceToDevice, c.stream));
    if (n_local) {
      HOVL_CUDA(cudaMemcpyAsync(d_len_out, R.len.get(), (size_t)n_local * 4, cudaMemcpyDeviceToDevice, c.stream));
      HOVL_CUDA(cudaMemcpyAsync(d_bad_out, R.bad.get(), (size_t)n_local * 4, cudaMemcpyDeviceToDevice, c.stream));
    }
    sync(c);
    resolve_timers(c);
    *n_words_out = R.n_words;
    h->sh_st.gpu_launches = c.launches;
    h->sh_st.ns_total = wall_ns() - t0;
    return HOVL_OK;
  });
}

extern "C" int hovl_shard_scatter(hovl_ctx *h, uint32_t n_parts, const uint64_t *peer_codes, const uint64_t *peer_vals,
                                  const uint64_t *peer_base) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(peer_codes && peer_vals && peer_base, HOVL_E_ARG, "NULL pointer");
    Ctx &c = h->c;
    const uint64_t t0 = wall_ns();
    {
      StageTimer tm(c, &h->sh_st.ns_encode, "peer_scatter");
      scatter_tuples_peer(c, h->sh_T.k(), h->sh_T.v(), h->sh_T.n, n_parts, h->sh_range_bits, peer_codes, peer_vals, peer_base);
    }
    h->sh_T = Tuples();
    resolve_timers(c);
    h->sh_st.gpu_launches = c.launches;
    h->sh_st.ns_total += wall_ns() - t0;
    return HOVL_OK;
  });
}

// Device buffers other ranks of the node may write into (plain cudaMalloc, outside the arena: an IPC handle
// names a whole allocation).
extern "C" int hovl_peer_alloc(hovl_ctx *h, uint64_t bytes, void **d_ptr, uint8_t *handle64) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(d_ptr && handle64, HOVL_E_ARG, "NULL pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void *p = nullptr;
    if (cudaMalloc(&p, bytes ? bytes : 256) != cudaSuccess) { (void)cudaGetLastError(); throw Error(HOVL_E_NOMEM, "cudaMalloc of a peer buffer failed"); }
    cudaIpcMemHandle_t hd;
    const cudaError_t e = cudaIpcGetMemHandle(&hd, p);
    if (e != cudaSuccess) { (void)cudaGetLastError(); cudaFree(p); throw Error(HOVL_E_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e)); }
    memcpy(handle64, &hd, 64);
    *d_ptr = p;
    return HOVL_OK;
  });
}
extern "C" int hovl_peer_open(hovl_ctx *h, const uint8_t *handle64, void **d_ptr) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(d_ptr && handle64, HOVL_E_ARG, "NULL pointer");
    cudaIpcMemHandle_t hd; memcpy(&hd, handle64, 64);
    void *p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { (void)cudaGetLastError(); throw Error(HOVL_E_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e)); }
    *d_ptr = p;
    return HOVL_OK;
  });
}
extern "C" int hovl_peer_close(hovl_ctx *h, void *d_ptr) {
  return guarded(h, [&]() -> int { if (d_ptr && cudaIpcCloseMemHandle(d_ptr) != cudaSuccess) (void)cudaGetLastError(); return HOVL_OK; });
}
extern "C" int hovl_peer_free(hovl_ctx *h, void *d_ptr) {
  return guarded(h, [&]() -> int { if (d_ptr && cudaFree(d_ptr) != cudaSuccess) (void)cudaGetLastError(); return HOVL_OK; });
}

// ---- alignment re-balance by DP cells (rebalance.cu): hovl_shard_align split in two, with an exchange of pairs in between
extern "C" int hovl_shard_select(hovl_ctx *h, const hovl_settings *s, uint64_t *d_keys, uint32_t *d_counts, uint64_t n_partial,
                                 const uint32_t *d_len_all, uint32_t n_reads_total, uint64_t *n_pairs_out, uint64_t *cost_out) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(s && d_len_all && n_pairs_out && cost_out && (n_partial == 0 || (d_keys && d_counts)), HOVL_E_ARG, "NULL pointer");
    Ctx &c = h->c;
    const uint64_t launches0 = c.launches, t0 = wall_ns();
    rebalance_select(h, *s, d_keys, d_counts, (size_t)n_partial, d_len_all, n_reads_total, n_pairs_out, cost_out);
    DevBuf<unsigned long long> nl(c, 1);                 // DispatchData.size of this rank's leads (before pairs move)
    HOVL_CUDA(cudaMemsetAsync(nl.get(), 0, 8, c.stream));
    if (h->sh_npairs) HOVL_LAUNCH(c, lead_count_kernel, (unsigned)div_up(h->sh_npairs, 256), 256, 0, h->sh_da.get(), h->sh_npairs, nl.get());
    unsigned long long hl = 0;
    HOVL_CUDA(cudaMemcpyAsync(&hl, nl.get(), 8, cudaMemcpyDeviceToHost, c.stream));
    sync(c);
    resolve_timers(c);
    h->sh_st.n_leads = hl;
    h->sh_st.gpu_launches += c.launches - launches0;
    h->sh_st.ns_total += wall_ns() - t0;
    return HOVL_OK;
  });
}
extern "C" int hovl_shard_split(hovl_ctx *h, uint64_t cost_before, uint64_t cost_total, uint32_t n_parts, uint64_t *counts_out) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(counts_out && n_parts >= 1 && n_parts <= 256, HOVL_E_ARG, "bad argument");
    rebalance_split(h, cost_before, cost_total, n_parts, counts_out);
    return HOVL_OK;
  });
}
// (a, b, count) triplets, 12 bytes per pair, in (a, b) order: one buffer so that one all-to-all carries them
extern "C" int hovl_shard_fetch_selected(hovl_ctx *h, uint32_t *d_abc_out, uint64_t cap_pairs) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(cap_pairs >= h->sh_npairs && (h->sh_npairs == 0 || d_abc_out), HOVL_E_ARG, "output buffer too small");
    rebalance_fetch(h, d_abc_out);
    return HOVL_OK;
  });
}
extern "C" int hovl_shard_align_pairs(hovl_ctx *h, const hovl_settings *s, const uint32_t *d_abc, uint64_t n_pairs,
                                      const uint32_t *d_packed_all, uint64_t n_words_all, const uint32_t *d_len_all,
                                      const uint32_t *d_bad_all, uint32_t n_reads_total, hovl_record *d_records_out,
                                      uint64_t records_cap, hovl_result **out) {
  if (out) *out = nullptr;
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(s && out && d_len_all && d_bad_all && (n_pairs == 0 || d_abc), HOVL_E_ARG, "NULL pointer");
    Ctx &c = h->c;
    const uint64_t launches0 = c.launches, t0 = wall_ns();
    hovl_result *res = new hovl_result();
    res->st = h->sh_st;
    int rc;
    try {
      Reads R;
      R.packed.borrow(const_cast<uint32_t *>(d_packed_all), n_words_all);
      R.len.borrow(const_cast<uint32_t *>(d_len_all), n_reads_total);
      R.bad.borrow(const_cast<uint32_t *>(d_bad_all), n_reads_total);
      shard_read_layout(c, d_len_all, n_reads_total, R);
      HOVL_REQUIRE(R.n_words <= n_words_all, HOVL_E_ARG, "packed reads shorter than their lengths say");
      DevBuf<uint32_t> da, db, dc;
      rebalance_unpack(c, d_abc, (size_t)n_pairs, da, db, dc);
      rc = align_and_collect(c, *s, R, da.get(), db.get(), dc.get(), (size_t)n_pairs, *res, d_records_out, records_cap);
      res->st.n_leads = h->sh_st.n_leads;
      sync(c);
      resolve_timers(c);
    } catch (...) { delete res; throw; }
    res->st.gpu_launches += c.launches - launches0;
    res->st.ns_total += wall_ns() - t0;
    c.arena.consolidate();
    *out = res;
    return rc;
  });
}

// ---- replicated-table variant: ranges of codes are sorted where they land, the sorted slices are gathered on every
// rank (their concatenation in rank order IS the sorted table), and each rank counts and aligns the pairs whose fst
// is one of ITS reads.  No partial counts travel, so the 7..222 filter stays fused with the counting.
extern "C" int hovl_shard_sort(hovl_ctx *h, const hovl_settings *s, uint32_t *d_codes, uint64_t *d_vals, uint64_t n_tuples) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(s && (n_tuples == 0 || (d_codes && d_vals)), HOVL_E_ARG, "NULL pointer");
    HOVL_REQUIRE(n_tuples < 0xfffffff0ull, HOVL_E_UNSUPPORTED, "more than 2^32 k-mers on one shard");
    Ctx &c = h->c;
    const uint64_t launches0 = c.launches, t0 = wall_ns();
    {
      Tuples T;
      T.n = n_tuples; T.cur = 0;
      T.key[0].borrow(d_codes, n_tuples); T.val[0].borrow(d_vals, n_tuples);
      T.key[1].alloc(c, n_tuples); T.val[1].alloc(c, n_tuples);
      sort_kmers(c, *s, T, h->sh_st);
      if (T.cur == 1 && n_tuples) {                       // the caller's buffers hold the result
        HOVL_CUDA(cudaMemcpyAsync(d_codes, T.k(), n_tuples * 4, cudaMemcpyDeviceToDevice, c.stream));
        HOVL_CUDA(cudaMemcpyAsync(d_vals, T.v(), n_tuples * 8, cudaMemcpyDeviceToDevice, c.stream));
      }
      sync(c);
    }
    resolve_timers(c);
    h->sh_st.gpu_launches += c.launches - launches0;
    h->sh_st.ns_total += wall_ns() - t0;
    return HOVL_OK;
  });
}

// Copies every rank's sorted slice (peer-mapped buffers, hovl_peer_open) into one local array: the all-gather of the
// table as n_parts device-to-device copies over NVLink.  counts[p] = tuples of rank p; the caller's barrier before this
// call says that every slice is sorted, the one after it that the slices may be overwritten.
extern "C" int hovl_shard_gather_peer(hovl_ctx *h, uint32_t n_parts, const uint64_t *peer_codes, const uint64_t *peer_vals,
                                      const uint64_t *counts, uint32_t *d_codes_all, uint64_t *d_vals_all) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(peer_codes && peer_vals && counts && d_codes_all && d_vals_all, HOVL_E_ARG, "NULL pointer");
    Ctx &c = h->c;
    const uint64_t t0 = wall_ns();
    {
      // one stream per peer: a single stream's peer copies run one after the other on one copy engine (75 ms for the 14.6 GB
      // a rank receives on K4 at 8 GPUs, 200 GB/s); side by side they use the NVSwitch fabric's width
      StageTimer tm(c, &h->sh_st.ns_kmer_sort, "table_gather");
      sync(c);
      while (h->sh_streams.size() < n_parts) {
        cudaStream_t st = nullptr;
        HOVL_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        h->sh_streams.push_back(st);
      }
      uint64_t at = 0;
      for (uint32_t p = 0; p < n_parts; p++) {
        if (counts[p]) {
          HOVL_CUDA(cudaMemcpyAsync(d_codes_all + at, reinterpret_cast<const void *>(peer_codes[p]), counts[p] * 4, cudaMemcpyDeviceToDevice, h->sh_streams[p]));
          HOVL_CUDA(cudaMemcpyAsync(d_vals_all + at, reinterpret_cast<const void *>(peer_vals[p]), counts[p] * 8, cudaMemcpyDeviceToDevice, h->sh_streams[p]));
        }
        at += counts[p];
      }
      for (uint32_t p = 0; p < n_parts; p++) HOVL_CUDA(cudaStreamSynchronize(h->sh_streams[p]));
    }
    sync(c);
    resolve_timers(c);
    h->sh_st.ns_total += wall_ns() - t0;
    return HOVL_OK;
  });
}

extern "C" int hovl_shard_pairs_local(hovl_ctx *h, const hovl_settings *s, const uint32_t *d_codes_all, const uint64_t *d_vals_all,
                                      uint64_t n_all, const uint32_t *d_packed_all, uint64_t n_words_all, const uint32_t *d_len_all,
                                      const uint32_t *d_bad_all, uint32_t n_reads_total, uint32_t read_lo, uint32_t n_local,
                                      hovl_result **out) {
  if (out) *out = nullptr;
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(s && out && d_len_all && d_bad_all && (n_all == 0 || (d_codes_all && d_vals_all)), HOVL_E_ARG, "NULL pointer");
    HOVL_REQUIRE(s->id_mode == 1 && s->order == 1, HOVL_E_UNSUPPORTED, "the sharded path needs id_mode=1, order=1");
    HOVL_REQUIRE((uint64_t)read_lo + n_local <= n_reads_total, HOVL_E_ARG, "read slice outside 1..n_reads_total");
    HOVL_REQUIRE(n_all < 0xfffffff0ull, HOVL_E_UNSUPPORTED, "more than 2^32 k-mers in the replicated table");
    Ctx &c = h->c;
    const uint64_t launches0 = c.launches, t0 = wall_ns();
    hovl_result *res = new hovl_result();
    res->st = h->sh_st;
    int rc;
    try {
      hovl_stats_t &st = res->st;
      Reads R;
      R.packed.borrow(const_cast<uint32_t *>(d_packed_all), n_words_all);
      R.len.borrow(const_cast<uint32_t *>(d_len_all), n_reads_total);
      R.bad.borrow(const_cast<uint32_t *>(d_bad_all), n_reads_total);
      shard_read_layout(c, d_len_all, n_reads_total, R);
      HOVL_REQUIRE(R.n_words <= n_words_all, HOVL_E_ARG, "packed reads shorter than their lengths say");
      shard_read_koff(c, d_len_all, n_reads_total, s->k, R);
      HOVL_REQUIRE(R.n_kmers == n_all, HOVL_E_ARG, "the gathered table does not hold every k-mer of the reads");
      Table B; Pairs P;
      {
        Tuples T;
        T.n = n_all; T.cur = 0; T.compact = false;
        T.key[0].borrow(const_cast<uint32_t *>(d_codes_all), n_all); T.val[0].borrow(const_cast<uint64_t *>(d_vals_all), n_all);
        group_kmers(c, *s, R, T, B, true, st);           // the whole table on every rank (replicated work)
      }
      if (read_lo != 0) st.n_buckets = 0;                // counted once in the job's sum
      GatherSrc src;
      src.n_reads = R.n_reads; src.koff = R.koff.get(); src.len = R.len.get(); src.ent = nullptr; src.n_tuples = n_all;
      src.first = read_lo; src.count = n_local;
      src.max_kmers = R.max_len >= (uint32_t)s->k ? R.max_len - (uint32_t)s->k + 1u : 0u;
      if (n_local) count_pairs_gather(c, *s, src, B, true, true, P, st);
      else { P.id_bits = 1; while (((uint64_t)1 << P.id_bits) <= (uint64_t)R.n_reads) P.id_bits++; P.key.alloc(c, 0); P.count.alloc(c, 0); }
      B = Table();
      DevBuf<uint32_t> da, db, dc;
      const size_t nd = select_pairs(c, *s, P, da, db, dc);
      P = Pairs();
      DevBuf<unsigned long long> nl(c, 1);
      HOVL_CUDA(cudaMemsetAsync(nl.get(), 0, 8, c.stream));
      if (nd) HOVL_LAUNCH(c, lead_count_kernel, (unsigned)div_up(nd, 256), 256, 0, da.get(), nd, nl.get());
      unsigned long long hl = 0;
      HOVL_CUDA(cudaMemcpyAsync(&hl, nl.get(), 8, cudaMemcpyDeviceToHost, c.stream));
      h->sh_recs.release(); h->sh_nrec = 0;
      rc = align_and_collect(c, *s, R, da.get(), db.get(), dc.get(), nd, *res, nullptr, 0, &h->sh_recs);
      h->sh_nrec = (size_t)st.n_records;
      st.n_leads = hl;
      sync(c);
      resolve_timers(c);
    } catch (...) { delete res; throw; }
    res->st.gpu_launches += c.launches - launches0;
    res->st.ns_total += wall_ns() - t0;
    *out = res;
    return rc;
  });
}

// The records hovl_shard_pairs_local left on the device -> `dst` (device memory, or page-locked host memory such as a
// node-shared landing buffer), once.  cap in records.
extern "C" int hovl_shard_fetch_records(hovl_ctx *h, hovl_record *dst, uint64_t cap) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(cap >= h->sh_nrec && (h->sh_nrec == 0 || dst), HOVL_E_ARG, "record buffer too small");
    Ctx &c = h->c;
    if (h->sh_nrec) HOVL_CUDA(cudaMemcpyAsync(dst, h->sh_recs.get(), h->sh_nrec * sizeof(hovl_record), cudaMemcpyDefault, c.stream));
    sync(c);
    h->sh_recs.release(); h->sh_nrec = 0;
    c.arena.consolidate();
    return HOVL_OK;
  });
}

extern "C" int hovl_shard_pairs(hovl_ctx *h, const hovl_settings *s, uint32_t *d_codes, uint64_t *d_vals, uint64_t n_tuples,
                                uint32_t n_reads_total, uint32_t n_parts, uint64_t *counts_out, uint64_t *n_partial_out) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(s && counts_out && n_partial_out && (n_tuples == 0 || (d_codes && d_vals)), HOVL_E_ARG, "NULL pointer");
    HOVL_REQUIRE(n_tuples < 0xfffffff0ull, HOVL_E_UNSUPPORTED, "more than 2^32 k-mers on one shard");
    Ctx &c = h->c;
    const uint64_t launches0 = c.launches;
    const uint64_t t0 = wall_ns();
    hovl_stats_t &st = h->sh_st;
    Pairs P;
    {
      Tuples T; Table B; Reads none;
      T.n = n_tuples; T.cur = 0;
      T.key[0].borrow(d_codes, n_tuples); T.val[0].borrow(d_vals, n_tuples);
      T.key[1].alloc(c, n_tuples); T.val[1].alloc(c, n_tuples);
      sort_kmers(c, *s, T, st);
      DevBuf<uint64_t> ekoff, ent(c, n_tuples);
      shard_read_offsets(c, T.v(), n_tuples, n_reads_total, ekoff);
      DevBuf<uint32_t> ecursor(c, (size_t)n_reads_total + 1);
      HOVL_CUDA(cudaMemsetAsync(ecursor.get(), 0, ecursor.bytes(), c.stream));
      const EntIndex ei{ent.get(), ekoff.get(), ecursor.get()};
      group_kmers(c, *s, none, T, B, false, st, &ei);
      T = Tuples();
      GatherSrc src;
      src.n_reads = n_reads_total; src.n_tuples = n_tuples; src.koff = ekoff.get(); src.ent = ent.get();
      count_pairs_gather(c, *s, src, B, false, false, P, st);
    }
    h->sh_keys.alloc(c, P.n); h->sh_counts.alloc(c, P.n); h->sh_n = P.n;
    partition_pairs(c, P.key.get(), P.count.get(), P.n, P.id_bits, n_reads_total, n_parts, h->sh_keys.get(),
                    h->sh_counts.get(), counts_out);
    resolve_timers(c);
    *n_partial_out = P.n;
    st.gpu_launches += c.launches - launches0;
    st.ns_total += wall_ns() - t0;
    return HOVL_OK;
  });
}

extern "C" int hovl_shard_fetch_pairs(hovl_ctx *h, uint64_t *d_keys_out, uint32_t *d_counts_out, uint64_t cap) {
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(cap >= h->sh_n && (h->sh_n == 0 || (d_keys_out && d_counts_out)), HOVL_E_ARG, "output buffer too small");
    Ctx &c = h->c;
    if (h->sh_n) {
      HOVL_CUDA(cudaMemcpyAsync(d_keys_out, h->sh_keys.get(), h->sh_n * 8, cudaMemcpyDeviceToDevice, c.stream));
      HOVL_CUDA(cudaMemcpyAsync(d_counts_out, h->sh_counts.get(), h->sh_n * 4, cudaMemcpyDeviceToDevice, c.stream));
    }
    sync(c);
    h->sh_keys.release(); h->sh_counts.release(); h->sh_n = 0;
    return HOVL_OK;
  });
}

extern "C" int hovl_shard_align(hovl_ctx *h, const hovl_settings *s, uint64_t *d_keys, uint32_t *d_counts, uint64_t n_partial,
                                const uint32_t *d_packed_all, uint64_t n_words_all, const uint32_t *d_len_all,
                                const uint32_t *d_bad_all, uint32_t n_reads_total, hovl_record *d_records_out,
                                uint64_t records_cap, hovl_result **out) {
  if (out) *out = nullptr;
  return guarded(h, [&]() -> int {
    HOVL_REQUIRE(s && out && d_len_all && d_bad_all, HOVL_E_ARG, "NULL pointer");
    Ctx &c = h->c;
    const uint64_t launches0 = c.launches;
    const uint64_t t0 = wall_ns();
    hovl_result *res = new hovl_result();
    res->st = h->sh_st;
    int rc;
    try {
      hovl_stats_t &st = res->st;
      Pairs P;
      merge_partial_pairs(c, d_keys, d_counts, n_partial, n_reads_total, P, st);
      DevBuf<uint32_t> da, db, dc;
      const size_t nd = select_pairs(c, *s, P, da, db, dc);
      P = Pairs();
      Reads R;
      R.packed.borrow(const_cast<uint32_t *>(d_packed_all), n_words_all);
      R.len.borrow(const_cast<uint32_t *>(d_len_all), n_reads_total);
      R.bad.borrow(const_cast<uint32_t *>(d_bad_all), n_reads_total);
      shard_read_layout(c, d_len_all, n_reads_total, R);
      HOVL_REQUIRE(R.n_words <= n_words_all, HOVL_E_ARG, "packed reads shorter than their lengths say");
      DevBuf<unsigned long long> nl(c, 1);
      HOVL_CUDA(cudaMemsetAsync(nl.get(), 0, 8, c.stream));
      if (nd) HOVL_LAUNCH(c, lead_count_kernel, (unsigned)div_up(nd, 256), 256, 0, da.get(), nd, nl.get());
      unsigned long long hl = 0;
      HOVL_CUDA(cudaMemcpyAsync(&hl, nl.get(), 8, cudaMemcpyDeviceToHost, c.stream));
      rc = align_and_collect(c, *s, R, da.get(), db.get(), dc.get(), nd, *res, d_records_out, records_cap);
      st.n_leads = hl;
      sync(c);
      resolve_timers(c);
    } catch (...) { delete res; throw; }
    res->st.gpu_launches += c.launches - launches0;
    res->st.ns_total += wall_ns() - t0;
    c.arena.consolidate();
    *out = res;
    return rc;
  });
}

// ---------------------------------------------------------------- no-allocation entry point
// hovl_filter_records_noalloc: the record compaction of align_and_collect (scan_*<Quad, RecIn, RecOut>: keep
// Alignment.valid records, count duds / valid alignments / valid overlaps) for callers that may not allocate
// (CUDA graph capture, serving loops).  The scan's scratch (tile sums + the status word) lives in a file-scope
// __device__ array instead of the context's arena; nothing here calls cudaMalloc / cudaFree or an allocating
// library routine.  No counterpart in the reference.
//
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current (below).  Every call on a device shares the
// array, so calls that overlap on the GPU race; a call clears what must start at zero (the status word) and
// overwrites every tile sum it reads, so nothing depends on the previous call's contents.
constexpr uint64_t NOALLOC_MAX_RECORDS = HOVL_NOALLOC_MAX_RECORDS;
constexpr size_t NOALLOC_TILES = (size_t)((NOALLOC_MAX_RECORDS + SCAN_TILE - 1) / SCAN_TILE);
constexpr size_t NOALLOC_STATUS_BYTES = 256;                                   // keeps the tile sums 256 B aligned
constexpr size_t NOALLOC_BYTES = NOALLOC_STATUS_BYTES + (NOALLOC_TILES + 1) * sizeof(Quad);
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

static std::mutex g_noalloc_mu;
static void *g_noalloc_addr[64];                     // device address of g_noalloc_scratch per device ordinal

__global__ void __launch_bounds__(32) noalloc_total_kernel(const Quad *__restrict__ total, const uint32_t *__restrict__ status,
                                                           uint64_t *__restrict__ out5) {
  if (threadIdx.x == 0) { out5[0] = total->a; out5[1] = total->b; out5[2] = total->c; out5[3] = total->d; out5[4] = *status; }
}

extern "C" int hovl_filter_records_noalloc(int device, void *stream, const hovl_record *d_records_in, uint64_t n,
                                           int filter_valid, hovl_record *d_records_out, uint64_t *d_counts_out5) {
  try {
    HOVL_REQUIRE(device >= 0 && device < 64, HOVL_E_ARG, "hovl_filter_records_noalloc: device ordinal out of range (0..63)");
    HOVL_REQUIRE(n <= NOALLOC_MAX_RECORDS, HOVL_E_UNSUPPORTED,
                 "hovl_filter_records_noalloc: more than HOVL_NOALLOC_MAX_RECORDS (2^24) records; the static scratch holds "
                 "the tile sums of at most that many - split the call");
    HOVL_REQUIRE(d_counts_out5 != nullptr && (n == 0 || (d_records_in && d_records_out)), HOVL_E_ARG,
                 "hovl_filter_records_noalloc: NULL device pointer");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) { cudaGetLastError(); throw Error(HOVL_E_NO_DEVICE, "no CUDA device (libhovl has no CPU fallback)"); }
    HOVL_REQUIRE(device < count, HOVL_E_ARG, "hovl_filter_records_noalloc: device index out of range");
    HOVL_CUDA(cudaSetDevice(device));
    void *base = nullptr;
    {
      std::lock_guard<std::mutex> g(g_noalloc_mu);
      if (!g_noalloc_addr[device]) {                 // set-up, once per device, with that device current
        void *p = nullptr;
        HOVL_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        HOVL_REQUIRE(p != nullptr, HOVL_E_CUDA, "cudaGetSymbolAddress returned NULL");
        g_noalloc_addr[device] = p;
      }
      base = g_noalloc_addr[device];
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *status = (uint32_t *)base;
    Quad *partials = (Quad *)((unsigned char *)base + NOALLOC_STATUS_BYTES);
    const size_t nt = (size_t)div_up(n, SCAN_TILE);
    HOVL_CUDA(cudaMemsetAsync(status, 0, NOALLOC_STATUS_BYTES, st));
    RecIn in{d_records_in, filter_valid, status};
    RecOut out{d_records_in, d_records_out};
    if (nt) scan_reduce_kernel<Quad, RecIn><<<(unsigned)nt, SCAN_THREADS, 0, st>>>((size_t)n, in, partials);
    scan_partials_kernel<Quad><<<1, SCAN_THREADS, 0, st>>>(nt, partials);
    if (nt) scan_apply_kernel<Quad, RecIn, RecOut><<<(unsigned)nt, SCAN_THREADS, 0, st>>>((size_t)n, in, out, partials);
    noalloc_total_kernel<<<1, 32, 0, st>>>(partials + nt, status, d_counts_out5);
    HOVL_CUDA(cudaGetLastError());
    return HOVL_OK;
  } catch (const Error &e) {
    g_create_error = e.msg;
    return e.code;
  } catch (...) {
    g_create_error = "unexpected exception";
    return HOVL_E_CUDA;
  }
}
