// hovl_cli — the Project4 command line (src/Project4.scala:48-268) over the C ABI of include/hovl.h.
//
// The north star keeps the option parser, FASTA reader and OVL writer on the JVM; there is no JVM in
// this image, so this C++ host stands where `scala Project4` stands (the Scala host is given as source
// in INTEGRATION.md).  Same flags, same defaults, "last flag wins", an unknown flag prints
// "Invalid Argument : <flag>" and exits 1, a missing -i prints "No input file specified" and exits -1.
// Everything that computes is one hovl_run call: there is no CPU path in here.
//
// Modes served: --calc-overlaps (default), --test-fasta-read, --bench-fasta-read, --test-kmer-cover,
// --test-dispatch-collisions, --test-block-dispatch, --test-overlaps, --bench-kmer-gen, --bench-kmer-analysis, --bench-align.
// The JVM's threading and block / single variants are one native path, so the bench modes print one line per native run.
// --bench-align-quick is served too: in the reference its debugStop guards never fire (src/Project4.scala:611, :642), so it
// aligns nothing and prints eight zero counts.  The one remaining probe mode prints gapped alignment strings
// (--test-alignment); it exits 2 with a message.
// Flags that do not exist in the reference start with --hovl-.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <map>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "../../include/hovl.h"

namespace {

[[noreturn]] void die_uncaught(const std::string &what) {   // an exception nobody catches on the JVM: trace, exit 1
  fprintf(stderr, "Exception in thread \"main\" %s\n", what.c_str());
  exit(1);
}
uint32_t f32_bits(float f) { uint32_t b; memcpy(&b, &f, 4); return b; }

int32_t to_int(int argc, char **argv, int i) {               // args(i + 1).toInt
  if (i + 1 >= argc) die_uncaught("java.lang.ArrayIndexOutOfBoundsException: " + std::to_string(i + 1));
  int32_t v;
  if (hovl_parse_int(argv[i + 1], &v) != HOVL_OK)
    die_uncaught(std::string("java.lang.NumberFormatException: For input string: \"") + argv[i + 1] + "\"");
  return v;
}
int32_t jabs(int32_t v) { return v < 0 ? (int32_t)(0u - (uint32_t)v) : v; }     // math.abs(Int): MIN_VALUE stays
float to_float(int argc, char **argv, int i) {               // args(i + 1).toFloat = Float.parseFloat
  if (i + 1 >= argc) die_uncaught("java.lang.ArrayIndexOutOfBoundsException: " + std::to_string(i + 1));
  std::string t = argv[i + 1];
  size_t a = 0, b = t.size();
  while (a < b && (unsigned char)t[a] <= ' ') a++;
  while (b > a && (unsigned char)t[b - 1] <= ' ') b--;
  t = t.substr(a, b - a);
  std::string num = t;
  if (!num.empty() && strchr("fFdD", num.back())) num.pop_back();
  char *end = nullptr;
  const float v = num.empty() ? 0.f : strtof(num.c_str(), &end);   // correctly rounded, like the JVM
  if (num.empty() || !end || *end != 0)
    die_uncaught("java.lang.NumberFormatException: For input string: \"" + t + "\"");
  return v;
}
const char *str_arg(int argc, char **argv, int i) {
  if (i + 1 >= argc) die_uncaught("java.lang.ArrayIndexOutOfBoundsException: " + std::to_string(i + 1));
  return argv[i + 1];
}

void print_help() {                                           // printHelp, src/Project4.scala:262-268
  printf(" Rohit Ramesh :         Cmsc 423 \n Project 4 : Sequence Overlapper \n                                 \n"
         " [Usage]                         \n   See README file for details   \n\n");
}

struct Options {
  hovl_settings s;
  std::string action = "calc-overlaps", input, output, hoxd;
  bool debug = false, wide_ids = false, stats = false, print_settings = false;
  int device = 0;
  std::vector<int> devices;                                    // --hovl-devices 0,1,..: one handle over several GPUs
};

Options read_args(int argc, char **argv) {                    // readArgs, src/Project4.scala:100-259
  Options o;
  hovl_settings_default(&o.s);
  int32_t k = 12, min_overlap = 40, max_ignore = 90, gap_open = -200, gap_extend = -20, min_coll = 7, max_coll = 222;
  float min_identity = 0.98f, kmer_center = 0.4f, kmer_edge = 0.4f;
  bool fd_align = true;
  int i = 1;
  auto is = [&](const char *a, const char *b = nullptr, const char *c = nullptr, const char *d = nullptr) {
    const char *x = argv[i];
    return !strcmp(x, a) || (b && !strcmp(x, b)) || (c && !strcmp(x, c)) || (d && !strcmp(x, d));
  };
  static const char *modes[] = {"--calc-overlaps", "--test-overlaps", "--test-alignment", "--test-dispatch-collisions",
                                "--test-block-dispatch", "--test-kmer-cover", "--test-fasta-read", "--bench-fasta-read",
                                "--bench-kmer-gen", "--bench-kmer-analysis", "--bench-align-quick", "--bench-align"};
  while (i < argc) {
    if (is("-h", "--help")) { print_help(); exit(0); }
    else if (is("-m", "--matrix", "-H", "--HOXD-matrix")) { o.hoxd = str_arg(argc, argv, i); i += 2; }
    else if (is("-k", "--kmer-size")) { k = to_int(argc, argv, i); i += 2; }
    else if (is("-i", "--input")) { o.input = str_arg(argc, argv, i); i += 2; }
    else if (is("-o", "--output")) { o.output = str_arg(argc, argv, i); i += 2; }
    else if (is("--match")) { (void)jabs(to_int(argc, argv, i)); i += 2; }            // parsed, never used (:133)
    else if (is("--mismatch")) { (void)jabs(to_int(argc, argv, i)); i += 2; }         // parsed, never used (:136)
    else if (is("--min-overlap")) { min_overlap = jabs(to_int(argc, argv, i)); i += 2; }
    else if (is("--min-identity")) {
      min_identity = to_float(argc, argv, i);
      if (min_identity >= 1) { volatile float t = min_identity * .01f; min_identity = t; }
      i += 2;
    }
    else if (is("--min-collisions")) { min_coll = jabs(to_int(argc, argv, i)); i += 2; }
    else if (is("--max-collisions")) { max_coll = jabs(to_int(argc, argv, i)); i += 2; }
    else if (is("--kmer-center")) { kmer_center = fabsf(to_float(argc, argv, i)); i += 2; }
    else if (is("--kmer-edge")) { kmer_edge = fabsf(to_float(argc, argv, i)); i += 2; }
    else if (is("-gO", "--gap-open")) { gap_open = (int32_t)(0u - (uint32_t)jabs(to_int(argc, argv, i))); i += 2; }
    else if (is("-gE", "--gap-extend")) { gap_extend = (int32_t)(0u - (uint32_t)jabs(to_int(argc, argv, i))); i += 2; }
    else if (is("--max-ignore")) { max_ignore = jabs(to_int(argc, argv, i)); i += 2; }
    else if (is("--st-hash", "--mt-hash", "--st-align", "--mt-align")) { i += 1; }   // threading of the JVM path: no effect on results
    else if (is("--block-align", "--single-align")) { i += 1; }                       // same pairs, same order (KmerTable.scala:225-273)
    else if (is("--quadratic-align")) { fd_align = false; i += 1; }
    else if (is("--linear-align")) { fd_align = true; i += 1; }
    else if (is("--debug")) { o.debug = true; i += 1; }
    else if (is("--sleep-for-debug")) { printf("Sleeping so debugger can connect.\n"); i += 1; }   // no JVM debugger to wait for
    else if (is("--hovl-wide-ids")) { o.wide_ids = true; i += 1; }
    else if (is("--hovl-stats")) { o.stats = true; i += 1; }
    else if (is("--hovl-print-settings")) { o.print_settings = true; i += 1; }
    else if (is("--hovl-device")) { o.device = to_int(argc, argv, i); i += 2; }
    else if (is("--hovl-devices")) {                                                  // e.g. 0,1,2,3: hovl_create(devices, n)
      const std::string list = str_arg(argc, argv, i);
      o.devices.clear();
      size_t at = 0;
      while (at <= list.size()) {
        size_t end = list.find(',', at);
        if (end == std::string::npos) end = list.size();
        int32_t d = 0;
        if (hovl_parse_int(list.substr(at, end - at).c_str(), &d) != HOVL_OK || d < 0)
          die_uncaught("java.lang.NumberFormatException: For input string: \"" + list + "\"");
        o.devices.push_back((int)d);
        at = end + 1;
      }
      i += 2;
    }
    else {
      bool mode = false;
      for (const char *m : modes) if (!strcmp(argv[i], m)) { o.action = m + 2; mode = true; }
      if (!mode) { printf("Invalid Argument : %s\nExiting Program.\n", argv[i]); exit(1); }
      i += 1;
    }
  }
  if (!o.hoxd.empty() && hovl_read_hoxd(o.hoxd.c_str(), o.s.cost) != HOVL_OK) die_uncaught(hovl_io_last_error());
  if (o.input.empty()) { printf("No input file specified\n"); exit(255); }          // System.exit(-1)
  o.s.k = k; o.s.gap_open = gap_open; o.s.gap_extend = gap_extend; o.s.min_overlap = min_overlap;
  o.s.min_identity_f32 = f32_bits(min_identity);
  o.s.max_ignore_f32 = f32_bits((float)max_ignore);                                 // Int -> Float, ObjectStore.scala:25
  o.s.min_coll = min_coll; o.s.max_coll = max_coll;
  hovl_settings_set_sections(&o.s, f32_bits(kmer_edge), f32_bits(kmer_center));
  o.s.aligner = fd_align ? 0 : 1;
  if (o.devices.size() > 1) o.wide_ids = true;                 // the sharded path returns canonical (a, b) order
  o.s.order = o.wide_ids ? 1 : 0; o.s.id_mode = o.wide_ids ? 1 : 0;
  o.s.filter_valid = 1;
  return o;
}

struct Run {
  hovl_ctx *ctx = nullptr; hovl_reads *reads = nullptr; hovl_result *res = nullptr;
  const hovl_record *recs = nullptr; uint64_t n = 0; hovl_stats_t st;
};

hovl_reads *open_reads(const Options &o) {
  hovl_reads *r = nullptr;
  if (hovl_reads_open(o.input.c_str(), &r) != HOVL_OK) die_uncaught(std::string("java.lang.Exception: ") + hovl_io_last_error());
  return r;
}

Run run_path(const Options &o, bool filter) {                 // generateKmerTable + genAlignment
  Run R;
  R.reads = open_reads(o);
  const int dev = o.device;
  int rc = o.devices.size() > 1 ? hovl_create(&R.ctx, o.devices.data(), (int)o.devices.size())
                                : hovl_create(&R.ctx, o.devices.size() == 1 ? o.devices.data() : &dev, 1);
  if (rc != HOVL_OK) { fprintf(stderr, "hovl: %s\n", hovl_last_error(nullptr)); exit(1); }
  hovl_settings s = o.s;
  s.filter_valid = filter ? 1 : 0;
  uint32_t n_reads = 0;
  const uint64_t *offs = hovl_reads_offsets(R.reads, &n_reads);
  rc = hovl_run(R.ctx, &s, hovl_reads_bases(R.reads, nullptr), offs, n_reads, &R.res);
  if (rc != HOVL_OK) { fprintf(stderr, "hovl: error %d: %s\n", rc, hovl_last_error(R.ctx)); exit(1); }
  R.recs = hovl_records(R.res, &R.n);
  hovl_stats(R.res, &R.st);
  if (o.stats)
    fprintf(stderr, "hovl: %llu reads, %llu k-mers, %llu buckets, %llu pair instances, %llu dispatched, %llu duds, %llu overlaps; "
            "%.3f ms on the device (%llu kernel launches)\n",
            (unsigned long long)R.st.n_reads, (unsigned long long)R.st.n_kmers, (unsigned long long)R.st.n_buckets,
            (unsigned long long)R.st.n_pair_instances, (unsigned long long)R.st.n_dispatched, (unsigned long long)R.st.n_duds,
            (unsigned long long)R.st.n_ovl_valid, R.st.ns_total / 1e6, (unsigned long long)R.st.gpu_launches);
  return R;
}
void finish(Run &R) { hovl_result_free(R.res); hovl_destroy(R.ctx); hovl_reads_free(R.reads); }

std::string java_float(float x) {                             // Float.toString, csrc/hostio.cpp
  char buf[64];
  hovl_format_float(x, buf, sizeof buf);
  return buf;
}

long long now_ms() {
  return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

int main(int argc, char **argv) {
  Options o = read_args(argc, argv);
  const std::string &a = o.action;
  if (o.print_settings) {                                      // what crosses the ABI, for the parser tests (no device needed)
    const unsigned char *b = reinterpret_cast<const unsigned char *>(&o.s);
    printf("action=%s input=%s output=%s devices=", a.c_str(), o.input.c_str(), o.output.c_str());
    if (o.devices.empty()) printf("%d", o.device);
    for (size_t j = 0; j < o.devices.size(); j++) printf("%s%d", j ? "," : "", o.devices[j]);
    printf(" settings=");
    for (size_t j = 0; j < sizeof o.s; j++) printf("%02x", b[j]);
    printf("\n");
    return 0;
  }
  if (a == "calc-overlaps") {                                  // src/Project4.scala:54-58
    Run R = run_path(o, true);
    if (hovl_write_ovl(o.output.c_str(), R.recs, R.n) != HOVL_OK) die_uncaught(std::string("java.io.IOException: ") + hovl_io_last_error());
    finish(R);
    return 0;
  }
  if (a == "test-overlaps") {                                  // testOverlaps, :484-504, over genAlignment(table, settings, false)
    Run R = run_path(o, false);
    uint32_t n_reads = 0; const uint64_t *off = hovl_reads_offsets(R.reads, &n_reads); const uint8_t *b = hovl_reads_bases(R.reads, nullptr);
    const uint64_t need = hovl_format_test_overlaps(R.recs, R.n, b, off, n_reads, nullptr, 0);
    std::vector<char> text((size_t)need);
    if (need) hovl_format_test_overlaps(R.recs, R.n, b, off, n_reads, text.data(), need);
    if (need) fwrite(text.data(), 1, (size_t)need, stdout);
    finish(R);
    return 0;
  }
  if (a == "test-fasta-read") {                                // testFastaRead, :271-284
    hovl_reads *r = open_reads(o);
    uint32_t n = 0; const uint64_t *off = hovl_reads_offsets(r, &n); const uint8_t *b = hovl_reads_bases(r, nullptr);
    printf("\n");
    for (uint32_t i = 0; i < n && i < 10; i++)
      printf("id : %u\nseq: %.*s\n\n", i + 1, (int)(off[i + 1] - off[i]), (const char *)b + off[i]);
    hovl_reads_free(r);
    return 0;
  }
  if (a == "bench-fasta-read") {                               // benchFastaRead, :287-296
    const long long t0 = now_ms();
    hovl_reads *r = open_reads(o);
    uint32_t n = 0; hovl_reads_offsets(r, &n);
    printf(" Read %u sequences from %s in %lld milliseconds.\n", n, o.input.c_str(), now_ms() - t0);
    hovl_reads_free(r);
    return 0;
  }
  if (a == "test-dispatch-collisions" || a == "test-block-dispatch") {   // :376-421
    Run R = run_path(o, false);                                // every dispatched pair, in dispatch order
    std::set<std::pair<int32_t, int32_t>> seen;
    std::map<uint64_t, uint64_t> hist;
    uint64_t i = 0, block = 0;
    for (uint64_t p = 0; p < R.n; p++) {
      const hovl_record &r = R.recs[p];
      i++;
      if (!seen.insert({r.a, r.b}).second) printf("!!!! Collission %d<->%d Dispatched more than once. \n", r.a, r.b);
      printf(" Dispatched Coll : %llu - %d <-> %d\n", (unsigned long long)i, r.a, r.b);
      block++;
      if (p + 1 == R.n || R.recs[p + 1].a != r.a) { hist[block]++; block = 0; }   // a lead is iterated once: one block per run of a
    }
    if (a == "test-block-dispatch") {
      printf("\n Histogram Of Relations : [Number of Aligns -> Number of Seqs w/ that many Aligns]\n");
      for (auto &kv : hist) printf("          [%llu -> %llu]\n", (unsigned long long)kv.first, (unsigned long long)kv.second);
      printf("\n");
    }
    finish(R);
    return 0;
  }
  if (a == "bench-kmer-analysis") {                            // benchKmerAnalysis, :353-373, from the device stage times
    printf("Starting kmer gen.\n");
    Run R = run_path(o, false);
    printf("Finished kmer gen.\n");
    printf("\nCalculated pair data in %llu milliseconds.\n\n",
           (unsigned long long)((R.st.ns_group + R.st.ns_expand + R.st.ns_pair_sort + R.st.ns_pair_reduce) / 1000000));
    printf("Calculated dispatch data in %llu milliseconds.\n\n", (unsigned long long)(R.st.ns_order / 1000000));
    finish(R);
    return 0;
  }
  if (a == "test-kmer-cover") {                                // testKmerCover, :299-320: k = 0..25
    hovl_reads *rd = open_reads(o);
    uint32_t n_reads = 0;
    const uint64_t *offs = hovl_reads_offsets(rd, &n_reads);
    const uint8_t *bases = hovl_reads_bases(rd, nullptr);
    hovl_ctx *ctx = nullptr;
    const int dev = o.device;
    if (hovl_create(&ctx, &dev, 1) != HOVL_OK) { fprintf(stderr, "hovl: %s\n", hovl_last_error(nullptr)); return 1; }
    for (int k = 0; k <= 25; k++) {
      uint64_t uniques = 0, n_sizes = 0, *sizes = nullptr, *counts = nullptr;
      uint64_t one_size = 0, one_count = 1;
      if (k == 0) {
        // every window of length 0 hashes to 0 (ObjectStore.scala:48-67): one bucket holding len + 1 entries per read
        // (BioLibs.scala:55: i <- 0 to len - k); no kernel runs a 0-base window
        for (uint32_t r = 0; r < n_reads; r++) one_size += offs[r + 1] - offs[r] + 1;
        uniques = n_reads ? 1 : 0; n_sizes = uniques;
      } else {
        hovl_settings s = o.s;
        s.k = k;
        if (n_reads > 32767) { s.order = 1; s.id_mode = 1; }
        const int rc = hovl_stage_cover(ctx, &s, bases, offs, n_reads, &uniques, &sizes, &counts, &n_sizes);
        if (rc != HOVL_OK) { fprintf(stderr, "hovl: error %d: %s\n", rc, hovl_last_error(ctx)); return 1; }
      }
      const float ratio = (float)(int32_t)uniques / (float)pow(4.0, (double)k);    // uniques.toFloat / possible.toFloat
      printf("Kmer Size : %d\n  uniques : %llu\n  ratio   : %s\n", k, (unsigned long long)uniques, java_float(ratio).c_str());
      printf("  [ number of collisions -> count of seqs with that many collisions ] :\n");
      for (uint64_t i = 0; i < n_sizes; i++)
        printf("          [%llu -> %llu]\n", (unsigned long long)(sizes ? sizes[i] : one_size), (unsigned long long)(counts ? counts[i] : one_count));
      printf("\n");
      hovl_free(sizes); hovl_free(counts);
    }
    hovl_destroy(ctx); hovl_reads_free(rd);
    return 0;
  }
  if (a == "bench-kmer-gen") {                                 // benchKmerGen, :324-349: read + split + table, twice
    // the JVM's single- and multi-threaded table builds are one native path: each line is one native run (file read included)
    hovl_ctx *ctx = nullptr;
    const int dev = o.device;
    if (hovl_create(&ctx, &dev, 1) != HOVL_OK) { fprintf(stderr, "hovl: %s\n", hovl_last_error(nullptr)); return 1; }
    for (int v = 0; v < 2; v++) {
      const long long t0 = now_ms();
      hovl_reads *rd = open_reads(o);
      uint32_t n_reads = 0;
      const uint64_t *offs = hovl_reads_offsets(rd, &n_reads);
      uint64_t uniques = n_reads ? 1 : 0, n_sizes = 0, *sizes = nullptr, *counts = nullptr;
      if (o.s.k > 0) {
        hovl_settings s = o.s;
        if (n_reads > 32767) { s.order = 1; s.id_mode = 1; }
        const int rc = hovl_stage_cover(ctx, &s, hovl_reads_bases(rd, nullptr), offs, n_reads, &uniques, &sizes, &counts, &n_sizes);
        if (rc != HOVL_OK) { fprintf(stderr, "hovl: error %d: %s\n", rc, hovl_last_error(ctx)); return 1; }
        hovl_free(sizes); hovl_free(counts);
      }
      printf("%sGenerated %llu unique kmers from %u sequences from %s %s in %lld milliseconds.\n\n", v == 0 ? "\n" : "",
             (unsigned long long)uniques, n_reads, o.input.c_str(), v == 0 ? "sequentially" : "in parellel", now_ms() - t0);
      hovl_reads_free(rd);
    }
    hovl_destroy(ctx);
    return 0;
  }
  if (a == "bench-align" || a == "bench-align-quick") {        // benchAlign, :469-481: quad x4 then linear x4, filter = false
    static const char *names[] = {"single threaded quad single", "single threaded quad block", "multi threaded quad single",
                                  "multi threaded quad block", "single threaded linear single", "single threaded linear block",
                                  "multi threaded linear single", "multi threaded linear block"};
    for (int v = 0; v < 8; v++) {
      // the JVM's threading and block/single variants are one native path (same pairs, same order): each line is a
      // run of its own; the time is what aligner(tab, s, false) covers: pair data + dispatch + alignment, no table build
      Options ov = o;
      ov.s.aligner = v < 4 ? 1 : 0; ov.stats = false;
      if (a == "bench-align-quick") {
        // benchAlignQuick (:461-465) sets debugStop = 500, and every guard `aligns.size > debugStop` / `i > debugStop`
        // (:611, :642, :694, :732) tests a counter that only grows INSIDE the guarded body: the reference walks the dispatch
        // and aligns nothing.  The native run is therefore the table + pair stage alone and the count is 0.
        hovl_reads *rd = open_reads(ov);
        hovl_ctx *ctx = nullptr;
        const int dev = ov.device;
        if (hovl_create(&ctx, &dev, 1) != HOVL_OK) { fprintf(stderr, "hovl: %s\n", hovl_last_error(nullptr)); return 1; }
        uint32_t n_reads = 0;
        const uint64_t *offs = hovl_reads_offsets(rd, &n_reads);
        if (n_reads > 32767) { ov.s.order = 1; ov.s.id_mode = 1; }
        uint64_t *keys = nullptr, n_pairs = 0, n_inst = 0; uint32_t *counts = nullptr;
        const long long t0 = now_ms();
        const int rc = hovl_stage_pairs(ctx, &ov.s, hovl_reads_bases(rd, nullptr), offs, n_reads, &keys, &counts, &n_pairs, &n_inst);
        if (rc != HOVL_OK) { fprintf(stderr, "hovl: error %d: %s\n", rc, hovl_last_error(ctx)); return 1; }
        printf("\nCalculated 0 %s alignments in %lld milliseconds.\n\n", names[v], now_ms() - t0);
        hovl_free(keys); hovl_free(counts);
        hovl_destroy(ctx); hovl_reads_free(rd);
        continue;
      }
      Run R = run_path(ov, false);
      const uint64_t ns = R.st.ns_group + R.st.ns_expand + R.st.ns_pair_sort + R.st.ns_pair_reduce + R.st.ns_order + R.st.ns_dp + R.st.ns_d2h;
      printf("\nCalculated %llu %s alignments in %llu milliseconds.\n\n", (unsigned long long)R.n, names[v], (unsigned long long)(ns / 1000000));
      finish(R);
    }
    return 0;
  }
  fprintf(stderr, "hovl_cli: mode --%s is not served by the native path (it prints gapped alignment strings, which the "
          "native path never builds); use --calc-overlaps, --test-fasta-read, --bench-fasta-read, --test-kmer-cover, "
          "--test-dispatch-collisions, --test-block-dispatch, --test-overlaps, --bench-kmer-gen, --bench-kmer-analysis, "
          "--bench-align or --bench-align-quick\n", a.c_str());
  return 2;
}
