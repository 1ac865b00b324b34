// hostio.cpp — the steps either side of the path, for hosts without the JVM reader/writer
// (SURVEY 8(f) n3): FASTA ingest with readSeq's semantics, the HOXD matrix reader, and the
// calcOverlaps file writer.  Host only; nothing here touches the device except that the read
// buffers are page-locked when a CUDA runtime is present, so hovl_run's H2D copy of them runs
// at full link speed with no staging copy.
//   readSeq      src/BioLibs.scala:26-50
//   readHOXD     src/BioLibs.scala:66-114
//   calcOverlaps src/Project4.scala:795-825
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../../include/hovl.h"

struct hovl_reads {
  uint8_t *bases = nullptr; uint64_t *offsets = nullptr;
  uint64_t n_bases = 0; uint32_t n_reads = 0;
  bool pinned_b = false, pinned_o = false;
};

namespace {

thread_local std::string g_io_error;

void *host_alloc(size_t bytes, bool *pinned) {
  *pinned = false;
  if (bytes == 0) bytes = 8;
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess) { *pinned = true; return p; }
  (void)cudaGetLastError();            // no device / no driver: plain memory, the readers still work
  return malloc(bytes);
}
void host_free(void *p, bool pinned) {
  if (!p) return;
  if (pinned) cudaFreeHost(p); else free(p);
}

// whole file into memory; returns false with g_io_error set
bool slurp(const char *path, std::vector<char> &buf) {
  const int fd = open(path, O_RDONLY);
  if (fd < 0) { g_io_error = std::string(path) + " (" + strerror(errno) + ")"; return false; }
  struct stat st;
  if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) { close(fd); g_io_error = std::string(path) + " (not a regular file)"; return false; }
  buf.resize((size_t)st.st_size);
  size_t got = 0;
  while (got < buf.size()) {
    const ssize_t r = read(fd, buf.data() + got, buf.size() - got);
    if (r < 0) { if (errno == EINTR) continue; close(fd); g_io_error = std::string(path) + " (" + strerror(errno) + ")"; return false; }
    if (r == 0) break;
    got += (size_t)r;
  }
  close(fd);
  buf.resize(got);
  return true;
}

// BufferedReader.readLine: a line ends at "\n", "\r" or "\r\n"; the terminator is not part of it;
// the last line needs no terminator; an empty file has no line at all.
struct LineReader {
  const char *p, *end;
  bool next(const char *&b, const char *&e) {
    if (p >= end) return false;
    b = p;
    const char *q = p;
    while (q < end && *q != '\n' && *q != '\r') q++;
    e = q;
    if (q < end) { if (*q == '\r' && q + 1 < end && q[1] == '\n') q += 2; else q += 1; }
    p = q;
    return true;
  }
};

}  // namespace

extern "C" const char *hovl_io_last_error(void) { return g_io_error.c_str(); }

// The file as a read-only mapping: the two passes below walk the page cache, no second copy of a 20 GB FASTA
// (K5) sits in memory beside the page-locked buffer.  Falls back to reading it whole where mmap is refused.
struct FileView {
  const char *data = nullptr; size_t size = 0; bool mapped = false; std::vector<char> owned;
  bool open_file(const char *path) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { g_io_error = std::string(path) + " (" + strerror(errno) + ")"; return false; }
    struct stat st;
    if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) { close(fd); g_io_error = std::string(path) + " (not a regular file)"; return false; }
    size = (size_t)st.st_size;
    if (size) {
      void *p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (p != MAP_FAILED) { data = (const char *)p; mapped = true; (void)madvise(p, size, MADV_SEQUENTIAL); }
    }
    close(fd);
    if (size && !mapped) { if (!slurp(path, owned)) return false; data = owned.data(); size = owned.size(); }
    return true;
  }
  ~FileView() { if (mapped) munmap(const_cast<char *>(data), size); }
};

extern "C" int hovl_reads_open(const char *path, hovl_reads **out) {
  if (!path || !out) { g_io_error = "null argument"; return HOVL_E_ARG; }
  *out = nullptr;
  FileView buf;
  if (!buf.open_file(path)) return HOVL_E_ARG;
  LineReader lr{buf.data, buf.data + buf.size};
  const char *b, *e;
  // `line.startsWith(">")` on the first line; an empty file makes it a NullPointerException
  if (!lr.next(b, e) || b == e || *b != '>') { g_io_error = std::string("Invalid Sequence File: ") + path; return HOVL_E_ARG; }
  // pass 1: sizes (so that the page-locked buffers are allocated once, at their final size)
  uint64_t n_bases = 0, n_reads = 1;
  { LineReader l2 = lr; while (l2.next(b, e)) { if (b < e && *b == '>') n_reads++; else n_bases += (uint64_t)(e - b); } }
  if (n_reads > 0xffffffffull) { g_io_error = "more than 2^32-1 records"; return HOVL_E_UNSUPPORTED; }
  hovl_reads *r = new hovl_reads();
  r->n_bases = n_bases; r->n_reads = (uint32_t)n_reads;
  r->bases = (uint8_t *)host_alloc(n_bases, &r->pinned_b);
  r->offsets = (uint64_t *)host_alloc((n_reads + 1) * 8, &r->pinned_o);
  if (!r->bases || !r->offsets) { host_free(r->bases, r->pinned_b); host_free(r->offsets, r->pinned_o); delete r; g_io_error = "out of host memory"; return HOVL_E_NOMEM; }
  // pass 2: every later '>' line closes a record (ids 1,2,3.. in file order, header text ignored),
  // body lines are concatenated and upper-cased, the last record is emitted at EOF
  uint64_t at = 0; uint32_t rec = 0;
  r->offsets[0] = 0;
  while (lr.next(b, e)) {
    if (b < e && *b == '>') { r->offsets[++rec] = at; continue; }
    for (const char *q = b; q < e; q++) { const unsigned char c = (unsigned char)*q; r->bases[at++] = (c >= 'a' && c <= 'z') ? (uint8_t)(c - 32) : c; }
  }
  r->offsets[++rec] = at;
  *out = r;
  return HOVL_OK;
}
extern "C" const uint8_t *hovl_reads_bases(const hovl_reads *r, uint64_t *n_bases) {
  if (!r) return nullptr;
  if (n_bases) *n_bases = r->n_bases;
  return r->bases;
}
extern "C" const uint64_t *hovl_reads_offsets(const hovl_reads *r, uint32_t *n_reads) {
  if (!r) return nullptr;
  if (n_reads) *n_reads = r->n_reads;
  return r->offsets;
}
extern "C" void hovl_reads_free(hovl_reads *r) {
  if (!r) return;
  host_free(r->bases, r->pinned_b); host_free(r->offsets, r->pinned_o);
  delete r;
}

// ---------------------------------------------------------------- readHOXD
namespace {
int acgt(char c) { switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; } }
std::string trim(const std::string &s) {                 // String.trim: strips chars <= ' ' on both ends
  size_t a = 0, b = s.size();
  while (a < b && (unsigned char)s[a] <= ' ') a++;
  while (b > a && (unsigned char)s[b - 1] <= ' ') b--;
  return s.substr(a, b - a);
}
std::vector<std::string> split_commas(const std::string &s) {   // String.split(","): trailing empty strings removed
  std::vector<std::string> v; size_t a = 0;
  for (;;) { const size_t c = s.find(',', a); if (c == std::string::npos) { v.push_back(s.substr(a)); break; } v.push_back(s.substr(a, c - a)); a = c + 1; }
  while (!v.empty() && v.back().empty()) v.pop_back();
  return v;
}
bool parse_java_int(const std::string &t, int32_t *out) {       // Integer.parseInt
  if (t.empty()) return false;
  size_t i = 0; bool neg = false;
  if (t[0] == '-' || t[0] == '+') { neg = t[0] == '-'; i = 1; if (t.size() == 1) return false; }
  int64_t v = 0;
  for (; i < t.size(); i++) { if (t[i] < '0' || t[i] > '9') return false; v = v * 10 + (t[i] - '0'); if (v > 2147483648ll) return false; }
  if (neg) v = -v;
  if (v > 2147483647ll || v < -2147483648ll) return false;
  *out = (int32_t)v; return true;
}
}  // namespace

extern "C" int hovl_parse_int(const char *text, int32_t *out) {
  int32_t v;
  if (!text || !out || !parse_java_int(text, &v)) return HOVL_E_ARG;
  *out = v; return HOVL_OK;
}

extern "C" int hovl_read_hoxd(const char *path, int32_t cost[4][4]) {
  if (!path || !cost) { g_io_error = "null argument"; return HOVL_E_ARG; }
  std::vector<char> buf;
  if (!slurp(path, buf)) return HOVL_E_ARG;
  LineReader lr{buf.data(), buf.data() + buf.size()};
  const char *b, *e;
  if (!lr.next(b, e)) { g_io_error = std::string("bad HOXD file: ") + path; return HOVL_E_ARG; }          // title line
  if (!lr.next(b, e)) { g_io_error = std::string("bad HOXD file: ") + path; return HOVL_E_ARG; }          // column header
  std::vector<std::string> cols = split_commas(std::string(b, e));
  for (auto &c : cols) c = trim(c);
  int32_t m[4][4]; memset(m, 0, sizeof m);
  while (lr.next(b, e)) {
    if (b == e) break;                                                       // an empty line ends the table
    const std::vector<std::string> row = split_commas(std::string(b, e));
    if (row.empty()) break;
    const std::string lab = trim(row[0]);
    const int a = lab.empty() ? -1 : acgt((char)toupper((unsigned char)lab[0]));
    for (size_t i = 1; i < row.size(); i++) {
      if (i >= cols.size()) { g_io_error = std::string("HOXD row longer than its header in ") + path; return HOVL_E_ARG; }   // ArrayIndexOutOfBounds
      const int bb = cols[i].empty() ? -1 : acgt((char)toupper((unsigned char)cols[i][0]));
      if (a < 0 || bb < 0) { g_io_error = std::string("bad HOXD base label in ") + path; return HOVL_E_ARG; }   // MatchError
      int32_t v;
      if (!parse_java_int(row[i], &v)) { g_io_error = "NumberFormatException: For input string: \"" + row[i] + "\""; return HOVL_E_ARG; }   // not trimmed, :91
      m[a][bb] = v;
    }
  }
  memcpy(cost, m, sizeof m);
  return HOVL_OK;
}

// ---------------------------------------------------------------- Float.toString and testOverlaps
namespace {
// Float.toString: shortest decimal that reads back as the same float, plain notation for 1e-3 <= |x| < 1e7,
// "d.dddE-n" otherwise (JDK 19+ prints exactly this; older JDKs print one digit more for some values).
std::string java_float(float x) {
  if (x != x) return "NaN";
  if (x == 0) return (1.0f / x < 0) ? "-0.0" : "0.0";
  if (x - x != 0) return x < 0 ? "-Infinity" : "Infinity";
  char buf[64];
  for (int prec = 1; prec <= 9; prec++) { snprintf(buf, sizeof buf, "%.*e", prec - 1, (double)x); if (strtof(buf, nullptr) == x) break; }
  std::string digits; int exp10 = 0;
  {
    const std::string t = buf;
    const size_t e = t.find('e');
    exp10 = atoi(t.c_str() + e + 1);
    for (char ch : t.substr(0, e)) if (ch >= '0' && ch <= '9') digits.push_back(ch);
  }
  std::string out = x < 0 ? "-" : "";
  const float ax = x < 0 ? -x : x;
  if (ax >= 1e-3f && ax < 1e7f) {
    if (exp10 >= 0) {
      std::string ip = digits.substr(0, (size_t)((int)digits.size() < exp10 + 1 ? (int)digits.size() : exp10 + 1));
      while ((int)ip.size() < exp10 + 1) ip.push_back('0');
      const std::string fp = (int)digits.size() > exp10 + 1 ? digits.substr((size_t)exp10 + 1) : "0";
      out += ip + "." + fp;
    } else {
      out += "0." + std::string((size_t)(-exp10 - 1), '0') + digits;
    }
  } else {
    out += digits.substr(0, 1) + "." + (digits.size() > 1 ? digits.substr(1) : "0") + "E" + std::to_string(exp10);
  }
  return out;
}
}  // namespace

extern "C" uint64_t hovl_format_float(float x, char *buf, uint64_t cap) {
  const std::string t = java_float(x);
  if (buf && cap) { const size_t l = t.size() < cap - 1 ? t.size() : (size_t)cap - 1; memcpy(buf, t.data(), l); buf[l] = 0; }
  return t.size();
}

// testOverlaps (src/Project4.scala:484-504) over genAlignment(table, settings, false): every dispatched pair in dispatch
// order, duds included (BioLibs.dud, src/BioLibs.scala:22: ids 0, empty sequences, errRatio 0f / (0f + 1f)).
extern "C" uint64_t hovl_format_test_overlaps(const hovl_record *recs, uint64_t n, const uint8_t *bases, const uint64_t *offsets,
                                               uint32_t n_reads, char *buf, uint64_t cap) {
  uint64_t need = 0;
  auto put = [&](const std::string &t) {
    if (buf && need + t.size() <= cap) memcpy(buf + need, t.data(), t.size());
    need += t.size();
  };
  auto dashes = [](int64_t k) { return std::string((size_t)(k > 0 ? k : 0), '-'); };   // ("").padTo(k, '-')
  for (uint64_t i = 0; i < n; i++) {
    const hovl_record &r = recs[i];
    const bool dud = (r.flags & HOVL_F_DUD) != 0;
    std::string sa, sb;
    int32_t ida = 0, idb = 0, ahg = 0, bhg = 0, si = 0, sj = 0, ei = 0, ej = 0;
    float ratio = 0.0f / (0.0f + 1.0f);
    if (!dud) {
      if (r.a < 1 || r.b < 1 || (uint32_t)r.a > n_reads || (uint32_t)r.b > n_reads) return 0;       // not a record of these reads
      ida = r.a; idb = r.b; ahg = r.ahg; bhg = r.bhg; si = r.start_i; sj = r.start_j; ei = r.end_i; ej = r.end_j;
      sa.assign((const char *)bases + offsets[r.a - 1], (size_t)(offsets[r.a] - offsets[r.a - 1]));
      sb.assign((const char *)bases + offsets[r.b - 1], (size_t)(offsets[r.b] - offsets[r.b - 1]));
      ratio = (float)r.correct / ((float)r.correct + (float)r.error);                               // ObjectStore.scala:98
    }
    put(" Overlap " + std::to_string(i + 1) + " : " + std::to_string(ida) + " <-> " + std::to_string(idb) + "\n");
    if (ahg >= 0) {
      put("   Seq A   : " + sa + dashes(bhg) + "\n");
      put("   Seq B   : " + dashes(ahg) + sb + "\n");
    } else {
      put("   Seq A   : " + dashes(-(int64_t)ahg) + sa + "\n");
      put("   Seq B   : " + sb + dashes(-(int64_t)bhg) + "\n");
    }
    put("   Ahg     : " + std::to_string(ahg) + "\n");
    put("   Bhg     : " + std::to_string(bhg) + "\n");
    put("   Start   : (" + std::to_string(si) + "," + std::to_string(sj) + ")\n");
    put("   End     : (" + std::to_string(ei) + "," + std::to_string(ej) + ")\n");
    put("   Error   : " + java_float(ratio) + "\n");
    put(std::string("   Valid?  : ") + ((!dud && (r.flags & HOVL_F_OVL_VALID)) ? "true" : "false") + "\n");
  }
  return need;
}

// ---------------------------------------------------------------- calcOverlaps
extern "C" int hovl_write_ovl(const char *path, const hovl_record *recs, uint64_t n) {
  const uint64_t need = hovl_format_ovl(recs, n, nullptr, 0);
  std::vector<char> text((size_t)need);
  if (need) hovl_format_ovl(recs, n, text.data(), need);
  if (!path || !*path) {                                   // output == "": println to stdout
    if (need && fwrite(text.data(), 1, (size_t)need, stdout) != need) { g_io_error = "short write to stdout"; return HOVL_E_ARG; }
    fflush(stdout);
    return HOVL_OK;
  }
  unlink(path);                                            // file.exists -> delete, createNewFile
  FILE *f = fopen(path, "wb");
  if (!f) { g_io_error = std::string(path) + " (" + strerror(errno) + ")"; return HOVL_E_ARG; }
  const bool ok = !need || fwrite(text.data(), 1, (size_t)need, f) == need;
  if (fclose(f) != 0 || !ok) { g_io_error = std::string(path) + " (write failed)"; return HOVL_E_ARG; }
  return HOVL_OK;
}
