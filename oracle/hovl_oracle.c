/* hovl_oracle.c — CPU oracle (TEST INFRASTRUCTURE ONLY; see hovl_oracle.h).
 *
 * Restates, function by function, the hash-overlap path of the reference:
 *   readSeq                 src/BioLibs.scala:26-50
 *   generateKmerSet/seqHash src/BioLibs.scala:54-61, src/ObjectStore.scala:40-67
 *   addKmerSet              src/KmerTable.scala:41-53
 *   addKmerPair/calcPairData src/KmerTable.scala:57-80, :85-149
 *   calcDispatchData        src/KmerTable.scala:155-187
 *   dispatchCollisionBlocks src/KmerTable.scala:246-273
 *   generateFastDovetailAlignment[Set] src/BioLibs.scala:373-591, :596-822
 *   Alignment.valid / Overlap src/ObjectStore.scala:89-142
 *   calcOverlaps            src/Project4.scala:795-825
 * plus a model of gnu.trove 3.0.3 TIntObjectHashMap (the reference's lib/trove.jar,
 * bytecode only) because its iteration order is the reference's output order.
 *
 * JVM semantics kept: Int = wrapping int32, Float = IEEE binary32 (build with
 * -ffp-contract=off, no -ffast-math), >> arithmetic.
 */
#define _POSIX_C_SOURCE 200809L
#include "hovl_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static uint64_t now_ns(void) {
  struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
  return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}
static float bits_f(uint32_t b) { float f; memcpy(&f, &b, 4); return f; }
static uint32_t f_bits(float f) { uint32_t b; memcpy(&b, &f, 4); return b; }

/* ===================== Trove 3.0.3 TIntObjectHashMap model ========================== *
 * gnu/trove/impl/hash/{THash,TIntHash}.class, impl/PrimeFinder.class,
 * map/hash/TIntObjectHashMap.class: open addressing, identity hash, double hashing,
 * prime capacities, load factor 0.5, rehash re-inserts old slots in DESCENDING slot order,
 * iterator walks slots from capacity-1 down to 0.  No key is ever removed by the
 * reference (src/KmerTable.scala only calls contains/put/get/clear-on-empty).          */
static const int32_t k_primes[] = {
  3,5,7,11,17,23,31,37,43,47,67,79,89,97,137,163,179,197,277,311,331,359,379,397,433,557,599,631,
  673,719,761,797,877,953,1039,1117,1201,1277,1361,1439,1523,1597,1759,1907,2081,2237,2411,2557,
  2729,2879,3049,3203,3527,3821,4177,4481,4831,5119,5471,5779,6101,6421,7057,7643,8363,8963,9677,
  10243,10949,11579,12203,12853,14143,15287,16729,17929,19373,20507,21911,23159,24407,25717,28289,
  30577,33461,35863,38747,41017,43853,46327,48817,51437,56591,61169,66923,71741,77509,82037,87719,
  92657,97649,102877,113189,122347,133853,143483,155027,164089,175447,185323,195311,205759,226379,
  244703,267713,286973,310081,328213,350899,370661,390647,411527,452759,489407,535481,573953,
  620171,656429,701819,741337,781301,823117,905551,978821,1070981,1147921,1240361,1312867,1403641,
  1482707,1562611,1646237,1811107,1957651,2141977,2295859,2480729,2625761,2807303,2965421,3125257,
  3292489,3622219,3915341,4283963,4591721,4961459,5251529,5614657,5930887,6250537,6584983,7244441,
  7830701,8567929,9183457,9922933,10503061,11229331,11861791,12501169,13169977,14488931,15661423,
  17135863,18366923,19845871,21006137,22458671,23723597,25002389,26339969,28977863,31322867,
  34271747,36733847,39691759,42012281,44917381,47447201,50004791,52679969,57955739,62645741,
  68543509,73467739,79383533,84024581,89834777,94894427,100009607,105359939,115911563,125291483,
  137087021,146935499,158767069,168049163,179669557,189788857,200019221,210719881,231823147,
  250582987,274174111,293871013,317534141,336098327,359339171,379577741,400038451,421439783,
  463646329,501165979,548348231,587742049,635068283,672196673,718678369,759155483,800076929,
  842879579,927292699,1002331963,1096696463,1175484103,1270136683,1344393353,1437356741,1518310967,
  1600153859,1685759167,1854585413,2004663929,2147483647 };
#define N_PRIMES ((int)(sizeof(k_primes) / sizeof(k_primes[0])))

static int32_t next_prime(int64_t n) {
  if (n > 2147483647) return 2147483647;
  int lo = 0, hi = N_PRIMES - 1;
  while (lo < hi) { int mid = (lo + hi) / 2; if (k_primes[mid] >= n) hi = mid; else lo = mid + 1; }
  return k_primes[lo];
}

struct trove_map {
  int32_t cap, size, free_, max_size;
  int32_t *keys, *vals; uint8_t *full;
};
static void trove_alloc(trove_map *m, int32_t cap) {
  m->cap = cap;
  m->keys = (int32_t *)calloc((size_t)cap, 4);
  m->vals = (int32_t *)calloc((size_t)cap, 4);
  m->full = (uint8_t *)calloc((size_t)cap, 1);
  if (!m->keys || !m->vals || !m->full) { fprintf(stderr, "oracle: out of memory\n"); abort(); }
}
static void trove_max(trove_map *m) {            /* THash.computeMaxSize */
  int32_t a = m->cap - 1, b = (int32_t)((float)m->cap * 0.5f);
  m->max_size = a < b ? a : b;
  m->free_ = m->cap - m->size;
}
trove_map *trove_new(void) {
  trove_map *m = (trove_map *)calloc(1, sizeof *m);
  trove_alloc(m, next_prime(20));                 /* ceil(10 / 0.5f) -> 23 */
  trove_max(m);
  return m;
}
void trove_free(trove_map *m) { if (m) { free(m->keys); free(m->vals); free(m->full); free(m); } }
/* TIntHash.insertKey: slot index; *found=1 when the key was already stored */
static int32_t trove_slot(trove_map *m, int32_t key, int *found) {
  int32_t h = key & 0x7fffffff, cap = m->cap, idx = h % cap;
  *found = 0;
  if (!m->full[idx]) return idx;
  if (m->keys[idx] == key) { *found = 1; return idx; }
  int32_t probe = 1 + h % (cap - 2);
  for (;;) {
    idx -= probe; if (idx < 0) idx += cap;
    if (!m->full[idx]) return idx;
    if (m->keys[idx] == key) { *found = 1; return idx; }
  }
}
static void trove_rehash(trove_map *m, int32_t new_cap) {
  int32_t old_cap = m->cap, *ok = m->keys, *ov = m->vals; uint8_t *of = m->full;
  trove_alloc(m, new_cap);
  for (int32_t i = old_cap; i-- > 0;)
    if (of[i]) { int f; int32_t s = trove_slot(m, ok[i], &f); m->full[s] = 1; m->keys[s] = ok[i]; m->vals[s] = ov[i]; }
  free(ok); free(ov); free(of);
}
int32_t *trove_get(trove_map *m, int32_t key) {
  int f; int32_t s = trove_slot(m, key, &f);
  return f ? &m->vals[s] : NULL;
}
int32_t *trove_put_if_absent(trove_map *m, int32_t key, int *is_new) {
  int f; int32_t s = trove_slot(m, key, &f);
  if (f) { if (is_new) *is_new = 0; return &m->vals[s]; }
  m->full[s] = 1; m->keys[s] = key; m->vals[s] = 0;
  if (is_new) *is_new = 1;
  /* THash.postInsertHook(usedFreeSlot = true) */
  m->free_--; m->size++;
  if (m->size > m->max_size || m->free_ == 0) {
    int32_t nc = m->size > m->max_size ? next_prime((int64_t)m->cap << 1) : m->cap;
    trove_rehash(m, nc);
    trove_max(m);
    s = trove_slot(m, key, &f);
  }
  return &m->vals[s];
}
int32_t trove_size(const trove_map *m) { return m->size; }
int32_t trove_capacity(const trove_map *m) { return m->cap; }
void trove_iterate(const trove_map *m, int32_t *keys, int32_t *vals) {
  int32_t n = 0;
  for (int32_t i = m->cap; i-- > 0;)
    if (m->full[i]) { if (keys) keys[n] = m->keys[i]; if (vals) vals[n] = m->vals[i]; n++; }
}

/* ================================ readers ============================================ */
/* readSeq (src/BioLibs.scala:26-50): first line must start with '>'; every later '>'
 * line closes a record; ids are 1,2,3.. in file order; body lines are concatenated and
 * upper-cased; the last record is emitted at EOF (even when empty).                     */
int orc_read_fasta(const char *path, uint8_t **bases, uint64_t **offsets, uint32_t *n_reads) {
  FILE *f = fopen(path, "rb");
  if (!f) return HOVL_E_ARG;
  size_t cap = 1 << 20, len = 0, ocap = 1024, nr = 0;
  uint8_t *b = (uint8_t *)malloc(cap);
  uint64_t *o = (uint64_t *)malloc(ocap * 8);
  char *line = NULL; size_t lcap = 0; ssize_t ll;
  ll = getline(&line, &lcap, f);
  if (ll <= 0 || line[0] != '>') { free(b); free(o); free(line); fclose(f); return HOVL_E_ARG; }
  o[0] = 0;
  while ((ll = getline(&line, &lcap, f)) >= 0) {
    while (ll > 0 && (line[ll - 1] == '\n' || line[ll - 1] == '\r')) ll--;  /* readLine strips EOL */
    if (ll > 0 && line[0] == '>') {
      nr++;
      if (nr + 1 >= ocap) { ocap *= 2; o = (uint64_t *)realloc(o, ocap * 8); }
      o[nr] = len;
    } else {
      if (len + (size_t)ll > cap) { while (len + (size_t)ll > cap) cap *= 2; b = (uint8_t *)realloc(b, cap); }
      for (ssize_t i = 0; i < ll; i++) {
        uint8_t c = (uint8_t)line[i];
        b[len++] = (c >= 'a' && c <= 'z') ? (uint8_t)(c - 32) : c;      /* toUpperCase */
      }
    }
  }
  nr++;
  if (nr + 1 >= ocap) { ocap += 2; o = (uint64_t *)realloc(o, ocap * 8); }
  o[nr] = len;
  free(line); fclose(f);
  *bases = b; *offsets = o; *n_reads = (uint32_t)nr;
  return HOVL_OK;
}

static int acgt_index(int c) {                    /* readHOXD / defaultHOXD order */
  switch (c) { case 'A': case 'a': return 0; case 'C': case 'c': return 1;
               case 'G': case 'g': return 2; case 'T': case 't': return 3; default: return -1; }
}
/* readHOXD (src/BioLibs.scala:66-114): skip the title line, read the column header, then
 * one CSV row per base until EOF or an empty line.                                       */
int orc_read_hoxd(const char *path, int32_t cost[4][4]) {
  FILE *f = fopen(path, "rb");
  if (!f) return HOVL_E_ARG;
  char buf[4096], *cols[64]; int ncol = 0;
  memset(cost, 0, 16 * sizeof(int32_t));
  if (!fgets(buf, sizeof buf, f)) { fclose(f); return HOVL_E_ARG; }          /* title */
  if (!fgets(buf, sizeof buf, f)) { fclose(f); return HOVL_E_ARG; }          /* header */
  char hdr[4096]; strcpy(hdr, buf);
  for (char *p = strtok(hdr, ",\r\n"); p && ncol < 64; p = strtok(NULL, ",\r\n")) cols[ncol++] = p;
  int colidx[64];
  for (int i = 0; i < ncol; i++) { char *p = cols[i]; while (*p == ' ') p++; colidx[i] = acgt_index(*p); }
  while (fgets(buf, sizeof buf, f)) {
    size_t l = strlen(buf); while (l && (buf[l - 1] == '\n' || buf[l - 1] == '\r')) buf[--l] = 0;
    if (l == 0) break;
    char *row[64]; int nrow = 0;
    for (char *p = strtok(buf, ","); p && nrow < 64; p = strtok(NULL, ",")) row[nrow++] = p;
    char *r0 = row[0]; while (*r0 == ' ') r0++;
    int a = acgt_index(*r0);
    for (int i = 1; i < nrow && i < ncol; i++) {
      if (a < 0 || colidx[i] < 0) { fclose(f); return HOVL_E_ARG; }           /* MatchError */
      char *end; long v = strtol(row[i], &end, 10);
      if (end == row[i] || *end != 0) { fclose(f); return HOVL_E_ARG; }       /* NumberFormatException */
      cost[a][colidx[i]] = (int32_t)v;
    }
  }
  fclose(f);
  return HOVL_OK;
}

/* ================================ settings =========================================== */
void orc_settings_set_sections(hovl_settings *s, float e, float c) {
  /* AlignSettings (src/ObjectStore.scala:32-35); every operation rounds to binary32 */
  volatile float half_c = c * 0.5f;
  volatile float tail = 1.0f - e, lead = 0.5f - half_c, mtail = 0.5f + half_c;
  s->head_f32 = f_bits(e); s->tail_f32 = f_bits(tail);
  s->mid_lead_f32 = f_bits(lead); s->mid_tail_f32 = f_bits(mtail);
}
void orc_settings_default(hovl_settings *s) {
  static const int32_t hoxd[4][4] = {              /* defaultHOXD, src/BioLibs.scala:122-140 */
    {  91, -114,  -31, -123}, {-114,  100, -125,  -31},
    { -31, -125,  100, -114}, {-123,  -31, -114,   91} };
  memset(s, 0, sizeof *s);
  s->k = 12;                                       /* src/Project4.scala:41  */
  memcpy(s->cost, hoxd, sizeof hoxd);
  s->gap_open = -200; s->gap_extend = -20;         /* :109-110 */
  s->min_overlap = 40;                             /* :106 */
  s->min_identity_f32 = f_bits(0.98f);             /* :107 */
  s->max_ignore_f32 = f_bits((float)90);           /* :108 -> Float at ObjectStore.scala:25 */
  s->min_coll = 7; s->max_coll = 222;              /* :111-112 */
  orc_settings_set_sections(s, 0.4f, 0.4f);        /* :113-114, tuple order :257 */
  s->aligner = 0; s->order = 0; s->id_mode = 0; s->filter_valid = 1;
}

/* ================================ k-mer tuples ======================================= */
/* Kmer.seqHash (src/ObjectStore.scala:48-67): first min(16,k) chars, h = (h<<2) ^ v with
 * A0 C1 T2 G3; any other char contributes 0 (the reference also prints a warning).      */
static inline int32_t kmer_code(const uint8_t *p, int32_t k) {
  int32_t h = 0, n = k < 16 ? k : 16;
  for (int32_t i = 0; i < n; i++) {
    uint32_t v;
    switch (p[i]) { case 'A': v = 0; break; case 'C': v = 1; break;
                    case 'T': v = 2; break; case 'G': v = 3; break; default: v = 0; }
    h = (int32_t)(((uint32_t)h << 2) ^ v);
  }
  return h;
}
/* generateKmerSet (src/BioLibs.scala:54-61): i = 0..L-k inclusive, loc = i.toFloat / d  */
static inline float kmer_loc(int64_t i, int64_t L, int32_t k) {
  volatile float d = (float)(L - k);
  volatile float fi = (float)i;
  volatile float q = fi / d;
  if (q != q) {          /* L == k: 0f/0f.  Stored with Java's canonical NaN bits (Float.floatToIntBits) */
    const uint32_t nan_bits = 0x7fc00000u; float f; memcpy(&f, &nan_bits, 4); return f;
  }
  return q;
}

/* ================================ aligner ============================================ */
static inline int32_t max2(int32_t a, int32_t b) { return a > b ? a : b; }
static inline int32_t max3(int32_t a, int32_t b, int32_t c) { return max2(a, max2(b, c)); }

/* width (src/BioLibs.scala:389-390 / :619-620):
 *   max(kmerSize, floor(A.size * (1 - minIdentity)).toInt + 1), product in binary32      */
static int32_t band_width(const hovl_settings *s, int32_t n) {
  volatile float omi = 1.0f - bits_f(s->min_identity_f32);
  volatile float prod = (float)n * omi;
  int32_t w = (int32_t)floor((double)prod) + 1;
  return s->k > w ? s->k : w;
}

static void set_dud(hovl_record *r, int extra) {    /* BioLibs.dud, src/BioLibs.scala:22 */
  r->flags = HOVL_F_DUD | extra;
  r->start_i = r->start_j = r->end_i = r->end_j = 0;
  r->correct = 0; r->error = 1; r->score = 0; r->ahg = 0; r->bhg = 0;
}

/* Alignment.valid (src/ObjectStore.scala:99-107) and Overlap (:124-125, :137-141)       */
static void finish_record(const hovl_settings *s, int32_t n, int32_t m, hovl_record *r) {
  volatile float fc = (float)r->correct, fe = (float)r->error;
  volatile float den = fc + fe;
  volatile float ratio = fc / den;
  int av = (ratio >= bits_f(s->min_identity_f32)) &&
           (r->correct + r->error >= s->min_overlap) &&
           ((r->start_i == 0 && m == r->end_j) || (r->start_j == 0 && n == r->end_i));
  r->ahg = r->start_i - r->start_j;
  r->bhg = m - n + r->ahg;
  float mi = bits_f(s->max_ignore_f32);
  int32_t aa = r->ahg < 0 ? -r->ahg : r->ahg, ab = r->bhg < 0 ? -r->bhg : r->bhg;
  int ov = av && ((float)aa < mi) && ((float)ab < mi);
  if (av) r->flags |= HOVL_F_ALIGN_VALID;
  if (ov) r->flags |= HOVL_F_OVL_VALID;
}

/* per-thread scratch for the three DP matrices */
typedef struct { int32_t *M, *X, *Y; size_t cap; } dp_scratch;
static void dp_reserve(dp_scratch *d, size_t cells) {
  if (cells <= d->cap) return;
  free(d->M); free(d->X); free(d->Y);
  d->M = (int32_t *)malloc(cells * 4); d->X = (int32_t *)malloc(cells * 4); d->Y = (int32_t *)malloc(cells * 4);
  d->cap = cells;
  if (!d->M || !d->X || !d->Y) { fprintf(stderr, "oracle: out of memory\n"); abort(); }
}

static inline int idx_acgt(uint8_t c) {            /* costFunc, src/BioLibs.scala:142-160 */
  switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}

/* generateFastDovetailAlignment (src/BioLibs.scala:373-591; the Set variant :596-822 runs
 * the same statements per B with fresh arrays).  Full matrices and the backward greedy
 * trace are kept exactly as the reference has them — the GPU's forward-carried trace is
 * checked against THIS.                                                                  */
static uint64_t align_pair(const hovl_settings *s, const uint8_t *A, int32_t n,
                           const uint8_t *B, int32_t m, hovl_record *r, dp_scratch *d) {
  const int32_t gO = s->gap_open, gE = s->gap_extend;
  const int32_t width = band_width(s, n);
  uint64_t cells = 0;
  r->flags = 0;
  for (int32_t i = 0; i < n; i++) if (idx_acgt(A[i]) < 0) { set_dud(r, HOVL_F_BAD_BASE); return 0; }
  for (int32_t j = 0; j < m; j++) if (idx_acgt(B[j]) < 0) { set_dud(r, HOVL_F_BAD_BASE); return 0; }
  if (m < width) { set_dud(r, HOVL_F_SHORT_B); return 0; }   /* charAt(j-1) would throw, :418/:648 */
  const int32_t W1 = width + 1;
  dp_reserve(d, (size_t)(n + 1) * (size_t)W1);
  int32_t *M = d->M, *X = d->X, *Y = d->Y;
  memset(M, 0, (size_t)(n + 1) * W1 * 4);          /* Array.ofDim: zero-filled, :392-394 */
  memset(X, 0, (size_t)(n + 1) * W1 * 4);
  memset(Y, 0, (size_t)(n + 1) * W1 * 4);
#define AT(P, i, j) P[(size_t)(i) * W1 + (j)]
  for (int32_t i = 0; i < n; i++) { AT(M, i, 0) = 0; AT(X, i, 0) = 0; AT(Y, i, 0) = gO + i * gE; }   /* :399-403 */
  for (int32_t j = 0; j < width; j++) { AT(M, 0, j) = 0; AT(X, 0, j) = gO + j * gE; AT(Y, 0, j) = 0; } /* :405-409 */

  int32_t best = 0, bi = 0, bj = 0;
  for (int32_t i = 1; i <= n; i++) {               /* phase 1, :415-438 */
    const int32_t *crow = s->cost[idx_acgt(A[i - 1])];
    for (int32_t j = 1; j <= width; j++) {
      int32_t mv = crow[idx_acgt(B[j - 1])] +
                   max2(max2(AT(M, i - 1, j - 1), AT(Y, i - 1, j - 1)), max2(AT(X, i - 1, j - 1), 0));
      int32_t xv = gE + max2(max2(AT(M, i, j - 1) + gO, AT(Y, i, j - 1) + gO), max2(AT(X, i, j - 1), 0));
      int32_t yv = gE + max2(max2(AT(M, i - 1, j) + gO, AT(Y, i - 1, j)), max2(AT(X, i - 1, j) + gO, 0));
      AT(M, i, j) = mv; AT(X, i, j) = xv; AT(Y, i, j) = yv;
      int32_t t = max3(mv, xv, yv);
      if (t > best) { best = t; bi = i; bj = j; }
    }
  }
  cells += (uint64_t)n * (uint64_t)width;
  if (best <= 0) { set_dud(r, 0); return cells; }  /* reference would index [-1]; extension: dud */

  int32_t i = bi, j = bj;                          /* trace 1, :443-459 */
  int32_t cur = max3(AT(M, i, j), AT(X, i, j), AT(Y, i, j));
  do {
    if (AT(M, i, j) == cur) { i--; j--; }
    else if (AT(X, i, j) == cur) { j--; }
    else if (AT(Y, i, j) == cur) { i--; }
    cur = max3(AT(M, i, j), AT(X, i, j), AT(Y, i, j));
  } while (cur > 0);
  if (j != 0) { set_dud(r, 0); return cells; }     /* :464-466 */

  const int32_t ds = i, dl = n - ds, z = width / 2; /* :473-475 */
  best = 0; int32_t bu = 0, bk = 0;
  for (int32_t u = 0; u <= dl; u++) {              /* phase 2, :495-534 */
    for (int32_t kk = 0; kk <= width; kk++) {
      int32_t ii = u + ds, jj = kk - z + u;
      int32_t mv, xv, yv;
      if (ii <= ds || jj <= 0 || jj > m) { mv = xv = yv = 0; }
      else {
        mv = (u != 0) ? s->cost[idx_acgt(A[ii - 1])][idx_acgt(B[jj - 1])] +
                        max2(max2(AT(M, u - 1, kk), AT(Y, u - 1, kk)), max2(AT(X, u - 1, kk), 0)) : 0;
        xv = (kk != 0) ? gE + max2(max2(AT(M, u, kk - 1) + gO, AT(Y, u, kk - 1) + gO),
                                   max2(AT(X, u, kk - 1), 0)) : 0;
        yv = (u != 0 && kk != width) ?
             gE + max2(max2(AT(M, u - 1, kk + 1) + gO, AT(Y, u - 1, kk + 1)),
                       max2(AT(X, u - 1, kk + 1) + gO, 0)) : 0;
      }
      AT(M, u, kk) = mv; AT(X, u, kk) = xv; AT(Y, u, kk) = yv;
      int32_t t = max3(mv, xv, yv);
      if (t > best) { best = t; bu = u; bk = kk; }
    }
  }
  cells += (uint64_t)(dl + 1) * (uint64_t)W1;
  if (best <= 0) { set_dud(r, 0); return cells; }  /* unreachable after a non-dud phase 1 */

  int32_t u = bu, kk = bk, c = 0, e = 0;            /* trace 2, :538-579 */
  cur = max3(AT(M, u, kk), AT(X, u, kk), AT(Y, u, kk));
  do {
    int32_t ii = u + ds, jj = kk - z + u;
    int mismatch;
    if (AT(M, u, kk) == cur) { mismatch = A[ii - 1] != B[jj - 1]; u--; }
    else if (AT(X, u, kk) == cur) { mismatch = 1; kk--; }
    else { mismatch = 1; u--; kk++; }               /* Y(u)(k) == max */
    if (mismatch) e++; else c++;
    cur = max3(AT(M, u, kk), AT(X, u, kk), AT(Y, u, kk));
  } while (cur > 0);
#undef AT
  r->start_i = u + ds; r->start_j = kk - z + u;     /* :582-584 */
  r->end_i = bu + ds;  r->end_j = bk - z + bu;
  r->correct = c; r->error = e; r->score = best;
  finish_record(s, n, m, r);
  return cells;
}

/* generateLocalAlignment (src/BioLibs.scala:171-263; the Set variant :267-368 runs the same
 * statements per B on matrices it allocates once as (lenA+1) x (maxL+1): cells right of lenB
 * keep stale values, but no recurrence reads them, and row 0 / column 0 are never rewritten).
 * Selected by --quadratic-align (src/Project4.scala:187-192, :585-601): three full matrices,
 * no band, the same three recurrences, arg-max with strict '>' in row-major order, greedy
 * backward trace.  Border quirks kept literally: the init loops run `until` the length, so
 * the last border entries stay at the 0 of Array.ofDim (:181-191).                        */
static uint64_t align_pair_quadratic(const hovl_settings *s, const uint8_t *A, int32_t n,
                                     const uint8_t *B, int32_t m, hovl_record *r, dp_scratch *d) {
  const int32_t gO = s->gap_open, gE = s->gap_extend;
  r->flags = 0;
  for (int32_t i = 0; i < n; i++) if (idx_acgt(A[i]) < 0) { set_dud(r, HOVL_F_BAD_BASE); return 0; }
  for (int32_t j = 0; j < m; j++) if (idx_acgt(B[j]) < 0) { set_dud(r, HOVL_F_BAD_BASE); return 0; }
  const int32_t W1 = m + 1;
  dp_reserve(d, (size_t)(n + 1) * (size_t)W1);
  int32_t *M = d->M, *X = d->X, *Y = d->Y;
  memset(M, 0, (size_t)(n + 1) * W1 * 4);
  memset(X, 0, (size_t)(n + 1) * W1 * 4);
  memset(Y, 0, (size_t)(n + 1) * W1 * 4);
#define AT(P, i, j) P[(size_t)(i) * W1 + (j)]
  for (int32_t i = 0; i < n; i++) { AT(M, i, 0) = 0; AT(X, i, 0) = 0; AT(Y, i, 0) = gO + i * gE; }   /* :181-185 */
  for (int32_t j = 0; j < m; j++) { AT(M, 0, j) = 0; AT(X, 0, j) = gO + j * gE; AT(Y, 0, j) = 0; }   /* :187-191 */
  int32_t best = 0, bi = 0, bj = 0;
  for (int32_t i = 1; i <= n; i++) {                  /* :197-219 */
    const int32_t *crow = s->cost[idx_acgt(A[i - 1])];
    for (int32_t j = 1; j <= m; j++) {
      int32_t mv = crow[idx_acgt(B[j - 1])] +
                   max2(max2(AT(M, i - 1, j - 1), AT(Y, i - 1, j - 1)), max2(AT(X, i - 1, j - 1), 0));
      int32_t xv = gE + max2(max2(AT(M, i, j - 1) + gO, AT(Y, i, j - 1) + gO), max2(AT(X, i, j - 1), 0));
      int32_t yv = gE + max2(max2(AT(M, i - 1, j) + gO, AT(Y, i - 1, j)), max2(AT(X, i - 1, j) + gO, 0));
      AT(M, i, j) = mv; AT(X, i, j) = xv; AT(Y, i, j) = yv;
      int32_t t = max3(mv, xv, yv);
      if (t > best) { best = t; bi = i; bj = j; }
    }
  }
  const uint64_t cells = (uint64_t)n * (uint64_t)m;
  /* maxLoc stays (0,0): the trace would read A.charAt(-1) (:236) — StringIndexOutOfBounds
   * in the reference; recorded as a dud here, like the fast aligner's same case.          */
  if (best <= 0) { set_dud(r, 0); return cells; }
  int32_t i = bi, j = bj, c = 0, e = 0;               /* :221-259 */
  int32_t cur = max3(AT(M, i, j), AT(X, i, j), AT(Y, i, j));
  do {
    int mismatch;
    if (AT(M, i, j) == cur) { mismatch = A[i - 1] != B[j - 1]; i--; j--; }
    else if (AT(X, i, j) == cur) { mismatch = 1; j--; }      /* pa = A char, pb = '-' */
    else { mismatch = 1; i--; }                              /* Y(i)(j) == max: pa = '-' */
    if (mismatch) e++; else c++;
    cur = max3(AT(M, i, j), AT(X, i, j), AT(Y, i, j));
  } while (cur > 0);
#undef AT
  r->start_i = i; r->start_j = j; r->end_i = bi; r->end_j = bj;   /* Alignment(.., (i,j), opt, c, e), :261 */
  r->correct = c; r->error = e; r->score = best;
  finish_record(s, n, m, r);
  return cells;
}

uint64_t orc_align_pair(const hovl_settings *s, const uint8_t *A, int32_t n,
                        const uint8_t *B, int32_t m, hovl_record *rec) {
  dp_scratch d = {0, 0, 0, 0};
  uint64_t c = s->aligner == 1 ? align_pair_quadratic(s, A, n, B, m, rec, &d) : align_pair(s, A, n, B, m, rec, &d);
  free(d.M); free(d.X); free(d.Y);
  return c;
}

int orc_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                    uint32_t n_reads, const uint32_t *pa, const uint32_t *pb, uint64_t np,
                    int threads, hovl_record *out, uint64_t *cells_out) {
  uint64_t cells = 0; int bad = 0;
  if (threads < 1) threads = 1;
#pragma omp parallel num_threads(threads) reduction(+ : cells)
  {
    dp_scratch d = {0, 0, 0, 0};
#pragma omp for schedule(dynamic, 64)
    for (int64_t p = 0; p < (int64_t)np; p++) {
      uint32_t a = pa[p], b = pb[p];
      if (a < 1 || a > n_reads || b < 1 || b > n_reads) { bad = 1; continue; }
      int32_t cnt = out[p].count;
      hovl_record r; memset(&r, 0, sizeof r);
      cells += (s->aligner == 1 ? align_pair_quadratic : align_pair)(
                          s, bases + offsets[a - 1], (int32_t)(offsets[a] - offsets[a - 1]),
                          bases + offsets[b - 1], (int32_t)(offsets[b] - offsets[b - 1]), &r, &d);
      r.a = (int32_t)a; r.b = (int32_t)b; r.count = cnt;
      out[p] = r;
    }
    free(d.M); free(d.X); free(d.Y);
  }
  if (cells_out) *cells_out = cells;
  return bad ? HOVL_E_ARG : HOVL_OK;
}

/* Overlap.print (src/ObjectStore.scala:127-135) + calcOverlaps (src/Project4.scala:806-820):
 * one record per Overlap.valid alignment, each followed by "\n".                         */
uint64_t orc_format_ovl(const hovl_record *recs, uint64_t n, char *buf, uint64_t cap) {
  uint64_t need = 0; char tmp[160];
  for (uint64_t i = 0; i < n; i++) {
    if (!(recs[i].flags & HOVL_F_OVL_VALID)) continue;
    int l = snprintf(tmp, sizeof tmp, "{OVL\nadj:N\nrds:%d,%d\nscr:0\nahg:%d\nbhg:%d\n}\n",
                     recs[i].a, recs[i].b, recs[i].ahg, recs[i].bhg);
    if (buf && need + (uint64_t)l <= cap) memcpy(buf + need, tmp, (size_t)l);
    need += (uint64_t)l;
  }
  return need;
}

/* ============================ table, pairs, dispatch ================================== */
/* wide-mode pair counter (extension for ids >= 32768): open addressing u64 -> u32        */
typedef struct { uint64_t *keys; uint32_t *vals; uint64_t cap, size; } wmap;
static void wmap_init(wmap *w, uint64_t cap) {
  w->cap = cap; w->size = 0;
  w->keys = (uint64_t *)malloc(cap * 8); w->vals = (uint32_t *)calloc(cap, 4);
  if (!w->keys || !w->vals) { fprintf(stderr, "oracle: out of memory\n"); abort(); }
  memset(w->keys, 0xff, cap * 8);
}
static inline uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x;
}
static void wmap_add(wmap *w, uint64_t key, uint32_t inc);
static void wmap_grow(wmap *w) {
  wmap n; wmap_init(&n, w->cap * 2);
  for (uint64_t i = 0; i < w->cap; i++) if (w->keys[i] != ~0ull) wmap_add(&n, w->keys[i], w->vals[i]);
  free(w->keys); free(w->vals); *w = n;
}
static void wmap_add(wmap *w, uint64_t key, uint32_t inc) {
  if ((w->size + 1) * 2 > w->cap) wmap_grow(w);
  uint64_t i = mix64(key) & (w->cap - 1);
  while (w->keys[i] != ~0ull && w->keys[i] != key) i = (i + 1) & (w->cap - 1);
  if (w->keys[i] == ~0ull) { w->keys[i] = key; w->size++; }
  w->vals[i] += inc;
}

typedef struct { uint64_t key; uint32_t count; } kc_t;
static int cmp_kc(const void *a, const void *b) {
  uint64_t x = ((const kc_t *)a)->key, y = ((const kc_t *)b)->key;
  return x < y ? -1 : x > y;
}

/* stable LSD sort of tuple indices by unsigned code (stage product for parity tests) */
static void sort_by_code(const uint32_t *code, uint64_t n, uint32_t *perm) {
  uint32_t *tmp = (uint32_t *)malloc((n ? n : 1) * 4);
  for (uint64_t i = 0; i < n; i++) perm[i] = (uint32_t)i;
  for (int pass = 0; pass < 2; pass++) {
    uint64_t *cnt = (uint64_t *)calloc(65537, 8);
    int sh = pass * 16;
    for (uint64_t i = 0; i < n; i++) cnt[((code[perm[i]] >> sh) & 0xffff) + 1]++;
    for (int d = 0; d < 65536; d++) cnt[d + 1] += cnt[d];
    for (uint64_t i = 0; i < n; i++) tmp[cnt[(code[perm[i]] >> sh) & 0xffff]++] = perm[i];
    memcpy(perm, tmp, n * 4);
    free(cnt);
  }
  free(tmp);
}

void orc_result_free(orc_result *r) {
  if (!r) return;
  free(r->recs); free(r->kmer_code); free(r->kmer_rid); free(r->kmer_loc);
  free(r->pair_key); free(r->pair_count); free(r->disp_a); free(r->disp_b); free(r->disp_count);
  free(r);
}

int orc_run(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
            uint32_t n_reads, int threads, int keep_stages, orc_result **out) {
  if (!s || !offsets || !out || (!bases && offsets[n_reads] > 0)) return HOVL_E_ARG;
  if (s->k < 1 || s->gap_open > 0 || s->gap_extend > 0) return HOVL_E_ARG;
  if (s->aligner != 0 && s->aligner != 1) return HOVL_E_UNSUPPORTED;  /* 1 = --quadratic-align */
  if (s->id_mode == 0 && n_reads > 32767) return HOVL_E_UNSUPPORTED;   /* (fst<<16)^snd, KmerTable.scala:73 */
  if (s->id_mode != 0 && s->order == 0) return HOVL_E_UNSUPPORTED;      /* Trove order is defined on the 32-bit key */
  if (threads < 1) threads = 1;
  orc_result *R = (orc_result *)calloc(1, sizeof *R);
  const int32_t k = s->k;
  const float head = bits_f(s->head_f32), tail = bits_f(s->tail_f32);
  const float mlead = bits_f(s->mid_lead_f32), mtail = bits_f(s->mid_tail_f32);
  uint64_t t0 = now_ns(), t1;

  /* ---- a1: generateKmerSet for every read, in read order (genMTKmerTable joins its
   *      futures in submission order, src/Project4.scala:550-560)                       */
  uint64_t *koff = (uint64_t *)malloc(((size_t)n_reads + 1) * 8);
  koff[0] = 0;
  for (uint32_t r = 0; r < n_reads; r++) {
    int64_t L = (int64_t)(offsets[r + 1] - offsets[r]);
    koff[r + 1] = koff[r] + (L >= k ? (uint64_t)(L - k + 1) : 0);
  }
  const uint64_t nk = koff[n_reads];
  uint32_t *code = (uint32_t *)malloc((nk ? nk : 1) * 4);
  uint32_t *rid = (uint32_t *)malloc((nk ? nk : 1) * 4);
  float *loc = (float *)malloc((nk ? nk : 1) * 4);
#pragma omp parallel for num_threads(threads) schedule(static)
  for (int64_t r = 0; r < (int64_t)n_reads; r++) {
    int64_t L = (int64_t)(offsets[r + 1] - offsets[r]);
    const uint8_t *p = bases + offsets[r];
    uint64_t o = koff[r];
    for (int64_t i = 0; i + k <= L; i++, o++) {
      code[o] = (uint32_t)kmer_code(p + i, k); rid[o] = (uint32_t)(r + 1); loc[o] = kmer_loc(i, L, k);
    }
  }
  t1 = now_ns(); R->ns_kmer = t1 - t0; t0 = t1;

  /* ---- a2: addKmerSet (src/KmerTable.scala:41-53): KmerData keyed by code, buckets in
   *      arrival order                                                                   */
  trove_map *kd = trove_new();
  uint32_t *bidx = (uint32_t *)malloc((nk ? nk : 1) * 4);     /* tuple -> bucket (first-occurrence rank) */
  uint64_t nb = 0;
  for (uint64_t i = 0; i < nk; i++) {
    int isnew; int32_t *v = trove_put_if_absent(kd, (int32_t)code[i], &isnew);
    if (isnew) *v = (int32_t)nb++;
    bidx[i] = (uint32_t)*v;
  }
  uint64_t *bstart = (uint64_t *)calloc(nb + 2, 8);
  for (uint64_t i = 0; i < nk; i++) bstart[bidx[i] + 1]++;
  uint64_t maxb = 0;
  for (uint64_t b = 0; b < nb; b++) { if (bstart[b + 1] > maxb) maxb = bstart[b + 1]; bstart[b + 1] += bstart[b]; }
  uint32_t *bel = (uint32_t *)malloc((nk ? nk : 1) * 4);       /* bucket contents: tuple indices */
  {
    uint64_t *fill = (uint64_t *)malloc((nb + 1) * 8);
    memcpy(fill, bstart, (nb + 1) * 8);
    for (uint64_t i = 0; i < nk; i++) bel[fill[bidx[i]]++] = (uint32_t)i;
    free(fill);
  }
  t1 = now_ns(); R->ns_table = t1 - t0; t0 = t1;

  /* ---- a3/a4: calcPairData (src/KmerTable.scala:85-149) over KmerData.iterator()      */
  int32_t *kd_keys = (int32_t *)malloc((nb ? nb : 1) * 4), *kd_vals = (int32_t *)malloc((nb ? nb : 1) * 4);
  trove_iterate(kd, kd_keys, kd_vals);
  trove_map *pd = NULL; wmap wm; memset(&wm, 0, sizeof wm);
  if (s->id_mode == 0) pd = trove_new(); else wmap_init(&wm, 1 << 16);
  uint32_t *st = (uint32_t *)malloc((maxb ? maxb : 1) * 4), *md = (uint32_t *)malloc((maxb ? maxb : 1) * 4),
           *en = (uint32_t *)malloc((maxb ? maxb : 1) * 4);
  uint64_t n_inst = 0;
  for (uint64_t it = 0; it < nb; it++) {
    uint64_t b = (uint64_t)kd_vals[it], nst = 0, nmd = 0, nen = 0;
    for (uint64_t q = bstart[b]; q < bstart[b + 1]; q++) {     /* :105-116 */
      uint32_t t = bel[q]; float l = loc[t];
      if (l <= head) st[nst++] = t;
      if (mlead <= l && l <= mtail) md[nmd++] = t;
      if (tail <= l) en[nen++] = t;
    }
    for (int phase = 0; phase < 2; phase++) {                  /* :118-128 */
      const uint32_t *ed = phase ? en : st; uint64_t ned = phase ? nen : nst;
      for (uint64_t x = 0; x < ned; x++) for (uint64_t y = 0; y < nmd; y++) {
        uint32_t a = ed[x], bb = md[y];                        /* addKmerPair(a, b), :57-80 */
        if (rid[a] == rid[bb]) continue;
        uint32_t fst, snd;
        if (loc[a] > loc[bb]) { fst = rid[a]; snd = rid[bb]; } else { fst = rid[bb]; snd = rid[a]; }
        n_inst++;
        if (pd) { int32_t key = (int32_t)((fst << 16) ^ snd); (*trove_put_if_absent(pd, key, NULL))++; }
        else wmap_add(&wm, ((uint64_t)fst << 32) | snd, 1);
      }
    }
  }
  free(st); free(md); free(en);
  t1 = now_ns(); R->ns_pairs = t1 - t0; t0 = t1;

  /* ---- a5/a6: calcDispatchData (:155-187) + dispatchCollisionBlocks order (:246-273)   */
  uint64_t np = pd ? (uint64_t)trove_size(pd) : wm.size;
  kc_t *sorted = (kc_t *)malloc((np ? np : 1) * sizeof(kc_t));     /* canonical (key asc) view */
  uint32_t *da = (uint32_t *)malloc((np ? np : 1) * 4), *db = (uint32_t *)malloc((np ? np : 1) * 4),
           *dc = (uint32_t *)malloc((np ? np : 1) * 4);
  uint64_t nd = 0, n_leads = 0;
  if (pd) {
    int32_t *pk = (int32_t *)malloc((np ? np : 1) * 4), *pv = (int32_t *)malloc((np ? np : 1) * 4);
    trove_iterate(pd, pk, pv);
    for (uint64_t i = 0; i < np; i++) {
      int32_t a = pk[i] >> 16, b = (int32_t)((uint32_t)pk[i] << 16) >> 16;   /* :169-170 */
      sorted[i].key = ((uint64_t)(uint32_t)a << 32) | (uint32_t)b; sorted[i].count = (uint32_t)pv[i];
    }
    if (s->order == 0) {
      /* DispatchData keyed by lead a; per-lead list in PairData iteration order */
      trove_map *dd = trove_new();
      uint32_t *lead_of = (uint32_t *)malloc((np ? np : 1) * 4);
      uint64_t *lcnt = NULL; uint64_t lcap = 0;
      for (uint64_t i = 0; i < np; i++) {
        int32_t cnt = pv[i];
        if (!(s->min_coll <= cnt && cnt <= s->max_coll)) { lead_of[i] = 0xffffffffu; continue; }
        int isnew; int32_t *v = trove_put_if_absent(dd, pk[i] >> 16, &isnew);
        if (isnew) {
          *v = (int32_t)n_leads++;
          if (n_leads > lcap) { lcap = lcap ? lcap * 2 : 1024; lcnt = (uint64_t *)realloc(lcnt, (lcap + 1) * 8); }
        }
        lead_of[i] = (uint32_t)*v;
      }
      uint64_t *lstart = (uint64_t *)calloc(n_leads + 2, 8);
      for (uint64_t i = 0; i < np; i++) if (lead_of[i] != 0xffffffffu) lstart[lead_of[i] + 1]++;
      for (uint64_t l = 0; l < n_leads; l++) lstart[l + 1] += lstart[l];
      uint64_t *lfill = (uint64_t *)malloc((n_leads + 1) * 8); memcpy(lfill, lstart, (n_leads + 1) * 8);
      uint32_t *lb = (uint32_t *)malloc((np ? np : 1) * 4), *lc = (uint32_t *)malloc((np ? np : 1) * 4);
      for (uint64_t i = 0; i < np; i++) if (lead_of[i] != 0xffffffffu) {
        uint64_t q = lfill[lead_of[i]]++;
        lb[q] = (uint32_t)((int32_t)((uint32_t)pk[i] << 16) >> 16); lc[q] = (uint32_t)pv[i];
      }
      int32_t *dk = (int32_t *)malloc((n_leads ? n_leads : 1) * 4), *dv = (int32_t *)malloc((n_leads ? n_leads : 1) * 4);
      trove_iterate(dd, dk, dv);
      for (uint64_t it = 0; it < n_leads; it++) {
        uint64_t l = (uint64_t)dv[it];
        for (uint64_t q = lstart[l]; q < lstart[l + 1]; q++) { da[nd] = (uint32_t)dk[it]; db[nd] = lb[q]; dc[nd] = lc[q]; nd++; }
      }
      free(dk); free(dv); free(lb); free(lc); free(lfill); free(lstart); free(lcnt); free(lead_of);
      trove_free(dd);
    }
    free(pk); free(pv);
  } else {
    uint64_t q = 0;
    for (uint64_t i = 0; i < wm.cap; i++) if (wm.keys[i] != ~0ull) { sorted[q].key = wm.keys[i]; sorted[q].count = wm.vals[i]; q++; }
  }
  qsort(sorted, np, sizeof(kc_t), cmp_kc);
  if (s->order != 0) {                                /* canonical order: (a, b) ascending */
    uint32_t last_a = 0;
    for (uint64_t i = 0; i < np; i++) {
      int64_t cnt = (int64_t)sorted[i].count;
      if (!(s->min_coll <= cnt && cnt <= s->max_coll)) continue;
      da[nd] = (uint32_t)(sorted[i].key >> 32); db[nd] = (uint32_t)sorted[i].key; dc[nd] = sorted[i].count;
      if (nd == 0 || da[nd] != last_a) { n_leads++; last_a = da[nd]; }
      nd++;
    }
  }
  t1 = now_ns(); R->ns_dispatch = t1 - t0; t0 = t1;

  /* ---- a7-a9: one calcLocalAlignmentSet per lead (src/Project4.scala:748-760), joined
   *      in submission order (:764-789)                                                  */
  hovl_record *recs = (hovl_record *)calloc(nd ? nd : 1, sizeof(hovl_record));
  for (uint64_t i = 0; i < nd; i++) recs[i].count = (int32_t)dc[i];
  uint64_t cells = 0;
  orc_align_pairs(s, bases, offsets, n_reads, da, db, nd, threads, recs, &cells);
  t1 = now_ns(); R->ns_align = t1 - t0;

  uint64_t n_dud = 0, n_av = 0, n_ov = 0, n_out = 0; int err = HOVL_OK;
  for (uint64_t i = 0; i < nd; i++) {
    /* the reference throws at whichever pair a worker reaches first; here BAD_BASE wins */
    if ((recs[i].flags & HOVL_F_SHORT_B) && err != HOVL_E_BAD_BASE) err = HOVL_E_SHORT_READ;
    if (recs[i].flags & HOVL_F_BAD_BASE) err = HOVL_E_BAD_BASE;
    if (recs[i].flags & HOVL_F_DUD) n_dud++;
    if (recs[i].flags & HOVL_F_ALIGN_VALID) n_av++;
    if (recs[i].flags & HOVL_F_OVL_VALID) n_ov++;
    if (!s->filter_valid || (recs[i].flags & HOVL_F_ALIGN_VALID)) recs[n_out++] = recs[i];
  }
  R->recs = recs; R->n_recs = n_out;
  R->st.n_reads = n_reads; R->st.n_bases = offsets[n_reads]; R->st.n_kmers = nk; R->st.n_buckets = nb;
  R->st.max_bucket = maxb; R->st.n_pair_instances = n_inst; R->st.n_distinct_pairs = np;
  R->st.n_dispatched = nd; R->st.n_leads = n_leads; R->st.n_duds = n_dud; R->st.n_align_valid = n_av;
  R->st.n_ovl_valid = n_ov; R->st.dp_cells = cells; R->st.n_records = n_out;
  R->ns_total = R->ns_kmer + R->ns_table + R->ns_pairs + R->ns_dispatch + R->ns_align;
  R->st.ns_total = R->ns_total;

  if (keep_stages) {
    uint32_t *perm = (uint32_t *)malloc((nk ? nk : 1) * 4);
    sort_by_code(code, nk, perm);
    R->kmer_code = (uint32_t *)malloc((nk ? nk : 1) * 4); R->kmer_rid = (uint32_t *)malloc((nk ? nk : 1) * 4);
    R->kmer_loc = (uint32_t *)malloc((nk ? nk : 1) * 4);
    for (uint64_t i = 0; i < nk; i++) {
      R->kmer_code[i] = code[perm[i]]; R->kmer_rid[i] = rid[perm[i]]; R->kmer_loc[i] = f_bits(loc[perm[i]]);
    }
    free(perm);
    R->n_kmers = nk;
    R->pair_key = (uint64_t *)malloc((np ? np : 1) * 8); R->pair_count = (uint32_t *)malloc((np ? np : 1) * 4);
    for (uint64_t i = 0; i < np; i++) { R->pair_key[i] = sorted[i].key; R->pair_count[i] = sorted[i].count; }
    R->n_pairs = np;
    R->disp_a = da; R->disp_b = db; R->disp_count = dc; R->n_disp = nd;
    da = db = dc = NULL;
  }
  free(da); free(db); free(dc); free(sorted);
  if (pd) trove_free(pd); else { free(wm.keys); free(wm.vals); }
  free(kd_keys); free(kd_vals); free(bel); free(bstart); free(bidx); trove_free(kd);
  free(code); free(rid); free(loc); free(koff);
  *out = R;
  return err;
}
