/* hovl.h — C ABI of the B200-native hash-overlap engine.
 *
 * This is the drop-in boundary for the hash-overlap hot path of
 * rohit507/Sequence-Aligner.  The reference has no FFI; the seam it replaces is
 * the pair of Scala calls
 *     generateKmerTable(file)                 src/Project4.scala:508
 *     genAlignment(table, settings, filter)   src/Project4.scala:567
 * whose result (Seq[Alignment]) is consumed by calcOverlaps (src/Project4.scala:795).
 * The JVM keeps readArgs / readSeq / readHOXD / calcOverlaps and replaces those two
 * calls with one hovl_run().  Plain pointers and sizes only; no C++ or torch types.
 *
 * There is NO CPU fallback behind these entry points: every compute call needs a
 * CUDA device (sm_100a) and fails with HOVL_E_NO_DEVICE / HOVL_E_CUDA otherwise.
 */
#ifndef HOVL_H
#define HOVL_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HOVL_ABI_VERSION 1

/* ---- status codes (0 = OK, negative = error; no exception crosses the ABI) ---- */
#define HOVL_OK              0
#define HOVL_E_ARG          -1  /* null pointer, bad size, bad settings value           */
#define HOVL_E_UNSUPPORTED  -2  /* e.g. n_reads > 32767 in id_mode 0 (reference key is
                                   (fst<<16)^snd, src/KmerTable.scala:73), band too wide */
#define HOVL_E_SHORT_READ   -3  /* a dispatched B is shorter than the band width: the
                                   reference throws StringIndexOutOfBounds at
                                   src/BioLibs.scala:648                                */
#define HOVL_E_BAD_BASE     -4  /* a dispatched read holds a non-ACGT base: the reference
                                   throws MatchError in costFunc, src/BioLibs.scala:144 */
#define HOVL_E_CUDA         -5
#define HOVL_E_NOMEM        -6
#define HOVL_E_NCCL         -7
#define HOVL_E_NO_DEVICE    -8

/* ---- settings: mirrors AlignSettings, src/ObjectStore.scala:17-36 ------------- */
typedef struct hovl_settings {
  int32_t  k;                 /* kmerSize (Project4.scala:41, default 12); codes use the
                                 first min(k,16) bases (ObjectStore.scala:52)            */
  int32_t  cost[4][4];        /* substitution matrix, rows/cols in A,C,G,T order
                                 (BioLibs.scala:119-141)                                 */
  int32_t  gap_open;          /* <= 0 (Project4.scala:160-162)                           */
  int32_t  gap_extend;        /* <= 0 (Project4.scala:163-165)                           */
  int32_t  min_overlap;       /* default 40                                              */
  uint32_t min_identity_f32;  /* IEEE-754 binary32 bits of minIdentity (default 0.98f)   */
  uint32_t max_ignore_f32;    /* bits of (float)maxIgnore (ObjectStore.scala:25)         */
  int32_t  min_coll;          /* default 7                                               */
  int32_t  max_coll;          /* default 222                                             */
  uint32_t head_f32;          /* kmerHeadEdge    = e            (ObjectStore.scala:32)   */
  uint32_t tail_f32;          /* kmerTailEdge    = 1.0f - e     (:33)                    */
  uint32_t mid_lead_f32;      /* kmerMidLeadEdge = 0.5f-(c*0.5f) (:34)                   */
  uint32_t mid_tail_f32;      /* kmerMidTailEdge = 0.5f+(c*0.5f) (:35)                   */
  int32_t  aligner;           /* 0 = fast dovetail (BioLibs.scala:596), 1 = quadratic
                                 local, --quadratic-align (BioLibs.scala:267-368)        */
  int32_t  order;             /* 0 = reference (Trove iteration) order,
                                 1 = canonical: records sorted by (a, b)                 */
  int32_t  id_mode;           /* 0 = reference 16-bit pair key, n_reads <= 32767 enforced
                                 1 = wide 64-bit pair key (extension)                    */
  int32_t  filter_valid;      /* genAlignment's `filter` (Project4.scala:567): 1 keeps
                                 only Alignment.valid records, 0 keeps every dispatched
                                 pair (duds included)                                    */
} hovl_settings;

/* One dispatched pair = the integer content of Alignment + Overlap
 * (src/ObjectStore.scala:89-142).  13 int32, 52 bytes.                                  */
#define HOVL_F_DUD         1   /* phase-1 trace did not reach column 0 (BioLibs.scala:694) */
#define HOVL_F_ALIGN_VALID 2   /* Alignment.valid  (ObjectStore.scala:102-107)             */
#define HOVL_F_OVL_VALID   4   /* Overlap.valid    (ObjectStore.scala:137-141)             */
#define HOVL_F_SHORT_B     8   /* len(B) < width: pair not aligned (-> HOVL_E_SHORT_READ)  */
#define HOVL_F_BAD_BASE   16   /* non-ACGT base in A or B (-> HOVL_E_BAD_BASE)             */
typedef struct hovl_record {
  int32_t a, b;               /* read ids, 1-based (BioLibs.scala:27,38)                  */
  int32_t count;              /* PairData collision count (KmerTable.scala:79)            */
  int32_t flags;              /* HOVL_F_*                                                 */
  int32_t start_i, start_j;   /* Alignment.start                                          */
  int32_t end_i, end_j;       /* Alignment.end                                            */
  int32_t correct, error;     /* Alignment.correct / .error                               */
  int32_t score;              /* phase-2 DP maximum (internal to the reference)           */
  int32_t ahg, bhg;           /* Overlap.ahg / .bhg (ObjectStore.scala:124-125)           */
} hovl_record;

typedef struct hovl_stats_t {
  uint64_t n_reads, n_bases;
  uint64_t n_kmers;            /* tuples produced by the splitter (a1)                    */
  uint64_t n_buckets;          /* distinct k-mer codes (a2)                               */
  uint64_t max_bucket;
  uint64_t n_pair_instances;   /* addKmerPair calls that reach the count (a3/a4)          */
  uint64_t n_distinct_pairs;   /* PairData.size                                           */
  uint64_t n_dispatched;       /* pairs passing min/max collisions (a5)                   */
  uint64_t n_leads;            /* DispatchData.size                                       */
  uint64_t n_duds, n_align_valid, n_ovl_valid;
  uint64_t dp_cells;           /* cells of the reference's phase-1 + phase-2 matrices for
                                  the aligned pairs (evaluated = dp_cells - dp_cells_skipped) */
  uint64_t n_records;          /* records returned (after filter_valid)                   */
  /* device time per stage in nanoseconds (CUDA events), host replay in wall ns */
  uint64_t ns_h2d, ns_pack, ns_encode, ns_kmer_sort, ns_group, ns_expand,
           ns_pair_sort, ns_pair_reduce, ns_order, ns_dp, ns_d2h, ns_total;
  uint64_t gpu_launches;       /* kernels launched by this run                            */
  uint64_t n_list_entries;     /* 8-byte section-list entries the pair gather read        */
  uint64_t kmer_tuple_bytes;   /* bytes per k-mer tuple through encode / sort / group: 12, or 8 in the
                                  compact (code, read-major index) form for reads of one length  */
  uint64_t dp_cells_skipped;   /* phase-1 cells of pairs whose doveStart came from the exact-
                                  prefix probe instead of the DP (dp.cu, tp_probe_kernel)  */
  uint64_t dp_cells_phase1;    /* the phase-1 part of dp_cells (lenA x width per aligned pair); phase-1 cells
                                  evaluated = dp_cells_phase1 - dp_cells_skipped              */
} hovl_stats_t;

typedef struct hovl_ctx    hovl_ctx;
typedef struct hovl_result hovl_result;

/* Fill `s` with the Project4 defaults (src/Project4.scala:103-114, :41) and the built-in
 * HOXD matrix (src/BioLibs.scala:119-141).                                               */
void hovl_settings_default(hovl_settings *s);
/* Derive the four section thresholds exactly as AlignSettings does from
 * (kmerEdge, kmerCenter) float32 bits (src/ObjectStore.scala:30-35, Project4.scala:257). */
void hovl_settings_set_sections(hovl_settings *s, uint32_t kmer_edge_f32, uint32_t kmer_center_f32);

/* A context owns its CUDA device(s): streams and scratch memory.  `devices` may be NULL (device 0) when n_devices is 1.
 * n_devices > 1: one handle over several GPUs of this machine that can reach each other's memory (NVLink / NVSwitch peer
 * access).  hovl_run on such a handle shards the work inside this process, one host thread per device, every exchange
 * through peer memory (csrc/multi.cu: the sharded path below without NCCL or a launcher); it needs id_mode = 1 and
 * order = 1 and returns exactly the records of a one-device run.  The other compute entry points (hovl_run_device,
 * hovl_stage_*, hovl_shard_*) need a one-device context.  One process per GPU with the stage entry points below
 * (hovl/dist.py) remains the way to span processes.                                              */
int  hovl_create(hovl_ctx **ctx, const int *devices, int n_devices);
void hovl_destroy(hovl_ctx *ctx);
const char *hovl_last_error(const hovl_ctx *ctx);

/* The hot path in one call.  `bases` = upper-case ASCII of all reads concatenated (host
 * memory, borrowed); read i (id i+1, BioLibs.scala:27) = bases[offsets[i] .. offsets[i+1]).
 * Replaces generateKmerTable + genAlignment (src/Project4.scala:57-59).                  */
int  hovl_run(hovl_ctx *ctx, const hovl_settings *s,
              const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads,
              hovl_result **out);
/* Same, with `bases`/`offsets` already resident on the context's device (device pointers).
 * Used by bench.py for the HBM-resident figure; results are still returned on the host. */
int  hovl_run_device(hovl_ctx *ctx, const hovl_settings *s,
                     const uint8_t *d_bases, const uint64_t *d_offsets, uint32_t n_reads,
                     uint64_t n_bases, hovl_result **out);

const hovl_record *hovl_records(const hovl_result *r, uint64_t *n);
int  hovl_stats(const hovl_result *r, hovl_stats_t *st);
void hovl_result_free(hovl_result *r);

/* Format the Overlap.valid records of `r` exactly as calcOverlaps + Overlap.print do
 * (src/Project4.scala:806-820, src/ObjectStore.scala:127-135).  Returns the number of
 * bytes needed; writes at most `cap` bytes to `buf` (may be NULL to size).  Host only.  */
uint64_t hovl_format_ovl(const hovl_record *recs, uint64_t n, char *buf, uint64_t cap);
/* The text of testOverlaps (src/Project4.scala:484-504, mode --test-overlaps) for records made with filter_valid = 0,
 * i.e. every dispatched pair in dispatch order, duds printed as BioLibs.dud (src/BioLibs.scala:22).  `bases` / `offsets`
 * are the reads the records came from (upper case, as readSeq leaves them).  Same sizing protocol as hovl_format_ovl;
 * returns 0 if a record names a read that does not exist.  Host only.                                               */
uint64_t hovl_format_test_overlaps(const hovl_record *recs, uint64_t n, const uint8_t *bases, const uint64_t *offsets,
                                   uint32_t n_reads, char *buf, uint64_t cap);
/* java.lang.Float.toString (what `println(" ... " + errRatio)` prints): writes a NUL-terminated string of at most
 * cap - 1 characters and returns the full length.  Host only.                                                       */
uint64_t hovl_format_float(float x, char *buf, uint64_t cap);

/* ---- the steps either side of the path, for hosts without the JVM reader / writer (host only) ----
 * hovl_reads_open = BioLibs.readSeq (src/BioLibs.scala:26-50): the first line must start with '>';
 * every later '>' line closes a record; header text is ignored, ids are 1,2,3.. in file order; body
 * lines are concatenated and upper-cased; lines end at "\n", "\r" or "\r\n" (BufferedReader.readLine);
 * the last record is emitted at EOF.  The buffers are page-locked when a CUDA runtime is present, so
 * hovl_run copies them to the device without staging.  Owned by the handle until hovl_reads_free.    */
typedef struct hovl_reads hovl_reads;
int  hovl_reads_open(const char *fasta_path, hovl_reads **out);
const uint8_t  *hovl_reads_bases(const hovl_reads *r, uint64_t *n_bases);
const uint64_t *hovl_reads_offsets(const hovl_reads *r, uint32_t *n_reads);   /* n_reads + 1 entries */
void hovl_reads_free(hovl_reads *r);
/* BioLibs.readHOXD (src/BioLibs.scala:66-114): title line, comma-separated column header, one row per
 * base until EOF or an empty line; values go through Integer.parseInt untrimmed.  cost is A,C,G,T.  */
int  hovl_read_hoxd(const char *path, int32_t cost[4][4]);
/* Integer.parseInt, as `args(i+1).toInt` uses it (src/Project4.scala:124-166).                       */
int  hovl_parse_int(const char *text, int32_t *out);
/* Project4.calcOverlaps (src/Project4.scala:795-825): the Overlap.valid records, formatted by
 * hovl_format_ovl, to `path` (an existing file is replaced) or to stdout when path is NULL or "".    */
int  hovl_write_ovl(const char *path, const hovl_record *recs, uint64_t n);
/* Message of the last failed call of this group on the calling thread.                               */
const char *hovl_io_last_error(void);

/* Iteration order of a gnu.trove 3.0.3 TIntObjectHashMap (the reference's lib/trove.jar,
 * used at src/KmerTable.scala:26-37) after inserting the distinct `keys` in the given order:
 * order_out[j] = index into keys[] of the j-th entry its iterator yields.  Host only.  This
 * is what makes the record order of hovl_run(order=0) equal to the reference's.            */
int  hovl_trove_order(const int32_t *keys, uint64_t n, uint32_t *order_out);
/* The same order computed on the context's device (csrc/trove.cu: every key claims the first slot of its probe path that
 * no earlier-inserted key holds, all keys at once; rehashes are further rounds).  This is what hovl_run(order = 0) uses
 * for PairData (27.4 M keys on c_ruddii); exported so that it can be checked against the host replay above.          */
int  hovl_trove_order_device(hovl_ctx *ctx, const int32_t *keys, uint64_t n, uint32_t *order_out);

/* Tuning knobs that do not change results (the table of names and defaults is in INTEGRATION.md section 5b; e.g.
 * "stream": the cudaStream_t to enqueue on, "kernel_timing": per-kernel times for hovl_kernel_times, "pair_chunk":
 * how many pair instances (addKmerPair calls) are materialised per expansion chunk, "sort_tma" / "sort_rank_ballot" /
 * "sort_early_counts": variants of the radix pass, "dp_wide" / "dp_pack16" / "dp_probe": variants of the aligner).
 * Unknown names and out-of-range values return HOVL_E_ARG.  On a handle over several devices the option goes to every
 * device.  The environment variable HOVL_OPTS="name=value,name=value" applies the same list to every handle the process
 * creates, inside hovl_create (a bad entry makes hovl_create fail with HOVL_E_ARG).                                    */
int  hovl_set_option(hovl_ctx *ctx, const char *name, int64_t value);
/* More options: "stream" = a cudaStream_t of the context's device (as an integer) that every
 * later call launches on instead of the context's own stream (borrowed, never destroyed);
 * "kernel_timing" = 1 brackets every kernel launch with CUDA events on that stream.
 * hovl_kernel_times reads the per-kernel totals of the last hovl_run* on `ctx`: it returns the
 * number of rows and, when i is in range, copies row i ("stage:kernel", launches, device ns). */
int  hovl_kernel_times(const hovl_ctx *ctx, uint32_t i, char *name, uint32_t cap,
                       uint64_t *launches, uint64_t *ns);

/* ---- stage entry points (host arrays in, host arrays out) ---------------------------
 * These expose the intermediate products of the path so that each stage can be checked
 * against the oracle: the counterpart of the reference's --test-kmer-cover /
 * --test-dispatch-collisions / --test-alignment probes (src/Project4.scala:299-440).
 * All output arrays are malloc'ed by the library and released with hovl_free().          */

/* a1+a2: splitter + KmerTable.  Returns the k-mer tuples sorted by (code, read id,
 * position) = the bucket contents of KmerData in addKmerSet order
 * (src/KmerTable.scala:41-53).                                                           */
int  hovl_stage_table(hovl_ctx *ctx, const hovl_settings *s,
                      const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads,
                      uint32_t **codes, uint32_t **read_ids, uint32_t **loc_f32,
                      uint64_t *n_kmers);
/* a3+a4: pair instances reduced by key.  keys[i] = (uint64)fst << 32 | snd, ascending;
 * counts[i] = PairData value (src/KmerTable.scala:57-80, :85-149).                       */
int  hovl_stage_pairs(hovl_ctx *ctx, const hovl_settings *s,
                      const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads,
                      uint64_t **keys, uint32_t **counts, uint64_t *n_pairs,
                      uint64_t *n_instances);
/* a7-a9 alone: align the given (a,b) id pairs (1-based ids) and fill records in order.  */
int  hovl_stage_align(hovl_ctx *ctx, const hovl_settings *s,
                      const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads,
                      const uint32_t *pair_a, const uint32_t *pair_b, uint64_t n_pairs,
                      hovl_record *records_out);
/* a2 statistics: KmerTable.uniqueKmers + kmerCollisionHistogram (src/KmerTable.scala:189-221), what
 * --test-kmer-cover prints per k (src/Project4.scala:299-320).  sizes[i] ascending = a bucket size that occurs,
 * counts[i] = how many distinct k-mers have that many entries; n_buckets = KmerData.size.                  */
int  hovl_stage_cover(hovl_ctx *ctx, const hovl_settings *s,
                      const uint8_t *bases, const uint64_t *offsets, uint32_t n_reads,
                      uint64_t *n_buckets, uint64_t **sizes, uint64_t **counts, uint64_t *n_sizes);
void hovl_free(void *p);

/* ---- sharded path: the hot path over several GPUs (north_star) ------------
 * One context per rank / GPU.  Every pointer below named d_* is a DEVICE pointer on the
 * context's device; counts_out arrays are host arrays of n_parts entries.  The library lays the
 * data out with one contiguous segment per destination rank and the caller performs the two
 * exchanges between the calls (NCCL all-to-all-v; see sequence-aligner_b200/hovl/dist.py and
 * INTEGRATION.md).  The reference has no counterpart: its KmerTable is one JVM heap
 * (src/KmerTable.scala:26-37).  Needs id_mode = 1 and order = 1; the ranks' record lists
 * concatenated in rank order are the canonical (a, b) order of hovl_run(order = 1).
 *
 *   rank r owns reads id_base+1 .. id_base+n_local          hovl_shard_encode
 *   all-to-all #1: tuples to hovl_shard_of_code(code)        (caller)
 *   KmerTable slice + partial PairData                       hovl_shard_pairs, hovl_shard_fetch_pairs
 *   all-to-all #2: (key, count) to hovl_shard_of_read(fst)   (caller)
 *   all-gather of the 2-bit packed reads                     (caller)
 *   sum, count filter, dovetail DP on the owner              hovl_shard_align
 */
uint32_t hovl_shard_of_code(uint32_t code, uint32_t n_parts);
uint32_t hovl_shard_of_read(uint32_t read_id /* 1-based */, uint32_t n_reads_total, uint32_t n_parts);

/* Splitter on the rank's reads (a1): tuples (code, (global read id << 32) | loc bits) written to
 * d_codes_out / d_vals_out grouped by destination rank, counts_out[p] = tuples for rank p.
 * Also hands out the rank's packed reads (2 bit/base, word-aligned per read), lengths and
 * bad-base flags for replication.  Capacities: tuple_cap >= sum(max(0, len-k+1)),
 * packed_cap >= sum(ceil(len/16)).                                                          */
int  hovl_shard_encode(hovl_ctx *ctx, const hovl_settings *s, const uint8_t *d_bases, const uint64_t *d_offsets,
                       uint32_t n_local, uint64_t n_bases_local, uint32_t id_base, uint32_t n_reads_total,
                       uint32_t n_parts, uint32_t *d_codes_out, uint64_t *d_vals_out, uint64_t tuple_cap,
                       uint32_t *d_packed_out, uint64_t packed_cap, uint32_t *d_len_out, uint32_t *d_bad_out,
                       uint64_t *counts_out, uint64_t *n_words_out);
/* All-to-all #1 fused with its partition, for ranks of one node.  hovl_shard_encode with d_codes_out ==
 * d_vals_out == NULL keeps the tuples in the context and only counts them per owner.  The caller agrees on
 * the layout of every rank's receive buffer (segments in source-rank order), maps the peers' buffers
 * (hovl_peer_alloc / hovl_peer_open: cudaMalloc + CUDA IPC), and hovl_shard_scatter then stores every tuple
 * straight into its owner's buffer over NVLink: peer_codes[p] / peer_vals[p] are the device addresses of
 * rank p's buffers in THIS process (the own buffers for p = this rank), peer_base[p] is where this rank's
 * segment starts in them.  The caller's barrier after it orders the ranks.                              */
int  hovl_shard_scatter(hovl_ctx *ctx, uint32_t n_parts, const uint64_t *peer_codes, const uint64_t *peer_vals,
                        const uint64_t *peer_base);
int  hovl_peer_alloc(hovl_ctx *ctx, uint64_t bytes, void **d_ptr, uint8_t *handle64);
int  hovl_peer_open(hovl_ctx *ctx, const uint8_t *handle64, void **d_ptr);
int  hovl_peer_close(hovl_ctx *ctx, void *d_ptr);
int  hovl_peer_free(hovl_ctx *ctx, void *d_ptr);
/* KmerTable + pair counting (a2-a4) over the tuples this rank received (clobbered).  The partial
 * (key = fst << id_bits | snd, count) list stays in the context, grouped by the rank that owns
 * fst; counts_out[p] = entries for rank p; hovl_shard_fetch_pairs copies it out (once).      */
int  hovl_shard_pairs(hovl_ctx *ctx, const hovl_settings *s, uint32_t *d_codes, uint64_t *d_vals, uint64_t n_tuples,
                      uint32_t n_reads_total, uint32_t n_parts, uint64_t *counts_out, uint64_t *n_partial_out);
int  hovl_shard_fetch_pairs(hovl_ctx *ctx, uint64_t *d_keys_out, uint32_t *d_counts_out, uint64_t cap);
/* Sum the partial counts received (clobbered), count filter (a5), dovetail DP + validity (a7-a9)
 * against the replicated reads.  d_records_out == NULL: the rank's records come back on the host
 * like hovl_run's.  Otherwise they are left in the caller's DEVICE buffer (records_cap >= n_partial
 * always suffices) for a device-side gather, and stats.n_records says how many there are.    */
int  hovl_shard_align(hovl_ctx *ctx, const hovl_settings *s, uint64_t *d_keys, uint32_t *d_counts, uint64_t n_partial,
                      const uint32_t *d_packed_all, uint64_t n_words_all, const uint32_t *d_len_all,
                      const uint32_t *d_bad_all, uint32_t n_reads_total, hovl_record *d_records_out,
                      uint64_t records_cap, hovl_result **out);

/* ---- no-allocation entry point -----------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It is the record compaction at the end of hovl_run (keep the Alignment.valid records when
 * filter_valid != 0, every record otherwise; count duds / valid alignments / valid overlaps), on DEVICE
 * buffers the caller owns, enqueued on `stream` (a cudaStream_t of `device`, NULL = the default stream)
 * and not synchronised.  The scan's scratch is a file-scope __device__ array of the library (about
 * 129 KiB, held by every GPU that loads libhovl.so until the process exits), so the call makes no
 * cudaMalloc / cudaFree / allocating library call.  Limits: n <= HOVL_NOALLOC_MAX_RECORDS, else
 * HOVL_E_UNSUPPORTED; device ordinal < 64.  All calls on one device share the scratch: calls that overlap
 * on the GPU (two streams, two host threads) race - serialise them.  A call clears what it needs zeroed.
 *   d_records_out : room for n records (d_counts_out5[0] of them are written, input order kept)
 *   d_counts_out5 : 5 x uint64 on the device: kept, duds, align-valid, ovl-valid, status bits
 *                   (1 = HOVL_F_SHORT_B seen, 2 = HOVL_F_BAD_BASE seen, 4 = a record with negative flags)
 * Errors: message through hovl_last_error(NULL) on the calling thread.                              */
#define HOVL_NOALLOC_MAX_RECORDS (1ull << 24)
int  hovl_filter_records_noalloc(int device, void *stream, const hovl_record *d_records_in, uint64_t n,
                                 int filter_valid, hovl_record *d_records_out, uint64_t *d_counts_out5);

/* ---- sharded path: alignment re-balance by DP cells --------------------------------------------
 * north_star: "the deduplicated pairs are re-balanced across GPUs for alignment".  hovl_shard_align in two halves with an
 * exchange of pairs between them (csrc/rebalance.cu).  hovl_shard_select sums the partial counts the rank received and
 * applies the count filter; the dispatched pairs stay in the context, sorted by (a, b), with the prefix of their cost
 * (lenA * width + (lenA + 1) * (width + 1) cells, src/BioLibs.scala:619-624); *cost_out is the rank's total.  With the
 * totals of all ranks known, hovl_shard_split cuts the rank's pairs into n_parts contiguous segments (counts_out): a pair
 * at position x of the job's cost line goes to rank floor(x * n_parts / cost_total).  hovl_shard_fetch_selected hands the
 * pairs out as (a, b, count) uint32 triplets (one all-to-all), hovl_shard_align_pairs aligns the triplets a rank received
 * (in source-rank order they are sorted by (a, b), so rank order remains the canonical record order).               */
int  hovl_shard_select(hovl_ctx *ctx, const hovl_settings *s, uint64_t *d_keys, uint32_t *d_counts, uint64_t n_partial,
                       const uint32_t *d_len_all, uint32_t n_reads_total, uint64_t *n_pairs_out, uint64_t *cost_out);
int  hovl_shard_split(hovl_ctx *ctx, uint64_t cost_before, uint64_t cost_total, uint32_t n_parts, uint64_t *counts_out);
int  hovl_shard_fetch_selected(hovl_ctx *ctx, uint32_t *d_abc_out, uint64_t cap_pairs);
int  hovl_shard_align_pairs(hovl_ctx *ctx, const hovl_settings *s, const uint32_t *d_abc, uint64_t n_pairs,
                            const uint32_t *d_packed_all, uint64_t n_words_all, const uint32_t *d_len_all,
                            const uint32_t *d_bad_all, uint32_t n_reads_total, hovl_record *d_records_out,
                            uint64_t records_cap, hovl_result **out);

/* ---- sharded path, replicated-table variant ---------------------------------------------------
 * For tables in which most bucket entries are chance matches (12-mers of a 100 Mbp genome: 2.6 x 10^10 distinct
 * pairs of which 0.5 % reach 7 collisions), exchanging partial pair counts is the wrong cut: the 7..222 filter can
 * only run after the exchange.  Here the TABLE is exchanged instead.  With option "shard_by_range" = 1
 * (hovl_set_option) hovl_shard_encode / hovl_shard_scatter give every rank a contiguous range of k-mer codes
 * (hovl_shard_of_code_range; balanced on uniform genomes, use the hashed default on skewed ones), the rank sorts its
 * range where it landed (hovl_shard_sort), every rank copies all sorted slices out of the peers' buffers
 * (hovl_shard_gather_peer, or the caller's all-gather) — in rank order they ARE the sorted KmerTable — and then groups
 * the whole table (replicated work) but counts, filters and aligns only the pairs whose fst is one of ITS reads
 * (hovl_shard_pairs_local).  The records stay on the device until hovl_shard_fetch_records copies them to `dst`
 * (device memory or page-locked host memory).  Rank order of the records == canonical (a, b) order.
 * The reference has no counterpart (one JVM heap, src/KmerTable.scala:26-37).                                     */
uint32_t hovl_shard_of_code_range(uint32_t code, int32_t k, uint32_t n_parts);
int  hovl_shard_sort(hovl_ctx *ctx, const hovl_settings *s, uint32_t *d_codes, uint64_t *d_vals, uint64_t n_tuples);
int  hovl_shard_gather_peer(hovl_ctx *ctx, uint32_t n_parts, const uint64_t *peer_codes, const uint64_t *peer_vals,
                            const uint64_t *counts, uint32_t *d_codes_all, uint64_t *d_vals_all);
int  hovl_shard_pairs_local(hovl_ctx *ctx, const hovl_settings *s, const uint32_t *d_codes_all, const uint64_t *d_vals_all,
                            uint64_t n_all, const uint32_t *d_packed_all, uint64_t n_words_all, const uint32_t *d_len_all,
                            const uint32_t *d_bad_all, uint32_t n_reads_total, uint32_t read_lo, uint32_t n_local,
                            hovl_result **out);
int  hovl_shard_fetch_records(hovl_ctx *ctx, hovl_record *dst, uint64_t cap);

#ifdef __cplusplus
}
#endif
#endif /* HOVL_H */
