/* hovl_oracle.h — CPU oracle for the hash-overlap hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the algorithm in /root/reference/src/{BioLibs,KmerTable,
 * ObjectStore,Project4}.scala (Scala 2.9; cannot be compiled here: no JVM/scalac).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (libhovl.so) never links or calls it.
 *
 * PARITY STATUS: "parity unpinned" for the Scala program's exact record set and order —
 * the reference ships no output of its own code and no JVM exists on this box.  What IS
 * pinned (tests/test_oracle_golden.py): every record the oracle emits on
 * amos/small/crp177.seq and on the reconstructed c_ruddii reads is contained, with equal
 * ahg/bhg, in the AMOS golden overlap sets the reference ships
 * (amos/small/crp177.ovl, amos/c_ruddii.bnk), and the oracle's .ovl assembles through the
 * reference's own AMOS tools to the shipped contigs.  Output ORDER follows a model of
 * gnu.trove 3.0.3 TIntObjectHashMap (lib/trove.jar, bytecode only) — see trove_* below.
 */
#ifndef HOVL_ORACLE_H
#define HOVL_ORACLE_H
#include "../include/hovl.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_result {
  hovl_record *recs;  uint64_t n_recs;        /* after filter_valid, in output order       */
  hovl_stats_t st;
  /* stage products (filled when keep_stages != 0) */
  uint32_t *kmer_code, *kmer_rid, *kmer_loc;  /* sorted by (code, read id, position)       */
  uint64_t  n_kmers;
  uint64_t *pair_key;  uint32_t *pair_count;  /* sorted by key = (u64)fst<<32 | snd        */
  uint64_t  n_pairs;
  uint32_t *disp_a, *disp_b, *disp_count;     /* every dispatched pair, in dispatch order  */
  uint64_t  n_disp;
  /* wall-clock per stage, ns */
  uint64_t ns_kmer, ns_table, ns_pairs, ns_dispatch, ns_align, ns_total;
} orc_result;

/* readSeq: src/BioLibs.scala:26-50.  Returns 0 or a negative code; arrays are malloc'ed. */
int  orc_read_fasta(const char *path, uint8_t **bases, uint64_t **offsets, uint32_t *n_reads);
/* readHOXD: src/BioLibs.scala:66-114 (CSV form only, like the reference). */
int  orc_read_hoxd(const char *path, int32_t cost[4][4]);
/* Project4 defaults: src/Project4.scala:103-114; sections: src/ObjectStore.scala:30-35. */
void orc_settings_default(hovl_settings *s);
void orc_settings_set_sections(hovl_settings *s, float kmer_edge, float kmer_center);

/* The whole path: genMTKmerTable + genBlockMTAlign (src/Project4.scala:531-563,:725-790). */
int  orc_run(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
             uint32_t n_reads, int threads, int keep_stages, orc_result **out);
void orc_result_free(orc_result *r);

/* generateFastDovetailAlignment for one pair: src/BioLibs.scala:373-591.  Fills every
 * field of `rec` except a,b,count.  Returns cells evaluated.                              */
uint64_t orc_align_pair(const hovl_settings *s, const uint8_t *A, int32_t n,
                        const uint8_t *B, int32_t m, hovl_record *rec);
/* Align a list of (a,b) id pairs (1-based).  */
int  orc_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                     uint32_t n_reads, const uint32_t *pa, const uint32_t *pb, uint64_t np,
                     int threads, hovl_record *out, uint64_t *cells);

/* Overlap.print + calcOverlaps: src/ObjectStore.scala:127-135, src/Project4.scala:806-820 */
uint64_t orc_format_ovl(const hovl_record *recs, uint64_t n, char *buf, uint64_t cap);

/* Trove 3.0.3 TIntObjectHashMap model, exposed for unit tests. */
typedef struct trove_map trove_map;
trove_map *trove_new(void);
void  trove_free(trove_map *m);
/* returns the value slot for `key`, inserting it (value 0) when absent; *is_new says which */
int32_t *trove_put_if_absent(trove_map *m, int32_t key, int *is_new);
int32_t *trove_get(trove_map *m, int32_t key);
int32_t  trove_size(const trove_map *m);
int32_t  trove_capacity(const trove_map *m);
/* iteration in TIntObjectHashMap.iterator() order: fills keys[]/vals[] (size entries) */
void  trove_iterate(const trove_map *m, int32_t *keys, int32_t *vals);

#ifdef __cplusplus
}
#endif
#endif
