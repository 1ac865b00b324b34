// engine.cuh — device-side stage products of the hash-overlap path and the functions that
// make them.  Stage map (reference file:line -> here):
//   generateKmerSet / Kmer.seqHash  src/BioLibs.scala:54-61, src/ObjectStore.scala:40-67 -> build_reads + encode_kmers
//   KmerTable.addKmerSet            src/KmerTable.scala:41-53                             -> sort_kmers + group_kmers
//   calcPairData / addKmerPair      src/KmerTable.scala:57-149                            -> count_pairs
//   calcDispatchData                src/KmerTable.scala:155-187                           -> select_pairs (+ host order replay)
//   generateFastDovetailAlignmentSet, Alignment.valid, Overlap  src/BioLibs.scala:596-822,
//                                   src/ObjectStore.scala:89-142                          -> align_pairs
#pragma once
#include "common.cuh"

namespace hovl {

// Reads resident on the device: ASCII as given, plus the 2-bit packed copy the kernels use.
struct Reads {
  uint32_t n_reads = 0;
  uint64_t n_bases = 0, n_words = 0, n_kmers = 0;
  uint32_t max_len = 0, min_len = 0;
  const uint8_t *d_bases = nullptr;       // borrowed or = own_bases
  const uint64_t *d_offsets = nullptr;    // n_reads + 1
  DevBuf<uint8_t> own_bases; DevBuf<uint64_t> own_offsets;
  DevBuf<uint64_t> woff;                  // n_reads + 1: first packed word of each read
  DevBuf<uint64_t> koff;                  // n_reads + 1: first k-mer tuple of each read
  DevBuf<uint32_t> len;                   // n_reads
  DevBuf<uint32_t> packed;                // n_words (+2 slack): 16 bases / word, first base in the top bits
  DevBuf<uint32_t> bad;                   // n_reads: 1 when the read holds a non-ACGT byte
};

// k-mer tuples (code | read id, loc) as sort key / value, double-buffered.
struct Tuples {
  size_t n = 0; int cur = 0;
  DevBuf<uint32_t> key[2];                // code (ObjectStore.scala:48-67)
  DevBuf<uint64_t> val[2];                // (read id << 32) | float32 bits of loc
  uint32_t *k() { return key[cur].get(); }
  uint64_t *v() { return val[cur].get(); }
};

// KmerData as a sorted array cut into buckets, with the three section lists compacted.
struct Table {
  size_t n_buckets = 0, n_st = 0, n_md = 0, n_en = 0; uint64_t max_bucket = 0;
  DevBuf<uint64_t> bucket_start;          // n_buckets + 1 (index into the sorted tuples)
  DevBuf<uint32_t> st_base, md_base, en_base;   // n_buckets + 1 (index into the lists below)
  DevBuf<uint64_t> st_list, md_list, en_list;   // tuple values, bucket order, arrival order inside
};

// PairData: distinct (fst,snd) keys ascending with their collision counts.
struct Pairs {
  size_t n = 0; uint64_t n_instances = 0;
  int id_bits = 0;                        // key = (fst << id_bits) | snd
  DevBuf<uint64_t> key; DevBuf<uint32_t> count;
  DevBuf<uint32_t> seq;                   // order mode 0: rank of the pair's first instance in
                                          // the reference's traversal (KmerTable.scala:94-128)
};

void build_reads(Ctx &c, const hovl_settings &s, Reads &R, hovl_stats_t &st);
void encode_kmers(Ctx &c, const hovl_settings &s, const Reads &R, Tuples &T, hovl_stats_t &st);
void sort_kmers(Ctx &c, const hovl_settings &s, Tuples &T, hovl_stats_t &st);
void group_kmers(Ctx &c, const hovl_settings &s, Tuples &T, Table &B, hovl_stats_t &st);
// perm: optional device array of n_buckets bucket ids = the order buckets are traversed in
void count_pairs(Ctx &c, const hovl_settings &s, const Table &B, const uint32_t *d_perm, uint32_t n_reads,
                 bool want_seq, Pairs &P, hovl_stats_t &st);
// keeps min_coll <= count <= max_coll; outputs a,b,count arrays (canonical order) and returns how many
size_t select_pairs(Ctx &c, const hovl_settings &s, const Pairs &P, DevBuf<uint32_t> &a, DevBuf<uint32_t> &b,
                    DevBuf<uint32_t> &cnt);
void align_pairs(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                 const uint32_t *d_cnt, size_t np, hovl_record *d_out, hovl_stats_t &st);

// host-side replay of gnu.trove 3.0.3 TIntObjectHashMap iteration order (order.cpp)
void trove_iteration_order(const int32_t *keys_in_insertion_order, size_t n, std::vector<uint32_t> &order_out);

}  // namespace hovl
