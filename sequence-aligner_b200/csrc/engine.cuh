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
  uint32_t id_base = 0;                   // ids of these reads are id_base + 1 .. id_base + n_reads (a rank's slice)
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
  // compact form, for read sets of one common length: the value is the tuple's read-major index
  // (read * windows-per-read + window start), from which read id and loc are recomputed; 8 B per tuple
  // instead of 12 B through the encode, the radix passes and the group stage
  bool compact = false;
  DevBuf<uint32_t> idx[2];
  uint32_t *k() { return key[cur].get(); }
  uint64_t *v() { return val[cur].get(); }
  uint32_t *ix() { return idx[cur].get(); }
};

// KmerData as a sorted array cut into buckets, with the three section lists compacted.
struct Table {
  size_t n_buckets = 0, n_st = 0, n_md = 0, n_en = 0; uint64_t max_bucket = 0;
  DevBuf<uint64_t> bucket_start;          // n_buckets + 1 (index into the sorted tuples)
  DevBuf<uint32_t> st_base, md_base, en_base;   // n_buckets + 1 (index into the lists below)
  DevBuf<uint64_t> st_list, md_list, en_list;   // tuple values, bucket order, arrival order inside
  DevBuf<uint4> base3;                    // optional: (st_base, md_base, en_base) of a bucket in one 16 B entry
  DevBuf<uint32_t> bucket_of;             // optional: bucket of tuple koff[read] + window start (read-major order)
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
void encode_kmers(Ctx &c, const hovl_settings &s, const Reads &R, Tuples &T, hovl_stats_t &st, bool allow_compact = false);
void sort_kmers(Ctx &c, const hovl_settings &s, Tuples &T, hovl_stats_t &st);
struct EntIndex { uint64_t *ent; const uint64_t *ekoff; uint32_t *ecursor; };   // k-mer shard: per-read tuple entries
void group_kmers(Ctx &c, const hovl_settings &s, const Reads &R, Tuples &T, Table &B, bool want_bucket_of, hovl_stats_t &st,
                 const EntIndex *ei = nullptr);
// perm: optional device array of n_buckets bucket ids = the order buckets are traversed in
void count_pairs(Ctx &c, const hovl_settings &s, const Reads &R, const Table &B, const uint32_t *d_perm,
                 bool want_seq, bool filter, Pairs &P, hovl_stats_t &st);
// Read-major view of the k-mer tuples for the per-read pair gather: either every window of
// every read (tuple koff[r] + start has bucket Table::bucket_of[...]), or, on a k-mer shard,
// explicit (bucket, loc) entries per read.
struct GatherSrc {
  uint32_t n_reads = 0; size_t n_tuples = 0;
  uint32_t max_kmers = 0;                 // most k-mers a read has (window form; 0 = unknown): sizes the gather's CTAs
  uint32_t first = 0, count = 0;          // gather only reads first .. first + count - 1 (count 0 = all): the pairs whose fst they are
  const uint64_t *koff = nullptr;         // n_reads + 1
  const uint32_t *len = nullptr;          // n_reads (window form only)
  const uint64_t *ent = nullptr;          // entry form: (bucket << 32) | float32 bits of loc
};
// filter: write out only the pairs that pass minCollisions <= count <= maxCollisions
void count_pairs_gather(Ctx &c, const hovl_settings &s, const GatherSrc &R, const Table &B, bool sort_result, bool filter,
                        Pairs &P, hovl_stats_t &st);
void merge_partial_pairs(Ctx &c, uint64_t *keys, uint32_t *counts, size_t m, uint32_t n_reads, Pairs &P, hovl_stats_t &st);
// keeps min_coll <= count <= max_coll; outputs a,b,count arrays (canonical order) and returns how many
size_t select_pairs(Ctx &c, const hovl_settings &s, const Pairs &P, DevBuf<uint32_t> &a, DevBuf<uint32_t> &b,
                    DevBuf<uint32_t> &cnt);
void align_pairs(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                 const uint32_t *d_cnt, size_t np, hovl_record *d_out, hovl_stats_t &st);

// aligner = 1 (--quadratic-align): generateLocalAlignmentSet, quad.cu; align_pairs forwards to it
void align_pairs_quadratic(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                           const uint32_t *d_cnt, size_t np, hovl_record *d_out, hovl_stats_t &st);

// ---- sharded path (one context per rank; see include/hovl.h "sharded path") ----
__host__ __device__ inline uint32_t shard_mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x85ebca6bu; x ^= x >> 13; x *= 0xc2b2ae35u; x ^= x >> 16;
  return x;
}
// owner of a k-mer code: high bits of a mixed hash (raw top bits are skewed on AT-rich genomes)
__host__ __device__ inline uint32_t shard_of_code(uint32_t code, uint32_t parts) {
  return (uint32_t)(((uint64_t)shard_mix32(code) * parts) >> 32);
}
// range mode (replicated-table variant): contiguous code ranges, so the ranks' sorted slices concatenate into the sorted
// table; code_bits = 2 * min(k, 16).  Balanced on uniform genomes only (the mixed hash above is the default).
__host__ __device__ inline uint32_t shard_of_code_range(uint32_t code, int code_bits, uint32_t parts) {
  return (uint32_t)(((uint64_t)code * parts) >> code_bits);
}
__host__ __device__ inline uint32_t shard_owner(uint32_t code, int range_bits, uint32_t parts) {
  return range_bits ? shard_of_code_range(code, range_bits, parts) : shard_of_code(code, parts);
}
// owner of a pair = contiguous range of its fst read id, so that the ranks' sorted outputs concatenate
__host__ __device__ inline uint32_t shard_of_read(uint32_t read_id, uint32_t n_reads, uint32_t parts) {
  return (uint32_t)(((uint64_t)(read_id - 1u) * parts) / (n_reads ? n_reads : 1u));
}
// scatter (a[i], b[i]) into `parts` contiguous segments by owner; counts_host[p] = segment sizes
void partition_tuples(Ctx &c, const uint32_t *codes, const uint64_t *vals, size_t n, uint32_t parts, int range_bits,
                      uint32_t *codes_out, uint64_t *vals_out, uint64_t *counts_host);
void partition_pairs(Ctx &c, const uint64_t *keys, const uint32_t *counts, size_t n, int id_bits, uint32_t n_reads,
                     uint32_t parts, uint64_t *keys_out, uint32_t *counts_out, uint64_t *counts_host);
// all-to-all #1 fused with its partition: count per owner, then store every tuple into its owner's (peer-mapped) buffer
void count_owners(Ctx &c, const uint32_t *codes, size_t n, uint32_t parts, int range_bits, uint64_t *counts_host);
void scatter_tuples_peer(Ctx &c, const uint32_t *codes, const uint64_t *vals, size_t n, uint32_t parts, int range_bits,
                         const uint64_t *peer_codes, const uint64_t *peer_vals, const uint64_t *peer_base);
// per-read tuple counts of a k-mer shard -> ekoff (n_reads + 1, exclusive prefix)
void shard_read_offsets(Ctx &c, const uint64_t *vals, size_t n, uint32_t n_reads, DevBuf<uint64_t> &ekoff);
// woff (n_reads + 1) and max/min length from a replicated len array
void shard_read_layout(Ctx &c, const uint32_t *len, uint32_t n_reads, Reads &R);
// koff (n_reads + 1: first k-mer tuple of each read in read-major order) and n_kmers from a replicated len array
void shard_read_koff(Ctx &c, const uint32_t *len, uint32_t n_reads, int32_t k, Reads &R);

// host-side replay of gnu.trove 3.0.3 TIntObjectHashMap iteration order (order.cpp)
void trove_iteration_order(const int32_t *keys_in_insertion_order, size_t n, std::vector<uint32_t> &order_out);
// the same on the device (trove.cu): d_order[j] = index into d_keys of the j-th entry the iterator yields
void trove_iteration_order_device(Ctx &c, const int32_t *d_keys, size_t n, DevBuf<uint32_t> &d_order);

}  // namespace hovl
