// trove.cu — the iteration order of a gnu.trove 3.0.3 TIntObjectHashMap, computed on the device.
//
// The reference keeps PairData in a Trove map (src/KmerTable.scala:26-37) and dispatches in its iteration order
// (:162-168), so byte-identical output needs that order for every distinct pair: 27.4 M keys on c_ruddii (K1).
// order.cpp replays the inserts one by one on the host (2.7 s there: every insert is a cache miss in a 230 MB
// table); here the same slot table is built by all keys at once.
//
// Why that is possible.  TIntHash is open addressing with double hashing and no removals: a key lands on the first
// slot of ITS probe path that no EARLIER-inserted key occupies.  That makes the final layout a function of the
// insertion ranks alone: give every key its rank as a priority, let every key walk its probe path and claim a slot
// with atomicMin(rank); a key that finds a better-ranked occupant moves on, a key that displaces a worse-ranked
// occupant takes over that occupant's walk from the next slot of the occupant's path.  A slot's occupant only ever
// improves, so no key has to go back, and the fixed point is exactly the sequential layout (induction on the rank).
// A rehash (load factor 0.5, capacity -> next PrimeFinder prime >= 2 * capacity) re-inserts the old table's keys from
// the top slot down and then continues with the new arrivals: one more round of the same insertion with new ranks —
// old keys by their position in that walk, new keys by their index.  The walk order of a table is a compaction of its
// slots from the top, i.e. a scan.  Work: sum of the capacities, about 4 slots per key in total.
#include "engine.cuh"
#include "scan.cuh"

namespace hovl {

int32_t trove_prime_at_least(int64_t want);          // order.cpp (PrimeFinder table)

constexpr unsigned long long TROVE_EMPTY = ~0ull;

__global__ void __launch_bounds__(256) trove_iota_kernel(uint32_t *__restrict__ rank, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rank[i] = (uint32_t)i;
}

// keys [0, n_g) into an empty table of `cap` slots; slot = (rank << 32) | key index
__global__ void __launch_bounds__(256) trove_insert_kernel(const int32_t *__restrict__ keys, const uint32_t *__restrict__ rank,
                                                           size_t n_g, int32_t cap, unsigned long long *__restrict__ tab) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_g) return;
  unsigned long long mine = ((unsigned long long)rank[i] << 32) | (unsigned long long)i;
  int32_t h = keys[i] & 0x7fffffff;                      // TIntHash: HashFunctions.hash(int) = value, & 0x7fffffff
  int32_t at = h % cap, step = 1 + h % (cap - 2);
  for (;;) {
    const unsigned long long old = atomicMin(&tab[at], mine);
    if (old == TROVE_EMPTY) return;                      // a free slot: placed (until a better rank comes by)
    if (old > mine) {                                    // displaced a later key: carry on with ITS walk
      mine = old;
      h = keys[(uint32_t)old] & 0x7fffffff;
      step = 1 + h % (cap - 2);
    }
    at -= step;
    if (at < 0) at += cap;
  }
}

// slots from the top down: flag = occupied
struct TroveWalkIn {
  const unsigned long long *tab; size_t cap;
  __device__ uint32_t operator()(size_t j) const { return tab[cap - 1 - j] != TROVE_EMPTY ? 1u : 0u; }
};
// ordinal of a key in that walk -> its rank in the next table (rank_out) or the iteration order itself (order_out)
struct TroveWalkOut {
  const unsigned long long *tab; size_t cap; uint32_t *rank_out, *order_out;
  __device__ void operator()(size_t j, const uint32_t &v, const uint32_t &p) const {
    if (!v) return;
    const uint32_t idx = (uint32_t)tab[cap - 1 - j];
    if (rank_out) rank_out[idx] = p;
    if (order_out) order_out[p] = idx;
  }
};

// d_order[j] = index (into d_keys, which is in insertion order) of the j-th entry the map's iterator yields.
void trove_iteration_order_device(Ctx &c, const int32_t *d_keys, size_t n, DevBuf<uint32_t> &d_order) {
  d_order.alloc(c, n);
  if (n == 0) return;
  HOVL_REQUIRE(n < 0x7fffffffull / 2, HOVL_E_UNSUPPORTED, "too many keys for a Trove table");
  DevBuf<uint32_t> rank(c, n);
  HOVL_LAUNCH(c, trove_iota_kernel, (unsigned)div_up(n, 256), 256, 0, rank.get(), n);
  int32_t cap = trove_prime_at_least(20);               // new TIntObjectHashMap(): 10 / 0.5f -> 23
  // the last table is the largest: allocate it once
  int32_t cap_last = cap;
  for (;;) {
    const int32_t half = (int32_t)((float)cap_last * 0.5f), limit = cap_last - 1 < half ? cap_last - 1 : half;
    if ((int64_t)n <= limit) break;
    cap_last = trove_prime_at_least((int64_t)cap_last << 1);
  }
  DevBuf<unsigned long long> tab(c, (size_t)cap_last);
  DevBuf<uint32_t> part;
  for (;;) {
    const int32_t half = (int32_t)((float)cap * 0.5f), limit = cap - 1 < half ? cap - 1 : half;
    const size_t n_g = n <= (size_t)limit + 1 ? n : (size_t)limit + 1;     // the insert that passes the limit is made, then the table grows
    const bool last = n <= (size_t)limit;
    HOVL_CUDA(cudaMemsetAsync(tab.get(), 0xff, (size_t)cap * 8, c.stream));
    HOVL_LAUNCH(c, trove_insert_kernel, (unsigned)div_up(n_g, 256), 256, 0, d_keys, rank.get(), n_g, cap, tab.get());
    TroveWalkIn in{tab.get(), (size_t)cap};
    TroveWalkOut out{tab.get(), (size_t)cap, last ? nullptr : rank.get(), last ? d_order.get() : nullptr};
    scan_act<uint32_t, TroveWalkIn, TroveWalkOut>(c, (size_t)cap, in, out, part);
    if (last) break;
    cap = trove_prime_at_least((int64_t)cap << 1);
  }
}

}  // namespace hovl
