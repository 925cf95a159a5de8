// host_logic.hpp — the sequential, parity-critical driver pieces of Spark's RandomForest.run that
// stay on the host inside emrf_fit: java.util.Random, XORShiftRandom,
// reservoir sampling of per-node feature subsets, Scala map iteration orders, the findSplits
// row sampler.  Product code: independent of oracle/.
//
// Upstream (spark 2.4.6 / scala 2.12, none vendored in the reference; pins: pom.xml:51-61):
//   core:util/random/XORShiftRandom.scala, core:util/random/SamplingUtils.scala,
//   core:util/random/RandomSampler.scala, core:rdd/PartitionwiseSampledRDD.scala,
//   scala/util/hashing/MurmurHash3.scala, scala/collection/mutable/HashTable.scala,
//   scala/collection/immutable/HashMap.scala, mllib:ml/tree/impl/DecisionTreeMetadata.scala
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/emrf.h"

namespace emrf {
namespace hostlogic {

// ---- bit generators -------------------------------------------------------------------------
// java.util.Random's LCG.  next(bits) = top `bits` of the 48-bit state, as a signed int.
struct LcgBits {
  uint64_t state;
  explicit LcgBits(int64_t seed) : state((static_cast<uint64_t>(seed) ^ 0x5DEECE66DULL) & kMask) {}
  static constexpr uint64_t kMask = (1ULL << 48) - 1;
  int32_t next(int bits) {
    state = (state * 0x5DEECE66DULL + 11ULL) & kMask;
    return static_cast<int32_t>(state >> (48 - bits));
  }
};

inline uint32_t murmur_mix_k(uint32_t k) {
  k *= 0xcc9e2d51u;
  k = (k << 15) | (k >> 17);
  return k * 0x1b873593u;
}
// MurmurHash3_x86_32 over `len` bytes (len must be a multiple of 4 here: 8 or 64).
inline uint32_t murmur3_words(const uint8_t* p, int len, uint32_t seed) {
  uint32_t h = seed;
  for (int i = 0; i + 4 <= len; i += 4) {
    uint32_t k;
    std::memcpy(&k, p + i, 4);  // little-endian host
    h ^= murmur_mix_k(k);
    h = (h << 13) | (h >> 19);
    h = h * 5u + 0xe6546b64u;
  }
  h ^= static_cast<uint32_t>(len);
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
// XORShiftRandom.hashSeed, Spark 2.4 flavour: ByteBuffer.allocate(java.lang.Long.SIZE) is a
// 64-BYTE buffer (big-endian seed then 56 zero bytes).  Version-sensitive (SPARK-23643).
inline uint64_t xorshift_seed_v24(int64_t seed) {
  uint8_t buf[64] = {0};
  uint64_t s = static_cast<uint64_t>(seed);
  for (int i = 7; i >= 0; --i) {
    buf[i] = static_cast<uint8_t>(s & 0xff);
    s >>= 8;
  }
  const uint32_t lo = murmur3_words(buf, 64, 0x3c074a61u);
  const uint32_t hi = murmur3_words(buf, 64, lo);
  return (static_cast<uint64_t>(hi) << 32) | lo;
}
struct XorShiftBits {
  uint64_t state;
  explicit XorShiftBits(int64_t seed) : state(xorshift_seed_v24(seed)) {}
  int32_t next(int bits) {
    uint64_t x = state;
    x ^= x << 21;
    x ^= x >> 35;
    x ^= x << 4;
    state = x;
    return static_cast<int32_t>(static_cast<uint32_t>(x & ((1ULL << bits) - 1ULL)));
  }
};
// The java.util.Random derived draws, shared by both generators.
template <class Bits>
struct JavaDraws : Bits {
  explicit JavaDraws(int64_t seed) : Bits(seed) {}
  int32_t next_int() { return this->next(32); }
  int64_t next_long() {
    const uint64_t hi = static_cast<uint64_t>(static_cast<int64_t>(this->next(32)));
    const uint64_t lo = static_cast<uint64_t>(static_cast<int64_t>(this->next(32)));
    return static_cast<int64_t>((hi << 32) + lo);
  }
  double next_double() {
    const int64_t a = this->next(26);
    const int64_t b = this->next(27);
    return static_cast<double>((a << 27) + b) * (1.0 / 9007199254740992.0);
  }
};
using JavaUtilRandom = JavaDraws<LcgBits>;
using SparkXorShift = JavaDraws<XorShiftBits>;

// ---- SamplingUtils.reservoirSampleAndCount over Range(0, n_items) -----------------------------
inline void reservoir_features(int n_items, int k, int64_t seed, int32_t* out, int* out_len) {
  const int fill = std::min(k, n_items);
  for (int i = 0; i < fill; ++i) out[i] = i;
  *out_len = fill;
  if (n_items <= k) return;
  SparkXorShift rnd(seed);
  for (int item = k; item < n_items; ++item) {
    const double seen = static_cast<double>(item + 1);  // l after increment
    const int64_t j = static_cast<int64_t>(rnd.next_double() * seen);
    if (j < k) out[j] = item;
  }
}

// ---- Scala 2.12 iteration orders for Int keys ------------------------------------------------
// mutable.HashMap (HashTable): table of chains, newest at the head; iteration from the highest
// bucket down.  Simulated with per-bucket singly linked lists in two flat arrays.
class ScalaMutableIntTable {
 public:
  ScalaMutableIntTable() : heads_(16, -1), count_(0), threshold_(12) {}
  void insert(int key) {
    const int b = bucket(key, static_cast<int>(heads_.size()));
    for (int e = heads_[b]; e >= 0; e = next_[e])
      if (keys_[e] == key) return;
    keys_.push_back(key);
    next_.push_back(heads_[b]);
    heads_[b] = static_cast<int>(keys_.size()) - 1;
    if (++count_ > threshold_) grow();
  }
  std::vector<int> iteration_order() const {
    std::vector<int> order;
    order.reserve(keys_.size());
    for (int b = static_cast<int>(heads_.size()) - 1; b >= 0; --b)
      for (int e = heads_[b]; e >= 0; e = next_[e]) order.push_back(keys_[e]);
    return order;
  }

 private:
  static uint32_t byteswap32(uint32_t v) {
    uint32_t hc = v * 0x9e3775cdu;
    hc = __builtin_bswap32(hc);
    return hc * 0x9e3775cdu;
  }
  static int bucket(int key, int table_len) {
    const uint32_t ones = static_cast<uint32_t>(table_len - 1);
    const int rotation = __builtin_popcount(ones) % 32;
    const uint32_t h = byteswap32(static_cast<uint32_t>(key));
    const uint32_t improved = rotation ? ((h >> rotation) | (h << (32 - rotation))) : h;
    return static_cast<int>((improved >> __builtin_clz(ones)) & ones);
  }
  void grow() {
    const int new_len = static_cast<int>(heads_.size()) * 2;
    std::vector<int> new_heads(new_len, -1);
    // HashTable.resize: old buckets from the top, each chain head→tail, head-inserting.
    for (int b = static_cast<int>(heads_.size()) - 1; b >= 0; --b) {
      int e = heads_[b];
      while (e >= 0) {
        const int following = next_[e];
        const int nb = bucket(keys_[e], new_len);
        next_[e] = new_heads[nb];
        new_heads[nb] = e;
        e = following;
      }
    }
    heads_.swap(new_heads);
    threshold_ = static_cast<int>((static_cast<int64_t>(new_len) * 750) / 1000);
  }
  std::vector<int> heads_, keys_, next_;
  int count_, threshold_;
};

// immutable.HashMap (HashTrieMap) order: by 5-bit digits of improve(hash), least significant first.
inline uint32_t scala_trie_hash(uint32_t h) {
  h = h + ~(h << 9);
  h ^= h >> 14;
  h += h << 4;
  return h ^ (h >> 10);
}
inline uint64_t scala_trie_sort_key(int key) {
  const uint32_t h = scala_trie_hash(static_cast<uint32_t>(key));
  uint64_t k = 0;  // digit 0 most significant in the sort key
  for (int d = 0; d < 7; ++d) k = (k << 5) | ((h >> (5 * d)) & 31u);
  return k;
}
// Order in which findBestSplits' `nodesForGroup.foreach` visits trees, given the trees in the
// order selectNodesToSplit first saw them: mutable.HashMap → .mapValues(_.toArray).toMap.
inline std::vector<int> group_tree_visit_order(const std::vector<int>& first_seen) {
  ScalaMutableIntTable table;
  for (int t : first_seen) table.insert(t);
  std::vector<int> order = table.iteration_order();
  if (order.size() > 4) {  // Map1..Map4 keep insertion order; beyond that a HashTrieMap
    std::sort(order.begin(), order.end(),
              [](int a, int b) { return scala_trie_sort_key(a) < scala_trie_sort_key(b); });
  }
  return order;
}

// ---- DecisionTreeMetadata.buildMetadata: numFeaturesPerNode -----------------------------------
inline int features_per_node(const emrf_params_t& p, int n_features) {
  int kind = p.feature_subset_kind;
  if (kind == 0) kind = (p.num_trees == 1) ? 1 : 2;
  const double F = static_cast<double>(n_features);
  switch (kind) {
    case 2: return static_cast<int>(std::ceil(std::sqrt(F)));
    case 3: return std::max(1, static_cast<int>(std::ceil(std::log(F) / std::log(2.0))));
    case 4: return static_cast<int>(std::ceil(F / 3.0));
    case 5: return std::min(static_cast<int>(p.feature_subset_value), n_features);
    case 6: return static_cast<int>(std::ceil(p.feature_subset_value * F));
    default: return n_features;
  }
}

// ---- findSplits: which rows RDD.sample keeps ---------------------------------------------------
inline double find_splits_fraction(const emrf_params_t& p, int64_t n_rows) {
  const int64_t required = std::max<int64_t>(static_cast<int64_t>(p.max_bins) * p.max_bins, 10000);
  return required < n_rows ? static_cast<double>(required) / static_cast<double>(n_rows) : 1.0;
}
// Returns false when every row is kept (fraction == 1.0; `rows` left empty).
inline bool find_splits_rows(const emrf_params_t& p, int64_t n_rows, const int64_t* part_off, int n_parts,
                             std::vector<int64_t>* rows) {
  rows->clear();
  const double fraction = find_splits_fraction(p, n_rows);
  if (fraction >= 1.0) return false;
  // input.sample(false, fraction, new XORShiftRandom(seed).nextInt()); per-partition seeds are
  // successive nextLong() of java.util.Random(sampleSeed) (PartitionwiseSampledRDD.getPartitions).
  SparkXorShift seeder(p.seed);
  JavaUtilRandom partition_seeds(static_cast<int64_t>(seeder.next_int()));
  const bool gap = fraction <= 0.4;  // RandomSampler.defaultMaxGapSamplingFraction
  const double lnq = std::log1p(-fraction);
  for (int part = 0; part < n_parts; ++part) {
    SparkXorShift rng(partition_seeds.next_long());
    const int64_t lo = part_off[part], hi = part_off[part + 1];
    if (gap) {
      // GapSampling: `skip` items are dropped, the next one is kept, then a new Geometric(fraction)
      // skip is drawn (the first one lazily, at the partition's first item).  Walked by jumps.
      if (lo < hi) {
        int64_t i = lo;
        while (true) {
          const int64_t skip = static_cast<int32_t>(std::log(std::max(rng.next_double(), 5e-11)) / lnq);
          i += skip;
          if (i >= hi) break;
          rows->push_back(i);
          ++i;
          if (i >= hi) break;
        }
      }
    } else {
      for (int64_t i = lo; i < hi; ++i)
        if (rng.next_double() <= fraction) rows->push_back(i);
    }
  }
  return true;
}

// ---- impurity on the host (driver side; plain IEEE, same operation order as
// mllib/tree/impurity/{Gini,Entropy}.scala) -------------------------------------------------------
inline double impurity_of(int kind, const uint32_t* counts, int n_classes) {
  double total = 0.0;
  for (int c = 0; c < n_classes; ++c) total += static_cast<double>(counts[c]);
  if (total == 0.0) return 0.0;
  if (kind == 0) {
    double imp = 1.0;
    for (int c = 0; c < n_classes; ++c) {
      const double f = static_cast<double>(counts[c]) / total;
      imp -= f * f;
    }
    return imp;
  }
  double imp = 0.0;
  for (int c = 0; c < n_classes; ++c) {
    if (counts[c] != 0) {
      const double f = static_cast<double>(counts[c]) / total;
      imp -= f * (std::log(f) / std::log(2.0));
    }
  }
  return imp;
}

// POLICY: how RandomForest.selectNodesToSplit ends a group.  Default (`groupDone` form, taken to be Spark
// 2.4.6's): the node that does not fit is examined — and spends a draw of the driver RNG — even when the group's memory
// already equals the limit.  policy_flags bit 0: the older loop tests `memUsage < maxMemoryUsage || memUsage == 0` first
// and stops without a draw in exactly that case.
inline bool select_nodes_examines_next(int policy_flags, int64_t mem_usage, int64_t max_mem) {
  return (policy_flags & 1) == 0 ? true : (mem_usage < max_mem || mem_usage == 0);
}
// POLICY: the pivot exp(-mean) of commons-math's PoissonDistribution.sample() (FastMath.exp there).
inline double poisson_exp_policy(double mean) { return std::exp(-mean); }

// Rows sharded over `world` ranks: where each node's partial histogram sits in the buffer that is reduce-scattered.
// A node is LIGHT when its weight over all ranks is <= 65535 (its cells then fit 16 bits, two per 32-bit word) and takes
// one half slot, a heavy node two.  Rank r receives the half slots [r*share, (r+1)*share): first its heavy nodes, then its
// light ones; heavy node number h goes to rank h / ceil(heavy/world), light node l to rank l / ceil(light/world).
// slot_out[i] = first half slot of node i, owner_out[i] = the rank that scans it; returns share.
inline int packed_slot_layout(const uint32_t* weight, int n, int world, int* slot_out, int* owner_out) {
  int n_light = 0;
  for (int i = 0; i < n; i++) n_light += (weight[i] <= 65535u);
  const int n_heavy = n - n_light;
  const int npr_h = (n_heavy + world - 1) / world, npr_l = (n_light + world - 1) / world;
  const int share = 2 * npr_h + npr_l;
  int h = 0, l = 0;
  for (int i = 0; i < n; i++) {
    int r;
    if (weight[i] <= 65535u) { r = l / npr_l; slot_out[i] = r * share + 2 * npr_h + l % npr_l; l++; }
    else { r = h / npr_h; slot_out[i] = r * share + 2 * (h % npr_h); h++; }
    owner_out[i] = r;
  }
  return share;
}

// Rows sharded, groups tiled on the device: the host knows a node's local entry count only as a bound `ub` (parent's
// bound, node's weight over all ranks).  The nodes of one tree that are pending together are disjoint, so together they
// hold at most the tree's local in-bag entries; a group's entries are therefore at most
//   sum over its trees of min(sum of the tree's node bounds, tree_total[tree]),
// and its tiles at most  entries / tile + nodes  (every node gets >= 1 tile: ceil(len/tile) <= len/tile + 1).
// `acc` is scratch of one zeroed word per tree (left zeroed).
inline uint64_t group_entries_bound(const uint32_t* ub, const int32_t* tree, int n, const uint64_t* tree_total, uint64_t* acc) {
  for (int i = 0; i < n; i++) acc[tree[i]] += ub[i];
  uint64_t total = 0;
  for (int i = 0; i < n; i++) {
    const int t = tree[i];
    if (acc[t]) { total += acc[t] < tree_total[t] ? acc[t] : tree_total[t]; acc[t] = 0; }
  }
  return total;
}
inline uint64_t tiles_bound(uint64_t entries_bound, uint32_t tile_len, int n_nodes) { return entries_bound / tile_len + (uint64_t)n_nodes; }

}  // namespace hostlogic
}  // namespace emrf
