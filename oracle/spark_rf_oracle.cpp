// =====================================================================================
// oracle/spark_rf_oracle.cpp — CPU restatement of Spark MLlib 2.4.6 RandomForestClassifier
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product (euromillioner_b200/, libemrf.so) may
// include, link, import or call this file.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs use it, and only as the checker / the
// timed CPU baseline.
//
// PARITY UNPINNED.  The algorithm restated here lives in a third-party dependency of the
// reference that is NOT vendored under /root/reference and is never called by it:
//   org.apache.spark:spark-mllib_2.12:2.4.6   (reference pin: pom.xml:56-61, runtime scope)
//   org.apache.spark:spark-core_2.12:2.4.6    (reference pin: pom.xml:51-55)
//   org.apache.commons:commons-math3:3.4.1    (transitive; PoissonDistribution, Well19937c)
//   org.scala-lang:scala-library:2.12.x       (MurmurHash3, mutable/immutable HashMap order)
// No JVM exists in this image, the reference holds no tests or golden vectors for this path
// (SURVEY.md §4, §8c), so this restatement follows the published upstream algorithm (SURVEY.md
// Appendix A) and is pinned only by the known answers of SURVEY.md Appendix B
// (tests/test_oracle_kat.py).  Every version-sensitive policy is an isolated, named function
//  so it can be corrected when someone diffs against a real Spark 2.4.6 run.
//
// Upstream paths cited below are relative to spark/mllib/src/main/scala/org/apache/spark/
// (prefix "core:" = spark/core/src/main/scala/org/apache/spark/, "cm:" = commons-math3).
// =====================================================================================
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

// -------------------------------------------------------------------------------------
// java.util.Random (JDK): 48-bit LCG.  KAT: SURVEY.md B.1.
// -------------------------------------------------------------------------------------
struct JavaRandom {
  uint64_t s = 0;
  JavaRandom() {}
  explicit JavaRandom(int64_t seed) { setSeed(seed); }
  virtual ~JavaRandom() {}
  virtual void setSeed(int64_t seed) { s = ((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
  virtual int32_t next(int bits) {
    s = (s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
    return (int32_t)(int64_t)(s >> (48 - bits));
  }
  int32_t nextInt() { return next(32); }
  int64_t nextLong() {
    int64_t hi = (int64_t)next(32);
    int64_t lo = (int64_t)next(32);
    return (int64_t)(((uint64_t)hi << 32) + (uint64_t)lo);
  }
  double nextDouble() {
    int64_t hi = (int64_t)next(26);
    int64_t lo = (int64_t)next(27);
    return (double)((hi << 27) + lo) * 0x1.0p-53;
  }
};

// -------------------------------------------------------------------------------------
// scala.util.hashing.MurmurHash3.bytesHash == MurmurHash3_x86_32.  KAT: SURVEY.md B.2.
// -------------------------------------------------------------------------------------
static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static uint32_t murmur3_bytes_hash(const uint8_t* data, int len, uint32_t seed) {
  uint32_t h = seed;
  int i = 0, rem = len;
  while (rem >= 4) {
    uint32_t k = (uint32_t)data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16) |
                 ((uint32_t)data[i + 3] << 24);
    k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u;
    h ^= k; h = rotl32(h, 13); h = h * 5u + 0xe6546b64u;
    i += 4; rem -= 4;
  }
  uint32_t k = 0;
  if (rem == 3) k ^= (uint32_t)data[i + 2] << 16;
  if (rem >= 2) k ^= (uint32_t)data[i + 1] << 8;
  if (rem >= 1) {
    k ^= (uint32_t)data[i];
    k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u;
    h ^= k;
  }
  h ^= (uint32_t)len;
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
static const uint32_t kScalaArraySeed = 0x3c074a61u;

// -------------------------------------------------------------------------------------
// core:util/random/XORShiftRandom.scala (Spark 2.4.x).
// POLICY hashSeed_v24 [SURVEY A.9]: 2.4 hashes a 64-BYTE buffer (ByteBuffer.allocate(
// java.lang.Long.SIZE)) holding the big-endian seed followed by 56 zero bytes.  Spark 3.0
// (SPARK-23643) changed it to 8 bytes; `hash_seed(seed, 8)` gives that variant.
// -------------------------------------------------------------------------------------
static int64_t xorshift_hash_seed(int64_t seed, int buf_len = 64) {
  uint8_t bytes[64];
  memset(bytes, 0, sizeof(bytes));
  for (int i = 0; i < 8; i++) bytes[i] = (uint8_t)((uint64_t)seed >> (56 - 8 * i));
  uint32_t lo = murmur3_bytes_hash(bytes, buf_len, kScalaArraySeed);
  uint32_t hi = murmur3_bytes_hash(bytes, buf_len, lo);
  return (int64_t)(((uint64_t)hi << 32) | (uint64_t)lo);
}
struct XORShiftRandom : JavaRandom {
  uint64_t x = 0;
  explicit XORShiftRandom(int64_t init) { x = (uint64_t)xorshift_hash_seed(init); }
  void setSeed(int64_t seed) override { x = (uint64_t)xorshift_hash_seed(seed); }
  int32_t next(int bits) override {
    uint64_t n = x ^ (x << 21);
    n ^= (n >> 35);
    n ^= (n << 4);
    x = n;
    return (int32_t)(uint32_t)(n & ((1ULL << bits) - 1));
  }
};

// -------------------------------------------------------------------------------------
// cm:random/AbstractWell.java + Well19937c.java + BitsStreamGenerator.nextDouble (3.4.1)
// -------------------------------------------------------------------------------------
struct Well19937c {
  static const int R = 624, M1 = 70, M2 = 179, M3 = 449;
  int32_t v[R];
  int index = 0;
  void setSeed(int64_t seed) {
    v[0] = (int32_t)((uint64_t)seed >> 32);
    v[1] = (int32_t)((uint64_t)seed & 0xffffffffULL);
    for (int i = 2; i < R; i++) {
      int64_t l = (int64_t)v[i - 2];
      v[i] = (int32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
    }
    index = 0;
  }
  uint32_t next(int bits) {
    const int iRm1 = (index + R - 1) % R, iRm2 = (index + R - 2) % R;
    const uint32_t v0 = (uint32_t)v[index];
    const uint32_t vM1 = (uint32_t)v[(index + M1) % R];
    const uint32_t vM2 = (uint32_t)v[(index + M2) % R];
    const uint32_t vM3 = (uint32_t)v[(index + M3) % R];
    const uint32_t z0 = (0x80000000u & (uint32_t)v[iRm1]) ^ (0x7FFFFFFFu & (uint32_t)v[iRm2]);
    const uint32_t z1 = (v0 ^ (v0 << 25)) ^ (vM1 ^ (vM1 >> 27));
    const uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
    const uint32_t z3 = z1 ^ z2;
    uint32_t z4 = z0 ^ (z1 ^ (z1 << 9)) ^ (z2 ^ (z2 << 21)) ^ (z3 ^ (z3 >> 21));
    v[index] = (int32_t)z3;
    v[iRm1] = (int32_t)z4;
    v[iRm2] = (int32_t)((uint32_t)v[iRm2] & 0x80000000u);
    index = iRm1;
    z4 ^= (z4 << 7) & 0xe46e1700u;
    z4 ^= (z4 << 15) & 0x9b868000u;
    return z4 >> (32 - bits);
  }
  double nextDouble() {
    const uint64_t high = (uint64_t)next(26) << 26;
    const uint64_t low = next(26);
    return (double)(high | low) * 0x1.0p-52;
  }
};

// cm:distribution/PoissonDistribution.sample() for mean < 40 (the only branch RF bagging hits:
// mean = subsamplingRate <= 1).  p = exp(-mean) (FastMath.exp; libm here — SURVEY A.5).
struct PoissonSampler {
  Well19937c rng;
  double mean, p;
  explicit PoissonSampler(double mean_);
  void reseed(int64_t seed) { rng.setSeed(seed); }
  int sample() {
    int64_t n = 0;
    double r = 1.0;
    while ((double)n < 1000.0 * mean) {
      double rnd = rng.nextDouble();
      r *= rnd;
      if (r >= p) n++; else return (int)n;
    }
    return (int)n;
  }
};

// (constructor defined below Params: the pivot is a policy function)
// -------------------------------------------------------------------------------------
// core:util/random/SamplingUtils.reservoirSampleAndCount.  Result NOT sorted.
// -------------------------------------------------------------------------------------
static std::vector<int> reservoir_sample_range(int n_items, int k, int64_t seed) {
  std::vector<int> res;
  int i = 0;
  while (i < k && i < n_items) { res.push_back(i); i++; }
  if (i < k) return res;  // input smaller than k: trimmed
  int64_t l = i;
  XORShiftRandom rand(seed);
  for (int item = i; item < n_items; item++) {
    l += 1;
    int64_t repl = (int64_t)(rand.nextDouble() * (double)l);
    if (repl < k) res[(int)repl] = item;
  }
  return res;
}

// -------------------------------------------------------------------------------------
// POLICY group_tree_iteration_order [SURVEY A.6/A.9]: iteration order of
//   mutableNodesForGroup.mapValues(_.toArray).toMap      (ml/tree/impl/RandomForest.scala,
//   selectNodesToSplit) for Int keys inserted in `keys` order (Scala 2.12):
//  (1) scala.collection.mutable.HashMap (HashTable): buckets high→low, newest first in a chain
//  (2) .toMap: <=4 entries keep that order (Map1..Map4); >4 → immutable.HashMap (HashTrieMap),
//      ordered by 5-bit groups of improve(hash) from the low bits up.
// -------------------------------------------------------------------------------------
static inline uint32_t bswap32(uint32_t v) {
  return (v >> 24) | ((v >> 8) & 0xff00u) | ((v << 8) & 0xff0000u) | (v << 24);
}
static inline uint32_t scala_byteswap32(uint32_t v) {
  uint32_t hc = v * 0x9e3775cdu;
  hc = bswap32(hc);
  return hc * 0x9e3775cdu;
}
static std::vector<int> scala_mutable_hashmap_order(const std::vector<int>& keys) {
  // HashTable: initial table 16, load factor 750/1000, chains head-inserted.
  std::vector<std::vector<int>> table(16);  // each chain: front = head
  int size = 0, threshold = 12;
  auto index_of = [&](int key, int tlen) {
    int ones = tlen - 1;
    int seedv = __builtin_popcount((unsigned)ones);
    int exponent = __builtin_clz((unsigned)ones);
    uint32_t i = scala_byteswap32((uint32_t)key);
    int rot = seedv % 32;
    uint32_t rotated = rot == 0 ? i : ((i >> rot) | (i << (32 - rot)));
    return (int)((rotated >> exponent) & (uint32_t)ones);
  };
  for (int key : keys) {
    int h = index_of(key, (int)table.size());
    bool found = false;
    for (int k2 : table[h]) if (k2 == key) { found = true; break; }
    if (found) continue;
    table[h].insert(table[h].begin(), key);
    size++;
    if (size > threshold) {
      int nlen = 2 * (int)table.size();
      std::vector<std::vector<int>> nt(nlen);
      for (int i = (int)table.size() - 1; i >= 0; i--)
        for (int e : table[i]) {  // head → tail
          int nh = index_of(e, nlen);
          nt[nh].insert(nt[nh].begin(), e);
        }
      table.swap(nt);
      threshold = (int)(((int64_t)nlen * 750) / 1000);
    }
  }
  std::vector<int> out;
  for (int i = (int)table.size() - 1; i >= 0; i--)
    for (int e : table[i]) out.push_back(e);
  return out;
}
static inline uint32_t scala_immutable_improve(uint32_t hcode) {
  uint32_t h = hcode + ~(hcode << 9);
  h = h ^ (h >> 14);
  h = h + (h << 4);
  return h ^ (h >> 10);
}
static std::vector<int> scala_immutable_map_order(const std::vector<int>& insertion) {
  if (insertion.size() <= 4) return insertion;  // Map1..Map4 keep insertion order
  std::vector<int> out = insertion;
  std::sort(out.begin(), out.end(), [](int a, int b) {
    uint32_t ha = scala_immutable_improve((uint32_t)a), hb = scala_immutable_improve((uint32_t)b);
    for (int lvl = 0; lvl < 35; lvl += 5) {
      uint32_t ca = (ha >> lvl) & 31u, cb = (hb >> lvl) & 31u;
      if (ca != cb) return ca < cb;
    }
    return false;
  });
  return out;
}
static std::vector<int> group_tree_iteration_order(const std::vector<int>& trees_in_first_seen_order) {
  return scala_immutable_map_order(scala_mutable_hashmap_order(trees_in_first_seen_order));
}

// -------------------------------------------------------------------------------------
// Parameters (mirrors Spark ML RandomForestClassifier params)
// -------------------------------------------------------------------------------------
struct Params {
  int32_t num_trees = 20, max_depth = 5, max_bins = 32, min_instances_per_node = 1, num_classes = 0;
  int32_t impurity = 0;             // 0 gini, 1 entropy
  int32_t feature_subset_kind = 0;  // 0 auto,1 all,2 sqrt,3 log2,4 onethird,5 n,6 fraction
  double feature_subset_value = 0.0, min_info_gain = 0.0, subsampling_rate = 1.0;
  int32_t max_memory_mb = 256;
  int32_t find_splits_policy = 0;   // 0: Spark 2.4 (zero filter + expected numSamples); 1: pre-2.4
  int64_t seed = 207336481LL;
  int32_t select_nodes_loop_policy = 0;   // 0: `groupDone` loop; 1: `while (memUsage < maxMemoryUsage || memUsage == 0)` loop
};

// POLICY: how RandomForest.selectNodesToSplit ends a group.  Both forms admit a node while
// memUsage + nodeMemUsage <= maxMemoryUsage (or the group is empty) and both spend an rng.nextLong() on the node that does
// not fit.  They differ only when a group's memory lands EXACTLY on the limit: the `groupDone` form (policy 0, what this
// restatement takes Spark 2.4.6 to do) then draws for one more node and rejects it; the older form (policy 1) tests
// `memUsage < maxMemoryUsage || memUsage == 0` before looking at the next node and stops without a draw.
static inline bool select_nodes_examines_next(int policy, int64_t mem_usage, int64_t max_mem) {
  return policy == 0 ? true : (mem_usage < max_mem || mem_usage == 0);
}

// POLICY: the pivot of commons-math 3.4.1 PoissonDistribution.sample() is FastMath.exp(-mean); libm's exp
// stands in for it here.  For subsamplingRate = 1.0 both are the double nearest to 1/e (0x3FD78B56362CEF38) as far as
// this restatement knows; if a JVM ever disagrees, this is the one place to change.
static inline double poisson_exp_policy(double mean) { return std::exp(-mean); }
inline PoissonSampler::PoissonSampler(double mean_) : mean(mean_), p(poisson_exp_policy(mean_)) {}

// ml/tree/impl/DecisionTreeMetadata.buildMetadata
static int num_features_per_node(const Params& p, int F) {
  int kind = p.feature_subset_kind;
  if (kind == 0) kind = (p.num_trees == 1) ? 1 : 2;  // auto → all | sqrt (classification)
  switch (kind) {
    case 1: return F;
    case 2: return (int)std::ceil(std::sqrt((double)F));
    case 3: return std::max(1, (int)std::ceil(std::log((double)F) / std::log(2.0)));
    case 4: return (int)std::ceil((double)F / 3.0);
    case 5: return std::min((int)p.feature_subset_value, F);
    case 6: return (int)std::ceil(p.feature_subset_value * (double)F);
  }
  return F;
}

// -------------------------------------------------------------------------------------
// findSplits: sampling.  ml/tree/impl/RandomForest.findSplits + core:rdd/RDD.sample →
// PartitionwiseSampledRDD + core:util/random/RandomSampler.scala BernoulliSampler/GapSampling.
// POLICY bernoulli_gap_sampling, partition_slicing [SURVEY A.3/A.9].
// Returns the sampled global row indices in partition order.
// -------------------------------------------------------------------------------------
static double samples_fraction_for_find_splits(const Params& p, int64_t num_examples) {
  int64_t required = std::max<int64_t>((int64_t)p.max_bins * p.max_bins, 10000);
  return required < num_examples ? (double)required / (double)num_examples : 1.0;
}
static std::vector<int64_t> find_splits_sample_rows(const Params& p, int64_t N, const int64_t* part_off, int P) {
  std::vector<int64_t> rows;
  double fraction = samples_fraction_for_find_splits(p, N);
  if (fraction >= 1.0) {
    rows.resize(N);
    for (int64_t i = 0; i < N; i++) rows[i] = i;
    return rows;
  }
  // input.sample(withReplacement=false, fraction, new XORShiftRandom(seed).nextInt())
  int64_t sample_seed = (int64_t)XORShiftRandom(p.seed).nextInt();
  JavaRandom part_seeds(sample_seed);  // PartitionwiseSampledRDD.getPartitions
  const double lnq = std::log1p(-fraction);
  for (int pi = 0; pi < P; pi++) {
    int64_t pseed = part_seeds.nextLong();
    XORShiftRandom rng(0);
    rng.setSeed(pseed);
    if (fraction <= 0.4) {
      // GapSampling: advance() at (lazy) construction, then per item.
      auto advance = [&]() {
        double u = std::max(rng.nextDouble(), 5e-11);
        return (int32_t)(std::log(u) / lnq);  // .toInt (saturating in Scala; never hit here)
      };
      int32_t drop = 0;
      bool constructed = false;
      for (int64_t i = part_off[pi]; i < part_off[pi + 1]; i++) {
        if (!constructed) { drop = advance(); constructed = true; }
        if (drop > 0) { drop--; }
        else { drop = advance(); rows.push_back(i); }
      }
    } else {
      for (int64_t i = part_off[pi]; i < part_off[pi + 1]; i++)
        if (rng.nextDouble() <= fraction) rows.push_back(i);
    }
  }
  return rows;
}

// ml/tree/impl/RandomForest.findSplitsForContinuousFeature (KAT B.3).
// `samples`: the feature's sampled values (after the zero filter when policy==0).
// `num_samples_expected`: (fraction*numExamples).toInt under policy 0; samples.size() under 1.
// POLICY findSplits_zeroFilter_and_numSamples [SURVEY A.9].
static std::vector<double> find_splits_for_continuous_feature(std::vector<double> samples, int num_splits,
                                                               int64_t num_samples_expected, int policy) {
  std::vector<double> out;
  if (samples.empty()) return out;
  int64_t part_num = (int64_t)samples.size();
  std::sort(samples.begin(), samples.end());
  std::vector<std::pair<double, int64_t>> vc;  // distinct value → count, ascending
  for (double x : samples) {
    if (!vc.empty() && vc.back().first == x) vc.back().second++;
    else vc.emplace_back(x, 1);
  }
  int64_t num_samples = part_num;
  if (policy == 0) {
    num_samples = num_samples_expected;
    if (num_samples - part_num > 0) {
      // valueCountMap + (0.0 -> (numSamples - partNumSamples)); zero values were filtered, so
      // 0.0 is not a key yet.  (−0.0 == 0.0 also filtered.)
      auto it = std::lower_bound(vc.begin(), vc.end(), std::make_pair(0.0, (int64_t)0),
                                 [](const std::pair<double, int64_t>& a, const std::pair<double, int64_t>& b) {
                                   return a.first < b.first;
                                 });
      vc.insert(it, std::make_pair(0.0, num_samples - part_num));
    }
  }
  int possible = (int)vc.size() - 1;
  if (possible == 0) return out;
  if (possible <= num_splits) {
    for (int i = 1; i <= possible; i++) out.push_back((vc[i - 1].first + vc[i].first) / 2.0);
    return out;
  }
  double stride = (double)num_samples / (double)(num_splits + 1);
  int64_t current = vc[0].second;
  double target = stride;
  for (size_t idx = 1; idx < vc.size(); idx++) {
    int64_t previous = current;
    current += vc[idx].second;
    double pgap = std::fabs((double)previous - target);
    double cgap = std::fabs((double)current - target);
    if (pgap < cgap) {
      out.push_back((vc[idx - 1].first + vc[idx].first) / 2.0);
      target += stride;
    }
  }
  return out;
}

// ml/tree/impl/TreePoint.findBin: java.util.Arrays.binarySearch semantics.
// Double.compare ordering: -0.0 < 0.0, NaN greatest.
static inline int java_double_compare(double a, double b) {
  if (a < b) return -1;
  if (a > b) return 1;
  int64_t ab, bb;
  // doubleToLongBits canonicalises NaN
  if (a != a) ab = 0x7ff8000000000000LL; else memcpy(&ab, &a, 8);
  if (b != b) bb = 0x7ff8000000000000LL; else memcpy(&bb, &b, 8);
  return ab == bb ? 0 : (ab < bb ? -1 : 1);
}
static inline int find_bin(const double* thr, int n, double x) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
    int c = java_double_compare(thr[mid], x);
    if (c < 0) lo = mid + 1;
    else if (c > 0) hi = mid - 1;
    else return mid;
  }
  return lo;  // -(lo+1) → -idx-1 = lo
}

// -------------------------------------------------------------------------------------
// mllib/tree/impurity/{Gini,Entropy}.scala calculate.  Plain IEEE ops, no FMA
// (compile with -ffp-contract=off).
// -------------------------------------------------------------------------------------
static double impurity_calculate(int kind, const double* counts, int C, double total) {
  if (total == 0) return 0.0;
  if (kind == 0) {
    double imp = 1.0;
    for (int c = 0; c < C; c++) {
      double f = counts[c] / total;
      imp -= f * f;
    }
    return imp;
  }
  double imp = 0.0;
  for (int c = 0; c < C; c++) {
    double cc = counts[c];
    if (cc != 0) {
      double f = cc / total;
      imp -= f * (std::log(f) / std::log(2.0));  // POLICY entropy_log: not bit-exact by nature
    }
  }
  return imp;
}
static double counts_sum(const double* counts, int C) {
  double s = 0.0;
  for (int c = 0; c < C; c++) s += counts[c];
  return s;
}
static int index_of_largest(const double* a, int C) {  // ImpurityCalculator.indexOfLargestArrayElement
  int mi = -1;
  double mv = -std::numeric_limits<double>::max();
  for (int c = 0; c < C; c++) if (a[c] > mv) { mi = c; mv = a[c]; }
  return mi;
}

// -------------------------------------------------------------------------------------
// Learning node / forest containers
// -------------------------------------------------------------------------------------
struct LNode {
  int id = 1;
  bool is_leaf = false;
  int left = -1, right = -1;      // pool indices
  bool has_split = false;
  int feature = -1, split_bin = -1;
  double threshold = 0.0;
  // ImpurityStats
  double gain = std::numeric_limits<double>::quiet_NaN();
  double impurity = 0.0;
  bool valid = true;
  bool has_stats = false;         // root before its first group has stats == null
  std::vector<double> calc;       // impurityCalculator.stats (class counts)
};
struct FlatNode {
  int32_t id, left, right, feature;
  double threshold, gain, impurity, prediction;
};
struct Forest {
  int num_trees = 0, num_classes = 0, num_features = 0;
  std::vector<int64_t> tree_off;       // T+1
  std::vector<FlatNode> nodes;         // pre-order per tree
  std::vector<double> stats;           // nodes.size()*C
  // fit by-products kept for stage-level parity tests
  std::vector<std::vector<double>> thresholds;  // per feature
  std::vector<int32_t> group_sizes;    // nodes per group, in scheduler order
  std::vector<int32_t> trace;          // per popped node: tree, id, then F_sub features
  int f_sub = 0;
};

// LearningNode.toNode(prune) (ml/tree/Node.scala).  POLICY prune_on_export.
static int64_t to_node(const std::vector<LNode>& pool, int ni, bool prune, int C, Forest& out) {
  const LNode& n = pool[ni];
  int64_t my = (int64_t)out.nodes.size();
  out.nodes.push_back(FlatNode());
  out.stats.resize(out.nodes.size() * (size_t)C);
  auto set_leaf = [&](int64_t at, double pred, double imp, const std::vector<double>& calc) {
    FlatNode& f = out.nodes[at];
    f.id = n.id; f.left = -1; f.right = -1; f.feature = -1;
    f.threshold = 0.0; f.gain = 0.0; f.impurity = imp; f.prediction = pred;
    for (int c = 0; c < C; c++) out.stats[at * C + c] = calc[c];
  };
  if (n.left >= 0 || n.right >= 0) {
    int64_t l = to_node(pool, n.left, prune, C, out);
    int64_t r = to_node(pool, n.right, prune, C, out);
    bool l_leaf = out.nodes[l].feature < 0 && out.nodes[l].left < 0;
    bool r_leaf = out.nodes[r].feature < 0 && out.nodes[r].left < 0;
    if (prune && l_leaf && r_leaf && out.nodes[l].prediction == out.nodes[r].prediction) {
      double pred = out.nodes[l].prediction;
      out.nodes.resize(my + 1);
      out.stats.resize(out.nodes.size() * (size_t)C);
      set_leaf(my, pred, n.impurity, n.calc);
    } else {
      FlatNode& f = out.nodes[my];
      f.id = n.id; f.left = (int32_t)(l - out.tree_off.back()); f.right = (int32_t)(r - out.tree_off.back());
      f.feature = n.feature; f.threshold = n.threshold; f.gain = n.gain; f.impurity = n.impurity;
      f.prediction = (double)index_of_largest(n.calc.data(), C);
      for (int c = 0; c < C; c++) out.stats[my * C + c] = n.calc[c];
    }
  } else {
    double pred = (double)index_of_largest(n.calc.data(), C);
    set_leaf(my, pred, n.valid ? n.impurity : -1.0, n.calc);
  }
  return my;
}

// -------------------------------------------------------------------------------------
// binsToBestSplit + calculateImpurityStats for ONE node (ml/tree/impl/RandomForest.scala;
// SURVEY A.8).  `hist` layout = DTStatsAggregator.allStats: [k][bin][class] with per-feature
// offsets scanLeft(numBins(f_k)*C).  hist is modified in place (prefix sums), as upstream.
// Outputs: best k (index into subset; -1 = no valid feature → dummy split), best split index,
// ImpurityStats.
// -------------------------------------------------------------------------------------
struct BestSplit {
  int k = -1, split = 0;
  double gain = -std::numeric_limits<double>::max();
  double impurity = 0.0, left_impurity = 0.0, right_impurity = 0.0;
  bool valid = false;
  std::vector<double> parent, left, right;
};
static BestSplit bins_to_best_split(double* hist, const int* subset, int f_sub, const int* num_bins_all,
                                    int C, int impurity_kind, int min_instances, double min_info_gain,
                                    bool have_node_stats, double node_impurity, const double* node_calc,
                                    const double* parent_stats /*C, aggregator parentStats*/) {
  BestSplit best;
  bool stats_null = !have_node_stats;  // gainAndImpurityStats == null (level 0)
  std::vector<double> parent_calc(C);
  double parent_imp = node_impurity;
  if (!stats_null) for (int c = 0; c < C; c++) parent_calc[c] = node_calc[c];
  bool any_feature = false;
  bool first_overall = true;
  std::vector<double> left(C), right(C);
  int64_t off = 0;
  for (int k = 0; k < f_sub; k++) {
    int f = subset[k];
    int nb = num_bins_all[f];
    int num_splits = nb - 1;
    double* fh = hist + off;
    off += (int64_t)nb * C;
    if (num_splits == 0) continue;
    any_feature = true;
    for (int s = 0; s < num_splits; s++)
      for (int c = 0; c < C; c++) fh[(s + 1) * C + c] += fh[s * C + c];
    // per-feature maxBy(gain): first maximum
    int fbest = -1;
    double fgain = 0.0, fli = 0.0, fri = 0.0;
    bool fvalid = false;
    std::vector<double> fl, fr;
    for (int s = 0; s < num_splits; s++) {
      for (int c = 0; c < C; c++) {
        left[c] = fh[s * C + c];
        right[c] = fh[num_splits * C + c] - left[c];
      }
      if (stats_null) {  // parent = left.copy.add(right); impurity = parent.calculate()
        for (int c = 0; c < C; c++) parent_calc[c] = left[c] + right[c];
        parent_imp = impurity_calculate(impurity_kind, parent_calc.data(), C, counts_sum(parent_calc.data(), C));
        stats_null = false;
      }
      int64_t lc = (int64_t)counts_sum(left.data(), C), rc = (int64_t)counts_sum(right.data(), C);
      int64_t tc = lc + rc;
      double gain, li = 0.0, ri = 0.0;
      bool valid;
      if (lc < min_instances || rc < min_instances) {
        gain = -std::numeric_limits<double>::max(); valid = false;
      } else {
        li = impurity_calculate(impurity_kind, left.data(), C, counts_sum(left.data(), C));
        ri = impurity_calculate(impurity_kind, right.data(), C, counts_sum(right.data(), C));
        double lw = (double)lc / (double)tc;
        double rw = (double)rc / (double)tc;
        gain = parent_imp - lw * li - rw * ri;
        valid = true;
        if (gain < min_info_gain) { gain = -std::numeric_limits<double>::max(); valid = false; }
      }
      if (fbest < 0 || gain > fgain) {
        fbest = s; fgain = gain; fli = li; fri = ri; fvalid = valid; fl = left; fr = right;
      }
    }
    if (first_overall || fgain > best.gain) {
      first_overall = false;
      best.k = k; best.split = fbest; best.gain = fgain; best.valid = fvalid;
      best.left_impurity = fli; best.right_impurity = fri; best.left = fl; best.right = fr;
    }
  }
  if (!any_feature) {
    // dummy split on subset.head, invalid stats built from the aggregator's parentStats
    best.k = -1; best.split = 0; best.valid = false; best.gain = -std::numeric_limits<double>::max();
    best.parent.assign(parent_stats, parent_stats + C);
    best.impurity = impurity_calculate(impurity_kind, best.parent.data(), C, counts_sum(best.parent.data(), C));
    return best;
  }
  best.parent = parent_calc;
  // ImpurityStats.getInvalidImpurityStats recomputes impurity from the calculator: same value.
  best.impurity = best.valid ? parent_imp
                             : impurity_calculate(impurity_kind, parent_calc.data(), C, counts_sum(parent_calc.data(), C));
  return best;
}

// -------------------------------------------------------------------------------------
// Stage functions
// -------------------------------------------------------------------------------------
static void compute_thresholds(const Params& p, const double* X, int64_t N, int F, const int64_t* part_off, int P,
                               std::vector<std::vector<double>>& thr) {
  // DecisionTreeMetadata: maxPossibleBins = min(maxBins, numExamples); numSplits = numBins-1
  int max_possible_bins = (int)std::min<int64_t>(p.max_bins, N);
  int num_splits = max_possible_bins - 1;
  double fraction = samples_fraction_for_find_splits(p, N);
  int64_t expected = (int64_t)(int32_t)(fraction * (double)N);
  std::vector<int64_t> rows = find_splits_sample_rows(p, N, part_off, P);
  thr.assign(F, {});
#pragma omp parallel for schedule(dynamic)
  for (int f = 0; f < F; f++) {
    std::vector<double> s;
    s.reserve(rows.size());
    for (int64_t r : rows) {
      double x = X[r * F + f];
      if (p.find_splits_policy == 0) { if (x != 0.0) s.push_back(x); }
      else s.push_back(x);
    }
    thr[f] = find_splits_for_continuous_feature(std::move(s), num_splits, expected, p.find_splits_policy);
  }
}

// ml/tree/impl/BaggedPoint.convertToBaggedRDD: weights[i*T + t]
static void bag_weights(const Params& p, int64_t N, const int64_t* part_off, int P, std::vector<uint8_t>& w,
                        int* max_weight) {
  const int T = p.num_trees;
  w.assign((size_t)N * T, 0);
  int gmax = 0;
  if (T > 1) {  // withReplacement
#pragma omp parallel for schedule(dynamic) reduction(max : gmax)
    for (int pi = 0; pi < P; pi++) {
      PoissonSampler ps(p.subsampling_rate);
      ps.reseed(p.seed + pi + 1);
      for (int64_t i = part_off[pi]; i < part_off[pi + 1]; i++)
        for (int t = 0; t < T; t++) {
          int s = ps.sample();
          if (s > gmax) gmax = s;
          w[(size_t)i * T + t] = (uint8_t)std::min(s, 255);
        }
    }
  } else if (p.subsampling_rate == 1.0) {
    std::fill(w.begin(), w.end(), 1);
    gmax = 1;
  } else {
#pragma omp parallel for schedule(dynamic)
    for (int pi = 0; pi < P; pi++) {
      XORShiftRandom rng(0);
      rng.setSeed(p.seed + pi + 1);
      for (int64_t i = part_off[pi]; i < part_off[pi + 1]; i++)
        for (int t = 0; t < T; t++) w[(size_t)i * T + t] = rng.nextDouble() < p.subsampling_rate ? 1 : 0;
    }
    gmax = 1;
  }
  if (max_weight) *max_weight = gmax;
}

// -------------------------------------------------------------------------------------
// RandomForest.run (SURVEY A.6–A.8)
// -------------------------------------------------------------------------------------
static Forest* fit(const Params& p, const double* X, const double* y, int64_t N, int F, const int64_t* part_off,
                   int P, std::string& err) {
  const int T = p.num_trees, C = p.num_classes;
  if (C <= 0 || C > 256) { err = "num_classes must be in 1..256"; return nullptr; }
  for (int64_t i = 0; i < N; i++) {
    double l = y[i];
    if (!(l >= 0 && l < C && std::floor(l) == l)) { err = "label out of range"; return nullptr; }
  }
  Forest* fo = new Forest();
  fo->num_trees = T; fo->num_classes = C; fo->num_features = F;
  // --- findSplits + metadata
  compute_thresholds(p, X, N, F, part_off, P, fo->thresholds);
  std::vector<int> num_bins(F);
  for (int f = 0; f < F; f++) num_bins[f] = (int)fo->thresholds[f].size() + 1;
  const int f_sub = num_features_per_node(p, F);
  fo->f_sub = f_sub;
  const bool subsampling = (f_sub != F);
  // --- TreePoint
  std::vector<uint8_t> bins((size_t)N * F);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < N; i++)
    for (int f = 0; f < F; f++)
      bins[(size_t)i * F + f] =
          (uint8_t)find_bin(fo->thresholds[f].data(), (int)fo->thresholds[f].size(), X[i * F + f]);
  std::vector<uint8_t> lab(N);
  for (int64_t i = 0; i < N; i++) lab[i] = (uint8_t)(int)y[i];
  // --- BaggedPoint
  std::vector<uint8_t> w;
  int maxw = 0;
  bag_weights(p, N, part_off, P, w, &maxw);
  if (maxw > 255) { err = "bag weight > 255"; delete fo; return nullptr; }
  // --- trees
  std::vector<std::vector<LNode>> pools(T);
  for (int t = 0; t < T; t++) { pools[t].reserve(64); pools[t].push_back(LNode()); }
  // node id cache (equivalent to re-walking predictImpl every pass): pool index per (tree,row)
  std::vector<int32_t> node_of((size_t)T * N, 0);
  std::vector<std::pair<int, int>> stack;  // (tree, pool index); back() = top
  for (int t = 0; t < T; t++) stack.emplace_back(t, 0);
  JavaRandom rng(p.seed);
  const int64_t max_mem = (int64_t)p.max_memory_mb * 1024LL * 1024LL;
  std::vector<int> all_features(F);
  for (int f = 0; f < F; f++) all_features[f] = f;

  struct GNode { int tree, pool; std::vector<int> subset; int64_t hist_off; };
  while (!stack.empty()) {
    // ---- selectNodesToSplit
    std::vector<GNode> group;
    std::vector<int> trees_seen;  // first-seen order = mutable HashMap insertion order
    std::map<int, std::vector<int>> nodes_for_tree;  // tree → group indices in pop order
    int64_t mem = 0, hist_total = 0;
    while (!stack.empty() && select_nodes_examines_next(p.select_nodes_loop_policy, mem, max_mem)) {
      auto top = stack.back();
      std::vector<int> subset;
      if (subsampling) subset = reservoir_sample_range(F, f_sub, rng.nextLong());
      else subset = all_features;
      int64_t total_bins = 0;
      for (int f : subset) total_bins += num_bins[f];
      int64_t node_mem = (int64_t)C * total_bins * 8;
      if (mem + node_mem <= max_mem || mem == 0) {
        stack.pop_back();
        if (!nodes_for_tree.count(top.first)) trees_seen.push_back(top.first);
        nodes_for_tree[top.first].push_back((int)group.size());
        GNode g; g.tree = top.first; g.pool = top.second; g.subset = subset; g.hist_off = hist_total;
        hist_total += (int64_t)C * total_bins;
        group.push_back(std::move(g));
        mem += node_mem;
      } else break;
    }
    fo->group_sizes.push_back((int32_t)group.size());
    for (auto& g : group) {
      fo->trace.push_back(g.tree);
      fo->trace.push_back(pools[g.tree][g.pool].id);
      for (int f : g.subset) fo->trace.push_back(f);
    }
    // ---- findBestSplits: aggregation (binSeqOp) — per tree independent, exact integer sums
    std::vector<double> hist((size_t)hist_total, 0.0);
    std::vector<double> parent_stats(group.size() * (size_t)C, 0.0);
    std::vector<int> tlist;
    for (auto& kv : nodes_for_tree) tlist.push_back(kv.first);
#pragma omp parallel for schedule(dynamic)
    for (int ti = 0; ti < (int)tlist.size(); ti++) {
      int t = tlist[ti];
      std::vector<int32_t> pool_to_g(pools[t].size(), -1);
      for (int gi : nodes_for_tree[t]) pool_to_g[group[gi].pool] = gi;
      const int32_t* no = &node_of[(size_t)t * N];
      for (int64_t i = 0; i < N; i++) {
        const int gidx = pool_to_g[no[i]];
        if (gidx < 0) continue;
        const GNode& g = group[gidx];
        // Spark adds zero weights too (harmless); they are skipped here.
        if (w[(size_t)i * T + t] == 0) continue;
        double wt = (double)w[(size_t)i * T + t];
        double* h = &hist[g.hist_off];
        const uint8_t* b = &bins[(size_t)i * F];
        int label = lab[i];
        int64_t off = 0;
        for (int f : g.subset) {
          h[off + (int64_t)b[f] * C + label] += wt;
          off += (int64_t)num_bins[f] * C;
        }
        parent_stats[(size_t)gidx * C + label] += wt;
      }
    }
    // ---- binsToBestSplit per node
    std::vector<BestSplit> bests(group.size());
#pragma omp parallel for schedule(dynamic)
    for (int gi = 0; gi < (int)group.size(); gi++) {
      const GNode& g = group[gi];
      const LNode& n = pools[g.tree][g.pool];
      bests[gi] = bins_to_best_split(&hist[g.hist_off], g.subset.data(), (int)g.subset.size(), num_bins.data(), C,
                                     p.impurity, p.min_instances_per_node, p.min_info_gain, n.has_stats,
                                     n.impurity, n.calc.data(), &parent_stats[(size_t)gi * C]);
    }
    // ---- driver: update nodes, create children, push (tree order = Scala map iteration order)
    std::vector<int> order = group_tree_iteration_order(trees_seen);
    struct Upd { int tree, pool, feature, split_bin, left, right; };
    std::vector<Upd> upds;
    for (int t : order) {
      for (int gi : nodes_for_tree[t]) {
        const GNode& g = group[gi];
        const BestSplit& bs = bests[gi];
        int ni = g.pool;
        int id = pools[t][ni].id;
        int level = 31 - __builtin_clz((unsigned)id);
        bool is_leaf = (bs.gain <= 0) || (level == p.max_depth);
        {
          LNode& n = pools[t][ni];
          n.is_leaf = is_leaf;
          n.gain = bs.gain; n.impurity = bs.impurity; n.valid = bs.valid; n.calc = bs.parent; n.has_stats = true;
        }
        if (!is_leaf) {
          int feat = g.subset[bs.k];
          bool child_is_leaf = (level + 1) == p.max_depth;
          // POLICY child_leaf_impurity_test: `== 0.0` in 2.4 (|x| < EPSILON in 3.x)
          bool l_leaf = child_is_leaf || (bs.left_impurity == 0.0);
          bool r_leaf = child_is_leaf || (bs.right_impurity == 0.0);
          LNode l, r;
          l.id = 2 * id; l.is_leaf = l_leaf; l.has_stats = true; l.calc = bs.left;
          l.impurity = impurity_calculate(p.impurity, l.calc.data(), C, counts_sum(l.calc.data(), C));
          r.id = 2 * id + 1; r.is_leaf = r_leaf; r.has_stats = true; r.calc = bs.right;
          r.impurity = impurity_calculate(p.impurity, r.calc.data(), C, counts_sum(r.calc.data(), C));
          int li = (int)pools[t].size();
          pools[t].push_back(l);
          pools[t].push_back(r);
          LNode& n = pools[t][ni];
          n.has_split = true; n.feature = feat; n.split_bin = bs.split;
          n.threshold = fo->thresholds[feat][bs.split];
          n.left = li; n.right = li + 1;
          if (!l_leaf) stack.emplace_back(t, li);
          if (!r_leaf) stack.emplace_back(t, li + 1);
          upds.push_back({t, ni, feat, bs.split, li, li + 1});
        }
      }
    }
    // ---- row → node (LearningNode.predictImpl one step; shouldGoLeft ≡ bin <= splitIndex)
    std::map<int, std::vector<int>> by_tree;  // tree → indices into upds
    for (size_t ui = 0; ui < upds.size(); ui++) by_tree[upds[ui].tree].push_back((int)ui);
    std::vector<int> ut;
    for (auto& kv : by_tree) ut.push_back(kv.first);
#pragma omp parallel for schedule(dynamic)
    for (int ti = 0; ti < (int)ut.size(); ti++) {
      int t = ut[ti];
      std::vector<int32_t> pool_to_u(pools[t].size(), -1);
      for (int ui : by_tree[t]) pool_to_u[upds[ui].pool] = ui;
      int32_t* no = &node_of[(size_t)t * N];
      for (int64_t i = 0; i < N; i++) {
        const int ui = pool_to_u[no[i]];
        if (ui < 0) continue;
        const Upd& u = upds[ui];
        no[i] = (bins[(size_t)i * F + u.feature] <= u.split_bin) ? u.left : u.right;
      }
    }
  }
  // ---- topNodes.map(_.toNode(prune = true))
  fo->tree_off.push_back(0);
  for (int t = 0; t < T; t++) {
    to_node(pools[t], 0, true, C, *fo);
    fo->tree_off.push_back((int64_t)fo->nodes.size());
  }
  return fo;
}

// RandomForestClassificationModel.predictRaw / raw2probabilityInPlace / argmax
static void predict(const Forest& fo, const double* X, int64_t N, double* raw, double* prob, double* pred,
                    const double* thresholds = nullptr) {
  const int C = fo.num_classes, F = fo.num_features;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < N; i++) {
    std::vector<double> votes(C, 0.0);
    const double* x = X + i * F;
    for (int t = 0; t < fo.num_trees; t++) {
      int64_t base = fo.tree_off[t];
      int64_t n = base;
      while (fo.nodes[n].left >= 0) {
        const FlatNode& nd = fo.nodes[n];
        n = base + ((x[nd.feature] <= nd.threshold) ? nd.left : nd.right);
      }
      const double* st = &fo.stats[n * C];
      double total = counts_sum(st, C);
      if (total != 0) for (int c = 0; c < C; c++) votes[c] += st[c] / total;
    }
    if (raw) for (int c = 0; c < C; c++) raw[i * C + c] = votes[c];
    if (pred) {  // raw2prediction = argmax of the raw vector (first max; Vector.argmax)
      int mi = 0;
      double mv = votes[0];
      for (int c = 1; c < C; c++) if (votes[c] > mv) { mv = votes[c]; mi = c; }
      if (thresholds) {  // raw2prediction with thresholds = probability2prediction(raw2probability(raw))
        double s = counts_sum(votes.data(), C);
        double mx = -std::numeric_limits<double>::infinity();
        mi = 0;
        for (int c = 0; c < C; c++) {
          const double pr = (s != 0) ? votes[c] / s : votes[c];
          const double scaled = pr / thresholds[c];
          if (scaled > mx) { mx = scaled; mi = c; }
        }
      }
      pred[i] = (double)mi;
    }
    if (prob) {
      double s = counts_sum(votes.data(), C);
      for (int c = 0; c < C; c++) prob[i * C + c] = (s != 0) ? votes[c] / s : votes[c];
    }
  }
}

}  // namespace orc

// =====================================================================================
// C interface (ctypes) — prefix orc_
// =====================================================================================
extern "C" {

typedef struct {
  int32_t num_trees, max_depth, max_bins, min_instances_per_node, num_classes;
  int32_t impurity, feature_subset_kind;
  double feature_subset_value, min_info_gain, subsampling_rate;
  int32_t max_memory_mb, find_splits_policy;
  int64_t seed;
  int32_t select_nodes_loop_policy, pad;
} orc_params_t;

static orc::Params to_params(const orc_params_t* p) {
  orc::Params q;
  q.num_trees = p->num_trees; q.max_depth = p->max_depth; q.max_bins = p->max_bins;
  q.min_instances_per_node = p->min_instances_per_node; q.num_classes = p->num_classes;
  q.impurity = p->impurity; q.feature_subset_kind = p->feature_subset_kind;
  q.feature_subset_value = p->feature_subset_value; q.min_info_gain = p->min_info_gain;
  q.subsampling_rate = p->subsampling_rate; q.max_memory_mb = p->max_memory_mb;
  q.find_splits_policy = p->find_splits_policy; q.seed = p->seed;
  q.select_nodes_loop_policy = p->select_nodes_loop_policy;
  return q;
}

static thread_local std::string g_err;
const char* orc_last_error() { return g_err.c_str(); }
int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// ---- primitives (KATs)
void orc_java_random(int64_t seed, int32_t* next_int, int64_t* next_long, double* next_double) {
  { orc::JavaRandom r(seed); *next_int = r.nextInt(); }
  { orc::JavaRandom r(seed); *next_long = r.nextLong(); }
  { orc::JavaRandom r(seed); *next_double = r.nextDouble(); }
}
void orc_java_random_longs(int64_t seed, int32_t n, int64_t* out) {
  orc::JavaRandom r(seed);
  for (int i = 0; i < n; i++) out[i] = r.nextLong();
}
uint32_t orc_murmur3(const uint8_t* data, int32_t len, uint32_t seed) { return orc::murmur3_bytes_hash(data, len, seed); }
int64_t orc_xorshift_hash_seed(int64_t seed, int32_t buf_len) { return orc::xorshift_hash_seed(seed, buf_len); }
void orc_xorshift(int64_t seed, int32_t* first_int, double* first_double) {
  { orc::XORShiftRandom r(seed); *first_int = r.nextInt(); }
  { orc::XORShiftRandom r(seed); *first_double = r.nextDouble(); }
}
void orc_xorshift_doubles(int64_t seed, int32_t n, double* out) {
  orc::XORShiftRandom r(seed);
  for (int i = 0; i < n; i++) out[i] = r.nextDouble();
}
void orc_well19937c_doubles(int64_t seed, int32_t n, double* out) {
  orc::Well19937c w;
  w.setSeed(seed);
  for (int i = 0; i < n; i++) out[i] = w.nextDouble();
}
void orc_well19937c_bits32(int64_t seed, int32_t n, uint32_t* out) {
  orc::Well19937c w;
  w.setSeed(seed);
  for (int i = 0; i < n; i++) out[i] = w.next(32);
}
void orc_poisson(int64_t seed, double mean, int32_t n, int32_t* out) {
  orc::PoissonSampler ps(mean);
  ps.reseed(seed);
  for (int i = 0; i < n; i++) out[i] = ps.sample();
}
int32_t orc_reservoir(int32_t n_items, int32_t k, int64_t seed, int32_t* out) {
  std::vector<int> r = orc::reservoir_sample_range(n_items, k, seed);
  for (size_t i = 0; i < r.size(); i++) out[i] = r[i];
  return (int32_t)r.size();
}
void orc_scala_mutable_order(const int32_t* keys, int32_t n, int32_t* out) {
  std::vector<int> k(keys, keys + n);
  std::vector<int> o = orc::scala_mutable_hashmap_order(k);
  for (size_t i = 0; i < o.size(); i++) out[i] = o[i];
}
void orc_scala_immutable_order(const int32_t* keys, int32_t n, int32_t* out) {
  std::vector<int> k(keys, keys + n);
  std::vector<int> o = orc::scala_immutable_map_order(k);
  for (size_t i = 0; i < o.size(); i++) out[i] = o[i];
}
void orc_group_tree_order(const int32_t* keys, int32_t n, int32_t* out) {
  std::vector<int> k(keys, keys + n);
  std::vector<int> o = orc::group_tree_iteration_order(k);
  for (size_t i = 0; i < o.size(); i++) out[i] = o[i];
}
int32_t orc_find_splits_for_continuous_feature(const double* samples, int32_t n, int32_t num_splits,
                                               int64_t num_samples_expected, int32_t policy, double* out) {
  std::vector<double> s(samples, samples + n);
  std::vector<double> r = orc::find_splits_for_continuous_feature(s, num_splits, num_samples_expected, policy);
  for (size_t i = 0; i < r.size(); i++) out[i] = r[i];
  return (int32_t)r.size();
}
int32_t orc_num_features_per_node(const orc_params_t* p, int32_t F) { return orc::num_features_per_node(to_params(p), F); }

// ---- stages
// thresholds_out: F*(max_bins) doubles (row f holds n_splits_out[f] values)
int orc_find_splits(const orc_params_t* p, const double* X, int64_t N, int32_t F, const int64_t* part_off,
                    int32_t P, double* thresholds_out, int32_t* n_splits_out) {
  std::vector<std::vector<double>> thr;
  orc::compute_thresholds(to_params(p), X, N, F, part_off, P, thr);
  for (int f = 0; f < F; f++) {
    n_splits_out[f] = (int32_t)thr[f].size();
    if ((int)thr[f].size() > p->max_bins) { g_err = "more splits than max_bins"; return -1; }
    for (size_t i = 0; i < thr[f].size(); i++) thresholds_out[(size_t)f * p->max_bins + i] = thr[f][i];
  }
  return 0;
}
int64_t orc_find_splits_sample_rows(const orc_params_t* p, int64_t N, const int64_t* part_off, int32_t P,
                                    int64_t* rows_out, int64_t cap) {
  std::vector<int64_t> r = orc::find_splits_sample_rows(to_params(p), N, part_off, P);
  for (size_t i = 0; i < r.size() && (int64_t)i < cap; i++) rows_out[i] = r[i];
  return (int64_t)r.size();
}
// bins_out row-major N×F
void orc_bin(const double* X, int64_t N, int32_t F, const double* thresholds, const int32_t* n_splits,
             int32_t thr_stride, uint8_t* bins_out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < N; i++)
    for (int f = 0; f < F; f++)
      bins_out[(size_t)i * F + f] = (uint8_t)orc::find_bin(thresholds + (size_t)f * thr_stride, n_splits[f], X[i * F + f]);
}
// weights_out row-major N×T
int orc_bag_weights(const orc_params_t* p, int64_t N, const int64_t* part_off, int32_t P, uint8_t* weights_out) {
  std::vector<uint8_t> w;
  int maxw = 0;
  orc::bag_weights(to_params(p), N, part_off, P, w, &maxw);
  memcpy(weights_out, w.data(), w.size());
  return maxw;
}
// one node's best split from an integer histogram laid out [k][bin][class] with per-feature
// numBins (num_bins_all indexed by feature id).  node_calc==NULL → root semantics (stats==null).
int orc_best_split(const double* hist_in, int64_t hist_len, const int32_t* subset, int32_t f_sub,
                   const int32_t* num_bins_all, int32_t C, int32_t impurity_kind, int32_t min_instances,
                   double min_info_gain, const double* node_calc, int32_t* out_k, int32_t* out_split,
                   double* out_gain, double* out_impurity, double* out_left /*C*/, double* out_right /*C*/,
                   int32_t* out_valid) {
  std::vector<double> h(hist_in, hist_in + hist_len);
  std::vector<double> parent(C, 0.0);
  {  // aggregator parentStats = per-class totals
    int nb0 = num_bins_all[subset[0]];
    for (int b = 0; b < nb0; b++) for (int c = 0; c < C; c++) parent[c] += h[(size_t)b * C + c];
  }
  double node_imp = 0.0;
  if (node_calc) node_imp = orc::impurity_calculate(impurity_kind, node_calc, C, orc::counts_sum(node_calc, C));
  orc::BestSplit bs = orc::bins_to_best_split(h.data(), subset, f_sub, num_bins_all, C, impurity_kind, min_instances,
                                              min_info_gain, node_calc != nullptr, node_imp, node_calc, parent.data());
  *out_k = bs.k; *out_split = bs.split; *out_gain = bs.gain; *out_impurity = bs.impurity; *out_valid = bs.valid;
  for (int c = 0; c < C; c++) {
    out_left[c] = bs.left.empty() ? 0.0 : bs.left[c];
    out_right[c] = bs.right.empty() ? 0.0 : bs.right[c];
  }
  return 0;
}

// ---- fit / forest / predict
void* orc_fit(const orc_params_t* p, const double* X, const double* y, int64_t N, int32_t F, const int64_t* part_off,
              int32_t P) {
  std::string err;
  orc::Forest* f = orc::fit(to_params(p), X, y, N, F, part_off, P, err);
  if (!f) g_err = err;
  return f;
}
void orc_forest_free(void* h) { delete (orc::Forest*)h; }
int64_t orc_forest_total_nodes(const void* h) { return (int64_t)((const orc::Forest*)h)->nodes.size(); }
int32_t orc_forest_num_groups(const void* h) { return (int32_t)((const orc::Forest*)h)->group_sizes.size(); }
void orc_forest_group_sizes(const void* h, int32_t* out) {
  const orc::Forest* f = (const orc::Forest*)h;
  memcpy(out, f->group_sizes.data(), f->group_sizes.size() * 4);
}
int64_t orc_forest_trace_len(const void* h) { return (int64_t)((const orc::Forest*)h)->trace.size(); }
void orc_forest_trace(const void* h, int32_t* out) {
  const orc::Forest* f = (const orc::Forest*)h;
  memcpy(out, f->trace.data(), f->trace.size() * 4);
}
int32_t orc_forest_f_sub(const void* h) { return ((const orc::Forest*)h)->f_sub; }
void orc_forest_thresholds(const void* h, int32_t stride, double* thr_out, int32_t* n_splits_out) {
  const orc::Forest* f = (const orc::Forest*)h;
  for (int j = 0; j < f->num_features; j++) {
    n_splits_out[j] = (int32_t)f->thresholds[j].size();
    for (size_t i = 0; i < f->thresholds[j].size() && (int)i < stride; i++)
      thr_out[(size_t)j * stride + i] = f->thresholds[j][i];
  }
}
void orc_forest_export(const void* h, int64_t* tree_off, int32_t* node_id, int32_t* left, int32_t* right,
                       int32_t* feature, double* threshold, double* gain, double* impurity, double* prediction,
                       double* stats) {
  const orc::Forest* f = (const orc::Forest*)h;
  for (size_t t = 0; t < f->tree_off.size(); t++) tree_off[t] = f->tree_off[t];
  for (size_t i = 0; i < f->nodes.size(); i++) {
    const orc::FlatNode& n = f->nodes[i];
    node_id[i] = n.id; left[i] = n.left; right[i] = n.right; feature[i] = n.feature;
    threshold[i] = n.threshold; gain[i] = n.gain; impurity[i] = n.impurity; prediction[i] = n.prediction;
  }
  memcpy(stats, f->stats.data(), f->stats.size() * 8);
}
void orc_predict_thresholds(const void* h, const double* X, int64_t N, const double* thresholds, double* raw, double* prob,
                            double* pred) {
  orc::predict(*(const orc::Forest*)h, X, N, raw, prob, pred, thresholds);
}
void orc_predict(const void* h, const double* X, int64_t N, double* raw, double* prob, double* pred) {
  orc::predict(*(const orc::Forest*)h, X, N, raw, prob, pred);
}

}  // extern "C"
