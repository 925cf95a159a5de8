// Exercises every routine of euromillioner_b200/csrc/host_logic.hpp (the host scheduler's parity-critical pieces) under
// AddressSanitizer + UndefinedBehaviorSanitizer; see profiles/sanitize/run.sh.  Plain host code: no CUDA needed.
#include <cstdio>
#include <vector>

#include "../../euromillioner_b200/csrc/host_logic.hpp"

namespace hl = emrf::hostlogic;

int main() {
  unsigned long long acc = 0;
  for (int64_t seed : {0ll, 42ll, 207336481ll, -1ll, (long long)0x7fffffffffffffffll}) {
    hl::JavaDraws<hl::LcgBits> jr(seed);
    hl::SparkXorShift xr(seed);
    for (int i = 0; i < 1000; i++) { acc += (unsigned long long)jr.next_long(); acc += (unsigned long long)(xr.next_double() * 1e6); }
    for (int n_items : {1, 7, 64, 128, 1000}) {
      for (int k : {1, 8, 12, 64}) {
        std::vector<int32_t> out(k);
        int len = 0;
        hl::reservoir_features(n_items, k, seed, out.data(), &len);
        for (int i = 0; i < len; i++) acc += (unsigned)out[i];
      }
    }
  }
  for (int n : {1, 2, 3, 4, 5, 9, 33, 100, 500}) {
    std::vector<int> seen(n);
    for (int i = 0; i < n; i++) seen[i] = (i * 37) % 1009;
    for (int v : hl::group_tree_visit_order(seen)) acc += (unsigned)v;
  }
  emrf_params_t p{};
  p.num_trees = 20; p.max_depth = 5; p.max_bins = 32; p.min_instances_per_node = 1; p.subsampling_rate = 1.0;
  p.max_memory_mb = 256; p.seed = 207336481;
  for (int kind = 0; kind <= 6; kind++) {
    p.feature_subset_kind = kind; p.feature_subset_value = kind == 5 ? 10 : 0.5;
    for (int F : {1, 3, 64, 70, 128, 1000}) acc += (unsigned)hl::features_per_node(p, F);
  }
  for (int64_t n_rows : {1ll, 1490ll, 25000ll, 1000000ll}) {
    for (int n_parts : {1, 7, 64}) {
      std::vector<int64_t> off(n_parts + 1);
      for (int i = 0; i <= n_parts; i++) off[i] = i * n_rows / n_parts;
      std::vector<int64_t> rows;
      hl::find_splits_rows(p, n_rows, off.data(), n_parts, &rows);
      acc += rows.size();
    }
  }
  const uint32_t counts[5] = {3, 0, 7, 1, 0};
  acc += (unsigned long long)(hl::impurity_of(0, counts, 5) * 1e9) + (unsigned long long)(hl::impurity_of(1, counts, 5) * 1e9);
  acc += hl::select_nodes_examines_next(0, 10, 10) + 2 * hl::select_nodes_examines_next(1, 10, 10);
  acc += (unsigned long long)(hl::poisson_exp_policy(1.0) * 1e9);
  // rows sharded: packed slot layout and the device-tiling bounds
  for (int world : {1, 2, 3, 8}) {
    for (int n : {1, 2, 31, 2542}) {
      std::vector<uint32_t> w(n);
      for (int i = 0; i < n; i++) w[i] = (uint32_t)((i * 2654435761u) >> (i % 3 ? 8 : 20));
      std::vector<int> slot(n), owner(n);
      acc += (unsigned)hl::packed_slot_layout(w.data(), n, world, slot.data(), owner.data());
      for (int i = 0; i < n; i++) acc += (unsigned)slot[i] + (unsigned)owner[i];
      std::vector<int32_t> tree(n);
      std::vector<uint64_t> tot(7), scratch(7, 0);
      for (int i = 0; i < n; i++) tree[i] = i % 7;
      for (int t = 0; t < 7; t++) tot[t] = 1000u * (t + 1);
      acc += hl::tiles_bound(hl::group_entries_bound(w.data(), tree.data(), n, tot.data(), scratch.data()), 256, n);
      for (uint64_t v : scratch) acc += v;   // left zeroed
    }
  }
  hl::ScalaMutableIntTable tab;
  for (int i = 0; i < 300; i++) tab.insert((i * 7919) % 1000);
  for (int v : tab.iteration_order()) acc += (unsigned)v;
  printf("host_logic probe checksum %llu\n", acc);
  return 0;
}
