// Development aid (tools/walk_stats.py): the histogram sink behind RRT_STAT in csrc/rrt_trace.cuh when the traversal
// headers are compiled for the host with -DRRT_WALK_STATS.  Not product, not a test.
#include <atomic>
#include <cstdint>

namespace {
constexpr int kKinds = 8, kBins = 64;
std::atomic<uint64_t> g_hist[kKinds][kBins];
}

namespace rrt {
void rrt_walk_stat(int kind, unsigned value) {
    if (kind < 0 || kind >= kKinds) return;
    g_hist[kind][value < (unsigned)kBins ? value : (unsigned)kBins - 1].fetch_add(1, std::memory_order_relaxed);
}
}  // namespace rrt

extern "C" {
void walk_stats_reset() {
    for (auto& k : g_hist) for (auto& b : k) b.store(0);
}
void walk_stats_get(uint64_t* out) {   // [kKinds][kBins]
    for (int k = 0; k < kKinds; ++k) for (int b = 0; b < kBins; ++b) out[k * kBins + b] = g_hist[k][b].load();
}
}
