// shard.cpp -- how the path shards over the GPUs of one box.
//
// Site patterns are independent given tree and model (every kernel is a map over patterns followed by a sum,
// tree/phylokernelnew.h:2377, 2792, 3273; the reference's own "packets", :1114-1142), so
//   * one alignment: contiguous pattern ranges, one per GPU, rounded to the kernels' 32-pattern tile; each GPU
//     keeps its slice of every partial vector; the only exchange is the sum of {lnL | df, ddf};
//   * a supermatrix: whole partitions placed on GPUs by longest-processing-time-first on the reference's cost
//     T_p * N_p * n_p (tree/phylosupertree.cpp:664-667).
#include <algorithm>
#include <numeric>
#include <vector>

#include "b200host.h"

namespace b200 {

void shardPatterns(int64_t nptn, int world, int rank, int tile, int64_t *begin, int64_t *end) {
    if (world < 1 || rank < 0 || rank >= world || tile < 1) outError("shardPatterns: bad arguments");
    const int64_t tiles = (nptn + tile - 1) / tile;
    const int64_t base = tiles / world, extra = tiles % world;  // the first `extra` ranks take one more tile
    const int64_t t0 = rank * base + std::min<int64_t>(rank, extra);
    const int64_t t1 = t0 + base + (rank < extra ? 1 : 0);
    *begin = std::min(nptn, t0 * tile);
    *end = std::min(nptn, t1 * tile);
}

void placePartitions(const double *cost, int nparts, int world, int32_t *gpu_of_part) {
    if (world < 1) outError("placePartitions: bad world size");
    std::vector<int> order(nparts);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
    std::vector<double> load(world, 0.0);
    for (int p : order) {
        int g = (int)(std::min_element(load.begin(), load.end()) - load.begin());
        gpu_of_part[p] = g;
        load[g] += cost[p];
    }
}

}  // namespace b200

extern "C" {
int b200host_shard_patterns(int64_t nptn, int world, int rank, int tile, int64_t *begin, int64_t *end) {
    try {
        b200::shardPatterns(nptn, world, rank, tile, begin, end);
        return 0;
    } catch (...) {
        return -1;
    }
}
int b200host_place_partitions(const double *cost, int nparts, int world, int32_t *gpu_of_part) {
    try {
        b200::placePartitions(cost, nparts, world, gpu_of_part);
        return 0;
    } catch (...) {
        return -1;
    }
}
}
