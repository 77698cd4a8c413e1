// host_graph.h — host-side HNSW graph construction ("stays on CPU in the first build").
// Follows the insert semantics of the Java classes (HnswPqIndex.java:391-539, HnswIndex.java:139-248): exact distances,
// M nearest neighbours (no diversity heuristic), back-links pruned to the M nearest. The query path never runs here.
#pragma once
#include <cstdint>
#include <vector>

struct HostGraph {
    int dim = 0, M = 32, max_level = 1, metric = 0;
    uint64_t level_seed = 0, level_ctr = 0;
    int entry = -1;
    bool exact_order = false;        // build-time distances in Java's sequential fp32 order (HnswPqIndex.java:741-750, Vector.java:77-102):
                                     // the graph is then the oracle's graph link for link (strict mode); default = 8-lane sums (fast)
    const float* vecs = nullptr;     // [n][dim] (normalised rows for cosine), owned by the index
    const float* norms = nullptr;    // [n] for cosine
    std::vector<int32_t> level;      // [n]
    std::vector<std::vector<int32_t>> adj;   // [max_level] flat [n][M], -1 padded
    std::vector<std::vector<uint8_t>> deg;   // [max_level][n]
    std::vector<std::vector<float>> adjd;    // [max_level] flat [n][M]: distance of each stored edge (< 0 = not known yet);
                                             // distances are symmetric bit for bit, so back-link pruning re-uses them
    std::vector<uint8_t> dead;       // [n] 1 = removed (removeVector, HnswPqIndex.java:542-596): skipped like `!vectors.containsKey`
    // rows whose adjacency changed since the last upload: (level << 32 | row), deduplicated through dirty_flag
    std::vector<uint64_t> dirty;
    std::vector<std::vector<uint8_t>> dirty_flag;   // [max_level][n]
    bool all_dirty = true;           // everything must be uploaded (after clear / import / bulk build)
    // scratch
    std::vector<uint32_t> visit; uint32_t epoch = 0;

    void init(int dim_, int M_, int max_elements, int metric_, uint64_t seed);
    void clear();
    int size() const { return (int)level.size(); }
    int draw_level();
    // insert row `index` (must equal size()); ef = construction beam; variant 1 = HnswPqIndex.searchLayer, 2 = HnswIndex.searchLayer
    void insert(int index, int ef, int variant);
    // import an externally built graph: levels[n], adj_in[max_level][n][M]
    void import_graph(const int32_t* levels, const int32_t* adj_in, int n, int entry_);
    // removeVector (HnswPqIndex.java:542-596 / HnswIndex.java:254-317): drop the node's own lists, delete it from the lists of
    // the nodes IT links to (links pointing at it from elsewhere stay and are skipped by every walk), mark it dead.
    // new_entry = the replacement entry point the caller computed (first live key in ConcurrentHashMap order), used when
    // `row` is the entry point; -1 when no live node is left.
    void remove(int row, int new_entry);
    void mark_dirty(int l, int row) {
        if (all_dirty) return;
        if ((int)dirty_flag[l].size() <= row) dirty_flag[l].resize(level.size(), 0);
        if (!dirty_flag[l][row]) { dirty_flag[l][row] = 1; dirty.push_back(((uint64_t)l << 32) | (uint32_t)row); }
    }
    void clear_dirty() {
        for (uint64_t d : dirty) dirty_flag[(int)(d >> 32)][(uint32_t)d] = 0;
        dirty.clear(); all_dirty = false;
    }
};
