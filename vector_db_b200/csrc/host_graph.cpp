// host_graph.cpp — see host_graph.h. Product code (not the oracle): own implementation of the Java insert algorithm.
#include "host_graph.h"

#include <algorithm>
#include <cmath>
#include <queue>

namespace {
inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// build-time distances need no particular summation order: 8 independent accumulators vectorise well
inline float l2sq_fast(const float* a, const float* b, int d) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i = 0;
    for (; i + 8 <= d; i += 8)
        for (int j = 0; j < 8; j++) { float t = a[i + j] - b[i + j]; acc[j] += t * t; }
    float s = ((acc[0] + acc[4]) + (acc[1] + acc[5])) + ((acc[2] + acc[6]) + (acc[3] + acc[7]));
    for (; i < d; i++) { float t = a[i] - b[i]; s += t * t; }
    return s;
}
inline float dot_fast(const float* a, const float* b, int d) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i = 0;
    for (; i + 8 <= d; i += 8)
        for (int j = 0; j < 8; j++) acc[j] += a[i + j] * b[i + j];
    float s = ((acc[0] + acc[4]) + (acc[1] + acc[5])) + ((acc[2] + acc[6]) + (acc[3] + acc[7]));
    for (; i < d; i++) s += a[i] * b[i];
    return s;
}
// Java's order: sum = sum + round(diff * diff), one accumulator, no fused multiply-add (javac never contracts)
__attribute__((optimize("-ffp-contract=off"))) float l2sq_java(const float* a, const float* b, int d) {
    float s = 0.0f;
    for (int i = 0; i < d; i++) { const float t = a[i] - b[i]; const float m = t * t; s = s + m; }
    return s;
}
__attribute__((optimize("-ffp-contract=off"))) float dot_java(const float* a, const float* b, int d) {
    float s = 0.0f;
    for (int i = 0; i < d; i++) { const float m = a[i] * b[i]; s = s + m; }
    return s;
}
struct Nb { float d; int idx; };
struct NbLess { bool operator()(const Nb& a, const Nb& b) const { return a.d < b.d || (a.d == b.d && a.idx < b.idx); } };
struct NbGreater { bool operator()(const Nb& a, const Nb& b) const { return NbLess()(b, a); } };
}  // namespace

void HostGraph::init(int dim_, int M_, int max_elements, int metric_, uint64_t seed) {
    dim = dim_; M = M_; metric = metric_; level_seed = seed; level_ctr = 0;
    // maxLevel = (int)(ln(maxElements)/ln(m)) + 1   (HnswPqIndex.java:89, HnswIndex.java:49)
    max_level = (int)(std::log((double)max_elements) / std::log((double)M)) + 1;
    if (max_level < 1) max_level = 1;
    adj.assign(max_level, {}); deg.assign(max_level, {}); adjd.assign(max_level, {}); dirty_flag.assign(max_level, {});
    clear();
}
void HostGraph::clear() {
    level.clear(); dead.clear(); entry = -1;      // level_ctr is NOT reset: the seeded level stream continues across a rebuild
                                                  // (ThreadLocalRandom has no reset either, HnswPqIndex.java:752-759)
    dirty.clear(); all_dirty = true;
    for (auto& f : dirty_flag) f.clear();
    for (auto& a : adj) a.clear();
    for (auto& d : deg) d.clear();
    for (auto& d : adjd) d.clear();
}
// seeded stand-in for ThreadLocalRandom (HnswPqIndex.java:752-759): P(level >= l) = e^-l, capped at max_level-1
int HostGraph::draw_level() {
    int lv = 0;
    const double p = 1.0 / M_E;
    for (;;) {
        double u = (double)(mix64(level_seed * 0x9E3779B97F4A7C15ull + (level_ctr++)) >> 11) * (1.0 / 9007199254740992.0);
        if (u < p && lv < max_level - 1) lv++; else break;
    }
    return lv;
}

void HostGraph::import_graph(const int32_t* levels, const int32_t* adj_in, int n, int entry_) {
    clear();
    level.assign(levels, levels + n);
    dead.assign(n, 0);
    for (int l = 0; l < max_level; l++) {
        adj[l].assign(adj_in + (size_t)l * n * M, adj_in + (size_t)(l + 1) * n * M);
        deg[l].resize(n);
        adjd[l].assign((size_t)n * M, -1.0f);
        for (int i = 0; i < n; i++) {
            int c = 0;
            while (c < M && adj[l][(size_t)i * M + c] >= 0) c++;
            deg[l][i] = (uint8_t)c;
        }
    }
    entry = entry_;
}

void HostGraph::insert(int index, int ef, int variant) {
    auto dist = [&](int a, int b) -> float {
        const float* x = vecs + (size_t)a * dim; const float* y = vecs + (size_t)b * dim;
        if (exact_order) return metric == 0 ? l2sq_java(x, y, dim) : 1.0f - dot_java(x, y, dim) / (norms[a] * norms[b]);
        if (metric == 0) return l2sq_fast(x, y, dim);
        return 1.0f - dot_fast(x, y, dim) / (norms[a] * norms[b]);
    };
    const int lv = draw_level();
    level.push_back(lv);
    dead.push_back(0);
    for (int l = 0; l < max_level; l++) { adj[l].resize((size_t)(index + 1) * M, -1); adjd[l].resize((size_t)(index + 1) * M, -1.0f); deg[l].push_back(0); }
    if ((int)visit.size() <= index) visit.resize(std::max<size_t>(1024, (size_t)(index + 1) * 2), 0);
    if (entry < 0) { entry = index; return; }

    int cur = entry;
    // greedy descent above the node's level: first improving neighbour wins (HnswPqIndex.java:761-792)
    for (int l = max_level - 1; l > lv; l--) {
        float curd = dist(cur, index);
        bool changed = true;
        while (changed) {
            changed = false;
            const int c = deg[l][cur];
            for (int j = 0; j < c; j++) {
                int nb = adj[l][(size_t)cur * M + j];
                if (dead[nb]) continue;                             // `!vectors.containsKey(neighborId)` (:778)
                float d = dist(nb, index);
                if (d < curd) { curd = d; cur = nb; changed = true; break; }
            }
        }
    }
    std::vector<Nb> found, tmp;
    std::vector<int> sel;
    for (int l = std::min(lv, max_level - 1); l >= 0; l--) {
        // searchLayer with the reference's stale-furthest behaviour (HnswPqIndex.java:829-895 / HnswIndex.java:557-644)
        std::priority_queue<Nb, std::vector<Nb>, NbLess> results;
        std::priority_queue<Nb, std::vector<Nb>, NbGreater> cands;
        if (++epoch == 0) { std::fill(visit.begin(), visit.end(), 0); epoch = 1; }
        float d0 = dist(cur, index);
        results.push({d0, cur}); cands.push({d0, cur});
        visit[cur] = epoch;
        float furthest = d0;
        while (!cands.empty()) {
            Nb c = cands.top(); cands.pop();
            if (variant == 2) { if (c.d > furthest) break; }
            else if (c.d > furthest && (int)results.size() >= ef) break;   // `continue` in Java; equivalent (min-heap)
            const int cn = deg[l][c.idx];
            // pull the unvisited neighbours' rows towards the core before the distance loop walks them one by one
            for (int j = 0; j < cn; j++) {
                const int nb = adj[l][(size_t)c.idx * M + j];
                if (visit[nb] != epoch && !dead[nb]) {
                    const char* row = reinterpret_cast<const char*>(vecs + (size_t)nb * dim);
                    __builtin_prefetch(row); __builtin_prefetch(row + 64); __builtin_prefetch(row + 128); __builtin_prefetch(row + 192);
                }
            }
            for (int j = 0; j < cn; j++) {
                int nb = adj[l][(size_t)c.idx * M + j];
                if (visit[nb] == epoch || dead[nb]) continue;      // :869 visited or no longer stored
                visit[nb] = epoch;
                float d = dist(nb, index);
                if ((int)results.size() < ef || d < furthest) {
                    cands.push({d, nb}); results.push({d, nb});
                    if ((int)results.size() > ef) { results.pop(); furthest = results.top().d; }
                }
            }
        }
        found.resize(results.size());
        for (int i = (int)results.size() - 1; i >= 0; i--) { found[i] = results.top(); results.pop(); }
        // selectNeighbors: the M nearest (HnswPqIndex.java:1043-1064); `found` is already ascending
        sel.clear();
        for (int i = 0; i < (int)found.size() && i < M; i++) sel.push_back(found[i].idx);
        for (int j = 0; j < (int)sel.size(); j++) { adj[l][(size_t)index * M + j] = sel[j]; adjd[l][(size_t)index * M + j] = found[j].d; }
        deg[l][index] = (uint8_t)sel.size();
        mark_dirty(l, index);
        for (int j = 0; j < (int)sel.size(); j++) {
            const int nb = sel[j];
            mark_dirty(l, nb);
            const float dnew = found[j].d;                      // dist(nb, index) == dist(index, nb) bit for bit
            int c = deg[l][nb];
            if (c < M) { adj[l][(size_t)nb * M + c] = index; adjd[l][(size_t)nb * M + c] = dnew; deg[l][nb] = (uint8_t)(c + 1); continue; }
            // M+1 connections: keep the M nearest to nb, stable in list order (HnswPqIndex.java:512-523)
            tmp.clear();
            for (int t = 0; t < c; t++) {
                const int nn = adj[l][(size_t)nb * M + t];
                if (dead[nn]) continue;                             // `nnVector != null` (:435-438): dangling links drop out here (the
                                                                    // compressed path :514-517 has no null check and would abort the
                                                                    // insert with an NPE — not reproduced)
                float dd = adjd[l][(size_t)nb * M + t];
                if (dd < 0.0f) dd = dist(nb, nn);               // imported graph: edge distance not cached yet
                tmp.push_back({dd, nn});
            }
            tmp.push_back({dnew, index});
            // selectNeighbors (:1043-1064): at most M candidates are returned in the given order, more are stable-sorted
            if ((int)tmp.size() > M) std::stable_sort(tmp.begin(), tmp.end(), [](const Nb& a, const Nb& b) { return a.d < b.d; });
            const int keep = std::min<int>(M, (int)tmp.size());
            for (int t = 0; t < M; t++) {
                adj[l][(size_t)nb * M + t] = t < keep ? tmp[t].idx : -1;
                adjd[l][(size_t)nb * M + t] = t < keep ? tmp[t].d : -1.0f;
            }
            deg[l][nb] = (uint8_t)keep;
        }
        cur = sel.empty() ? index : sel[0];
    }
    if (lv > level[entry]) entry = index;
}

void HostGraph::remove(int row, int new_entry) {
    const int lv = level[row];
    for (int l = 0; l <= lv && l < max_level; l++) {
        const int c = deg[l][row];
        for (int j = 0; j < c; j++) {
            const int nb = adj[l][(size_t)row * M + j];
            if (nb < 0 || nb == row) continue;
            // neighborConnections.remove(Integer.valueOf(id)): first occurrence, later entries shift down
            const int cn = deg[l][nb];
            for (int t = 0; t < cn; t++) {
                if (adj[l][(size_t)nb * M + t] != row) continue;
                for (int u = t; u + 1 < cn; u++) { adj[l][(size_t)nb * M + u] = adj[l][(size_t)nb * M + u + 1]; adjd[l][(size_t)nb * M + u] = adjd[l][(size_t)nb * M + u + 1]; }
                adj[l][(size_t)nb * M + cn - 1] = -1; adjd[l][(size_t)nb * M + cn - 1] = -1.0f;
                deg[l][nb] = (uint8_t)(cn - 1);
                mark_dirty(l, nb);
                break;
            }
        }
        for (int j = 0; j < M; j++) { adj[l][(size_t)row * M + j] = -1; adjd[l][(size_t)row * M + j] = -1.0f; }   // currentGraph.remove(id)
        deg[l][row] = 0;
        mark_dirty(l, row);
    }
    dead[row] = 1;
    if (row == entry) entry = new_entry;
}
