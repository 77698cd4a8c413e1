// tests/probes/hostgraph_probe.cpp — TEST INFRASTRUCTURE: C-linkage handle around the product's host graph builder
// (vector_db_b200/csrc/host_graph.cpp) so that the CPU suite can compare it link for link with the oracle's insert / remove.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../vector_db_b200/csrc/host_graph.h"

struct Probe { HostGraph g; std::vector<float> vecs, norms; };
#define API extern "C" __attribute__((visibility("default")))
API void* hg_create(int dim, int M, int max_elements, int metric, uint64_t seed, int exact) {
    Probe* p = new Probe(); p->g.init(dim, M, max_elements, metric, seed); p->g.exact_order = exact != 0; return p;
}
API void hg_destroy(void* h) { delete (Probe*)h; }
API void hg_insert(void* h, const float* v, float norm, int ef, int variant) {
    Probe* p = (Probe*)h;
    p->vecs.insert(p->vecs.end(), v, v + p->g.dim); p->norms.push_back(norm);
    p->g.vecs = p->vecs.data(); p->g.norms = p->norms.data();
    p->g.insert(p->g.size(), ef, variant);
}
API void hg_remove(void* h, int row, int new_entry) { ((Probe*)h)->g.remove(row, new_entry); }
API int hg_size(void* h) { return ((Probe*)h)->g.size(); }
API int hg_entry(void* h) { return ((Probe*)h)->g.entry; }
API int hg_max_level(void* h) { return ((Probe*)h)->g.max_level; }
API void hg_export(void* h, int32_t* levels, int32_t* adj) {
    Probe* p = (Probe*)h; const int n = p->g.size();
    std::memcpy(levels, p->g.level.data(), (size_t)n * 4);
    for (int l = 0; l < p->g.max_level; l++) std::memcpy(adj + (size_t)l * n * p->g.M, p->g.adj[l].data(), (size_t)n * p->g.M * 4);
}
API int hg_dirty_count(void* h) { return (int)((Probe*)h)->g.dirty.size(); }
API void hg_clear_dirty(void* h) { ((Probe*)h)->g.clear_dirty(); }
