// snapshot.cu — vdb_save / vdb_load: binary index snapshot. The reference persists raw vectors
// only (storage/VectorStorage.java:49,185) and every native save()/load() is an empty TODO (HNSWIndex.cpp:535-541,
// HNSWPQIndex.cpp:774-780, PQIndex.cpp:299-300), so this format is this library's own:
//   "VDBSNAP1" | vdb_config | n | trained | max_level | entry | raw_rows
//   ids[n] i32 | vectors[n][dim] f32 | norms[n] f32 (cosine) | codebooks[M][256][sd] f32 + codes[n][M] u8 (if trained)
//   levels[n] i32 + adj[max_level][n][deg] i32 (graph indexes)            — little-endian, no padding.
#include <cstdio>
#include <cstring>

#include "kernels.h"
#include "vdb_index.h"

#define FAILS(code, msg) do { vdb_set_error(msg); return (code); } while (0)

extern "C" int vdb_get_graph(vdb_index* x, int32_t* levels, int32_t* adj, int32_t* entry);

namespace {
struct File {
    FILE* f = nullptr;
    ~File() { if (f) fclose(f); }
    bool w(const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
    bool r(void* p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }
};
}  // namespace

extern "C" int vdb_save(vdb_index* x, const char* path) {
    if (!x || !path) FAILS(VDB_ERR_INVALID, "vdb_save: bad arguments");
    if (x->device_only) FAILS(VDB_ERR_INVALID, "vdb_save: device-generated rows are not persisted");
    const bool graph = x->cfg.index_type == VDB_INDEX_HNSW || x->cfg.index_type == VDB_INDEX_HNSW_PQ;
    if (graph) { int rc = vdb_get_graph(x, nullptr, nullptr, nullptr); if (rc != VDB_OK) return rc; }   // forces a sync
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    File fh; fh.f = fopen(path, "wb");
    if (!fh.f) FAILS(VDB_ERR_IO, std::string("cannot open ") + path);
    const int64_t n = x->n;
    const int32_t trained = x->trained ? 1 : 0, max_level = x->g.max_level, entry = x->g.entry;
    const int64_t raw_rows = x->raw_rows;
    bool ok = fh.w("VDBSNAP1", 8) && fh.w(&x->cfg, sizeof(vdb_config)) && fh.w(&n, 8) && fh.w(&trained, 4) && fh.w(&max_level, 4) &&
              fh.w(&entry, 4) && fh.w(&raw_rows, 8);
    ok = ok && fh.w(x->h_ids.data(), (size_t)n * 4) && fh.w(x->h_vecs.data(), (size_t)n * x->cfg.dim * 4);
    if (x->cfg.metric == VDB_METRIC_COSINE) ok = ok && fh.w(x->h_norms.data(), (size_t)n * 4);
    if (trained && x->pq_m > 0) {
        std::vector<float> cb((size_t)x->pq_m * VDB_NCENT * x->sd);
        VDB_CUDA(cudaMemcpy(cb.data(), x->d_cb.p, cb.size() * 4, cudaMemcpyDeviceToHost));
        std::vector<uint8_t> codes((size_t)n * x->pq_m, 0);
        const int64_t have = std::min<int64_t>(x->n_coded, n);
        if (have > 0) VDB_CUDA(cudaMemcpy(codes.data(), x->d_codes.p, (size_t)have * x->pq_m, cudaMemcpyDeviceToHost));
        const int64_t nc = have;
        ok = ok && fh.w(cb.data(), cb.size() * 4) && fh.w(&nc, 8) && fh.w(codes.data(), codes.size());
    }
    if (graph) {
        ok = ok && fh.w(x->g.level.data(), (size_t)n * 4);
        for (int l = 0; l < max_level && ok; l++) ok = fh.w(x->g.adj[l].data(), (size_t)n * x->g.M * 4);
    }
    if (!ok) FAILS(VDB_ERR_IO, "short write");
    return VDB_OK;
}

extern "C" int vdb_load(vdb_index* x, const char* path) {
    if (!x || !path) FAILS(VDB_ERR_INVALID, "vdb_load: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    if (x->n != 0) FAILS(VDB_ERR_INVALID, "vdb_load: index must be empty");
    File fh; fh.f = fopen(path, "rb");
    if (!fh.f) FAILS(VDB_ERR_IO, std::string("cannot open ") + path);
    char magic[8]; vdb_config c; int64_t n = 0, raw_rows = 0; int32_t trained = 0, max_level = 0, entry = -1;
    if (!(fh.r(magic, 8) && fh.r(&c, sizeof(c)) && fh.r(&n, 8) && fh.r(&trained, 4) && fh.r(&max_level, 4) && fh.r(&entry, 4) && fh.r(&raw_rows, 8)))
        FAILS(VDB_ERR_IO, "short read (header)");
    if (std::memcmp(magic, "VDBSNAP1", 8) != 0) FAILS(VDB_ERR_IO, "not a vector_db_b200 snapshot");
    if (c.dim != x->cfg.dim || c.index_type != x->cfg.index_type || c.pq_m != x->cfg.pq_m || c.hnsw_m != x->cfg.hnsw_m)
        FAILS(VDB_ERR_INVALID, "snapshot was written with a different index configuration");
    if (n > x->cfg.max_elements) FAILS(VDB_ERR_FULL, "snapshot larger than max_elements");
    const bool graph = x->cfg.index_type == VDB_INDEX_HNSW || x->cfg.index_type == VDB_INDEX_HNSW_PQ;
    if (graph && max_level != x->g.max_level) FAILS(VDB_ERR_INVALID, "snapshot max_level differs (max_elements changed?)");
    x->h_ids.resize(n); x->h_vecs.resize((size_t)n * c.dim);
    bool ok = fh.r(x->h_ids.data(), (size_t)n * 4) && fh.r(x->h_vecs.data(), (size_t)n * c.dim * 4);
    if (x->cfg.metric == VDB_METRIC_COSINE) { x->h_norms.resize(n); ok = ok && fh.r(x->h_norms.data(), (size_t)n * 4); }
    x->n = n; x->n_dev = 0; x->n_trav = 0; x->n_fmirror = 0; x->n_coded = 0; x->raw_rows = raw_rows;
    if (ok && trained && x->pq_m > 0) {
        std::vector<float> cb((size_t)x->pq_m * VDB_NCENT * x->sd);
        int64_t nc = 0;
        x->h_codes_override.resize((size_t)n * x->pq_m);
        ok = fh.r(cb.data(), cb.size() * 4) && fh.r(&nc, 8) && fh.r(x->h_codes_override.data(), x->h_codes_override.size());
        if (ok) {
            VDB_TRY(x->d_cb.ensure(cb.size(), 0, x->stream));
            VDB_CUDA(cudaMemcpy(x->d_cb.p, cb.data(), cb.size() * 4, cudaMemcpyHostToDevice));
            x->trained = true;
            x->codes_overridden = (nc == n);
            if (!x->codes_overridden) x->h_codes_override.clear();
        }
    }
    if (ok && graph) {
        std::vector<int32_t> levels(n), adj((size_t)max_level * n * x->g.M);
        ok = fh.r(levels.data(), (size_t)n * 4) && fh.r(adj.data(), adj.size() * 4);
        if (ok) { x->g.import_graph(levels.data(), adj.data(), (int)n, entry); x->n_indexed = n; x->graph_uploaded = false; }
    }
    if (!ok) { x->n = 0; x->h_ids.clear(); x->h_vecs.clear(); FAILS(VDB_ERR_IO, "short read"); }
    return VDB_OK;
}
