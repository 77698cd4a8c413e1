// snapshot.cu — vdb_save / vdb_load: binary index snapshot. The reference persists raw vectors
// only (storage/VectorStorage.java:49,185) and every native save()/load() is an empty TODO (HNSWIndex.cpp:535-541,
// HNSWPQIndex.cpp:774-780, PQIndex.cpp:299-300, declared at native/index/VectorIndex.h:21-25), so this format is this
// library's own. It is a SHARED format: the test checker under oracle/ (snapshot_read / snapshot_write) reads and writes the
// same bytes, which is how the CPU side and the GPU path load "the same persisted graph and codebooks" in the parity tests.
//
//   "VDBSNAP2" | u32 cfg_bytes | vdb_config | i64 n | i32 trained | i32 max_level | i32 entry | i32 any_dead | i64 raw_rows |
//   u64 level_ctr | i64 n_train | i64 java_size | i64 n_coded
//   ids[n] i32 | vectors[n][dim] f32 | norms[n] f32 (cosine metric only)
//   PQ indexes, trained   : codebooks[M][256][sd] f32 | codes[n][M] u8 (rows >= n_coded are zero)
//   PQ indexes, untrained : samples[n_train][dim] f32 (collected by addVector, HnswPqIndex.java:170-175)
//   graph indexes         : levels[n] i32 | adj[max_level][n][deg] i32 (-1 padded) | dead[n] u8 | jindex[n] i64
// little-endian, no padding. Everything read from the file is validated before it reaches a kernel.
#include <cstdio>
#include <cstring>
#include <new>

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
struct Header {
    int64_t n; int32_t trained, max_level, entry, any_dead; int64_t raw_rows; uint64_t level_ctr; int64_t n_train, java_size, n_coded;
};
}  // namespace
int vdb_validate_graph(const int32_t* levels, const int32_t* adj, int64_t n, int max_level, int M, int32_t entry);

extern "C" int vdb_save(vdb_index* x, const char* path) try {
    if (!x || !path) FAILS(VDB_ERR_INVALID, "vdb_save: bad arguments");
    if (!x->shards.empty()) FAILS(VDB_ERR_INVALID, "vdb_save: not available on a multi-device index (n_devices > 1)");
    if (x->device_only) FAILS(VDB_ERR_INVALID, "vdb_save: device-generated rows are not persisted");
    const bool graph = x->cfg.index_type == VDB_INDEX_HNSW || x->cfg.index_type == VDB_INDEX_HNSW_PQ;
    if (graph) { int rc = vdb_get_graph(x, nullptr, nullptr, nullptr); if (rc != VDB_OK) return rc; }   // forces a sync
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    File fh; fh.f = fopen(path, "wb");
    if (!fh.f) FAILS(VDB_ERR_IO, std::string("cannot open ") + path);
    const int64_t n = x->n;
    Header h{};
    h.n = n; h.trained = x->trained ? 1 : 0; h.max_level = x->g.max_level; h.entry = x->g.entry; h.any_dead = x->any_dead ? 1 : 0;
    h.raw_rows = x->raw_rows; h.level_ctr = x->g.level_ctr; h.n_train = x->trained ? 0 : x->n_train; h.java_size = x->java_size;
    h.n_coded = std::min<int64_t>(x->n_coded, n);
    const uint32_t cfg_bytes = (uint32_t)sizeof(vdb_config);
    bool ok = fh.w("VDBSNAP2", 8) && fh.w(&cfg_bytes, 4) && fh.w(&x->cfg, sizeof(vdb_config)) && fh.w(&h, sizeof h);
    ok = ok && fh.w(x->h_ids.data(), (size_t)n * 4) && fh.w(x->h_vecs.data(), (size_t)n * x->cfg.dim * 4);
    if (x->cfg.metric == VDB_METRIC_COSINE) ok = ok && fh.w(x->h_norms.data(), (size_t)n * 4);
    if (x->pq_m > 0 && x->trained) {
        std::vector<float> cb((size_t)x->pq_m * VDB_NCENT * x->sd);
        VDB_CUDA(cudaMemcpy(cb.data(), x->d_cb.p, cb.size() * 4, cudaMemcpyDeviceToHost));
        std::vector<uint8_t> codes((size_t)n * x->pq_m, 0);
        if (h.n_coded > 0) VDB_CUDA(cudaMemcpy(codes.data(), x->d_codes.p, (size_t)h.n_coded * x->pq_m, cudaMemcpyDeviceToHost));
        ok = ok && fh.w(cb.data(), cb.size() * 4) && fh.w(codes.data(), codes.size());
    } else if (x->pq_m > 0) {
        ok = ok && fh.w(x->h_train.data(), (size_t)h.n_train * x->cfg.dim * 4);
    }
    if (graph) {
        ok = ok && fh.w(x->g.level.data(), (size_t)n * 4);
        for (int l = 0; l < h.max_level && ok; l++) ok = fh.w(x->g.adj[l].data(), (size_t)n * x->g.M * 4);
        std::vector<int64_t> ji(x->h_jindex);
        ji.resize(n, 0);
        ok = ok && fh.w(x->g.dead.data(), (size_t)n) && fh.w(ji.data(), (size_t)n * 8);
    }
    if (!ok) FAILS(VDB_ERR_IO, "short write");
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_save"); }

static int load_impl(vdb_index* x, const char* path) {
    File fh; fh.f = fopen(path, "rb");
    if (!fh.f) FAILS(VDB_ERR_IO, std::string("cannot open ") + path);
    char magic[8]; uint32_t cfg_bytes = 0; vdb_config c; Header h{};
    if (!(fh.r(magic, 8) && fh.r(&cfg_bytes, 4))) FAILS(VDB_ERR_IO, "short read (header)");
    if (std::memcmp(magic, "VDBSNAP2", 8) != 0) FAILS(VDB_ERR_IO, "not a vector_db_b200 snapshot (VDBSNAP2)");
    if (cfg_bytes != sizeof(vdb_config)) FAILS(VDB_ERR_IO, "snapshot was written by a different library version (vdb_config size)");
    if (!(fh.r(&c, sizeof c) && fh.r(&h, sizeof h))) FAILS(VDB_ERR_IO, "short read (header)");
    if (c.dim != x->cfg.dim || c.index_type != x->cfg.index_type || c.metric != x->cfg.metric || c.pq_m != x->cfg.pq_m ||
        c.hnsw_m != x->cfg.hnsw_m)
        FAILS(VDB_ERR_INVALID, "snapshot was written with a different index configuration (dim / type / metric / pq_m / hnsw_m)");
    const int64_t n = h.n;
    if (n < 0 || h.n_train < 0 || h.n_train > 10000 || h.n_coded < 0 || h.n_coded > n || h.raw_rows < 0 || h.raw_rows > n ||
        h.java_size < 0 || h.java_size > n)
        FAILS(VDB_ERR_IO, "snapshot header out of range");
    if (n > x->cfg.max_elements) FAILS(VDB_ERR_FULL, "snapshot larger than max_elements");
    const bool graph = x->cfg.index_type == VDB_INDEX_HNSW || x->cfg.index_type == VDB_INDEX_HNSW_PQ;
    if (graph && h.max_level != x->g.max_level) FAILS(VDB_ERR_INVALID, "snapshot max_level differs (max_elements changed?)");
    const int dim = c.dim;
    std::vector<int32_t> ids(n); std::vector<float> vecs((size_t)n * dim), norms;
    bool ok = fh.r(ids.data(), (size_t)n * 4) && fh.r(vecs.data(), vecs.size() * 4);
    if (x->cfg.metric == VDB_METRIC_COSINE) { norms.resize(n); ok = ok && fh.r(norms.data(), (size_t)n * 4); }
    std::vector<float> cb, train; std::vector<uint8_t> codes;
    if (ok && x->pq_m > 0 && h.trained) {
        cb.resize((size_t)x->pq_m * VDB_NCENT * x->sd); codes.resize((size_t)n * x->pq_m);
        ok = fh.r(cb.data(), cb.size() * 4) && fh.r(codes.data(), codes.size());
    } else if (ok && x->pq_m > 0) {
        train.resize((size_t)h.n_train * dim);
        ok = fh.r(train.data(), train.size() * 4);
    }
    std::vector<int32_t> levels, adj; std::vector<uint8_t> dead; std::vector<int64_t> jindex;
    if (ok && graph) {
        levels.resize(n); adj.resize((size_t)h.max_level * n * x->g.M); dead.resize(n); jindex.resize(n);
        ok = fh.r(levels.data(), (size_t)n * 4) && fh.r(adj.data(), adj.size() * 4) && fh.r(dead.data(), (size_t)n) && fh.r(jindex.data(), (size_t)n * 8);
    }
    if (!ok) FAILS(VDB_ERR_IO, "short read");
    if (graph) {
        VDB_TRY(vdb_validate_graph(levels.data(), adj.data(), n, h.max_level, x->g.M, n > 0 ? h.entry : -1));
        for (int64_t i = 0; i < n; i++)
            if (dead[i] > 1 || jindex[i] < 0 || jindex[i] >= std::max<int64_t>(n, 1)) FAILS(VDB_ERR_IO, "snapshot: dead / code-row table out of range");
        if (n > 0 && h.entry >= 0 && dead[h.entry]) FAILS(VDB_ERR_IO, "snapshot: entry point is a removed node");
    }
    // ---- commit ----
    x->h_ids.swap(ids); x->h_vecs.swap(vecs); x->h_norms.swap(norms);
    x->n = n; x->n_dev = 0; x->n_trav = 0; x->n_fmirror = 0; x->n_coded = 0; x->n_codesT = 0; x->raw_rows = h.raw_rows;
    x->trained = false; x->codes_overridden = false; x->h_codes_override.clear();
    if (x->pq_m > 0 && h.trained) {
        VDB_TRY(x->d_cb.ensure(cb.size(), 0, x->stream));
        VDB_CUDA(cudaMemcpy(x->d_cb.p, cb.data(), cb.size() * 4, cudaMemcpyHostToDevice));
        x->trained = true;
        if (h.n_coded == n) { x->h_codes_override.swap(codes); x->codes_overridden = true; }
        x->h_train.clear(); x->n_train = 0;
    } else if (x->pq_m > 0) {
        x->h_train.swap(train); x->n_train = h.n_train;
    }
    if (graph) {
        x->g.import_graph(levels.data(), adj.data(), (int)n, n > 0 ? h.entry : -1);
        x->g.level_ctr = h.level_ctr;
        x->g.dead.assign(dead.begin(), dead.end());
        x->n_indexed = n; x->graph_uploaded = false;
        x->h_jindex.swap(jindex); x->java_size = h.java_size;
        x->id2row.clear(); x->n_live = 0; x->any_dead = false; x->jindex_rows.clear();
        for (int64_t r = 0; r < n; r++) {
            if (x->g.dead[r]) { x->any_dead = true; continue; }
            x->id2row[x->h_ids[r]] = r; x->n_live++;
        }
        if (x->any_dead) for (int64_t r = 0; r < n; r++) x->jindex_rows.emplace(x->h_jindex[r], r);
        x->dead_dirty = x->any_dead;
    }
    return VDB_OK;
}

extern "C" int vdb_load(vdb_index* x, const char* path) try {
    if (!x || !path) FAILS(VDB_ERR_INVALID, "vdb_load: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    if (!x->shards.empty()) FAILS(VDB_ERR_INVALID, "vdb_load: not available on a multi-device index (n_devices > 1)");
    if (x->n != 0) FAILS(VDB_ERR_INVALID, "vdb_load: index must be empty");
    try { return load_impl(x, path); }
    catch (const std::bad_alloc&) { FAILS(VDB_ERR_IO, "vdb_load: out of memory (corrupt snapshot header?)"); }
    catch (const std::exception& e) { FAILS(VDB_ERR_IO, std::string("vdb_load: ") + e.what()); }
} catch (...) { return vdb_caught("vdb_load"); }
