// vdb_api.cu — C ABI (include/vdb.h) over the CUDA kernels. Host orchestration only: every distance, scan, traversal,
// top-k, encode and k-means step of the query / rebuild path runs in a kernel of this library (no CPU fallback).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include "kernels.h"
#include "vdb_index.h"

int vdb_validate_graph(const int32_t* levels, const int32_t* adj, int64_t n, int max_level, int M, int32_t entry);

static thread_local std::string g_err;
void vdb_set_error(const std::string& msg) { g_err = msg; }
#define FAIL(code, msg) do { vdb_set_error(msg); return (code); } while (0)

// Every extern "C" entry point is a function-try-block ending in `catch (...) { return vdb_caught(name); }`: a C++ exception
// (std::bad_alloc from a host vector, anything else) becomes an error code + message, never unwinds into the caller's frames
// (a JNI frame, ctypes). Called from inside a catch handler only.
int vdb_caught(const char* fn) {
    try { throw; }
    catch (const std::bad_alloc&) { vdb_set_error(std::string(fn) + ": out of host memory"); }
    catch (const std::exception& e) { vdb_set_error(std::string(fn) + ": " + e.what()); }
    catch (...) { vdb_set_error(std::string(fn) + ": unknown C++ exception"); }
    return VDB_ERR_INVALID;
}

extern "C" const char* vdb_last_error(void) { return g_err.c_str(); }
extern "C" const char* vdb_version(void) { return "vector_db_b200 0.1 (sm_100a)"; }

static int recommended_pq_m(int dim) {   // CompressionConfig.recommendedConfig (CompressionConfig.java:114-130)
    int m = std::max(8, dim / 8);
    while (dim % m != 0 && m > 1) m--;
    return m;
}

extern "C" int vdb_default_config(int32_t index_type, int32_t dim, int32_t max_elements, vdb_config* c) try {
    if (!c || dim <= 0 || max_elements <= 0) FAIL(VDB_ERR_INVALID, "vdb_default_config: bad arguments");
    std::memset(c, 0, sizeof(*c));
    c->dim = dim; c->max_elements = max_elements; c->index_type = index_type;
    c->hnsw_m = 32; c->pq_bits = 8; c->pq_iters = 25; c->level_seed = 1234;
    switch (index_type) {
        case VDB_INDEX_FLAT: c->metric = VDB_METRIC_SQL2; c->arith = VDB_ARITH_JAVA_SEQ; break;
        case VDB_INDEX_PQ:   // PQIndex.h:10-14 defaults M=8; callers normally pass M explicitly
            c->metric = VDB_METRIC_SQL2; c->arith = VDB_ARITH_NATIVE_AVX2; c->pq_m = 8; c->train_flavour = VDB_TRAIN_NATIVE_PQ; break;
        case VDB_INDEX_HNSW:   // HnswIndex.java:21-26
            c->metric = VDB_METRIC_COSINE; c->arith = VDB_ARITH_JAVA_SEQ; c->ef_construction = 400; c->ef_search = 400; break;
        case VDB_INDEX_HNSW_PQ:   // HnswPqIndex.java:27-32
            c->metric = VDB_METRIC_SQL2; c->arith = VDB_ARITH_JAVA_SEQ; c->ef_construction = 64; c->ef_search = 64;
            c->pq_m = recommended_pq_m(dim); c->train_flavour = VDB_TRAIN_JAVA; break;
        default: FAIL(VDB_ERR_INVALID, "unknown index_type");
    }
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_default_config"); }

extern "C" int vdb_create(const vdb_config* cfg, vdb_index** out) try {
    if (!cfg || !out) FAIL(VDB_ERR_INVALID, "vdb_create: null argument");
    *out = nullptr;
    if (cfg->dim <= 0 || cfg->max_elements <= 0) FAIL(VDB_ERR_INVALID, "dim and max_elements must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) FAIL(VDB_ERR_CUDA, "no CUDA device: libvectordb has no CPU fallback");
    if (cfg->device < 0 || cfg->device >= ndev) FAIL(VDB_ERR_INVALID, "bad device ordinal");
    cudaDeviceProp prop;
    VDB_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10) FAIL(VDB_ERR_CUDA, "libvectordb is built for sm_100a (B200) only");
    VDB_CUDA(cudaSetDevice(cfg->device));
    vdb_index* x = new vdb_index();
    x->cfg = *cfg;
    x->num_sms = prop.multiProcessorCount;
    vdb_config& c = x->cfg;
    if (c.hnsw_m <= 0) c.hnsw_m = 32;
    if (c.pq_iters <= 0) c.pq_iters = 25;
    if (c.pq_bits == 0) c.pq_bits = 8;
    const bool uses_pq = c.index_type == VDB_INDEX_PQ || c.index_type == VDB_INDEX_HNSW_PQ;
    if (uses_pq) {
        if (c.pq_bits != 8) { delete x; FAIL(VDB_ERR_INVALID, "only pq_bits == 8 is supported"); }
        if (c.pq_m <= 0) c.pq_m = recommended_pq_m(c.dim);
        if (c.dim % c.pq_m != 0) {
            if (c.index_type == VDB_INDEX_PQ) { delete x; FAIL(VDB_ERR_INVALID, "Dimension must be divisible by M"); }  // PQIndex.cpp:19-21
            c.pq_m = recommended_pq_m(c.dim);
        }
        x->pq_m = c.pq_m; x->sd = c.dim / c.pq_m;
        if (x->sd > 32) { delete x; FAIL(VDB_ERR_INVALID, "sub-dimension > 32 not supported"); }
    }
    if (c.index_type == VDB_INDEX_HNSW) c.metric = VDB_METRIC_COSINE;
    if (uses_pq) c.metric = VDB_METRIC_SQL2;
    if (c.index_type == VDB_INDEX_HNSW || c.index_type == VDB_INDEX_HNSW_PQ) {
        if (c.hnsw_m > 64) { delete x; FAIL(VDB_ERR_INVALID, "hnsw_m > 64 not supported"); }
        if ((c.trav_store == VDB_TRAV_FP8 || c.trav_store == VDB_TRAV_INT8) && (c.dim % 128) != 0) { delete x; FAIL(VDB_ERR_INVALID, "fp8 / int8 traversal rows need dim % 128 == 0"); }
        if (c.ef_construction <= 0) c.ef_construction = c.index_type == VDB_INDEX_HNSW ? 400 : 64;
        if (c.ef_search <= 0) c.ef_search = c.index_type == VDB_INDEX_HNSW ? 400 : 64;
        x->g.init(c.dim, c.hnsw_m, c.max_elements, c.metric, c.level_seed);
        x->g.exact_order = c.trav_store == VDB_TRAV_F32 && c.trav_batch == 1;      // strict mode: bit parity for the graph too
    }
    if (cudaStreamCreateWithFlags(&x->stream, cudaStreamNonBlocking) != cudaSuccess) { delete x; FAIL(VDB_ERR_CUDA, "stream create failed"); }
    cudaEventCreate(&x->ev_total0); cudaEventCreate(&x->ev_total1); cudaEventCreate(&x->ev_k0); cudaEventCreate(&x->ev_k1);
    cudaEventCreateWithFlags(&x->ev_done, cudaEventDisableTiming);
    if (x->d_counters.ensure(24, 0, x->stream) != VDB_OK) { delete x; return VDB_ERR_CUDA; }
    cudaMemsetAsync(x->d_counters.p, 0, 24 * sizeof(unsigned long long), x->stream);
    if (c.n_devices > 1) {
        const int rc = vdb_multi_create(x);
        if (rc != VDB_OK) { const std::string keep = g_err; vdb_destroy(x); g_err = keep; return rc; }
    }
    *out = x;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_create"); }
#define MULTI(x) (!(x)->shards.empty())
#define NOT_MULTI(x, what) do { if (MULTI(x)) FAIL(VDB_ERR_INVALID, what ": not available on a multi-device index (n_devices > 1)"); } while (0)

extern "C" int vdb_destroy(vdb_index* x) try {
    if (!x) return VDB_OK;
    if (MULTI(x)) vdb_multi_destroy(x);
    cudaSetDevice(x->cfg.device);
    if (x->stream) cudaStreamSynchronize(x->stream);
    x->d_base.release(); x->d_norms.release(); x->d_cb.release(); x->d_ids.release(); x->d_adj.release(); x->d_codes.release();
    x->d_q.release(); x->d_out_d.release(); x->d_lut.release(); x->d_out_ids.release(); x->d_keys.release(); x->d_keys_tmp.release();
    x->d_rows.release(); x->d_rows_tmp.release(); x->d_visited.release(); x->d_counters.release(); x->d_trav.release(); x->d_tnorm.release(); x->d_tscale.release(); x->d_codesT.release(); x->d_lutg.release(); x->d_flag.release();
    x->d_fmirror.release(); x->d_fab.release(); x->d_fcolsum.release(); x->d_fq.release(); x->d_fqf.release(); x->d_fscores.release(); x->d_ftau.release();
    x->d_fqstat.release(); x->d_fcnt.release(); x->d_fcand.release(); x->d_fslab_d.release(); x->d_fslab_i.release();
    x->d_train.release(); x->d_sub.release(); x->d_mind.release(); x->d_asg.release();
    x->d_dead.release(); x->d_adj_payload.release(); x->d_adj_targets.release();
    x->p_q.release(); x->p_d.release(); x->p_ids.release();
    if (x->ev_done) cudaEventDestroy(x->ev_done);
    if (x->ev_total0) { cudaEventDestroy(x->ev_total0); cudaEventDestroy(x->ev_total1); cudaEventDestroy(x->ev_k0); cudaEventDestroy(x->ev_k1); }
    if (x->stream) cudaStreamDestroy(x->stream);
    delete x;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_destroy"); }

static inline bool is_graph(const vdb_index* x) { return x->cfg.index_type == VDB_INDEX_HNSW || x->cfg.index_type == VDB_INDEX_HNSW_PQ; }
extern "C" int64_t vdb_size(const vdb_index* x) { return x ? (is_graph(x) ? x->n_live : x->n) : -1; }   // Java: vectors.size()
extern "C" int64_t vdb_rows(const vdb_index* x) { return x ? x->n : -1; }
extern "C" int vdb_is_trained(const vdb_index* x) { return x && x->trained ? 1 : 0; }
extern "C" int32_t vdb_max_level(const vdb_index* x) { return x ? x->g.max_level : 0; }

// ---- device store upkeep --------------------------------------------------------------------------------------
static int upload_rows(vdb_index* x) {
    // copy host rows [n_dev, n) to the device (cosine rows were normalised on the device at add time)
    if (x->device_only || x->n_dev >= x->n) return VDB_OK;
    const int dim = x->cfg.dim;
    VDB_TRY(x->d_base.ensure((size_t)std::max<int64_t>(x->n, x->n_dev) * dim, (size_t)x->n_dev * dim, x->stream));
    VDB_TRY(x->d_ids.ensure((size_t)x->n, (size_t)x->n_dev, x->stream));
    const int64_t cnt = x->n - x->n_dev;
    VDB_CUDA(cudaMemcpyAsync(x->d_base.p + (size_t)x->n_dev * dim, x->h_vecs.data() + (size_t)x->n_dev * dim, (size_t)cnt * dim * 4,
                             cudaMemcpyHostToDevice, x->stream));
    VDB_CUDA(cudaMemcpyAsync(x->d_ids.p + x->n_dev, x->h_ids.data() + x->n_dev, (size_t)cnt * 4, cudaMemcpyHostToDevice, x->stream));
    if (x->cfg.metric == VDB_METRIC_COSINE) {
        VDB_TRY(x->d_norms.ensure((size_t)x->n, (size_t)x->n_dev, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->d_norms.p + x->n_dev, x->h_norms.data() + x->n_dev, (size_t)cnt * 4, cudaMemcpyHostToDevice, x->stream));
    }
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->n_dev = x->n;
    return VDB_OK;
}

// traversal copy of rows [n_trav, n_dev): zero-padded to 8 elements, bf16 (default) or fp32 (strict mode)
static inline int trav_dpad(const vdb_index* x) { return (x->cfg.dim + 7) & ~7; }
static inline int trav_mode(const vdb_index* x) { return x->trav_eff; }      // kernel tstore: 0 fp32, 1 bf16, 2 fp8, 3 int8
static inline int trav_stride(const vdb_index* x) {
    return (x->trav_eff == 2 || x->trav_eff == 3) ? x->cfg.dim : trav_dpad(x) * (x->trav_eff == 0 ? 4 : 2);    // 8-bit rows: dim % 128 == 0
}
// VDB_TRAV_AUTO: int8 rows need every row's quantisation step (max|x| / 127) to be small against its typical element: rows whose
// largest element exceeds 6 x rms would spend the 8 bits on the outlier. Checked on the rows about to be converted.
static int pick_trav_format(vdb_index* x, int64_t row0, int64_t cnt, bool* changed) {
    *changed = false;
    int want;
    switch (x->cfg.trav_store) {
        case VDB_TRAV_F32: want = 0; break;
        case VDB_TRAV_FP8: want = 2; break;
        case VDB_TRAV_INT8: want = 3; break;
        case VDB_TRAV_BF16: want = 1; break;
        default: {
            want = 1;
            if ((x->cfg.dim % 128) == 0 && !getenv("VDB_TRAV_NO_INT8") && (row0 == 0 || x->trav_eff == 3)) {
                VDB_CUDA(cudaMemsetAsync(x->d_counters.p + 16, 0, sizeof(unsigned long long), x->stream));
                launch_count_outlier_rows(x->d_base.p + (size_t)row0 * x->cfg.dim, cnt, x->cfg.dim, 6.0f, x->d_counters.p + 16, x->stream);
                unsigned long long bad = 0;
                VDB_CUDA(cudaMemcpyAsync(&bad, x->d_counters.p + 16, sizeof bad, cudaMemcpyDeviceToHost, x->stream));
                VDB_CUDA(cudaStreamSynchronize(x->stream));
                if (bad * 1000ull <= (unsigned long long)cnt) want = 3;
            }
        }
    }
    if (row0 == 0 || want != x->trav_eff) *changed = want != x->trav_eff;
    x->trav_eff = want;
    return VDB_OK;
}
static int build_traversal_rows(vdb_index* x) {
    if (x->n_trav >= x->n_dev) return VDB_OK;
    bool changed = false;
    VDB_TRY(pick_trav_format(x, x->n_trav, x->n_dev - x->n_trav, &changed));
    if (changed) x->n_trav = 0;                    // newly added rows fail the int8 check (or first build): convert everything
    const size_t stride = (size_t)trav_stride(x);
    VDB_TRY(x->d_trav.ensure((size_t)x->n_dev * stride, (size_t)x->n_trav * stride, x->stream));
    VDB_TRY(x->d_tnorm.ensure((size_t)x->n_dev, (size_t)x->n_trav, x->stream));
    VDB_TRY(x->d_tscale.ensure((size_t)x->n_dev, (size_t)x->n_trav, x->stream));
    launch_make_traversal_rows(x->d_base.p + (size_t)x->n_trav * x->cfg.dim, x->n_dev - x->n_trav, x->cfg.dim, trav_dpad(x),
                               trav_mode(x), x->d_trav.p + (size_t)x->n_trav * stride,
                               x->d_tnorm.p + x->n_trav, x->d_tscale.p + x->n_trav, x->stream);
    VDB_CUDA(cudaGetLastError());
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->n_trav = x->n_dev;
    return VDB_OK;
}

int vdb_upload_rows_and_traversal_copy(vdb_index* x) {
    VDB_TRY(upload_rows(x));
    return build_traversal_rows(x);
}

static int train_on_device(vdb_index* x, const float* h_samples, int64_t ns) {
    if (ns <= 0) FAIL(VDB_ERR_INVALID, "Invalid training samples");   // PQIndex.cpp:30-32
    const int dim = x->cfg.dim, M = x->pq_m;
    VDB_TRY(x->d_train.ensure((size_t)ns * dim, 0, x->stream));
    VDB_TRY(x->d_sub.ensure((size_t)ns * dim, 0, x->stream));
    VDB_TRY(x->d_mind.ensure((size_t)ns * M, 0, x->stream));
    VDB_TRY(x->d_asg.ensure((size_t)ns * M, 0, x->stream));
    VDB_TRY(x->d_cb.ensure((size_t)M * VDB_NCENT * x->sd, 0, x->stream));
    VDB_CUDA(cudaMemcpyAsync(x->d_train.p, h_samples, (size_t)ns * dim * 4, cudaMemcpyHostToDevice, x->stream));
    if (launch_kmeans_train(x->d_train.p, (int)ns, dim, M, x->cfg.pq_iters, x->cfg.train_flavour, x->d_cb.p, x->d_sub.p, x->d_mind.p,
                            x->d_asg.p, x->stream) != 0)
        FAIL(VDB_ERR_INVALID, "k-means launch rejected");
    VDB_CUDA(cudaGetLastError());
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->trained = true;
    return VDB_OK;
}

// encode rows [n_coded, n_dev) on the device; rows below raw_rows keep all-zero codes (HnswPqIndex.java:104)
static int encode_pending(vdb_index* x) {
    if (!x->trained || x->n_coded >= x->n_dev) return VDB_OK;
    const int M = x->pq_m, dim = x->cfg.dim;
    VDB_TRY(x->d_codes.ensure((size_t)x->n_dev * M, (size_t)x->n_coded * M, x->stream));
    int64_t from = x->n_coded;
    if (from < x->raw_rows) {
        const int64_t upto = std::min(x->raw_rows, x->n_dev);
        VDB_CUDA(cudaMemsetAsync(x->d_codes.p + (size_t)from * M, 0, (size_t)(upto - from) * M, x->stream));
        from = upto;
    }
    if (from < x->n_dev) {
        if (launch_pq_encode(x->d_base.p + (size_t)from * dim, x->n_dev - from, dim, M, x->d_cb.p, x->cfg.arith, x->d_codes.p + (size_t)from * M,
                             x->stream) != 0)
            FAIL(VDB_ERR_INVALID, "encode launch rejected");
        VDB_CUDA(cudaGetLastError());
    }
    if (x->codes_overridden && (int64_t)x->h_codes_override.size() == x->n_dev * M) {
        VDB_CUDA(cudaMemcpyAsync(x->d_codes.p, x->h_codes_override.data(), x->h_codes_override.size(), cudaMemcpyHostToDevice, x->stream));
        x->n_codesT = 0;
    }
    if (x->cfg.index_type == VDB_INDEX_PQ) {
        // scan layout: 16-row blocks, subspace-major (adc_fast.cu); blocks that were complete stay, the rest is rewritten
        if (x->n_codesT > x->n_coded) x->n_codesT = x->n_coded;
        const int64_t row0 = x->n_codesT & ~(int64_t)15;
        VDB_TRY(x->d_codesT.ensure((size_t)((x->n_dev + 15) / 16) * adc_mpad(M) * 16, (size_t)(row0 / 16) * adc_mpad(M) * 16, x->stream));
        launch_transpose_codes(x->d_codes.p, x->n_dev, M, row0, x->d_codesT.p, x->stream);
        VDB_CUDA(cudaGetLastError());
        x->n_codesT = x->n_dev;
    }
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->n_coded = x->n_dev;
    return VDB_OK;
}

static int upload_graph(vdb_index* x) {
    const int64_t n = x->g.size();
    const int L = x->g.max_level, M = x->g.M;
    const bool grow = n > x->adj_cap;
    if (grow || x->g.all_dirty || (int64_t)x->g.dirty.size() * 8 > n) {
        // full upload; capacity grows geometrically so the level offsets stay put across incremental adds
        if (grow) x->adj_cap = std::min<int64_t>(std::max<int64_t>(x->cfg.max_elements, n), std::max<int64_t>(1024, n + n / 2));
        VDB_TRY(x->d_adj.ensure((size_t)L * x->adj_cap * M, 0, x->stream));
        for (int l = 0; l < L; l++)
            VDB_CUDA(cudaMemcpyAsync(x->d_adj.p + (size_t)l * x->adj_cap * M, x->g.adj[l].data(), (size_t)n * M * 4, cudaMemcpyHostToDevice, x->stream));
    } else if (!x->g.dirty.empty()) {
        // only the rows an insert / remove touched (the new node's lists + <= M back-links per level): one payload, one scatter
        const size_t cnt = x->g.dirty.size();
        std::vector<int32_t> payload(cnt * M);
        for (size_t i = 0; i < cnt; i++) {
            const int l = (int)(x->g.dirty[i] >> 32); const uint32_t row = (uint32_t)x->g.dirty[i];
            std::memcpy(&payload[i * M], &x->g.adj[l][(size_t)row * M], (size_t)M * 4);
        }
        VDB_TRY(x->d_adj_payload.ensure(cnt * M, 0, x->stream));
        VDB_TRY(x->d_adj_targets.ensure(cnt, 0, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->d_adj_payload.p, payload.data(), payload.size() * 4, cudaMemcpyHostToDevice, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->d_adj_targets.p, x->g.dirty.data(), cnt * 8, cudaMemcpyHostToDevice, x->stream));
        launch_scatter_adj_rows(x->d_adj.p, x->d_adj_payload.p, x->d_adj_targets.p, (int)cnt, M, x->adj_cap, x->stream);
        VDB_CUDA(cudaGetLastError());
    }
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->g.clear_dirty();
    x->graph_uploaded = true;
    return VDB_OK;
}
static int upload_dead(vdb_index* x) {
    if (!x->dead_dirty) return VDB_OK;
    const int64_t n = x->g.size();
    std::vector<uint32_t> w((size_t)(n + 31) / 32, 0u);
    for (int64_t i = 0; i < n; i++) if (x->g.dead[i]) w[i >> 5] |= 1u << (i & 31);
    VDB_TRY(x->d_dead.ensure(w.size(), 0, x->stream));
    VDB_CUDA(cudaMemcpyAsync(x->d_dead.p, w.data(), w.size() * 4, cudaMemcpyHostToDevice, x->stream));
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->dead_dirty = false;
    return VDB_OK;
}
// rows that share a Java code-row index with a newly inserted (trained-phase) row get that row's code (vdb_index.h: h_jindex)
static int propagate_shared_codes(vdb_index* x, int64_t from_row) {
    if (!x->any_dead || !x->trained) return VDB_OK;
    const int M = x->pq_m;
    for (int64_t r = from_row; r < x->n; r++) {
        const int64_t j = x->h_jindex[r];
        if (r >= x->raw_rows) {
            auto range = x->jindex_rows.equal_range(j);
            for (auto it = range.first; it != range.second; ++it)
                if (it->second != r)
                    VDB_CUDA(cudaMemcpyAsync(x->d_codes.p + (size_t)it->second * M, x->d_codes.p + (size_t)r * M, (size_t)M, cudaMemcpyDeviceToDevice, x->stream));
        }
        x->jindex_rows.emplace(j, r);
    }
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    return VDB_OK;
}

// bring the searchable state up to date with everything added so far (Java indexes are searchable after every addVector)
static int sync_index(vdb_index* x) {
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    const int t = x->cfg.index_type;
    if (t == VDB_INDEX_HNSW || t == VDB_INDEX_HNSW_PQ) {
        if (x->device_only) FAIL(VDB_ERR_INVALID, "graph indexes need host rows (vdb_fill_synthetic is for FLAT/PQ)");
        const int64_t first_new = x->n_indexed;
        if (x->n_indexed < x->n) {
            x->g.vecs = x->h_vecs.data(); x->g.norms = x->h_norms.data();
            const int dim = x->cfg.dim;
            for (int64_t i = x->n_indexed; i < x->n; i++) {
                if ((int64_t)x->h_jindex.size() <= i) x->h_jindex.push_back(x->java_size++);     // index = currentSize++ (:393,:467)
                int ef = x->cfg.ef_construction, variant = 2;
                if (t == VDB_INDEX_HNSW_PQ) {
                    variant = 1;
                    // addVector: collect samples until trained; train at min(10000, maxElements/10) (HnswPqIndex.java:170-175)
                    if (!x->trained && x->n_train < 10000) {
                        x->h_train.insert(x->h_train.end(), x->h_vecs.begin() + i * dim, x->h_vecs.begin() + (i + 1) * dim);
                        x->n_train++;
                        if (x->n_train >= std::min<int64_t>(10000, x->cfg.max_elements / 10)) {
                            VDB_TRY(train_on_device(x, x->h_train.data(), x->n_train));
                            x->raw_rows = i;   // rows before this one were added raw
                        }
                    }
                    if (!x->trained) x->raw_rows = i + 1;
                    ef = x->trained ? std::max(x->cfg.ef_construction, 200) : x->cfg.ef_construction;   // :498 / :417
                }
                x->g.insert((int)i, ef, variant);
            }
            x->n_indexed = x->n;
            x->graph_uploaded = false;
        }
        VDB_TRY(upload_rows(x));
        VDB_TRY(build_traversal_rows(x));
        if (t == VDB_INDEX_HNSW_PQ) { VDB_TRY(encode_pending(x)); VDB_TRY(propagate_shared_codes(x, first_new)); }
        if (!x->graph_uploaded) VDB_TRY(upload_graph(x));
        VDB_TRY(upload_dead(x));
    } else {
        VDB_TRY(upload_rows(x));
        if (t == VDB_INDEX_PQ) VDB_TRY(encode_pending(x));
    }
    return VDB_OK;
}

// ---- add / train / build ---------------------------------------------------------------------------------------
extern "C" int vdb_add_batch(vdb_index* x, const int32_t* ids, const float* vecs, int64_t cnt) try {
    if (!x || (!vecs && cnt > 0) || cnt < 0) FAIL(VDB_ERR_INVALID, "vdb_add_batch: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    if (x->device_only) FAIL(VDB_ERR_INVALID, "index holds device-generated rows; host adds are not allowed");
    if (x->cfg.index_type == VDB_INDEX_PQ && !x->trained)
        FAIL(VDB_ERR_UNTRAINED, "PQ index must be trained before adding vectors");    // PQIndex.cpp:124-126
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    const int dim = x->cfg.dim;
    std::vector<float> kept_v; std::vector<int32_t> kept_i;
    if (is_graph(x)) {
        // addVector answers false for an id that is already stored (HnswPqIndex.java:157-159, HnswIndex.java:144-146): skipped
        std::vector<int64_t> keep;
        for (int64_t i = 0; i < cnt; i++) {
            const int32_t id = ids ? ids[i] : (int32_t)(x->cfg.id_offset + x->n + i);
            if (x->id2row.count(id)) continue;
            x->id2row[id] = x->n + (int64_t)keep.size();
            keep.push_back(i);
        }
        if (x->n_live + (int64_t)keep.size() > x->cfg.max_elements) {      // `vectors.size() >= maxElements` (:161-163)
            for (int64_t i : keep) x->id2row.erase(ids ? ids[i] : (int32_t)(x->cfg.id_offset + x->n + i));
            FAIL(VDB_ERR_FULL, "max_elements exceeded");
        }
        if ((int64_t)keep.size() != cnt || !ids) {
            kept_v.resize(keep.size() * (size_t)dim); kept_i.resize(keep.size());
            for (size_t t = 0; t < keep.size(); t++) {
                std::memcpy(&kept_v[t * dim], vecs + (size_t)keep[t] * dim, (size_t)dim * 4);
                kept_i[t] = ids ? ids[keep[t]] : (int32_t)(x->cfg.id_offset + x->n + keep[t]);
            }
            vecs = kept_v.data(); ids = kept_i.data(); cnt = (int64_t)keep.size();
        }
        x->n_live += cnt;
    } else if (x->n + cnt > x->cfg.max_elements) FAIL(VDB_ERR_FULL, "max_elements exceeded");   // VectorStore.cpp:40-43
    const size_t old = x->h_vecs.size();
    x->h_vecs.insert(x->h_vecs.end(), vecs, vecs + (size_t)cnt * dim);
    for (int64_t i = 0; i < cnt; i++) x->h_ids.push_back(ids ? ids[i] : (int32_t)(x->cfg.id_offset + x->n + i));
    if (x->cfg.metric == VDB_METRIC_COSINE && cnt > 0 && !MULTI(x)) {       // (a multi-device front hands raw rows to its shards)
        // normalise on insert (HnswIndex.java:154) with Java's arithmetic, on the device; keep the normalised copy on the host
        VDB_TRY(x->d_q.ensure((size_t)cnt * dim, 0, x->stream));
        VDB_TRY(x->d_out_d.ensure((size_t)cnt, 0, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->d_q.p, vecs, (size_t)cnt * dim * 4, cudaMemcpyHostToDevice, x->stream));
        launch_normalize_rows(x->d_q.p, cnt, dim, x->d_out_d.p, x->stream);
        x->h_norms.resize(x->n + cnt);
        VDB_CUDA(cudaMemcpyAsync(x->h_vecs.data() + old, x->d_q.p, (size_t)cnt * dim * 4, cudaMemcpyDeviceToHost, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->h_norms.data() + x->n, x->d_out_d.p, (size_t)cnt * 4, cudaMemcpyDeviceToHost, x->stream));
        VDB_CUDA(cudaStreamSynchronize(x->stream));
    }
    x->n += cnt;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_add_batch"); }

// first live key of Java's `vectors` map in iteration order, skipping `skip_id` (HnswPqIndex.java:577-584). The map is a
// ConcurrentHashMap<>(maxElements): table = tableSizeFor(1.5 * maxElements + 1) bins and never resizes (<= maxElements entries
// stay below the 0.75 load factor); bin = spread(hash) & (table - 1) with spread(h) = (h ^ (h >>> 16)) & 0x7fffffff and
// Integer.hashCode() = the value; bins are walked in ascending order, entries of one bin in insertion order.
static int64_t java_first_live_row(const vdb_index* x, int32_t skip_id) {
    uint64_t want = (uint64_t)x->cfg.max_elements + ((uint64_t)x->cfg.max_elements >> 1) + 1, table = 1;
    while (table < want && table < (1ull << 30)) table <<= 1;
    int64_t best_row = -1; uint64_t best_bin = ~0ull;
    for (const auto& kv : x->id2row) {
        if (kv.first == skip_id) continue;
        const uint32_t h = (uint32_t)kv.first;
        const uint64_t bin = ((h ^ (h >> 16)) & 0x7fffffffu) & (table - 1);
        if (bin < best_bin || (bin == best_bin && kv.second < best_row)) { best_bin = bin; best_row = kv.second; }
    }
    return best_row;
}

extern "C" int vdb_contains(vdb_index* x, int32_t id) try {
    if (!x) return 0;
    std::lock_guard<std::mutex> lk(x->mu);
    if (is_graph(x)) return x->id2row.count(id) ? 1 : 0;
    for (int64_t i = 0; i < (int64_t)x->h_ids.size(); i++) if (x->h_ids[i] == id) return 1;
    return 0;
} catch (...) { return vdb_caught("vdb_contains"); }

extern "C" int vdb_remove(vdb_index* x, int32_t id, int32_t* removed) try {
    if (!x) FAIL(VDB_ERR_INVALID, "vdb_remove: null index");
    if (removed) *removed = 0;
    std::lock_guard<std::mutex> lk(x->mu);
    if (!is_graph(x)) return VDB_OK;                                  // NativeIndex.removeVector -> false (NativeIndex.java:47-51)
    auto it = x->id2row.find(id);
    if (it == x->id2row.end()) return VDB_OK;                         // `!vectors.containsKey(id)` -> false (:545-547)
    VDB_TRY(sync_index(x));                                           // everything added before this call is in the graph
    const int64_t row = it->second;
    if (!x->any_dead) {                                               // first removal: index the Java code rows handed out so far
        x->jindex_rows.clear();
        for (int64_t r = 0; r < (int64_t)x->h_jindex.size(); r++) x->jindex_rows.emplace(x->h_jindex[r], r);
    }
    int64_t new_entry = -1;
    if (row == x->g.entry && x->n_live > 1) new_entry = java_first_live_row(x, id);
    x->g.remove((int)row, (int)new_entry);
    x->id2row.erase(it);
    x->n_live--;
    if (x->cfg.index_type == VDB_INDEX_HNSW_PQ) x->java_size--;       // currentSize-- (:591); HnswIndex keeps no such counter
    x->any_dead = true; x->dead_dirty = true; x->graph_uploaded = false;
    if (removed) *removed = 1;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_remove"); }

extern "C" int vdb_train(vdb_index* x, const float* samples, int64_t ns) try {
    if (!x || !samples) FAIL(VDB_ERR_INVALID, "Invalid training samples");
    std::lock_guard<std::mutex> lk(x->mu);
    if (x->pq_m == 0) FAIL(VDB_ERR_INVALID, "index has no PQ stage");
    if (MULTI(x)) return vdb_multi_train(x, samples, ns);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    VDB_TRY(train_on_device(x, samples, ns));
    x->n_coded = 0;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_train"); }

extern "C" int vdb_set_codebooks(vdb_index* x, const float* cb) try {
    if (!x || !cb || x->pq_m == 0) FAIL(VDB_ERR_INVALID, "vdb_set_codebooks: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    if (MULTI(x)) { for (auto* sh : x->shards) VDB_TRY(vdb_set_codebooks(sh, cb)); x->trained = true; return VDB_OK; }
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    const size_t cnt = (size_t)x->pq_m * VDB_NCENT * x->sd;
    VDB_TRY(x->d_cb.ensure(cnt, 0, x->stream));
    VDB_CUDA(cudaMemcpyAsync(x->d_cb.p, cb, cnt * 4, cudaMemcpyHostToDevice, x->stream));
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->trained = true; x->n_coded = 0;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_set_codebooks"); }
extern "C" int vdb_get_codebooks(vdb_index* x, float* out) try {
    if (!x || !out || !x->trained) FAIL(VDB_ERR_UNTRAINED, "no codebooks");
    std::lock_guard<std::mutex> lk(x->mu);
    if (MULTI(x)) return vdb_get_codebooks(x->shards[0], out);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    VDB_CUDA(cudaMemcpy(out, x->d_cb.p, (size_t)x->pq_m * VDB_NCENT * x->sd * 4, cudaMemcpyDeviceToHost));
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_get_codebooks"); }
extern "C" int vdb_set_codes(vdb_index* x, const uint8_t* codes) try {
    if (!x || !codes || x->pq_m == 0) FAIL(VDB_ERR_INVALID, "vdb_set_codes: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    NOT_MULTI(x, "vdb_set_codes");
    x->h_codes_override.assign(codes, codes + (size_t)x->n * x->pq_m);
    x->codes_overridden = true; x->n_coded = 0;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_set_codes"); }
extern "C" int vdb_get_codes(vdb_index* x, uint8_t* out) try {
    if (!x || !out || x->pq_m == 0) FAIL(VDB_ERR_INVALID, "vdb_get_codes: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    NOT_MULTI(x, "vdb_get_codes");
    VDB_TRY(sync_index(x));
    if (x->n_coded < x->n_dev) FAIL(VDB_ERR_UNTRAINED, "codes not available (untrained)");
    VDB_CUDA(cudaMemcpy(out, x->d_codes.p, (size_t)x->n_dev * x->pq_m, cudaMemcpyDeviceToHost));
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_get_codes"); }
extern "C" int vdb_encode_all(vdb_index* x) try {
    if (!x || x->pq_m == 0) FAIL(VDB_ERR_INVALID, "vdb_encode_all: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    if (!x->trained) FAIL(VDB_ERR_UNTRAINED, "train first");
    if (MULTI(x)) { VDB_TRY(vdb_multi_sync(x)); return vdb_multi_encode_all(x); }
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    VDB_TRY(upload_rows(x));
    x->n_coded = 0; x->raw_rows = 0;
    return encode_pending(x);
} catch (...) { return vdb_caught("vdb_encode_all"); }

// every level in [0, max_level), every neighbour and the entry point inside [-1, n): a bad id would become an out-of-bounds
// shared-memory atomic in the beam kernel (sticky CUDA fault), so imported graphs are checked on the host first
int vdb_validate_graph(const int32_t* levels, const int32_t* adj, int64_t n, int max_level, int M, int32_t entry) {
    if (n < 0 || entry < -1 || entry >= n) FAIL(VDB_ERR_INVALID, "graph: entry point out of range");
    for (int64_t i = 0; i < n; i++)
        if (levels[i] < 0 || levels[i] >= max_level) FAIL(VDB_ERR_INVALID, "graph: node level out of range");
    const size_t tot = (size_t)max_level * n * M;
    for (size_t i = 0; i < tot; i++)
        if (adj[i] < -1 || adj[i] >= n) FAIL(VDB_ERR_INVALID, "graph: neighbour id out of range");
    return VDB_OK;
}

extern "C" int vdb_set_graph(vdb_index* x, const int32_t* levels, const int32_t* adj, int32_t max_level, int32_t entry) try {
    if (!x || !levels || !adj) FAIL(VDB_ERR_INVALID, "vdb_set_graph: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    if (!is_graph(x)) FAIL(VDB_ERR_INVALID, "vdb_set_graph: not a graph index");
    if (max_level != x->g.max_level) FAIL(VDB_ERR_INVALID, "max_level mismatch");
    VDB_TRY(vdb_validate_graph(levels, adj, x->n, max_level, x->g.M, entry));
    try { x->g.import_graph(levels, adj, (int)x->n, entry); }
    catch (const std::exception& e) { FAIL(VDB_ERR_INVALID, std::string("vdb_set_graph: ") + e.what()); }
    x->n_indexed = x->n; x->graph_uploaded = false;
    x->h_jindex.resize(x->n);
    for (int64_t r = 0; r < x->n; r++) x->h_jindex[r] = r;
    x->java_size = x->n; x->any_dead = false; x->dead_dirty = false; x->jindex_rows.clear();
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_set_graph"); }
extern "C" int vdb_get_graph(vdb_index* x, int32_t* levels, int32_t* adj, int32_t* entry) try {
    if (!x) FAIL(VDB_ERR_INVALID, "vdb_get_graph: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_TRY(sync_index(x));
    const int64_t n = x->g.size();
    if (levels) std::memcpy(levels, x->g.level.data(), (size_t)n * 4);
    if (adj) for (int l = 0; l < x->g.max_level; l++) std::memcpy(adj + (size_t)l * n * x->g.M, x->g.adj[l].data(), (size_t)n * x->g.M * 4);
    if (entry) *entry = x->g.entry;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_get_graph"); }

// rebuildIndex (HnswPqIndex.java:676-718, HnswIndex.java:430-478): train if needed, clear, re-add in ascending id order
extern "C" int vdb_build(vdb_index* x) try {
    if (!x) FAIL(VDB_ERR_INVALID, "vdb_build: null index");
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    const int t = x->cfg.index_type;
    if (MULTI(x)) return vdb_multi_rebuild(x);
    if (t == VDB_INDEX_FLAT || t == VDB_INDEX_PQ) return sync_index(x);     // NativeIndex.buildIndex() is a no-op (NativeIndex.java:84-87)
    if (x->device_only) FAIL(VDB_ERR_INVALID, "graph indexes need host rows");
    const int dim = x->cfg.dim;
    if (t == VDB_INDEX_HNSW_PQ && !x->trained) {
        // samples collected so far = the first min(n, 10000) rows in arrival order
        if (x->n_indexed < x->n) {
            for (int64_t i = x->n_indexed; i < x->n && x->n_train < 10000; i++) {
                x->h_train.insert(x->h_train.end(), x->h_vecs.begin() + i * dim, x->h_vecs.begin() + (i + 1) * dim);
                x->n_train++;
            }
        }
        if (x->n_train > 0) VDB_TRY(train_on_device(x, x->h_train.data(), x->n_train));
    }
    // buildIndex re-adds `new HashMap<>(vectors)` (:687): removed vectors are gone for good — compact the store first
    if (x->any_dead) {
        int64_t w = 0;
        for (int64_t r = 0; r < x->n; r++) {
            if (x->g.dead[r]) continue;
            if (w != r) {
                std::memmove(&x->h_vecs[(size_t)w * dim], &x->h_vecs[(size_t)r * dim], (size_t)dim * 4);
                x->h_ids[w] = x->h_ids[r];
                if (!x->h_norms.empty()) x->h_norms[w] = x->h_norms[r];
            }
            w++;
        }
        x->n = w; x->h_vecs.resize((size_t)w * dim); x->h_ids.resize(w);
        if (!x->h_norms.empty()) x->h_norms.resize(w);
    }
    // reorder the store by id (stable), as buildIndex sorts the saved vectors by id before re-adding
    std::vector<int64_t> order(x->n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return x->h_ids[a] < x->h_ids[b]; });
    bool sorted = true;
    for (int64_t i = 0; i < x->n; i++) if (order[i] != i) { sorted = false; break; }
    if (!sorted) {
        std::vector<float> nv(x->h_vecs.size());
        std::vector<int32_t> ni(x->n);
        std::vector<float> nn(x->h_norms.size());
        for (int64_t i = 0; i < x->n; i++) {
            std::memcpy(&nv[(size_t)i * dim], &x->h_vecs[(size_t)order[i] * dim], (size_t)dim * 4);
            ni[i] = x->h_ids[order[i]];
            if (!nn.empty()) nn[i] = x->h_norms[order[i]];
        }
        x->h_vecs.swap(nv); x->h_ids.swap(ni); x->h_norms.swap(nn);
    }
    if (x->cfg.metric == VDB_METRIC_COSINE && x->n > 0) {
        // buildIndex re-adds the STORED (already normalised) vectors through addVector, which normalises again
        // (HnswIndex.java:437-460 -> :154): not an exact no-op in fp32, so it is done here too (same device kernel as the add path)
        VDB_TRY(x->d_q.ensure((size_t)x->n * dim, 0, x->stream));
        VDB_TRY(x->d_out_d.ensure((size_t)x->n, 0, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->d_q.p, x->h_vecs.data(), (size_t)x->n * dim * 4, cudaMemcpyHostToDevice, x->stream));
        launch_normalize_rows(x->d_q.p, x->n, dim, x->d_out_d.p, x->stream);
        x->h_norms.resize(x->n);
        VDB_CUDA(cudaMemcpyAsync(x->h_vecs.data(), x->d_q.p, (size_t)x->n * dim * 4, cudaMemcpyDeviceToHost, x->stream));
        VDB_CUDA(cudaMemcpyAsync(x->h_norms.data(), x->d_out_d.p, (size_t)x->n * 4, cudaMemcpyDeviceToHost, x->stream));
        VDB_CUDA(cudaStreamSynchronize(x->stream));
    }
    x->g.clear();
    x->n_indexed = 0; x->n_dev = 0; x->n_trav = 0; x->n_fmirror = 0; x->n_coded = 0; x->raw_rows = 0; x->codes_overridden = false;
    x->id2row.clear();
    for (int64_t r = 0; r < x->n; r++) x->id2row[x->h_ids[r]] = r;
    x->n_live = x->n; x->any_dead = false; x->dead_dirty = false; x->h_jindex.clear(); x->java_size = 0; x->jindex_rows.clear();
    // large rebuilds run on the GPU (graph_build.cu: batched inserts with the beam kernel); strict mode and small indexes keep the
    // host builder, whose graph is the oracle's link for link
    const bool pq_ready = t != VDB_INDEX_HNSW_PQ || x->trained;
    if (!x->g.exact_order && pq_ready && x->n >= 8192 && !getenv("VDB_HOST_GRAPH_BUILD")) {
        const int ef = t == VDB_INDEX_HNSW_PQ ? std::max(x->cfg.ef_construction, 200) : x->cfg.ef_construction;     // :498 / HnswIndex.java:22
        VDB_TRY(vdb_gpu_build_graph(x, ef, t == VDB_INDEX_HNSW_PQ ? 1 : 2));
        x->n_indexed = x->n;
        x->h_jindex.resize(x->n);
        std::iota(x->h_jindex.begin(), x->h_jindex.end(), (int64_t)0);
        x->java_size = x->n;
    }
    return sync_index(x);
} catch (...) { return vdb_caught("vdb_build"); }

extern "C" int vdb_fill_synthetic(vdb_index* x, uint64_t seed, int64_t cnt) try {
    if (!x || cnt <= 0) FAIL(VDB_ERR_INVALID, "vdb_fill_synthetic: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    if (x->cfg.index_type != VDB_INDEX_FLAT && x->cfg.index_type != VDB_INDEX_PQ) FAIL(VDB_ERR_INVALID, "synthetic fill is for FLAT/PQ indexes");
    if (x->n > 0 && !x->device_only) FAIL(VDB_ERR_INVALID, "index already holds host rows");
    if (x->n + cnt > x->cfg.max_elements) FAIL(VDB_ERR_FULL, "max_elements exceeded");
    if (MULTI(x)) return vdb_multi_fill_synthetic(x, seed, cnt);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    const int dim = x->cfg.dim;
    VDB_TRY(x->d_base.ensure((size_t)(x->n + cnt) * dim, (size_t)x->n * dim, x->stream));
    // generator index = (global row) * dim + j so that shards of one logical dataset agree with the oracle
    launch_fill_synth(x->d_base.p + (size_t)x->n * dim, seed, (int64_t)(x->cfg.id_offset + x->n) * dim, cnt * dim, x->stream);
    if (x->cfg.metric == VDB_METRIC_COSINE) {
        VDB_TRY(x->d_norms.ensure((size_t)(x->n + cnt), (size_t)x->n, x->stream));
        launch_normalize_rows(x->d_base.p + (size_t)x->n * dim, cnt, dim, x->d_norms.p + x->n, x->stream);
    }
    VDB_CUDA(cudaGetLastError());
    VDB_CUDA(cudaStreamSynchronize(x->stream));
    x->device_only = true;
    x->n += cnt; x->n_dev = x->n;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_fill_synthetic"); }

// ---- search ---------------------------------------------------------------------------------------------------
static inline int java_search_ef(int64_t n, int k, int ef, int cap) {     // HnswPqIndex.java:616-621 (cap = 5000 there)
    int byRatio = (int)((double)n * 0.15);
    int s = std::max(std::max(byRatio, k * 100), ef);
    s = (int)std::min<int64_t>(s, n);
    return std::min(s, cap > 0 ? cap : 5000);
}

static int search_graph(vdb_index* x, const float* d_q, int64_t nq, int k, int32_t* d_ids, float* d_dists, cudaStream_t st) {
    HnswSearchParams p{};
    const bool pq = x->cfg.index_type == VDB_INDEX_HNSW_PQ;
    p.base = x->d_base.p; p.norms = pq ? nullptr : x->d_norms.p; p.ids = x->d_ids.p; p.adj = x->d_adj.p;
    p.codes = pq && x->trained ? x->d_codes.p : nullptr; p.codebooks = pq && x->trained ? x->d_cb.p : nullptr;
    p.queries = d_q; p.out_ids = d_ids; p.out_dists = d_dists; p.counters = x->d_counters.p;
    p.n = (int)x->n; p.dim = x->cfg.dim; p.deg = x->g.M; p.max_level = x->g.max_level; p.entry = x->g.entry;
    p.adj_stride = x->adj_cap; p.dead = x->any_dead ? x->d_dead.p : nullptr;
    const int64_t live = x->n_live;                                             // Java: dataSize = vectors.size() (:615)
    p.pq_m = x->pq_m; p.trained = pq && x->trained ? 1 : 0;
    p.metric = pq ? 0 : 1;
    if (pq) {
        const int sef = java_search_ef(live, k, x->cfg.ef_search, x->cfg.search_ef_cap);
        if (x->trained) { p.variant = 0; p.ef = std::max(sef, k * 100); }     // :638-640 searchLayerWithSize
        else { p.variant = 1; p.ef = sef; }                                    // :643 searchLayer
    } else { p.variant = 2; p.ef = std::max(x->cfg.ef_search, k * 4); }        // HnswIndex.java:347
    p.ef = (int)std::min<int64_t>(p.ef, std::max<int64_t>(live, 1));
    p.k = k; p.nq = (int)nq;
    return vdb_graph_search_launch(x, p, st);
}

// fills the traversal-copy fields and the launch shape of a beam-kernel call and launches it (searches and graph construction)
int vdb_graph_search_launch(vdb_index* x, HnswSearchParams& p, cudaStream_t st) {
    p.tbase = x->d_trav.p; p.tnorm = x->d_tnorm.p; p.tscale = x->d_tscale.p; p.tstride = trav_stride(x); p.tstore = trav_mode(x);
    // candidates expanded per round: ~1.5 % of the beam keeps >= 99 % of the ids equal to the Java pop order (tests) and halves the
    // per-round bookkeeping of small beams (C1, ef = 1500: 11 -> 23 per round, 4.65 -> 2.9 ms per 1000 queries)
    if (p.batch <= 0)      // the graph builder presets 1 (the Java pop order)
        p.batch = x->cfg.trav_batch > 0 ? x->cfg.trav_batch : (p.ef >= 4096 ? 64 : std::max(4, std::min(32, p.ef / 64)));
    p.batch = std::max(1, std::min(p.batch, hnsw_max_batch(p.deg)));
    p.profile = getenv("VDB_HNSW_PROFILE") ? 1 : 0;
    if (const char* eb = getenv("VDB_HNSW_BATCH")) p.batch = std::max(1, std::min(atoi(eb), hnsw_max_batch(p.deg)));
    // L2 prefetch of discovered rows: a gain only when the traversal copy is well beyond the 126 MB L2 (rows come from HBM);
    // on a mostly L2-resident copy (C3: 102 MB) the prefetch instructions cost more L1TEX wavefronts than they save latency
    p.prefetch = (size_t)x->n * (size_t)p.tstride > ((size_t)192 << 20) ? 1 : 0;
    if (const char* ep = getenv("VDB_HNSW_PREFETCH")) p.prefetch = atoi(ep) ? 1 : 0;
    p.visited_in_smem = 1;
    size_t smem = hnsw_search_smem_bytes(p);
    if (smem > 112 * 1024) { p.visited_in_smem = 0; smem = hnsw_search_smem_bytes(p); }
    if (smem > 227 * 1024) FAIL(VDB_ERR_INVALID, "ef too large for the shared-memory pool");
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(3, (227 * 1024) / (smem + 1024)));
    if (getenv("VDB_HNSW_TWO_CTAS")) per_sm = std::min(per_sm, 2);
    p.ctas_per_sm = per_sm;
    int grid = (int)std::min<int64_t>(p.nq, (int64_t)x->num_sms * per_sm);
    if (!p.visited_in_smem) {
        VDB_TRY(x->d_visited.ensure((size_t)grid * ((p.n + 31) / 32), 0, st));
        p.visited_scratch = x->d_visited.p;
    }
    if (x->timed) cudaEventRecord(x->ev_k0, st);
    p.qcounter = reinterpret_cast<int*>(x->d_counters.p + 22);
    VDB_CUDA(cudaMemsetAsync(p.qcounter, 0, sizeof(unsigned long long), st));
    int rc = launch_hnsw_search(p, grid, st);
    if (rc != 0) FAIL(VDB_ERR_INVALID, rc == 1 ? "k or degree too large for the graph kernel" : "shared memory request rejected");
    if (x->timed) cudaEventRecord(x->ev_k1, st);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

static int pick_chunks(const vdb_index* x, int64_t nq, int64_t n, int R) {
    int want = (int)((2 * (int64_t)x->num_sms + nq - 1) / nq);
    int maxc = std::max(1, 8192 / next_pow2(R));
    int c = std::max(1, std::min(want, maxc));
    while (c > 1 && n / c < 4096) c--;
    return c;
}

// ---- PQ search in two phases, so that shards can agree on ONE global shortlist in between (multi_gpu.cu, vdb_pq_*_device):
//      phase 1  LUT build + ADC scan -> the shard's top-R keys (ordered ADC distance << 32 | id) in x->d_keys, rows in x->d_rows;
//      phase 2  entries whose key is above the caller's per-query threshold (the global R-th key) are masked, then the exact
//               re-score of the rest (or, without rerank, the keys themselves) gives the shard's top-k.
// shortlist of the PQ search: max(rerank, k); sharded callers may shorten a shard's LOCAL list (shortlist_override, >= k): the
// global boundary is then checked for certainty by select_kth_key_kernel
static inline int pq_shortlist_len(const vdb_index* x, int k) {
    const int R = x->cfg.rerank > 0 ? std::max(x->cfg.rerank, k) : k;
    return x->shortlist_override > 0 ? std::min(R, std::max(x->shortlist_override, k)) : R;
}
int vdb_pq_phase1(vdb_index* x, const float* d_q, int64_t qb, int k, cudaStream_t st) {
    const int M = x->pq_m, dim = x->cfg.dim;
    const int R = pq_shortlist_len(x, k);
    if (R > 4096) FAIL(VDB_ERR_INVALID, "rerank shortlist > 4096 not supported");
    if (qb > 1024) FAIL(VDB_ERR_INVALID, "PQ phase 1 handles at most 1024 queries per call");
    const int32_t* ids = x->device_only ? nullptr : x->d_ids.p;
    const bool fast = adc_fast_supported(M, R) && !getenv("VDB_ADC_EXACT_SCAN");
    const int R2 = fast ? adc_fast_r2(R) : R;
    const int nch = pick_chunks(x, qb, x->n, R2);
    VDB_TRY(x->d_lut.ensure((size_t)qb * M * VDB_NCENT, 0, st));
    VDB_TRY(x->d_keys.ensure((size_t)qb * R, 0, st));
    VDB_TRY(x->d_rows.ensure((size_t)qb * R, 0, st));
    VDB_TRY(x->d_keys_tmp.ensure((size_t)qb * R2 * (nch + 1), 0, st));
    VDB_TRY(x->d_rows_tmp.ensure((size_t)qb * R2 * (nch + 1), 0, st));
    launch_pq_lut(d_q, x->d_cb.p, (int)qb, dim, M, x->cfg.arith, x->d_lut.p, st);
    if (fast) {
        const int groups = (M + 63) / 64;
        VDB_TRY(x->d_lutg.ensure((size_t)qb * groups * 16384, 0, st));
        VDB_TRY(x->d_flag.ensure((size_t)qb, 0, st));
        launch_lut_to_gather_layout(x->d_lut.p, (int)qb, M, x->d_lutg.p, st);
        if (x->timed) cudaEventRecord(x->ev_k0, st);
        if (launch_adc_scan_fast(x->d_codesT.p, x->d_codes.p, x->n, M, x->d_lut.p, x->d_lutg.p, (int)qb, ids, x->cfg.id_offset, R, nch,
                                 x->d_keys_tmp.p, x->d_rows_tmp.p, x->d_keys.p, x->d_rows.p, x->d_flag.p, st) != 0)
            FAIL(VDB_ERR_INVALID, "fast ADC scan rejected the launch");
        if (x->timed) cudaEventRecord(x->ev_k1, st);
        // queries whose shortlist could not be certified (massive ties) are redone in the reference's own order
        launch_adc_scan_topr(x->d_codes.p, x->n, M, x->d_lut.p, (int)qb, ids, x->cfg.id_offset, R, 1, nullptr, nullptr, x->d_keys.p,
                             x->d_rows.p, st, x->d_flag.p);
    } else {
        if (x->timed) cudaEventRecord(x->ev_k0, st);
        launch_adc_scan_topr(x->d_codes.p, x->n, M, x->d_lut.p, (int)qb, ids, x->cfg.id_offset, R, nch, x->d_keys_tmp.p, x->d_rows_tmp.p,
                             x->d_keys.p, x->d_rows.p, st);
        if (x->timed) cudaEventRecord(x->ev_k1, st);
    }
    VDB_CUDA(cudaGetLastError());
    x->stats.adc_rows += qb * x->n;
    return VDB_OK;
}
int vdb_pq_phase2(vdb_index* x, const float* d_q, int64_t qb, int k, const uint64_t* d_kth, int32_t* d_ids, float* d_dists, cudaStream_t st) {
    const int R = pq_shortlist_len(x, k);
    const int32_t* ids = x->device_only ? nullptr : x->d_ids.p;
    if (d_kth) launch_mask_keys_above(x->d_keys.p, x->d_rows.p, (int)qb, R, d_kth, st);
    if (x->cfg.rerank > 0)
        launch_rerank(x->d_base.p, ids, x->cfg.id_offset, x->cfg.dim, d_q, (int)qb, x->d_rows.p, R,
                      x->cfg.arith == VDB_ARITH_NATIVE_AVX2 ? ORDER_R1 : ORDER_SEQ, k, d_ids, d_dists, st);
    else
        launch_keys_to_results(x->d_keys.p, R, k, qb, d_ids, d_dists, st);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}

static int search_pq(vdb_index* x, const float* d_q, int64_t nq, int k, int32_t* d_ids, float* d_dists, cudaStream_t st) {
    if (!x->trained) {   // PQIndex.cpp:147-150 / :259-270: untrained -> no results
        launch_keys_to_results(nullptr, 0, k, nq, d_ids, d_dists, st);
        return VDB_OK;
    }
    const int dim = x->cfg.dim;
    const int64_t QB = 1024;   // query sub-batch: bounds the LUT buffers (QB * M * 1 KB each)
    for (int64_t q0 = 0; q0 < nq; q0 += QB) {
        const int64_t qb = std::min(QB, nq - q0);
        VDB_TRY(vdb_pq_phase1(x, d_q + q0 * dim, qb, k, st));
        VDB_TRY(vdb_pq_phase2(x, d_q + q0 * dim, qb, k, nullptr, d_ids + q0 * k, d_dists + q0 * k, st));
    }
    return VDB_OK;
}

// bf16 mirror for the tensor-core flat search; rebuilt whenever rows were added (the centring mean changes with them)
static int build_flat_mirror(vdb_index* x, cudaStream_t st) {
    if (x->n_fmirror == x->n) return VDB_OK;
    const FlatGemmPlan p = flat_gemm_plan(x->n, x->cfg.dim, 1);
    VDB_TRY(x->d_fmirror.ensure((size_t)x->n * p.dpad, 0, st));
    VDB_TRY(x->d_fab.ensure((size_t)p.n_pad, 0, st));
    VDB_TRY(x->d_fcolsum.ensure((size_t)x->cfg.dim + 4, 0, st));
    if (launch_flat_mirror(x->d_base.p, x->n, x->cfg.dim, x->cfg.metric, x->d_fmirror.p, x->d_fab.p, x->d_fcolsum.p, st))
        FAIL(VDB_ERR_CUDA, "flat mirror launch failed");
    x->n_fmirror = x->n;
    return VDB_OK;
}

static int search_flat(vdb_index* x, const float* d_q, int64_t nq, int k, int32_t* d_ids, float* d_dists, cudaStream_t st) {
    if (k > 2048) FAIL(VDB_ERR_INVALID, "k > 2048 not supported");
    const int32_t* ids = x->device_only ? nullptr : x->d_ids.p;
    const int order = x->cfg.arith == VDB_ARITH_NATIVE_AVX2 ? ORDER_R1 : ORDER_SEQ;
    // tensor-core path (flat_gemm.cu): bf16 GEMM shortlist + exact re-score + certification; VDB_FLAT_EXACT_SCAN=1 forces
    // the order-exact scan (tests compare the two)
    const bool gemm = flat_gemm_supported(x->n, x->cfg.dim, k) && !getenv("VDB_FLAT_EXACT_SCAN");
    if (gemm) VDB_TRY(build_flat_mirror(x, st));
    const int64_t QB = gemm ? 4096 : 16384;
    for (int64_t q0 = 0; q0 < nq; q0 += QB) {
        const int64_t qb = std::min(QB, nq - q0);
        const int nch = pick_chunks(x, qb, x->n, k);
        VDB_TRY(x->d_keys.ensure((size_t)qb * k, 0, st));
        VDB_TRY(x->d_rows.ensure((size_t)qb * k, 0, st));
        if (nch > 1) { VDB_TRY(x->d_keys_tmp.ensure((size_t)qb * k * nch, 0, st)); VDB_TRY(x->d_rows_tmp.ensure((size_t)qb * k * nch, 0, st)); }
        const int* only = nullptr;
        if (gemm) {
            const int dim = x->cfg.dim;
            const int64_t slab = flat_gemm_slab_rows(x->n, dim, k);
            const int nslabs = (int)((x->n + slab - 1) / slab);
            if ((int64_t)nslabs * k > 8192) FAIL(VDB_ERR_INVALID, "flat search: too many slabs for this k");
            const FlatGemmPlan p = flat_gemm_plan(std::min(slab, x->n), dim, k);
            VDB_TRY(x->d_fq.ensure((size_t)qb * p.dpad, 0, st));
            if (x->cfg.metric == 1) VDB_TRY(x->d_fqf.ensure((size_t)qb * dim, 0, st));
            VDB_TRY(x->d_fqstat.ensure((size_t)qb, 0, st));
            VDB_TRY(x->d_fscores.ensure((size_t)qb * p.S_pad, 0, st));
            VDB_TRY(x->d_ftau.ensure((size_t)qb, 0, st));
            VDB_TRY(x->d_fcnt.ensure((size_t)qb, 0, st));
            VDB_TRY(x->d_fcand.ensure((size_t)qb * p.cap, 0, st));
            VDB_TRY(x->d_flag.ensure((size_t)qb, 0, st));
            if (nslabs > 1) {
                VDB_TRY(x->d_fslab_d.ensure((size_t)nslabs * qb * k, 0, st));
                VDB_TRY(x->d_fslab_i.ensure((size_t)nslabs * qb * k, 0, st));
            }
            if (x->timed && q0 == 0) cudaEventRecord(x->ev_k0, st);
            for (int sl = 0; sl < nslabs; sl++) {
                const int64_t r0 = (int64_t)sl * slab, rows = std::min(slab, x->n - r0);
                FlatGemmArgs a{};
                a.d_base = x->d_base.p + (size_t)r0 * dim; a.d_norms = x->d_norms.p ? x->d_norms.p + r0 : nullptr; a.d_ids = ids ? ids + r0 : nullptr;
                a.id_offset = x->cfg.id_offset + (int)r0; a.n = rows; a.n_total = x->n; a.flag_or = sl > 0; a.dim = dim;
                a.metric = x->cfg.metric; a.order = order; a.d_mirror = x->d_fmirror.p + (size_t)r0 * p.dpad; a.d_arow = x->d_fab.p + r0;
                a.d_colsum = x->d_fcolsum.p;
                a.d_queries = d_q + q0 * dim; a.nq = (int)qb; a.k = k;
                a.d_qmirror = x->d_fq.p; a.d_qf = x->d_fqf.p; a.d_qstat = x->d_fqstat.p; a.d_scores = x->d_fscores.p; a.d_tau = x->d_ftau.p;
                a.d_cnt = x->d_fcnt.p; a.d_cand = x->d_fcand.p; a.d_flag = x->d_flag.p; a.d_stats = x->d_counters.p + 10;
                a.d_out_ids = nslabs > 1 ? x->d_fslab_i.p + (size_t)sl * qb * k : d_ids + q0 * k;
                a.d_out_d = nslabs > 1 ? x->d_fslab_d.p + (size_t)sl * qb * k : d_dists + q0 * k;
                const int rc = launch_flat_gemm_search(a, st);
                if (rc) { char msg[64]; snprintf(msg, sizeof msg, "flat GEMM launch failed (%d)", rc); FAIL(VDB_ERR_CUDA, msg); }
            }
            if (nslabs > 1) launch_merge_topk(x->d_fslab_d.p, x->d_fslab_i.p, nslabs, qb, k, d_dists + q0 * k, d_ids + q0 * k, st);
            if (x->timed && q0 == 0) cudaEventRecord(x->ev_k1, st);
            only = x->d_flag.p;                      // uncertified queries fall through to the order-exact scan below
        }
        if (!gemm && x->timed && q0 == 0) cudaEventRecord(x->ev_k0, st);
        launch_flat_search_exact(x->d_base.p, x->d_norms.p, ids, x->cfg.id_offset, x->n, x->cfg.dim, d_q + q0 * x->cfg.dim, (int)qb, k,
                                 x->cfg.metric, order, nch, x->d_keys_tmp.p, x->d_rows_tmp.p, x->d_keys.p, x->d_rows.p, st, only);
        if (!gemm && x->timed && q0 == 0) cudaEventRecord(x->ev_k1, st);
        launch_keys_to_results(x->d_keys.p, k, k, qb, d_ids + q0 * k, d_dists + q0 * k, st, only);
        VDB_CUDA(cudaGetLastError());
    }
    return VDB_OK;
}

static int search_device_locked(vdb_index* x, const float* d_q, int64_t nq, int k, int32_t* d_ids, float* d_dists, cudaStream_t st) {
    VDB_TRY(sync_index(x));
    if (nq == 0) return VDB_OK;
    // all scratch (keys, LUTs, visited sets, flags, timing events) belongs to the index: a search issued on a different stream
    // than the previous one is ordered after it on the device instead of overlapping with it
    if (x->have_done && x->last_stream != st) VDB_CUDA(cudaStreamWaitEvent(st, x->ev_done, 0));
    x->timed = true;
    cudaEventRecord(x->ev_total0, st);
    int rc;
    if (x->n == 0 || (is_graph(x) && (x->g.entry < 0 || x->n_live == 0))) {
        launch_keys_to_results(nullptr, 0, k, nq, d_ids, d_dists, st);     // empty index -> empty lists (HnswIndexTest.java:120-129)
        cudaEventRecord(x->ev_k0, st); cudaEventRecord(x->ev_k1, st);
        rc = VDB_OK;
    } else {
        switch (x->cfg.index_type) {
            case VDB_INDEX_FLAT: rc = search_flat(x, d_q, nq, k, d_ids, d_dists, st); break;
            case VDB_INDEX_PQ: rc = search_pq(x, d_q, nq, k, d_ids, d_dists, st); break;
            default: rc = search_graph(x, d_q, nq, k, d_ids, d_dists, st); break;
        }
    }
    cudaEventRecord(x->ev_total1, st);
    cudaEventRecord(x->ev_done, st);
    x->have_done = true; x->last_stream = st;
    x->stats.queries += nq;
    return rc;
}

int vdb_search_device_nolock(vdb_index* x, const float* d_q, int64_t nq, int k, int32_t* d_ids, float* d_dists, cudaStream_t st) {
    return search_device_locked(x, d_q, nq, k, d_ids, d_dists, st);
}
int vdb_sync_index_nolock(vdb_index* x) { return sync_index(x); }

extern "C" int vdb_search_batch_device(vdb_index* x, const float* d_q, int64_t nq, int32_t k, int32_t* d_ids, float* d_dists, void* stream) try {
    if (!x || nq < 0 || k <= 0 || (nq > 0 && (!d_q || !d_ids || !d_dists))) FAIL(VDB_ERR_INVALID, "vdb_search_batch_device: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaStream_t st = (cudaStream_t)stream;   // verbatim cudaStream_t: NULL is CUDA's legacy default stream
    if (MULTI(x)) return vdb_multi_search(x, d_q, true, nq, k, d_ids, d_dists, st);      // buffers on the front device (cfg.device)
    return search_device_locked(x, d_q, nq, k, d_ids, d_dists, st);
} catch (...) { return vdb_caught("vdb_search_batch_device"); }

extern "C" int vdb_search_batch(vdb_index* x, const float* q, int64_t nq, int32_t k, int32_t* out_ids, float* out_d) try {
    if (!x || nq < 0 || k <= 0 || (nq > 0 && (!q || !out_ids || !out_d))) FAIL(VDB_ERR_INVALID, "vdb_search_batch: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    if (nq == 0) return VDB_OK;
    if (MULTI(x)) return vdb_multi_search(x, q, false, nq, k, out_ids, out_d, x->stream);
    const int dim = x->cfg.dim;
    cudaStream_t st = x->stream;
    VDB_TRY(x->d_q.ensure((size_t)nq * dim, 0, st));
    VDB_TRY(x->d_out_d.ensure((size_t)nq * k, 0, st));
    VDB_TRY(x->d_out_ids.ensure((size_t)nq * k, 0, st));
    // caller memory is pageable in general (Java direct ByteBuffers): stage through pinned buffers so the copies are async DMA
    VDB_TRY(x->p_q.ensure((size_t)nq * dim));
    VDB_TRY(x->p_d.ensure((size_t)nq * k));
    VDB_TRY(x->p_ids.ensure((size_t)nq * k));
    {   // in 1 MB pieces: the DMA of a piece runs while the next one is being staged
        const size_t total = (size_t)nq * dim * 4, piece = (size_t)1 << 20;
        const char* src = reinterpret_cast<const char*>(q);
        char* pin = reinterpret_cast<char*>(x->p_q.p);
        char* dst = reinterpret_cast<char*>(x->d_q.p);
        for (size_t off = 0; off < total; off += piece) {
            const size_t len = std::min(piece, total - off);
            std::memcpy(pin + off, src + off, len);
            VDB_CUDA(cudaMemcpyAsync(dst + off, pin + off, len, cudaMemcpyHostToDevice, st));
        }
    }
    VDB_TRY(search_device_locked(x, x->d_q.p, nq, k, x->d_out_ids.p, x->d_out_d.p, st));
    VDB_CUDA(cudaMemcpyAsync(x->p_ids.p, x->d_out_ids.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, st));
    VDB_CUDA(cudaMemcpyAsync(x->p_d.p, x->d_out_d.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, st));
    VDB_CUDA(cudaStreamSynchronize(st));
    std::memcpy(out_ids, x->p_ids.p, (size_t)nq * k * 4);
    std::memcpy(out_d, x->p_d.p, (size_t)nq * k * 4);
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_search_batch"); }

extern "C" int vdb_get_stats(vdb_index* x, vdb_stats* out) try {
    if (!x || !out) FAIL(VDB_ERR_INVALID, "vdb_get_stats: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    if (MULTI(x)) return vdb_multi_stats(x, out);
    VDB_CUDA(cudaSetDevice(x->cfg.device));
    if (x->timed) {
        VDB_CUDA(cudaEventSynchronize(x->ev_total1));
        cudaEventElapsedTime(&x->stats.last_total_ms, x->ev_total0, x->ev_total1);
        if (cudaEventElapsedTime(&x->stats.last_kernel_ms, x->ev_k0, x->ev_k1) != cudaSuccess) { x->stats.last_kernel_ms = 0; cudaGetLastError(); }
    }
    unsigned long long c[24] = {0};
    VDB_CUDA(cudaMemcpy(c, x->d_counters.p, sizeof(c), cudaMemcpyDeviceToHost));
    if (getenv("VDB_HNSW_PROFILE"))
        fprintf(stderr, "[vdb hnsw profile] cycles: setup %llu select %llu gather %llu dist %llu sort %llu merge %llu final %llu | rounds %llu\n",
                c[2], c[3], c[4], c[5], c[6], c[7], c[8], c[9]);
    if (getenv("VDB_HNSW_PROFILE")) {
        fprintf(stderr, "[vdb hnsw profile] select split: C0 folds %llu cycles in %llu folds, candidate scan %llu cycles\n", c[17], c[18], c[19]);
        fprintf(stderr, "[vdb hnsw profile] accepted keys: sum %llu | rounds with > 256: %llu (sort cycles %llu), with 129..256: %llu\n", c[13], c[12], c[15], c[14]);
    }
    x->stats.dist_evals = (int64_t)c[0]; x->stats.expansions = (int64_t)c[1];
    x->stats.flat_certified = (int64_t)c[10]; x->stats.flat_fallback = (int64_t)c[11];
    x->stats.trav_format = x->trav_eff == 0 ? VDB_TRAV_F32 : (x->trav_eff == 2 ? VDB_TRAV_FP8 : (x->trav_eff == 3 ? VDB_TRAV_INT8 : VDB_TRAV_BF16));
    *out = x->stats;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_get_stats"); }

// ---- device-side building blocks ------------------------------------------------------------------------------
extern "C" int vdb_pq_lut_device(vdb_index* x, const float* d_q, int64_t nq, float* d_lut, void* stream) try {
    if (!x || !d_q || !d_lut || x->pq_m == 0) FAIL(VDB_ERR_INVALID, "vdb_pq_lut_device: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    NOT_MULTI(x, "vdb_pq_lut_device");
    if (!x->trained) FAIL(VDB_ERR_UNTRAINED, "train first");
    cudaStream_t st = (cudaStream_t)stream;   // verbatim cudaStream_t: NULL is CUDA's legacy default stream
    launch_pq_lut(d_q, x->d_cb.p, (int)nq, x->cfg.dim, x->pq_m, x->cfg.arith, d_lut, st);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_pq_lut_device"); }
extern "C" int vdb_adc_scan_device(vdb_index* x, const float* d_lut, float* d_out, void* stream) try {
    if (!x || !d_lut || !d_out || x->pq_m == 0) FAIL(VDB_ERR_INVALID, "vdb_adc_scan_device: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    NOT_MULTI(x, "vdb_adc_scan_device");
    VDB_TRY(sync_index(x));
    if (x->n_coded < x->n) FAIL(VDB_ERR_UNTRAINED, "no codes");
    cudaStream_t st = (cudaStream_t)stream;   // verbatim cudaStream_t: NULL is CUDA's legacy default stream
    launch_adc_scan_plain(x->d_codes.p, x->n, x->pq_m, d_lut, d_out, st);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_adc_scan_device"); }
// ---- the two PQ phases and the boundary selection as device-level calls: one process per GPU (torch.distributed) runs phase 1,
//      all-gathers the keys (ONE collective of nq x R x 8 bytes), selects the global boundary, runs phase 2, all-gathers the top-k ----
extern "C" int vdb_pq_shortlist_device(vdb_index* x, const float* d_q, int64_t nq, int32_t k, uint64_t* d_keys_out, int32_t* shortlist_len, void* stream) try {
    if (!x || !d_q || !d_keys_out || nq <= 0 || k <= 0 || x->cfg.index_type != VDB_INDEX_PQ) FAIL(VDB_ERR_INVALID, "vdb_pq_shortlist_device: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    NOT_MULTI(x, "vdb_pq_shortlist_device");
    if (!x->trained) FAIL(VDB_ERR_UNTRAINED, "train first");
    VDB_TRY(sync_index(x));
    cudaStream_t st = (cudaStream_t)stream;
    // *shortlist_len on entry: 0 = the full shortlist, else the length of this shard's LOCAL list (kept until the next call; phase 2
    // uses the same)
    x->shortlist_override = (shortlist_len && *shortlist_len > 0) ? *shortlist_len : 0;
    const int R = pq_shortlist_len(x, k);
    if (shortlist_len) *shortlist_len = R;
    if (x->n == 0) { VDB_CUDA(cudaMemsetAsync(d_keys_out, 0xff, (size_t)nq * R * 8, st)); return VDB_OK; }
    VDB_TRY(vdb_pq_phase1(x, d_q, nq, k, st));
    VDB_CUDA(cudaMemcpyAsync(d_keys_out, x->d_keys.p, (size_t)nq * R * 8, cudaMemcpyDeviceToDevice, st));
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_pq_shortlist_device"); }
extern "C" int vdb_pq_rerank_device(vdb_index* x, const float* d_q, int64_t nq, int32_t k, const uint64_t* d_kth, int32_t* d_ids, float* d_dists, void* stream) try {
    if (!x || !d_q || !d_ids || !d_dists || nq <= 0 || k <= 0 || x->cfg.index_type != VDB_INDEX_PQ) FAIL(VDB_ERR_INVALID, "vdb_pq_rerank_device: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    NOT_MULTI(x, "vdb_pq_rerank_device");
    cudaStream_t st = (cudaStream_t)stream;
    if (x->n == 0) { launch_keys_to_results(nullptr, 0, k, nq, d_ids, d_dists, st); return VDB_OK; }
    return vdb_pq_phase2(x, d_q, nq, k, d_kth, d_ids, d_dists, st);
} catch (...) { return vdb_caught("vdb_pq_rerank_device"); }
extern "C" int vdb_select_kth_key_device(const uint64_t* d_keys, int32_t g, int64_t nq, int32_t R, uint64_t* d_kth, void* stream) try {
    if (!d_keys || !d_kth || g <= 0 || R <= 0 || nq <= 0 || (int64_t)g * R > 16384) FAIL(VDB_ERR_INVALID, "vdb_select_kth_key_device: bad arguments (g x R <= 16384)");
    launch_select_kth_key(d_keys, g, (int)nq, R, R, d_kth, nullptr, (cudaStream_t)stream);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_select_kth_key_device"); }
extern "C" int vdb_select_kth_key_ex_device(const uint64_t* d_keys, int32_t g, int64_t nq, int32_t list_len, int32_t rank, uint64_t* d_kth,
                                            int32_t* d_uncertain, void* stream) {
    if (!d_keys || !d_kth || g <= 0 || list_len <= 0 || rank <= 0 || nq <= 0 || (int64_t)g * list_len > 16384 || (int64_t)g * list_len < rank)
        FAIL(VDB_ERR_INVALID, "vdb_select_kth_key_ex_device: bad arguments (rank <= g x list_len <= 16384)");
    launch_select_kth_key(d_keys, g, (int)nq, list_len, rank, d_kth, d_uncertain, (cudaStream_t)stream);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}
extern "C" int vdb_merge_topk_device(const float* d_dists, const int32_t* d_ids, int32_t g, int64_t nq, int32_t k, float* d_out_d,
                                     int32_t* d_out_ids, void* stream) {
    if (!d_dists || !d_ids || !d_out_d || !d_out_ids || g <= 0 || k <= 0 || (int64_t)g * k > 8192) FAIL(VDB_ERR_INVALID, "vdb_merge_topk_device: bad arguments");
    launch_merge_topk(d_dists, d_ids, g, nq, k, d_out_d, d_out_ids, (cudaStream_t)stream);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}
extern "C" int vdb_distance_device(const float* d_a, const float* d_b, int64_t n, int32_t dim, int32_t metric, int32_t order, float* d_out,
                                   void* stream) {
    if (!d_a || !d_b || !d_out) FAIL(VDB_ERR_INVALID, "vdb_distance_device: bad arguments");
    launch_distance(d_a, d_b, n, dim, metric, order, d_out, (cudaStream_t)stream);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}
