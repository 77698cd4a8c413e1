// oracle/ref_driver.cpp — TEST INFRASTRUCTURE ONLY.
//
// A C-linkage driver around the UNMODIFIED reference classes (compiled from /root/reference/native where
// they lie, see oracle/Makefile `ref`). It exists because the reference's own gtest suites cannot be built
// offline (FetchContent, native/CMakeLists.txt:158-163). Nothing here is product code and nothing from the
// reference is copied: this file only calls the reference's public (and, for inspection, private) members.
//
// `#define private public` is used solely to READ codebooks_/codes_ back out of PQIndex / HNSWPQIndex so the
// oracle restatement and the GPU path can be compared on "the same codebooks" (the reference has no
// save/load: PQIndex.cpp:299-300, HNSWPQIndex.cpp:774-780).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <future>
#include <memory>
#include <mutex>
#include <queue>
#include <random>
#include <shared_mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#define private public
#include "index/PQIndex.h"
#include "index/HNSWPQIndex.h"
#include "index/HNSWIndex.h"
#undef private
#include "compute/ADCUtils.h"
#include "compute/DistanceUtils.h"

using namespace vectordb;

namespace vectordb {
float euclideanDistanceAVX2(const float*, const float*, size_t);
float cosineDistanceAVX2(const float*, const float*, size_t);
void batchEuclideanDistanceAVX2(const float*, const float*, size_t, size_t, float*);
float euclideanDistanceScalar(const float*, const float*, size_t);
}

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API const char* ref_isa_name() { return getISAName(); }
REF_API float ref_l2sq_dispatch(const float* a, const float* b, int d) { return getEuclideanDistanceFunc()(a, b, (size_t)d); }
REF_API float ref_l2sq_avx2(const float* a, const float* b, int d) { return euclideanDistanceAVX2(a, b, (size_t)d); }
REF_API float ref_l2sq_scalar(const float* a, const float* b, int d) { return euclideanDistanceScalar(a, b, (size_t)d); }
REF_API float ref_cosine_avx2(const float* a, const float* b, int d) { return cosineDistanceAVX2(a, b, (size_t)d); }
REF_API void ref_batch_l2(const float* q, const float* v, int n, int d, float* out) { getBatchEuclideanDistanceFunc()(q, v, (size_t)n, (size_t)d, out); }
REF_API float ref_adc_scalar(const float* table, const uint8_t* codes, int M, int ncent) { return adcDistanceScalar(table, codes, M, ncent); }
REF_API float ref_adc_avx2(const float* table, const uint8_t* codes, int M, int ncent) { return adcDistanceAVX2(table, codes, M, ncent); }

// ---- PQIndex (native/index/PQIndex.{h,cpp}) ----
REF_API void* ref_pq_create(int dim, int maxElements, int M, int nBits, int maxIter) {
    try { PQConfig c; c.M = M; c.nBits = nBits; c.maxIterations = maxIter; return new PQIndex(dim, maxElements, c); }
    catch (...) { return nullptr; }
}
REF_API void ref_pq_destroy(void* h) { delete (PQIndex*)h; }
REF_API int ref_pq_train(void* h, int n, const float* samples) { try { ((PQIndex*)h)->train(n, samples); return 0; } catch (...) { return 1; } }
REF_API int ref_pq_add_batch(void* h, const float* v, const int* ids, int n) { try { ((PQIndex*)h)->addBatch(v, ids, n); return 0; } catch (...) { return 1; } }
REF_API int ref_pq_add(void* h, int id, const float* v) { try { ((PQIndex*)h)->add(id, v); return 0; } catch (...) { return 1; } }
REF_API int ref_pq_search(void* h, const float* q, int k, int* ids, float* d) { int c = 0; ((PQIndex*)h)->search(q, k, ids, d, &c); return c; }
REF_API void ref_pq_search_batch(void* h, const float* q, int nq, int k, int* ids, float* d) { ((PQIndex*)h)->searchBatch(q, nq, k, ids, d); }
REF_API int ref_pq_size(void* h) { return ((PQIndex*)h)->size(); }
REF_API void ref_pq_get_codebooks(void* h, float* out) { auto* p = (PQIndex*)h; std::memcpy(out, p->codebooks_.data(), p->codebooks_.size() * sizeof(float)); }
REF_API void ref_pq_get_codes(void* h, uint8_t* out) { auto* p = (PQIndex*)h; std::memcpy(out, p->codes_.data(), p->codes_.size()); }

// ---- HNSWPQIndex (native/index/HNSWPQIndex.{h,cpp}) ----
REF_API void* ref_hnswpq_create(int dim, int maxElements, int M, int efC, int ef, int pqM, int pqBits, int pqIters) {
    try {
        HNSWPQConfig c; c.M = M; c.efConstruction = efC; c.efSearch = ef; c.pqM = pqM; c.pqBits = pqBits; c.pqIterations = pqIters;
        return new HNSWPQIndex(dim, maxElements, c);
    } catch (...) { return nullptr; }
}
REF_API void ref_hnswpq_destroy(void* h) { delete (HNSWPQIndex*)h; }
REF_API int ref_hnswpq_train(void* h, int n, const float* samples) { try { ((HNSWPQIndex*)h)->train(n, samples); return 0; } catch (...) { return 1; } }
REF_API void ref_hnswpq_get_codebooks(void* h, float* out) { auto* p = (HNSWPQIndex*)h; std::memcpy(out, p->codebooks_.data(), p->codebooks_.size() * sizeof(float)); }
REF_API void ref_hnswpq_encode(void* h, const float* v, int n, uint8_t* codes) {
    auto* p = (HNSWPQIndex*)h;
    for (int i = 0; i < n; i++) p->encode(v + (size_t)i * p->dimension_, codes + (size_t)i * p->config_.pqM);
}
REF_API int ref_hnswpq_add(void* h, int id, const float* v) { try { ((HNSWPQIndex*)h)->add(id, v); return 0; } catch (...) { return 1; } }
REF_API int ref_hnswpq_search(void* h, const float* q, int k, int* ids, float* d) { int c = 0; ((HNSWPQIndex*)h)->search(q, k, ids, d, &c); return c; }
REF_API void ref_hnswpq_search_batch(void* h, const float* q, int nq, int k, int* ids, float* d) { ((HNSWPQIndex*)h)->searchBatch(q, nq, k, ids, d); }
REF_API float ref_hnswpq_pq_distance(void* h, const float* q, int node) { return ((HNSWPQIndex*)h)->computeDistancePQ(q, node); }

// ---- the three BLAS calls of native/compute/BatchDistance.cpp (declared in oracle/shim/cblas.h): the reference's own fallback
//      loops (BatchDistance.cpp:36-46, 62-70, 98-105), row-major, the only argument combinations the reference uses ----
extern "C" {
#include "cblas.h"
void cblas_sgemm(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, enum CBLAS_TRANSPOSE, int m, int n, int k, float alpha, const float* A, int lda,
                 const float* B, int ldb, float beta, float* C, int ldc) {          // C = alpha * A * B^T + beta * C
    for (int i = 0; i < m; i++)
        for (int j = 0; j < n; j++) {
            float sum = 0.0f;
            for (int l = 0; l < k; l++) sum += A[(size_t)i * lda + l] * B[(size_t)j * ldb + l];
            C[(size_t)i * ldc + j] = alpha * sum + (beta == 0.0f ? 0.0f : beta * C[(size_t)i * ldc + j]);
        }
}
float cblas_sdot(int n, const float* x, int incx, const float* y, int incy) {
    float sum = 0.0f;
    for (int i = 0; i < n; i++) sum += x[(size_t)i * incx] * y[(size_t)i * incy];
    return sum;
}
void cblas_sgemv(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, int m, int n, float alpha, const float* A, int lda, const float* x, int incx,
                 float beta, float* y, int incy) {                                   // y = alpha * A * x + beta * y
    for (int i = 0; i < m; i++) {
        float sum = 0.0f;
        for (int j = 0; j < n; j++) sum += A[(size_t)i * lda + j] * x[(size_t)j * incx];
        y[(size_t)i * incy] = alpha * sum + (beta == 0.0f ? 0.0f : beta * y[(size_t)i * incy]);
    }
}
}

// ---- native HNSWIndex (native/index/HNSWIndex.cpp): the one beam search of the reference that runs on this image ----
REF_API void* ref_hnsw_create(int dim, int maxElements, int M, int efC, int ef) {
    HNSWConfig c; c.M = M; c.efConstruction = efC; c.efSearch = ef;
    try { return new HNSWIndex(dim, maxElements, c); } catch (...) { return nullptr; }
}
REF_API void ref_hnsw_destroy(void* h) { delete (HNSWIndex*)h; }
REF_API int ref_hnsw_add(void* h, int id, const float* v) { try { ((HNSWIndex*)h)->add(id, v); return 0; } catch (...) { return 1; } }
REF_API int ref_hnsw_add_batch(void* h, const float* v, const int* ids, int n) { try { ((HNSWIndex*)h)->addBatch(v, ids, n); return 0; } catch (...) { return 1; } }
REF_API int ref_hnsw_size(void* h) { return ((HNSWIndex*)h)->size(); }
REF_API int ref_hnsw_search(void* h, const float* q, int k, int* ids, float* d) { int c = 0; ((HNSWIndex*)h)->search(q, k, ids, d, &c); return c; }
REF_API void ref_hnsw_search_batch(void* h, const float* q, int nq, int k, int* ids, float* d) { ((HNSWIndex*)h)->searchBatch(q, nq, k, ids, d); }
REF_API int ref_hnsw_entry(void* h) { return ((HNSWIndex*)h)->entryPoint_.load(); }
REF_API int ref_hnsw_node_level(void* h, int node) { return ((HNSWIndex*)h)->nodes_[node].level; }
// neighbours of `node` on `level` into out[cap]; returns the count
REF_API int ref_hnsw_neighbors(void* h, int node, int level, int* out, int cap) {
    auto& nb = ((HNSWIndex*)h)->nodes_[node].neighbors;
    if (level < 0 || level >= (int)nb.size()) return 0;
    int n = std::min<int>(cap, (int)nb[level].size());
    for (int i = 0; i < n; i++) out[i] = nb[level][i];
    return n;
}
