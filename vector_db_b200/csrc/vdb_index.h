// vdb_index.h — the index object behind the C ABI (include/vdb.h).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/vdb.h"
#include "host_graph.h"

void vdb_set_error(const std::string& msg);

#define VDB_CUDA(call)                                                                          \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            vdb_set_error(std::string(#call) + ": " + cudaGetErrorString(e__));                 \
            return VDB_ERR_CUDA;                                                                \
        }                                                                                       \
    } while (0)
#define VDB_TRY(call)                    \
    do {                                 \
        int rc__ = (call);               \
        if (rc__ != VDB_OK) return rc__; \
    } while (0)

// growable device array; ensure() keeps the first `keep` elements
template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n, size_t keep, cudaStream_t st) {
        if (n <= cap) return VDB_OK;
        size_t ncap = n;
        T* np = nullptr;
        VDB_CUDA(cudaMalloc(&np, ncap * sizeof(T)));
        if (p && keep) VDB_CUDA(cudaMemcpyAsync(np, p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st));
        if (p) { VDB_CUDA(cudaStreamSynchronize(st)); cudaFree(p); }
        p = np; cap = ncap;
        return VDB_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
template <typename T>
struct PinnedBuf {
    T* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n) {
        if (n <= cap) return VDB_OK;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        VDB_CUDA(cudaMallocHost(&p, n * sizeof(T)));
        cap = n;
        return VDB_OK;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct vdb_index {
    vdb_config cfg;
    int pq_m = 0, sd = 0, num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_total0 = nullptr, ev_total1 = nullptr, ev_k0 = nullptr, ev_k1 = nullptr;
    cudaEvent_t ev_done = nullptr;  // end of the last search: the scratch buffers are shared, so a search on ANOTHER stream waits on it
    cudaStream_t last_stream = nullptr;
    bool have_done = false;
    bool timed = false;
    std::mutex mu;

    // host store (row = insertion order since the last rebuild)
    std::vector<float> h_vecs;      // normalised rows for cosine
    std::vector<float> h_norms;     // cosine only
    std::vector<int32_t> h_ids;
    int64_t n = 0;                  // rows stored
    bool device_only = false;       // rows were generated on the device (vdb_fill_synthetic): no host copy

    // device store
    DevBuf<float> d_base, d_norms, d_cb;
    DevBuf<int32_t> d_ids, d_adj;
    DevBuf<uint8_t> d_codes;
    DevBuf<uint8_t> d_codesT;       // PQ index: codes transposed in 16-row blocks [ceil(n/16)][ceil8(M)][16] (adc_fast.cu)
    int64_t n_codesT = 0;           // rows present in d_codesT
    DevBuf<unsigned char> d_trav;   // traversal rows (bf16 or padded fp32), rebuilt when rows are uploaded
    int64_t n_trav = 0;
    int trav_eff = 1;               // effective row format (kernel tstore: 0 fp32, 1 bf16, 2 fp8, 3 int8), decided when the copy is (re)built
    DevBuf<float> d_tnorm;          // squared norms of the traversal rows
    DevBuf<float> d_tscale;         // fp8 traversal rows: inverse of the per-row scale
    DevBuf<uint16_t> d_fmirror;     // flat index: bf16 mirror [n][dpad] for the tensor-core search (flat_gemm.cu)
    DevBuf<float> d_fab;            // per-row additive term a: approximate score = a + b * (q' . x')
    DevBuf<float> d_fcolsum;        // [dim] column sums, [dim] max squared row norm
    int64_t n_fmirror = 0;          // rows the mirror was built for (rebuilt whenever it differs from n)
    int64_t n_dev = 0;              // rows present in d_base / d_ids
    int64_t n_coded = 0;            // rows with valid codes in d_codes

    // PQ state
    bool trained = false;
    std::vector<float> h_train;     // samples collected by addVector (HnswPqIndex.java:170-175)
    int64_t n_train = 0;
    int64_t raw_rows = 0;           // rows inserted before training finished: their codes stay zero (HnswPqIndex.java:104,391)
    std::vector<uint8_t> h_codes_override;  // vdb_set_codes
    bool codes_overridden = false;

    // graph
    HostGraph g;
    int64_t n_indexed = 0;          // rows inserted into the graph
    bool graph_uploaded = false;
    int64_t adj_cap = 0;            // rows per level of d_adj (capacity: grows geometrically, levels keep their offsets in between)
    // live set of a graph index (Java: the `vectors` map). Rows are never reused; a removed row stays as a dead node.
    std::unordered_map<int32_t, int64_t> id2row;
    int64_t n_live = 0;
    bool dead_dirty = false;        // g.dead changed since d_dead was uploaded
    bool any_dead = false;
    DevBuf<uint32_t> d_dead;        // [ceil(n/32)] bitmap
    // Java `currentSize` / idToIndex (HnswPqIndex.java:393,467,589-591): the code row of a vector is currentSize at insert
    // time and currentSize is DEcremented by removeVector, so an insert after a removal overwrites the code row of an earlier
    // vector. Reproduced: h_jindex[row] = that index, rows sharing one index share the latest code.
    std::vector<int64_t> h_jindex;
    int64_t java_size = 0;
    std::unordered_multimap<int64_t, int64_t> jindex_rows;   // built at the first removal
    DevBuf<int32_t> d_adj_payload;  // incremental adjacency upload: changed rows + their (level,row) targets
    DevBuf<uint64_t> d_adj_targets;

    // search scratch
    DevBuf<float> d_q, d_out_d, d_lut, d_lutg;
    DevBuf<int> d_flag;
    DevBuf<int32_t> d_out_ids;
    DevBuf<uint64_t> d_keys, d_keys_tmp;
    DevBuf<uint32_t> d_rows, d_rows_tmp, d_visited;
    DevBuf<unsigned long long> d_counters;
    DevBuf<uint16_t> d_fq;          // flat GEMM scratch: bf16 query mirror, normalised queries, scores, thresholds, candidates
    DevBuf<float> d_fqf, d_fscores, d_ftau, d_fslab_d;
    DevBuf<int32_t> d_fslab_i;
    DevBuf<float2> d_fqstat;
    DevBuf<uint32_t> d_fcnt;
    DevBuf<uint64_t> d_fcand;
    PinnedBuf<float> p_q, p_d;
    PinnedBuf<int32_t> p_ids;
    // training scratch
    DevBuf<float> d_train, d_sub, d_mind;
    DevBuf<int> d_asg;

    vdb_stats stats{};

    // multi-device index (cfg.n_devices > 1, FLAT / PQ; multi_gpu.cu): this object is the front, the rows live in `shards`
    std::vector<vdb_index*> shards;         // one single-device index per GPU (device cfg.device + s), id-range shards
    std::vector<cudaEvent_t> shard_ev;      // per shard: "my part of the step is done"
    cudaEvent_t ev_front = nullptr;         // front (device cfg.device): "the gathered data is merged / the boundary is known"
    int64_t n_distributed = 0;              // host rows already handed to the shards
    DevBuf<uint64_t> g_keys, g_kth;         // front device: gathered shortlist keys [g][qb][R], global boundary [qb]
    DevBuf<uint64_t> d_kth;                 // shard: its copy of the boundary
    DevBuf<int> g_unc;                      // front device: number of queries whose boundary is uncertain with local shortlists
    int shortlist_override = 0;             // PQ shard: length of the LOCAL shortlist of the current sharded search (0 = max(rerank, k))
    DevBuf<float> g_d;                      // gathered results [g][qb][k]
    DevBuf<int32_t> g_i;
};

struct HnswSearchParams;
// vdb_api.cu: completes the launch shape of a beam-kernel call (traversal rows, round batch, shared memory, grid) and launches it
int vdb_graph_search_launch(vdb_index* x, HnswSearchParams& p, cudaStream_t st);
// graph_build.cu: rebuild the whole graph of a graph index on the GPU (batched inserts); the host graph adopts the result
int vdb_gpu_build_graph(vdb_index* x, int ef, int variant);
int vdb_pq_phase1(vdb_index* x, const float* d_q, int64_t qb, int k, cudaStream_t st);
int vdb_pq_phase2(vdb_index* x, const float* d_q, int64_t qb, int k, const uint64_t* d_kth, int32_t* d_ids, float* d_dists, cudaStream_t st);
// vdb_api.cu helpers the builder needs
int vdb_upload_rows_and_traversal_copy(vdb_index* x);
// multi_gpu.cu
int vdb_caught(const char* fn);      // catch (...) handler of the extern "C" entry points (vdb_api.cu)
int vdb_multi_create(vdb_index* front);
void vdb_multi_destroy(vdb_index* front);
int vdb_multi_train(vdb_index* front, const float* samples, int64_t ns);
int vdb_multi_sync(vdb_index* front);
int vdb_multi_rebuild(vdb_index* front);
int vdb_multi_fill_synthetic(vdb_index* front, uint64_t seed, int64_t cnt);
int vdb_multi_encode_all(vdb_index* front);
int vdb_multi_search(vdb_index* front, const float* q, bool q_on_device, int64_t nq, int k, int32_t* out_ids, float* out_d, cudaStream_t user_stream);
int vdb_multi_stats(vdb_index* front, vdb_stats* out);
// per-shard search without the handle's mutex juggling (vdb_api.cu)
int vdb_search_device_nolock(vdb_index* x, const float* d_q, int64_t nq, int k, int32_t* d_ids, float* d_dists, cudaStream_t st);
int vdb_sync_index_nolock(vdb_index* x);
