/* include/vdb.h — C ABI of libvectordb.so, the B200 (sm_100a) implementation of VectorDatabase's batched
 * query path. This is the drop-in boundary: everything the reference's JNI bridge
 * (/root/reference/native/bridge/VectorDBJNI.cpp:39-200) does against its C++ `vectordb::VectorIndex`
 * (/root/reference/native/index/VectorIndex.h:9-35) is expressed here with plain pointers and sizes.
 * INTEGRATION.md shows the JNI / ctypes stubs that bind it.
 *
 * Conventions
 *   - every call returns an int status (VDB_OK == 0); nothing throws across the boundary (the reference lets
 *     C++ exceptions cross JNI frames: VectorDBJNI.cpp:132-159). vdb_last_error() gives the message of the
 *     last failure on the calling thread.
 *   - all pointers are caller-owned and borrowed for the duration of the call (reference: GetDirectBufferAddress
 *     / Get..ArrayElements with JNI_ABORT, VectorDBJNI.cpp:136-156,170-198).
 *   - result slots that hold no hit are filled with id = -1, distance = +inf (the reference forgets to:
 *     HNSWIndex.cpp:557-566 vs jni/NativeIndex.java:222).
 *   - there is NO CPU fallback: if no CUDA device / wrong architecture is present, vdb_create fails.
 */
#ifndef VDB_H_
#define VDB_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define VDB_OK 0
#define VDB_ERR_INVALID 1   /* bad argument / state (IllegalArgumentException on the Java side) */
#define VDB_ERR_CUDA 2      /* CUDA runtime failure, or no sm_100 device */
#define VDB_ERR_FULL 3      /* max_elements reached (VectorStore.cpp:40-43 throws; Java returns false) */
#define VDB_ERR_UNTRAINED 4 /* PQ index used before train (PQIndex.cpp:124-126 throws) */
#define VDB_ERR_IO 5

typedef struct vdb_index vdb_index;

/* which reference class the index reproduces */
enum {
    VDB_INDEX_FLAT = 0,    /* brute force (ground truth of RecallOptimizationTest.java:151-164; BatchDistance.cpp:121-147) */
    VDB_INDEX_PQ = 1,      /* native PQIndex: LUT + ADC scan + (dist,id) top-k   (native/index/PQIndex.cpp:146-213) */
    VDB_INDEX_HNSW = 2,    /* Java HnswIndex: cosine on normalised vectors       (index/HnswIndex.java) */
    VDB_INDEX_HNSW_PQ = 3  /* Java HnswPqIndex: sq-L2, PQ upper layers, exact layer 0 + rerank (index/HnswPqIndex.java) */
};
enum { VDB_METRIC_SQL2 = 0, VDB_METRIC_COSINE = 1 };
enum { VDB_TRAV_AUTO = 0, VDB_TRAV_F32 = 1, VDB_TRAV_FP8 = 2, VDB_TRAV_INT8 = 3, VDB_TRAV_BF16 = 4 };
/* fp32 reduction order of the reference implementation being reproduced */
enum {
    VDB_ARITH_JAVA_SEQ = 0,    /* sequential, mul then add (HnswPqIndex.java:741-750, :343-350) */
    VDB_ARITH_NATIVE_AVX2 = 1  /* 8-lane FMA + horizontal tree (DistanceAVX2.cpp:9-33 / :59-89) */
};
/* k-means flavour for vdb_train (oracle: orc_pq_train) */
enum {
    VDB_TRAIN_JAVA = 0,          /* HnswPqIndex.java:194-338  k-means++ with java.util.Random(42+m), 25 Lloyd passes */
    VDB_TRAIN_NATIVE_HNSWPQ = 1, /* HNSWPQIndex.cpp:127-244   k-means++ with std::mt19937(42+m), stop when stable   */
    VDB_TRAIN_NATIVE_PQ = 2      /* PQIndex.cpp:29-105        256 uniform draws with std::mt19937(42+m)             */
};

typedef struct vdb_config {
    int32_t dim;
    int32_t max_elements;
    int32_t index_type;      /* VDB_INDEX_* */
    int32_t metric;          /* VDB_METRIC_*; forced by index_type for HNSW (cosine) and HNSW_PQ / PQ (sq-L2) */
    int32_t arith;           /* VDB_ARITH_*; default: JAVA_SEQ for HNSW*, NATIVE_AVX2 for PQ */
    int32_t hnsw_m;          /* 32  (HnswPqIndex.java:27, HnswIndex.java:21) */
    int32_t ef_construction; /* 64 / 400 */
    int32_t ef_search;       /* 64 / 400 */
    int32_t pq_m;            /* subspaces; 0 = CompressionConfig.recommendedConfig(dim) (CompressionConfig.java:114-130) */
    int32_t pq_bits;         /* only 8 is supported (256 centroids, HnswPqIndex.java:86) */
    int32_t pq_iters;        /* 25 */
    int32_t train_flavour;   /* VDB_TRAIN_* */
    int32_t rerank;          /* PQ index: exact re-score of the best `rerank` ADC candidates (0 = none, as PQIndex.cpp) */
    int32_t device;          /* CUDA ordinal */
    int32_t id_offset;       /* multi-GPU id-range sharding: this shard holds global rows [id_offset, id_offset+n) */
    int32_t search_ef_cap;   /* HNSW_PQ: cap of the layer-0 beam; 0 = 5000, the constant of HnswPqIndex.java:621 */
    uint64_t level_seed;     /* seeded stand-in for ThreadLocalRandom level draws (HnswPqIndex.java:752-759) */
    int32_t trav_store;      /* HNSW*: rows read during traversal (VDB_TRAV_*): 0 = auto: int8 with a per-row scale when dim % 128 == 0
                                and (all but 0.1 % of) the rows pass the outlier check max|x| <= 6 rms(x), else the bf16 copy;
                                1 = fp32 (strict order); 2 = fp8 e4m3 with a per-row power-of-two scale (dim % 128 == 0);
                                3 = int8 always (dim % 128 == 0); 4 = bf16 always. Whatever the rows, the answer is re-scored
                                from the fp32 rows in the reference's order, so returned distances are exact */
    int32_t trav_batch;      /* HNSW*: candidates expanded per round: 0 = auto (ef/128 in [4,32]; 64 from ef 4096), 1 = the Java pop order */
    int32_t n_devices;       /* FLAT / PQ: GPUs of this box the rows are spread over by contiguous id range (0 or 1 = one GPU,
                                `device`); devices device..device+n_devices-1 are used from ONE process: one vdb_search_batch
                                fans the batch out to every shard and merges the lists (a JVM cannot run torchrun) */
    int32_t reserved2;
} vdb_config;

typedef struct vdb_stats {
    int64_t queries;
    int64_t dist_evals;      /* V: exact distances computed during traversal */
    int64_t expansions;      /* E: nodes expanded on layer 0 */
    int64_t adc_rows;        /* PQ code rows scanned */
    float last_kernel_ms;    /* device time of the dominant kernel of the last search (CUDA events) */
    float last_total_ms;     /* device time of the whole last search */
    int64_t flat_certified;  /* flat search: queries answered by the tensor-core shortlist and certified exact */
    int64_t flat_fallback;   /* flat search: queries that failed certification and were re-run by the order-exact scan */
    int32_t trav_format;     /* HNSW*: row format the traversal copy currently has (VDB_TRAV_F32 / FP8 / INT8 / BF16; what AUTO picked) */
    int32_t shortlist_retries;   /* n_devices > 1, PQ: queries repeated with full-length shard shortlists (0 on evenly spread rows) */
} vdb_stats;

const char* vdb_last_error(void);
const char* vdb_version(void);
/* fills *cfg with the defaults of the given reference class */
int vdb_default_config(int32_t index_type, int32_t dim, int32_t max_elements, vdb_config* cfg);

/* Device contract: every entry point that takes an index handle makes the index's device (cfg.device) current in the calling
 * thread with cudaSetDevice and leaves it current on return; callers that use several devices re-select their own afterwards.
 * An index with cfg.n_devices = g > 1 (FLAT / PQ only) owns id-range shards on devices cfg.device .. cfg.device + g - 1 and
 * returns with cfg.device current. */
/* nativeCreateHNSW / nativeCreatePQ (VectorDBJNI.cpp:39-65) — one constructor for all index types */
int vdb_create(const vdb_config* cfg, vdb_index** out);
/* nativeDestroy (VectorDBJNI.cpp:161-164) */
int vdb_destroy(vdb_index* idx);
/* nativeAdd / nativeAddBatch (VectorDBJNI.cpp:132-139,167-182): ids[n] (NULL = 0..n-1 continuing), vecs[n][dim] */
int vdb_add_batch(vdb_index* idx, const int32_t* ids, const float* vecs, int64_t n);
/* NativePqIndex.nativeTrain (VectorDBJNI.cpp:67-75): k-means on samples[n][dim] on the GPU */
int vdb_train(vdb_index* idx, const float* samples, int64_t n);
/* VectorIndex.buildIndex / VectorDatabase.rebuildIndex (HnswPqIndex.java:676-718): train if needed,
 * re-encode, rebuild the graph in ascending id order, upload */
int vdb_build(vdb_index* idx);
/* nativeSearch / nativeSearchBatch (VectorDBJNI.cpp:141-159,184-200): HOST buffers; queries[nq][dim],
 * out_ids[nq][k], out_dists[nq][k] */
int vdb_search_batch(vdb_index* idx, const float* queries, int64_t nq, int32_t k, int32_t* out_ids, float* out_dists);
/* same with DEVICE buffers on `stream` (a cudaStream_t taken verbatim: NULL = CUDA's legacy default stream, which is
 * also torch's default stream): no host copies of queries or results. One handle serves one search at a time: the scratch
 * buffers belong to the index, so calls are serialised by the handle's mutex on the host and, when consecutive calls use
 * different streams, ordered on the device by an event (they do not overlap). The first call after rows were added uploads
 * them and may grow scratch buffers, which synchronises the handle's own stream. */
int vdb_search_batch_device(vdb_index* idx, const float* d_queries, int64_t nq, int32_t k, int32_t* d_ids,
                            float* d_dists, void* stream);
/* VectorIndex.removeVector (index/HnswPqIndex.java:542-596, HnswIndex.java:254-317): HNSW / HNSW_PQ only — the node's own
 * lists are dropped, it is deleted from the lists of the nodes it links to, links pointing at it from elsewhere stay and are
 * skipped by every walk (`!vectors.containsKey`), the entry point moves to the first live key in ConcurrentHashMap order.
 * *removed = 1 when the id was present. FLAT / PQ: *removed = 0 (NativeIndex.removeVector answers false, NativeIndex.java:47-51) */
int vdb_remove(vdb_index* idx, int32_t id, int32_t* removed);
/* 1 when the id is stored and live (graph indexes; `vectors.containsKey(id)`, HnswPqIndex.java:157) */
int vdb_contains(vdb_index* idx, int32_t id);
/* live vectors (Java: vectors.size()) */
int64_t vdb_size(const vdb_index* idx);
/* stored rows including removed ones: the row count of vdb_get_codes / vdb_get_graph / vdb_set_graph buffers (== vdb_size
 * until the first vdb_remove; back to it after vdb_build) */
int64_t vdb_rows(const vdb_index* idx);
int vdb_is_trained(const vdb_index* idx);
int vdb_get_stats(vdb_index* idx, vdb_stats* out);

/* ---- state import/export: "the same persisted graph and codebooks" for parity runs, and save/load
 *      (every reference save()/load() is an empty TODO: HNSWIndex.cpp:535-541, HNSWPQIndex.cpp:774-780) ---- */
int vdb_set_codebooks(vdb_index* idx, const float* codebooks /* [pq_m][256][dim/pq_m] */);
int vdb_get_codebooks(vdb_index* idx, float* out);
int vdb_get_codes(vdb_index* idx, uint8_t* out /* [n][pq_m] */);
int vdb_set_codes(vdb_index* idx, const uint8_t* codes /* [n][pq_m] */);
/* levels[n]; adj[max_level][n][hnsw_m] (-1 padded, node rows in insertion order); entry = node row */
int vdb_set_graph(vdb_index* idx, const int32_t* levels, const int32_t* adj, int32_t max_level, int32_t entry);
int vdb_get_graph(vdb_index* idx, int32_t* levels, int32_t* adj, int32_t* entry);
int32_t vdb_max_level(const vdb_index* idx);
int vdb_save(vdb_index* idx, const char* path);
int vdb_load(vdb_index* idx, const char* path);

/* ---- device-side building blocks (benchmarks, multi-GPU merge, parity tests of single kernels) ---- */
/* vdb_merge_topk_device and vdb_distance_device take no index handle: they launch on the calling thread's CURRENT device,
 * which must be the device the pointers live on (cudaSetDevice before the call). */
/* fill rows [row0,row0+n) of the index with the counter-based synthetic generator (seed, (row)*dim+j) directly
 * in HBM (no PCIe traffic); ids = id_offset+row. Oracle twin: orc_fill_uniform. */
int vdb_fill_synthetic(vdb_index* idx, uint64_t seed, int64_t n);
/* encode all stored rows with the current codebooks on the GPU (PQ / HNSW_PQ) */
int vdb_encode_all(vdb_index* idx);
/* LUT build only: d_lut[nq][pq_m][256] */
int vdb_pq_lut_device(vdb_index* idx, const float* d_queries, int64_t nq, float* d_lut, void* stream);
/* ADC distances of every stored row for ONE query's LUT: d_out[n] (test hook, order R3) */
int vdb_adc_scan_device(vdb_index* idx, const float* d_lut, float* d_out, void* stream);
/* PQ search in two phases for id-range shards in SEPARATE processes (one per GPU, torch.distributed; inside one process
 * vdb_config.n_devices does all of this): phase 1 = LUT + ADC scan -> this shard's top-R keys d_keys_out[nq][R] (ordered ADC
 * distance << 32 | id, ascending; R = max(rerank, k), or the local length passed IN *shortlist_len, see below; the length used is
 * returned in *shortlist_len; nq <= 1024); the caller all-gathers the keys,
 * vdb_select_kth_key_device gives the GLOBAL R-th key per query from the g lists [g][nq][R]; phase 2 (vdb_pq_rerank_device, same
 * queries, right after phase 1 on the same handle) drops the shard's entries above that boundary and re-scores the rest exactly
 * -> the shard's top-k; vdb_merge_topk_device merges the g lists. The merged answer equals the single-GPU answer bit for bit. */
int vdb_pq_shortlist_device(vdb_index* idx, const float* d_queries, int64_t nq, int32_t k, uint64_t* d_keys_out, int32_t* shortlist_len, void* stream);
int vdb_pq_rerank_device(vdb_index* idx, const float* d_queries, int64_t nq, int32_t k, const uint64_t* d_kth /* [nq] or NULL */,
                         int32_t* d_ids, float* d_dists, void* stream);
int vdb_select_kth_key_device(const uint64_t* d_keys /* [g][nq][R] */, int32_t g, int64_t nq, int32_t R, uint64_t* d_kth, void* stream);
/* Shorter LOCAL shortlists (no reference counterpart: the reference has no shards). The scan's candidate bookkeeping grows with the
 * list length, and a shard of an evenly spread index holds only ~R/g of the global top-R, so sharded callers may pass a local length
 * L (k <= L < R, e.g. 2R/g + 64) in *shortlist_len of vdb_pq_shortlist_device (0 = R; the length used comes back in it and stays in
 * force for the following vdb_pq_rerank_device). The boundary of rank R among the g lists [g][nq][L] is then only certain if no shard
 * ran out of keys below it: this call also counts the uncertain queries in *d_uncertain (device int, zeroed by the caller; NULL = do
 * not check). A caller that finds it non-zero repeats the batch with full-length lists. */
int vdb_select_kth_key_ex_device(const uint64_t* d_keys /* [g][nq][list_len] */, int32_t g, int64_t nq, int32_t list_len, int32_t rank,
                                 uint64_t* d_kth, int32_t* d_uncertain, void* stream);
/* merge g per-shard result lists [g][nq][k] (dist,id) into [nq][k] ordered by (dist,id) */
int vdb_merge_topk_device(const float* d_dists, const int32_t* d_ids, int32_t g, int64_t nq, int32_t k,
                          float* d_out_dists, int32_t* d_out_ids, void* stream);
/* standalone exact distance, for known-answer tests: order = VDB_ARITH_*; metric as above */
int vdb_distance_device(const float* d_a, const float* d_b, int64_t n, int32_t dim, int32_t metric, int32_t order,
                        float* d_out, void* stream);

/* measured ceilings of the access patterns the gather-bound kernels use (csrc/microbench.cu; bench.py quotes roofline fractions
 * against them): kind 0 = random row gathers with the beam kernel's load shape, 1 = the same with 8 lines per row in flight,
 * 2 = the ADC scan's conflict-free shared-memory gather. out[0] = ms per launch (median of reps), out[1] = bytes (kinds 0, 1) or
 * gathers (kind 2) per launch. Allocates and frees its own buffers on the current device. */
int vdb_microbench_device(int32_t kind, int64_t rows, int32_t row_bytes, int32_t work, int32_t reps, double* out);

/* ---- no-allocation variant (no counterpart in the reference) ----
 * Order-exact flat search (squared L2) over caller-owned DEVICE rows d_base[n][dim] for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it): same kernels and same answer as a VDB_INDEX_FLAT
 * index (ids and distance bits, ties by id; global id = row number), but the call allocates NOTHING — no cudaMalloc /
 * cudaFree / pinned memory, no allocating library routine. Its scratch (per-chunk candidate lists) lives in one fixed 8 MiB
 * `static __device__` array inside the library, held by every GPU that loads the library until the process exits.
 *   limits : nq * k <= ~340 000 (scratch), k <= 1024, dim <= 8192, nq <= 65535; larger requests fail with VDB_ERR_INVALID and
 *            a message that states the limit.
 *   sharing: every call on a GPU uses the same array and finds the previous call's contents in it (nothing needs clearing:
 *            every slot that is read is written first by the same call). Calls that OVERLAP on one GPU — two streams, two
 *            host threads — race on it: issue one call at a time per GPU, or order them on one stream.
 *   device : the rows' device must be current (cudaSetDevice) in the calling thread; `order` = VDB_ARITH_*; `stream` as above. */
int vdb_flat_search_noalloc_device(const float* d_base, int64_t n, int32_t dim, int32_t order, const float* d_queries, int64_t nq,
                                   int32_t k, int32_t* d_out_ids, float* d_out_dists, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VDB_H_ */
