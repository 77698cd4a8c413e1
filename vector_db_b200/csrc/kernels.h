// kernels.h — launch entry points shared between the kernel translation units and the C-ABI layer.
#pragma once
#include "common.cuh"

// pq_kernels.cu
void launch_pq_lut(const float* d_q, const float* d_cb, int nq, int dim, int M, int arith, float* d_lut, cudaStream_t st);
void launch_adc_scan_topr(const uint8_t* d_codes, int64_t n, int M, const float* d_luts, int nq, const int32_t* d_ids,
                          int id_offset, int R, int nchunks, uint64_t* d_keys_tmp, uint32_t* d_rows_tmp, uint64_t* d_keys,
                          uint32_t* d_rows, cudaStream_t st, const int* only_flag = nullptr);
// adc_fast.cu
static inline int adc_mpad(int M) { return (M + 7) & ~7; }    // chunks per 16-row block of the transposed code layout
void launch_transpose_codes(const uint8_t* d_codes, int64_t n, int M, int64_t row0, uint8_t* d_codesT, cudaStream_t st);
void launch_lut_to_gather_layout(const float* d_lut, int nq, int M, float* d_lutg, cudaStream_t st);
int adc_fast_supported(int M, int R);
int adc_fast_r2(int R);
int launch_adc_scan_fast(const uint8_t* d_codesT, const uint8_t* d_codes, int64_t n, int M, const float* d_luts, const float* d_lutg, int nq,
                         const int32_t* d_ids, int id_offset, int R, int nchunks, uint64_t* d_keys_tmp, uint32_t* d_rows_tmp,
                         uint64_t* d_keys, uint32_t* d_rows, int* d_flag, cudaStream_t st);
void launch_merge_keys(const uint64_t* d_in_keys, const uint32_t* d_in_rows, int nq, int nlists, int R, uint64_t* d_keys,
                       uint32_t* d_rows, cudaStream_t st, const int* only_flag = nullptr);
void launch_keys_to_results(const uint64_t* d_keys, int R, int k, int64_t nq, int32_t* d_ids, float* d_dists, cudaStream_t st,
                            const int* only_flag = nullptr);
// shortlist entries whose key is above the per-query threshold kth[q] are removed (key = KEY_MAX, row = 0xffffffff)
void launch_mask_keys_above(uint64_t* d_keys, uint32_t* d_rows, int nq, int R, const uint64_t* d_kth, cudaStream_t st);
// kth[q] = the R-th smallest of the g * R keys keys[s][q][0..R) (each list ascending): the global shortlist boundary
void launch_select_kth_key(const uint64_t* d_keys, int g, int nq, int L, int rank, uint64_t* d_kth, int* d_uncertain, cudaStream_t st);
void launch_adc_scan_plain(const uint8_t* d_codes, int64_t n, int M, const float* d_lut, float* d_out, cudaStream_t st);
void launch_rerank(const float* d_base, const int32_t* d_ids, int id_offset, int dim, const float* d_q, int nq,
                   const uint32_t* d_rows, int R, int order, int k, int32_t* d_out_ids, float* d_out_d, cudaStream_t st);
int launch_pq_encode(const float* d_x, int64_t n, int dim, int M, const float* d_cb, int arith, uint8_t* d_codes, cudaStream_t st);
int launch_kmeans_train(const float* d_samples, int n, int dim, int M, int iters, int flavour, float* d_codebooks,
                        float* d_sub, float* d_mind, int* d_asg, cudaStream_t st);
void launch_fill_synth(float* d_out, uint64_t seed, int64_t start, int64_t count, cudaStream_t st);
void launch_normalize_rows(float* d_x, int64_t n, int dim, float* d_norms, cudaStream_t st);
void launch_distance(const float* d_a, const float* d_b, int64_t n, int dim, int metric, int order, float* d_out, cudaStream_t st);

// hnsw_kernels.cu
struct HnswSearchParams {
    const float* base;        // [n][dim]; normalised rows for cosine (exact re-score, greedy descent)
    const float* tnorm;       // [n] squared norm of the stored traversal row (bf16 / fp8 sq-L2 form)
    const float* tscale;      // [n] fp8 rows: 1 / (power-of-two scale the row was stored with)
    const void* tbase;        // traversal rows [n][tstride bytes]: fp32 or bf16, zero-padded to a multiple of 8 elements
    const float* norms;       // [n] Java norm of the stored row (cosine), else nullptr
    const int32_t* ids;       // [n] external ids
    const int32_t* adj;       // [max_level][adj_stride][deg], -1 padded
    const uint32_t* dead;     // [ceil(n/32)] bitmap of removed nodes (removeVector), or nullptr: pre-loaded into the visited bitset
    int64_t adj_stride;       // rows per level of adj (>= n: capacity, so incremental adds do not move the levels)
    const uint8_t* codes;     // [n][pq_m] or nullptr
    const float* codebooks;   // [pq_m][256][sd] or nullptr
    const float* queries;     // [nq][dim]
    int32_t* out_ids;         // [nq][k]
    float* out_dists;         // [nq][k]
    unsigned long long* counters;  // [16]: dist evals, expansions, then 8 phase cycle counters (profile)
    uint32_t* visited_scratch;     // [grid][ceil(n/32)] when the bitset does not fit in shared memory
    int n, dim, deg, max_level, entry;
    int pq_m, trained;
    int metric;               // 0 sq-L2 (HnswPqIndex), 1 cosine (HnswIndex)
    int variant;              // 0 searchLayerWithSize, 1 HnswPqIndex.searchLayer, 2 HnswIndex.searchLayer
    int ef, k, nq;
    int tstride, tstore;      // bytes per traversal row (multiple of 16); 0 = fp32, 1 = bf16, 2 = fp8 e4m3, 3 = int8
    int batch;                // candidates expanded per round (1 = the Java order)
    int profile;              // accumulate thread-0 cycle counters per phase into counters[2..9]
    int ctas_per_sm;          // CTAs per SM the launch is sized for (2 or 3): selects the kernel build (128 / 80 registers per thread)
    int* qcounter;            // zeroed before the launch: CTAs take their next query with atomicAdd (dynamic balance); NULL = static stride
    int visited_in_smem;
    int prefetch;             // pull fresh rows towards L2 as soon as they are discovered (traversal copy well beyond L2 only)
    // graph construction (graph_build.cu) runs the same kernel for the inserts of a batch:
    int beam_level;           // adjacency level the beam walks (0 for searches; l for the level-l links of new nodes)
    const int* qrows;         // query q = base row qrows[q] (new nodes of a batch that reach this level), or nullptr: queries[q]
    const int32_t* qentry;    // per-query start node of the beam (>= 0: no greedy descent: the nearest node found on the level above,
                              // HnswPqIndex.java:526-531), or nullptr / -1: greedy descent from `entry`
};
int launch_hnsw_search(const HnswSearchParams& p, int grid, cudaStream_t st);
size_t hnsw_search_smem_bytes(const HnswSearchParams& p);
int hnsw_max_batch(int deg);
// write `cnt` changed adjacency rows (payload[cnt][deg]) to their (level, row) targets: target = level << 32 | row
void launch_scatter_adj_rows(int32_t* d_adj, const int32_t* d_payload, const uint64_t* d_targets, int cnt, int deg, int64_t adj_stride,
                             cudaStream_t st);
// rows of d_base[n][dim] whose largest |element| exceeds `ratio` x rms(row): *d_count += that number (int8 suitability check)
void launch_count_outlier_rows(const float* d_base, int64_t n, int dim, float ratio, unsigned long long* d_count, cudaStream_t st);
void launch_make_traversal_rows(const float* d_base, int64_t n, int dim, int dpad, int mode /* 0 fp32, 1 bf16, 2 fp8, 3 int8 */, void* d_out, float* d_tnorm,
                                float* d_tscale, cudaStream_t st);

// flat_kernels.cu
void launch_flat_search_exact(const float* d_base, const float* d_norms, const int32_t* d_ids, int id_offset, int64_t n, int dim,
                              const float* d_q, int nq, int k, int metric, int order, int nchunks, uint64_t* d_keys_tmp,
                              uint32_t* d_rows_tmp, uint64_t* d_keys, uint32_t* d_rows, cudaStream_t st, const int* only_flag = nullptr);
void launch_merge_topk(const float* d_dists, const int32_t* d_ids, int g, int64_t nq, int k, float* d_out_d, int32_t* d_out_ids,
                       cudaStream_t st);

// flat_gemm.cu — tensor-core flat search (tcgen05 + TMA), certified against the order-exact re-score
struct FlatGemmPlan {
    int dpad;            // dim padded to a multiple of 64 (one 128-byte swizzle row of bf16)
    int64_t n_pad;       // rows padded to a multiple of 256 (length of the (a,b) array)
    int64_t step;        // sample stride
    int S, S_pad, P;     // sample rows, padded to 256, next power of two
    int s;               // threshold rank inside the sample
    int k2, k2p;         // rows re-scored exactly per query, next power of two
    uint32_t cap;        // candidate-list capacity per query
    int capP;
};
struct FlatGemmArgs {
    const float* d_base; const float* d_norms; const int32_t* d_ids; int id_offset; int64_t n; int dim; int metric; int order;
    int64_t n_total;             // rows of the whole index (n = rows of this slab; all row pointers are the slab's)
    int flag_or;                 // 1 for slabs after the first: OR the certificate flags
    const void* d_mirror;        // bf16 [n][dpad]
    const float* d_arow;         // [n_pad] additive term of the approximate score
    const float* d_colsum;       // [dim] column sums, [dim] = max ||row'||^2 bits
    const float* d_queries; int nq; int k;
    // scratch
    void* d_qmirror;             // bf16 [nq][dpad]
    float* d_qf;                 // [nq][dim] (cosine: normalised queries)
    float2* d_qstat;             // [nq]
    float* d_scores;             // [nq][S_pad]
    float* d_tau;                // [nq]
    uint32_t* d_cnt;             // [nq]
    uint64_t* d_cand;            // [nq][cap]
    int* d_flag;                 // [nq] 1 = not certified, answer with the exact scan
    unsigned long long* d_stats; // [2] certified, flagged
    int32_t* d_out_ids; float* d_out_d;
};
int flat_gemm_supported(int64_t n, int dim, int k);
int64_t flat_gemm_slab_rows(int64_t n, int dim, int k);
FlatGemmPlan flat_gemm_plan(int64_t n, int dim, int k);
int launch_flat_mirror(const float* d_base, int64_t n, int dim, int metric, void* d_mirror, float* d_arow, float* d_colsum, cudaStream_t st);
int launch_flat_gemm_search(const FlatGemmArgs& a, cudaStream_t st);
