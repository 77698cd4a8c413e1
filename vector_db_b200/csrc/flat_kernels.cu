// flat_kernels.cu — exact brute-force search in the reference's own fp32 order (a6, a13) and the (dist,id) list merge used
// for chunked scans and the multi-GPU result merge (a8).
//
// This is the order-exact path: one thread per base row keeps the reduction sequential (R6) or 8-lane (R1), so ids AND
// distances are bit-identical to the CPU reference. The tensor-core shortlist path (flat_gemm.cu) feeds its candidates
// through the same re-score for large N.
#include "kernels.h"
#include <algorithm>

template <int NT>
__global__ void __launch_bounds__(NT)
flat_exact_kernel(const float* __restrict__ base, const float* __restrict__ norms, const int32_t* __restrict__ ids, int id_offset,
                  int64_t n, int dim, const float* __restrict__ queries, int k, int cap, int metric, int order,
                  int64_t rows_per_chunk, uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_rows,
                  const int* __restrict__ only_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (only_flag && !only_flag[blockIdx.y]) return;        // certified by the tensor-core path: nothing to do
    float* sq = reinterpret_cast<float*>(smem_raw);
    uint64_t* bkey = reinterpret_cast<uint64_t*>(sq + ((dim + 3) & ~3));
    uint32_t* brow = reinterpret_cast<uint32_t*>(bkey + cap);
    __shared__ int s_cnt;
    __shared__ uint64_t s_tau;
    __shared__ float s_qn;
    const int q = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
    for (int i = threadIdx.x; i < dim; i += NT) sq[i] = queries[(size_t)q * dim + i];
    for (int i = threadIdx.x; i < cap; i += NT) { bkey[i] = KEY_MAX; brow[i] = 0xffffffffu; }
    if (threadIdx.x == 0) { s_cnt = 0; s_tau = KEY_MAX; }
    __syncthreads();
    if (metric == 1) {
        // query normalised the Java way (HnswIndex.java:344)
        if (threadIdx.x == 0) s_qn = norm_java(sq, dim);
        __syncthreads();
        const float nrm = s_qn;
        __syncthreads();
        for (int i = threadIdx.x; i < dim; i += NT) sq[i] = __fdiv_rn(sq[i], nrm);
        __syncthreads();
        if (threadIdx.x == 0) s_qn = norm_java(sq, dim);
        __syncthreads();
    }
    const float qn = metric == 1 ? s_qn : 0.0f;
    const int64_t r0 = (int64_t)chunk * rows_per_chunk;
    const int64_t r1 = min(n, r0 + rows_per_chunk);
    for (int64_t b0 = r0; b0 < r1; b0 += NT) {
        const int64_t i = b0 + threadIdx.x;
        if (i < r1) {
            const float* row = base + i * dim;
            float d;
            if (metric == 1) d = cosine_dist_from_dot(dot_seq(row, sq, dim), norms[i], qn);
            else d = l2sq_order(sq, row, dim, order);
            const uint64_t key = make_key(d, ids ? ids[i] : (int32_t)(id_offset + i));
            if (key < s_tau) {
                int slot = atomicAdd(&s_cnt, 1);
                bkey[slot] = key; brow[slot] = (uint32_t)i;
            }
        }
        // one uniform decision for the whole CTA: a warp that raced ahead to the next iteration's atomicAdd must not be able to
        // change what a slower warp reads here (a divergent barrier inside the sort would follow)
        if (__syncthreads_or(s_cnt > cap - NT)) {
            block_bitonic_sort<true>(bkey, brow, cap);
            if (threadIdx.x == 0) { int c = min(s_cnt, k); s_cnt = c; if (c >= k) s_tau = bkey[k - 1]; }
            __syncthreads();
            for (int j = k + threadIdx.x; j < cap; j += NT) { bkey[j] = KEY_MAX; brow[j] = 0xffffffffu; }
            __syncthreads();
        }
    }
    block_bitonic_sort<true>(bkey, brow, cap);
    uint64_t* ok = out_keys + ((size_t)q * nchunks + chunk) * k;
    uint32_t* orow = out_rows + ((size_t)q * nchunks + chunk) * k;
    for (int j = threadIdx.x; j < k; j += NT) { ok[j] = bkey[j]; orow[j] = brow[j]; }
}

void launch_flat_search_exact(const float* d_base, const float* d_norms, const int32_t* d_ids, int id_offset, int64_t n, int dim,
                              const float* d_q, int nq, int k, int metric, int order, int nchunks, uint64_t* d_keys_tmp,
                              uint32_t* d_rows_tmp, uint64_t* d_keys, uint32_t* d_rows, cudaStream_t st, const int* only_flag) {
    constexpr int NT = 128;
    const int cap = next_pow2(k + NT);
    const size_t smem = (size_t)((dim + 3) & ~3) * 4 + (size_t)cap * 12;
    cudaFuncSetAttribute(flat_exact_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int64_t rpc = (n + nchunks - 1) / nchunks;
    dim3 grid(nchunks, nq);
    if (nchunks == 1) {
        flat_exact_kernel<NT><<<grid, NT, smem, st>>>(d_base, d_norms, d_ids, id_offset, n, dim, d_q, k, cap, metric, order, rpc, d_keys, d_rows, only_flag);
    } else {
        flat_exact_kernel<NT><<<grid, NT, smem, st>>>(d_base, d_norms, d_ids, id_offset, n, dim, d_q, k, cap, metric, order, rpc, d_keys_tmp, d_rows_tmp, only_flag);
        launch_merge_keys(d_keys_tmp, d_rows_tmp, nq, nchunks, k, d_keys, d_rows, st, only_flag);
    }
}

// merge g per-shard result lists [g][nq][k] into [nq][k] by (dist,id); id < 0 marks an empty slot. One CTA per query.
__global__ void merge_topk_kernel(const float* __restrict__ dists, const int32_t* __restrict__ ids, int g, int64_t nq, int k, int cap,
                                  float* __restrict__ out_d, int32_t* __restrict__ out_ids) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* bkey = reinterpret_cast<uint64_t*>(smem_raw);
    const int64_t q = blockIdx.x;
    const int total = g * k;
    for (int i = threadIdx.x; i < cap; i += blockDim.x) {
        uint64_t key = KEY_MAX;
        if (i < total) {
            const int s = i / k, j = i % k;
            const size_t off = ((size_t)s * nq + q) * k + j;
            const int32_t id = ids[off];
            if (id >= 0) key = make_key(dists[off], id);
        }
        bkey[i] = key;
    }
    block_bitonic_sort<false>(bkey, nullptr, cap);
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const uint64_t key = bkey[j];
        if (key == KEY_MAX) { out_ids[q * k + j] = -1; out_d[q * k + j] = __int_as_float(0x7f800000); }
        else { out_ids[q * k + j] = key_id(key); out_d[q * k + j] = key_dist(key); }
    }
}
void launch_merge_topk(const float* d_dists, const int32_t* d_ids, int g, int64_t nq, int k, float* d_out_d, int32_t* d_out_ids,
                       cudaStream_t st) {
    if (nq <= 0) return;
    const int cap = next_pow2(std::max(g * k, k));
    const size_t smem = (size_t)cap * 8;
    cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    merge_topk_kernel<<<(unsigned)nq, 128, smem, st>>>(d_dists, d_ids, g, nq, k, cap, d_out_d, d_out_ids);
}
