// pq_kernels.cu — product-quantisation kernels: LUT build (a1), ADC code scan + top-R (a2, a8), exact rerank (a7),
// encode (a10), k-means training (a9). SURVEY.md §8a row numbers in parentheses.
#include "kernels.h"
#include <algorithm>

// =====================================================================================================
// a1  LUT build:  lut[q][m][c] = || q[m*sd:(m+1)*sd] - codebook[m][c] ||^2
//     native order R2 (PQIndex.cpp:157-163 via batchEuclideanDistanceAVX2) or Java sequential (PqIndex.java:306-330)
// =====================================================================================================
__global__ void pq_lut_kernel(const float* __restrict__ queries, const float* __restrict__ codebooks, int dim, int M,
                              int arith, float* __restrict__ lut) {
    const int q = blockIdx.x;
    const int sd = dim / M;
    extern __shared__ float sq[];
    for (int i = threadIdx.x; i < dim; i += blockDim.x) sq[i] = queries[(size_t)q * dim + i];
    __syncthreads();
    const int total = M * VDB_NCENT;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
        const int m = e >> 8;
        const float* cen = codebooks + (size_t)e * sd;
        float v = arith == ARITH_NATIVE_AVX2 ? l2sq_r2(sq + m * sd, cen, sd) : l2sq_seq(sq + m * sd, cen, sd);
        lut[(size_t)q * total + e] = v;
    }
}
void launch_pq_lut(const float* d_q, const float* d_cb, int nq, int dim, int M, int arith, float* d_lut, cudaStream_t st) {
    if (nq <= 0) return;
    pq_lut_kernel<<<nq, 256, dim * sizeof(float), st>>>(d_q, d_cb, dim, M, arith, d_lut);
}

// =====================================================================================================
// a2  ADC scan, generic version (any M): one thread per code row, dist = dist + lut[m][code[m]] for m ascending —
//     exactly order R3 (PQIndex.cpp:173-198), so distances are bit-identical to the reference.
//     Per-CTA streaming top-R with a threshold + shared candidate buffer; keys order by (dist, id).
// =====================================================================================================
template <int NT>
__global__ void __launch_bounds__(NT)
adc_scan_topr_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, const float* __restrict__ luts,
                     const int32_t* __restrict__ ids, int id_offset, int R, int cap, int64_t rows_per_chunk,
                     uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_rows, const int* __restrict__ only_flag) {
    if (only_flag && !only_flag[blockIdx.y]) return;      // fallback launch: only queries the fast scan could not certify
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* lut = reinterpret_cast<float*>(smem_raw);
    uint64_t* bkey = reinterpret_cast<uint64_t*>(lut + (size_t)M * VDB_NCENT);
    uint32_t* brow = reinterpret_cast<uint32_t*>(bkey + cap);
    __shared__ int s_cnt;
    __shared__ uint64_t s_tau;
    const int q = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
    const float* gl = luts + (size_t)q * M * VDB_NCENT;
    for (int i = threadIdx.x; i < M * VDB_NCENT; i += NT) lut[i] = gl[i];
    for (int i = threadIdx.x; i < cap; i += NT) { bkey[i] = KEY_MAX; brow[i] = 0xffffffffu; }
    if (threadIdx.x == 0) { s_cnt = 0; s_tau = KEY_MAX; }
    __syncthreads();
    const int64_t r0 = (int64_t)chunk * rows_per_chunk;
    const int64_t r1 = min(n, r0 + rows_per_chunk);
    for (int64_t base = r0; base < r1; base += NT) {
        const int64_t i = base + threadIdx.x;
        if (i < r1) {
            const uint8_t* p = codes + i * M;
            float d = 0.0f;
            for (int m = 0; m < M; m++) d = __fadd_rn(d, lut[m * VDB_NCENT + p[m]]);
            const int32_t id = ids ? ids[i] : (int32_t)(id_offset + i);
            const uint64_t key = make_key(d, id);
            if (key < s_tau) {
                int slot = atomicAdd(&s_cnt, 1);
                bkey[slot] = key; brow[slot] = (uint32_t)i;
            }
        }
        __syncthreads();
        if (s_cnt > cap - NT) {                       // uniform: s_cnt only changes between barriers
            block_bitonic_sort<true>(bkey, brow, cap);
            if (threadIdx.x == 0) {
                int c = min(s_cnt, R);
                s_cnt = c;
                if (c >= R) s_tau = bkey[R - 1];
            }
            __syncthreads();
            for (int j = R + threadIdx.x; j < cap; j += NT) { bkey[j] = KEY_MAX; brow[j] = 0xffffffffu; }
            __syncthreads();
        }
    }
    block_bitonic_sort<true>(bkey, brow, cap);
    uint64_t* ok = out_keys + ((size_t)q * nchunks + chunk) * R;
    uint32_t* orow = out_rows + ((size_t)q * nchunks + chunk) * R;
    for (int j = threadIdx.x; j < R; j += NT) { ok[j] = bkey[j]; orow[j] = brow[j]; }
}

// merge `nlists` sorted-or-not lists of R (key,row) per query into the R best; one CTA per query.
__global__ void merge_keys_kernel(const uint64_t* __restrict__ in_keys, const uint32_t* __restrict__ in_rows, int nlists,
                                  int R, int cap, uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_rows,
                                  const int* __restrict__ only_flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* bkey = reinterpret_cast<uint64_t*>(smem_raw);
    uint32_t* brow = reinterpret_cast<uint32_t*>(bkey + cap);
    const int q = blockIdx.x;
    if (only_flag && !only_flag[q]) return;
    const int total = nlists * R;
    for (int i = threadIdx.x; i < cap; i += blockDim.x) {
        bkey[i] = i < total ? in_keys[(size_t)q * total + i] : KEY_MAX;
        brow[i] = i < total ? in_rows[(size_t)q * total + i] : 0xffffffffu;
    }
    block_bitonic_sort<true>(bkey, brow, cap);
    for (int j = threadIdx.x; j < R; j += blockDim.x) { out_keys[(size_t)q * R + j] = bkey[j]; out_rows[(size_t)q * R + j] = brow[j]; }
}

// keys -> (ids, dists), first k of each query's R
__global__ void keys_to_results_kernel(const uint64_t* __restrict__ keys, int R, int k, int64_t nq, int32_t* __restrict__ out_ids,
                                       float* __restrict__ out_d, const int* __restrict__ only_flag) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq * k) return;
    int64_t q = t / k; int j = (int)(t % k);
    if (only_flag && !only_flag[q]) return;
    uint64_t key = j < R ? keys[q * R + j] : KEY_MAX;
    if (key == KEY_MAX) { out_ids[t] = -1; out_d[t] = __int_as_float(0x7f800000); }
    else { out_ids[t] = key_id(key); out_d[t] = key_dist(key); }
}

void launch_merge_keys(const uint64_t* d_in_keys, const uint32_t* d_in_rows, int nq, int nlists, int R, uint64_t* d_keys,
                       uint32_t* d_rows, cudaStream_t st, const int* only_flag) {
    const int mcap = next_pow2(nlists * R);
    const size_t msmem = (size_t)mcap * 12;
    cudaFuncSetAttribute(merge_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem);
    merge_keys_kernel<<<nq, 256, msmem, st>>>(d_in_keys, d_in_rows, nlists, R, mcap, d_keys, d_rows, only_flag);
}
void launch_adc_scan_topr(const uint8_t* d_codes, int64_t n, int M, const float* d_luts, int nq, const int32_t* d_ids,
                          int id_offset, int R, int nchunks, uint64_t* d_keys_tmp, uint32_t* d_rows_tmp, uint64_t* d_keys,
                          uint32_t* d_rows, cudaStream_t st, const int* only_flag) {
    constexpr int NT = 256;
    if (only_flag) nchunks = 1;
    const int cap = next_pow2(R + NT);
    const size_t smem = (size_t)M * VDB_NCENT * 4 + (size_t)cap * 12;
    cudaFuncSetAttribute(adc_scan_topr_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int64_t rpc = (n + nchunks - 1) / nchunks;
    dim3 grid(nchunks, nq);
    if (nchunks == 1) {
        adc_scan_topr_kernel<NT><<<grid, NT, smem, st>>>(d_codes, n, M, d_luts, d_ids, id_offset, R, cap, rpc, d_keys, d_rows, only_flag);
    } else {
        adc_scan_topr_kernel<NT><<<grid, NT, smem, st>>>(d_codes, n, M, d_luts, d_ids, id_offset, R, cap, rpc, d_keys_tmp, d_rows_tmp, nullptr);
        launch_merge_keys(d_keys_tmp, d_rows_tmp, nq, nchunks, R, d_keys, d_rows, st);
    }
}
void launch_keys_to_results(const uint64_t* d_keys, int R, int k, int64_t nq, int32_t* d_ids, float* d_dists, cudaStream_t st,
                            const int* only_flag) {
    int64_t tot = nq * k;
    if (tot <= 0) return;
    keys_to_results_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(d_keys, R, k, nq, d_ids, d_dists, only_flag);
}

__global__ void mask_keys_above_kernel(uint64_t* __restrict__ keys, uint32_t* __restrict__ rows, int nq, int R, const uint64_t* __restrict__ kth) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)nq * R) return;
    if (keys[t] > kth[t / R]) { keys[t] = KEY_MAX; rows[t] = 0xffffffffu; }
}
void launch_mask_keys_above(uint64_t* d_keys, uint32_t* d_rows, int nq, int R, const uint64_t* d_kth, cudaStream_t st) {
    const int64_t tot = (int64_t)nq * R;
    if (tot > 0) mask_keys_above_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(d_keys, d_rows, nq, R, d_kth);
}
// one CTA per query: the g sorted lists (L keys each) are merged by sorting their g * L keys in shared memory; element rank-1 is
// the boundary. When the lists are SHORTER than the rank asked for (L < rank: shards keep a local shortlist of ~2 rank / g), the
// boundary is only certain if no shard ran out of keys below it: a full list whose last key is below the boundary may hide more.
// Such queries are counted in *uncertain (the caller repeats them with full-length lists).
__global__ void select_kth_key_kernel(const uint64_t* __restrict__ keys, int g, int nq, int L, int rank, int cap, uint64_t* __restrict__ kth,
                                      int* __restrict__ uncertain) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* b = reinterpret_cast<uint64_t*>(smem_raw);
    const int q = blockIdx.x, total = g * L;
    for (int i = threadIdx.x; i < cap; i += blockDim.x) b[i] = i < total ? keys[((size_t)(i / L) * nq + q) * L + (i % L)] : KEY_MAX;
    block_bitonic_sort<false>(b, nullptr, cap);
    if (threadIdx.x == 0) {
        const uint64_t bound = rank - 1 < cap ? b[rank - 1] : KEY_MAX;
        kth[q] = bound;
        if (uncertain && L < rank) {
            bool bad = false;
            for (int s = 0; s < g; s++) {
                const uint64_t last = keys[((size_t)s * nq + q) * L + (L - 1)];
                if (last != KEY_MAX && last < bound) bad = true;
            }
            if (bad) atomicAdd(uncertain, 1);
        }
    }
}
void launch_select_kth_key(const uint64_t* d_keys, int g, int nq, int L, int rank, uint64_t* d_kth, int* d_uncertain, cudaStream_t st) {
    const int cap = next_pow2(g * L);
    const size_t smem = (size_t)cap * 8;
    cudaFuncSetAttribute(select_kth_key_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    select_kth_key_kernel<<<nq, 256, smem, st>>>(d_keys, g, nq, L, rank, cap, d_kth, d_uncertain);
}

// plain ADC distances for one LUT (test hook; order R3)
__global__ void adc_scan_plain_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, const float* __restrict__ lut_g,
                                      float* __restrict__ out) {
    extern __shared__ float lut[];
    for (int i = threadIdx.x; i < M * VDB_NCENT; i += blockDim.x) lut[i] = lut_g[i];
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t* p = codes + i * M;
        float d = 0.0f;
        for (int m = 0; m < M; m++) d = __fadd_rn(d, lut[m * VDB_NCENT + p[m]]);
        out[i] = d;
    }
}
void launch_adc_scan_plain(const uint8_t* d_codes, int64_t n, int M, const float* d_lut, float* d_out, cudaStream_t st) {
    const size_t smem = (size_t)M * VDB_NCENT * 4;
    cudaFuncSetAttribute(adc_scan_plain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 4);
    if (blocks < 1) blocks = 1;
    adc_scan_plain_kernel<<<blocks, 256, smem, st>>>(d_codes, n, M, d_lut, d_out);
}

// =====================================================================================================
// a7  exact rerank: recompute the exact distance of each shortlisted row in the reference's order (one thread per
//     candidate keeps the order sequential), then (dist,id) top-k.   HnswPqIndex.java:646-664 / HNSWPQIndex.cpp:617-655
// =====================================================================================================
__global__ void __launch_bounds__(256)
rerank_kernel(const float* __restrict__ base, const int32_t* __restrict__ ids, int id_offset, int dim,
              const float* __restrict__ queries, const uint32_t* __restrict__ rows, int R, int cap, int order,
              int k, int32_t* __restrict__ out_ids, float* __restrict__ out_d) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sq = reinterpret_cast<float*>(smem_raw);
    uint64_t* bkey = reinterpret_cast<uint64_t*>(sq + ((dim + 3) & ~3));
    float* tiles = reinterpret_cast<float*>(bkey + cap);
    const int q = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < dim; i += blockDim.x) sq[i] = queries[(size_t)q * dim + i];
    __syncthreads();
    // 32 candidates per warp pass: coalesced row reads, per-row sums in the reference order (rescore32, common.cuh)
    float* tile = tiles + warp * RS_TILE;
    for (int g = warp; g * 32 < cap; g += 8) {
        const int j = g * 32 + lane;
        uint32_t row = j < R ? rows[(size_t)q * R + j] : 0xffffffffu;
        const unsigned live = __ballot_sync(0xffffffffu, row != 0xffffffffu);
        if (!live) { bkey[j] = KEY_MAX; continue; }
        const uint32_t any = __shfl_sync(0xffffffffu, row, __ffs(live) - 1);      // dead lanes walk a live row (result unused)
        const bool ok = row != 0xffffffffu;
        if (!ok) row = any;
        const float d = order == ORDER_R1 ? rescore32<1>(base, dim, sq, row, tile, lane) : rescore32<0>(base, dim, sq, row, tile, lane);
        bkey[j] = ok ? make_key(d, ids ? ids[row] : (int32_t)(id_offset + row)) : KEY_MAX;
    }
    block_bitonic_sort<false>(bkey, nullptr, cap);
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        uint64_t key = j < cap ? bkey[j] : KEY_MAX;
        if (key == KEY_MAX) { out_ids[(size_t)q * k + j] = -1; out_d[(size_t)q * k + j] = __int_as_float(0x7f800000); }
        else { out_ids[(size_t)q * k + j] = key_id(key); out_d[(size_t)q * k + j] = key_dist(key); }
    }
}
void launch_rerank(const float* d_base, const int32_t* d_ids, int id_offset, int dim, const float* d_q, int nq,
                   const uint32_t* d_rows, int R, int order, int k, int32_t* d_out_ids, float* d_out_d, cudaStream_t st) {
    const int cap = std::max(32, next_pow2(R));
    const size_t smem = (size_t)((dim + 3) & ~3) * 4 + (size_t)cap * 8 + (size_t)8 * RS_TILE * 4;
    cudaFuncSetAttribute(rerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    rerank_kernel<<<nq, 256, smem, st>>>(d_base, d_ids, id_offset, dim, d_q, d_rows, R, cap, order, k, d_out_ids, d_out_d);
}

// =====================================================================================================
// a10 PQ encode: code[i][m] = argmin_c subdist(x[i][m], codebook[m][c]), strict '<', lowest index wins.
//     grid (row tiles, M); the subspace's 256 centroids are staged in shared memory and broadcast.
// =====================================================================================================
__device__ __forceinline__ int nearest_centroid_dev(const float* __restrict__ cb, const float* __restrict__ x, int sd, int arith) {
    int best = 0;
    float bd = 3.402823466e+38f;   // FLT_MAX (HnswPqIndex.java:327, HNSWPQIndex.cpp:248)
    for (int c = 0; c < VDB_NCENT; c++) {
        float d = arith == ARITH_NATIVE_AVX2 ? l2sq_r1(x, cb + c * sd, sd) : l2sq_seq(x, cb + c * sd, sd);
        if (d < bd) { bd = d; best = c; }
    }
    return best;
}
#define ENC_MAX_SD 32
__global__ void pq_encode_kernel(const float* __restrict__ x, int64_t n, int dim, int M, const float* __restrict__ codebooks,
                                 int arith, uint8_t* __restrict__ codes) {
    extern __shared__ float cb[];
    const int m = blockIdx.y, sd = dim / M;
    for (int i = threadIdx.x; i < VDB_NCENT * sd; i += blockDim.x) cb[i] = codebooks[(size_t)m * VDB_NCENT * sd + i];
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float xv[ENC_MAX_SD];
        const float* xp = x + i * dim + (size_t)m * sd;
        for (int d = 0; d < sd; d++) xv[d] = xp[d];
        codes[i * M + m] = (uint8_t)nearest_centroid_dev(cb, xv, sd, arith);
    }
}
int launch_pq_encode(const float* d_x, int64_t n, int dim, int M, const float* d_cb, int arith, uint8_t* d_codes, cudaStream_t st) {
    const int sd = dim / M;
    if (sd > ENC_MAX_SD) return 1;
    if (n <= 0) return 0;
    int bx = (int)std::min<int64_t>((n + 255) / 256, 65535);
    dim3 grid(bx, M);
    pq_encode_kernel<<<grid, 256, (size_t)VDB_NCENT * sd * 4, st>>>(d_x, n, dim, M, d_cb, arith, d_codes);
    return 0;
}

// =====================================================================================================
// a9  k-means training, one CTA per subspace. Every floating-point reduction whose order is observable in the
//     reference (k-means++ totals / cumulative sums, centroid sums) is done sequentially in sample-index order by one
//     thread; the embarrassingly parallel parts (distances, argmin) are spread over the CTA.
//     flavour 0: HnswPqIndex.java:194-338   1: HNSWPQIndex.cpp:127-244   2: PQIndex.cpp:29-105
// =====================================================================================================
struct JavaRandomDev {
    uint64_t seed;
    __device__ void init(int64_t s) { seed = ((uint64_t)s ^ 0x5DEECE66Dull) & ((1ull << 48) - 1); }
    __device__ int32_t next(int bits) {
        seed = (seed * 0x5DEECE66Dull + 0xBull) & ((1ull << 48) - 1);
        return (int32_t)((int64_t)seed >> (48 - bits));
    }
    __device__ int32_t nextInt(int32_t bound) {
        int32_t r = next(31);
        int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        for (int32_t u = r; (int32_t)((uint32_t)u - (uint32_t)(r = u % bound) + (uint32_t)m) < 0; u = next(31)) {}
        return r;
    }
    __device__ float nextFloat() { return __fdiv_rn((float)next(24), 16777216.0f); }
};
// std::mt19937 with libstdc++'s uniform_int_distribution<int> (Lemire) and uniform_real_distribution<float>
struct Mt19937Dev {
    uint32_t* mt; int idx;
    __device__ void init(uint32_t* state, uint32_t s) {
        mt = state; mt[0] = s;
        for (int i = 1; i < 624; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
        idx = 624;
    }
    __device__ uint32_t gen() {
        if (idx >= 624) {
            for (int i = 0; i < 624; i++) {
                uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
                mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
        return y;
    }
    __device__ int uniform_int(int lo, int hi) {
        uint32_t range = (uint32_t)(hi - lo) + 1u;
        uint64_t product = (uint64_t)gen() * (uint64_t)range;
        uint32_t low = (uint32_t)product;
        if (low < range) {
            uint32_t threshold = (0u - range) % range;
            while (low < threshold) { product = (uint64_t)gen() * (uint64_t)range; low = (uint32_t)product; }
        }
        return lo + (int)(product >> 32);
    }
    __device__ float uniform_real(float a, float b) {
        float sum = __uint2float_rn(gen());
        float ret = __fdiv_rn(sum, 4294967296.0f);
        if (ret >= 1.0f) ret = __uint_as_float(0x3f7fffffu);
        return __fadd_rn(__fmul_rn(ret, __fsub_rn(b, a)), a);
    }
};

#define KM_NT 512
// scratch per subspace in global memory: sub[n][sd] (gathered sub-vectors), mind[n], asg[n] (int), newc[256][sd]
__global__ void __launch_bounds__(KM_NT)
kmeans_train_kernel(const float* __restrict__ samples, int n, int dim, int M, int iters, int flavour,
                    float* __restrict__ codebooks, float* __restrict__ g_sub, float* __restrict__ g_mind, int* __restrict__ g_asg) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int m = blockIdx.x, sd = dim / M, tid = threadIdx.x;
    float* cb = reinterpret_cast<float*>(smem_raw);            // [256][sd] current centroids
    float* acc = cb + VDB_NCENT * sd;                          // [256][sd] sums
    int* cnt = reinterpret_cast<int*>(acc + VDB_NCENT * sd);   // [256]
    uint32_t* mtstate = reinterpret_cast<uint32_t*>(cnt + VDB_NCENT);  // [624]
    __shared__ int s_sel, s_changed;
    float* sub = g_sub + (size_t)m * n * sd;
    float* mind = g_mind + (size_t)m * n;
    int* asg = g_asg + (size_t)m * n;
    const int arith = flavour == 0 ? ARITH_JAVA_SEQ : ARITH_NATIVE_AVX2;

    for (int i = tid; i < n * sd; i += KM_NT) { int r = i / sd, d = i % sd; sub[i] = samples[(size_t)r * dim + (size_t)m * sd + d]; }
    for (int i = tid; i < VDB_NCENT * sd; i += KM_NT) cb[i] = 0.0f;
    for (int i = tid; i < n; i += KM_NT) { mind[i] = 3.402823466e+38f; asg[i] = 0; }
    __syncthreads();

    JavaRandomDev jr; Mt19937Dev mt;
    if (tid == 0) {
        if (flavour == 0) { jr.init(42 + m); s_sel = jr.nextInt(n); }
        else { mt.init(mtstate, 42u + (uint32_t)m); s_sel = mt.uniform_int(0, n - 1); }
    }
    __syncthreads();
    if (flavour == 2) {
        // PQIndex.cpp:52-60: every centroid is an independent uniform draw
        for (int c = 0; c < VDB_NCENT; c++) {
            int sel = s_sel;
            for (int d = tid; d < sd; d += KM_NT) cb[c * sd + d] = sub[(size_t)sel * sd + d];
            __syncthreads();
            if (tid == 0 && c + 1 < VDB_NCENT) s_sel = mt.uniform_int(0, n - 1);
            __syncthreads();
        }
    } else {
        for (int d = tid; d < sd; d += KM_NT) cb[d] = sub[(size_t)s_sel * sd + d];
        __syncthreads();
        const int cmax = flavour == 0 ? min(VDB_NCENT, n) : VDB_NCENT;     // HnswPqIndex.java:266 bounds c by n too
        for (int c = 1; c < cmax; c++) {
            for (int i = tid; i < n; i += KM_NT) {
                float d = arith == ARITH_NATIVE_AVX2 ? l2sq_r1(sub + (size_t)i * sd, cb + (c - 1) * sd, sd)
                                                     : l2sq_seq(sub + (size_t)i * sd, cb + (c - 1) * sd, sd);
                if (d < mind[i]) mind[i] = d;
            }
            __syncthreads();
            if (tid == 0) {
                float total = 0.0f;
                for (int i = 0; i < n; i++) total = __fadd_rn(total, mind[i]);
                float target = flavour == 0 ? __fmul_rn(jr.nextFloat(), total) : mt.uniform_real(0.0f, total);
                float cum = 0.0f; int sel = 0;
                for (int i = 0; i < n; i++) { cum = __fadd_rn(cum, mind[i]); if (cum >= target) { sel = i; break; } }
                s_sel = sel;
            }
            __syncthreads();
            for (int d = tid; d < sd; d += KM_NT) cb[c * sd + d] = sub[(size_t)s_sel * sd + d];
            __syncthreads();
        }
    }
    // Lloyd
    for (int it = 0; it < iters; it++) {
        if (tid == 0) s_changed = 0;
        __syncthreads();
        int changed = 0;
        for (int i = tid; i < n; i += KM_NT) {
            float xv[ENC_MAX_SD];
            for (int d = 0; d < sd; d++) xv[d] = sub[(size_t)i * sd + d];
            int nn = nearest_centroid_dev(cb, xv, sd, arith);
            int prev = flavour == 0 ? 0 : asg[i];                 // Java re-zeroes assignments each pass (:234)
            if (prev != nn) changed = 1;
            asg[i] = nn;
        }
        if (changed) s_changed = 1;
        __syncthreads();
        if (flavour == 0) { if (!s_changed && it > 5) break; }
        else if (!s_changed) break;
        // M-step: thread (c,d) walks all samples in index order so each centroid component is summed sequentially
        for (int e = tid; e < VDB_NCENT * sd; e += KM_NT) {
            const int c = e / sd, d = e % sd;
            float s = 0.0f; int k = 0;
            for (int i = 0; i < n; i++) if (asg[i] == c) { s = __fadd_rn(s, sub[(size_t)i * sd + d]); k++; }
            acc[e] = s;
            if (d == 0) cnt[c] = k;
        }
        __syncthreads();
        for (int e = tid; e < VDB_NCENT * sd; e += KM_NT) {
            const int c = e / sd;
            if (cnt[c] > 0) cb[e] = __fmul_rn(acc[e], __fdiv_rn(1.0f, (float)cnt[c]));
            else if (flavour != 0) cb[e] = 0.0f;                 // native: empty cluster becomes the zero vector
        }
        __syncthreads();
    }
    for (int i = tid; i < VDB_NCENT * sd; i += KM_NT) codebooks[(size_t)m * VDB_NCENT * sd + i] = cb[i];
}
int launch_kmeans_train(const float* d_samples, int n, int dim, int M, int iters, int flavour, float* d_codebooks,
                        float* d_sub, float* d_mind, int* d_asg, cudaStream_t st) {
    const int sd = dim / M;
    if (sd > ENC_MAX_SD) return 1;
    const size_t smem = (size_t)2 * VDB_NCENT * sd * 4 + VDB_NCENT * 4 + 624 * 4;
    cudaFuncSetAttribute(kmeans_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kmeans_train_kernel<<<M, KM_NT, smem, st>>>(d_samples, n, dim, M, iters, flavour, d_codebooks, d_sub, d_mind, d_asg);
    return 0;
}

// =====================================================================================================
// synthetic fill / normalisation helpers
// =====================================================================================================
__global__ void fill_synth_kernel(float* __restrict__ out, uint64_t seed, int64_t start, int64_t count) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = vdb_synth(seed, (uint64_t)(start + i));
}
void launch_fill_synth(float* d_out, uint64_t seed, int64_t start, int64_t count, cudaStream_t st) {
    if (count <= 0) return;
    int blocks = (int)std::min<int64_t>((count + 255) / 256, 148 * 16);
    fill_synth_kernel<<<blocks, 256, 0, st>>>(d_out, seed, start, count);
}
// Java Vector.normalize + getNorm of the normalised copy (core/Vector.java:93-116), one thread per row (sequential sums)
__global__ void normalize_rows_kernel(float* __restrict__ x, int64_t n, int dim, float* __restrict__ norms_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float* r = x + i * dim;
    float nrm = norm_java(r, dim);
    for (int d = 0; d < dim; d++) r[d] = __fdiv_rn(r[d], nrm);
    if (norms_out) norms_out[i] = norm_java(r, dim);
}
void launch_normalize_rows(float* d_x, int64_t n, int dim, float* d_norms, cudaStream_t st) {
    if (n <= 0) return;
    normalize_rows_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_x, n, dim, d_norms);
}
__global__ void distance_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n, int dim, int metric, int order,
                                float* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* x = a + i * dim; const float* y = b + i * dim;
    if (metric == 1) out[i] = cosine_dist_from_dot(dot_seq(x, y, dim), norm_java(x, dim), norm_java(y, dim));
    else out[i] = l2sq_order(x, y, dim, order);
}
void launch_distance(const float* d_a, const float* d_b, int64_t n, int dim, int metric, int order, float* d_out, cudaStream_t st) {
    if (n <= 0) return;
    distance_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_a, d_b, n, dim, metric, order, d_out);
}
