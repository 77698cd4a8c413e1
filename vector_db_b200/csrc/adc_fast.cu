// adc_fast.cu — the ADC code scan at shared-memory-gather speed.
//
// Reference semantics: PQIndex::search second half (native/index/PQIndex.cpp:173-198): d[i] = sum_m lut[m][codes[i][m]],
// m ascending (order R3), then the k smallest by (dist, id) (:200-205).
//
// Why not one thread per row (pq_kernels.cu adc_scan_topr_kernel, kept as the exact fallback): 32 lanes gathering
// lut[m][c_lane] for the SAME m hit random banks of one 1 KB LUT row — ~3.5-way conflicts — and each lane reads its own
// 125-byte row (uncoalesced). This kernel instead
//   * stores codes TRANSPOSED in blocks of 16 rows: codesT[block][m][16 rows] — lane l loads 16 bytes = subspace
//     m = l + 32 j of 16 consecutive rows; a warp's load is 512 contiguous bytes (coalesced, 128-bit);
//   * lays the LUT out in shared memory as  offset(m, c) = (m >> 6) * 64 KB + c * 256 B + (m & 63) * 4 B, so lane l's
//     gathers all land in bank l: conflict-free BY CONSTRUCTION for any code values, and the whole address is ONE
//     byte-permute of (code word, lane constant): PRMT + LDS + FADD per gathered value;
//   * reduces the 16 per-lane partial sums across the warp with a transposing butterfly (16 shuffles per 16 rows);
//   * keeps a per-CTA candidate buffer with a lazily tightened threshold; after warm-up almost every row fails one compare.
// The sums are associated differently from R3 (within 2*M*2^-24 relative), so the scan returns a shortlist of R2 > R rows
// and adc_refine_kernel re-evaluates exactly those in order R3, sorts by (dist, id) and CERTIFIES that no row outside the
// shortlist can belong to the exact top R; a query that cannot be certified (massive ties) is redone by the exact kernel.
// Result: ids and distances bit-identical to the reference, at gather speed.
#include <algorithm>

#include "kernels.h"

#define AF_NT 512
#define AF_WARPS (AF_NT / 32)
#define AF_K 8                                  // 16-row blocks per warp between barriers
#define AF_SPAN (16 * AF_WARPS * AF_K)          // rows per super-iteration (= max insertions between barriers)

// ---- codes [n][M] -> codesT [ceil(n/16)][M][16]; rows past n are code 0 (never reported: row < n is checked).
//      Inside a chunk the 16 rows are stored in the order the reading lane's butterfly wants (see the scan kernel): byte a
//      = row a ^ p(lane), which removes every select from the transposing reduction. ----
__global__ void transpose_codes_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, int64_t row0, uint8_t* __restrict__ codesT) {
    // one thread per output byte of blocks covering rows [row0, n)
    const int64_t b0 = row0 >> 4;
    const int64_t total = (((n + 15) >> 4) - b0) * (int64_t)M * 16;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t bm = t >> 4;
        const int m = (int)(bm % M);
        const int l = m & 31;             // the lane that reads this 16-byte chunk; byte a holds row a ^ p(l), p = row_of_lane
        const int r = (int)(t & 15) ^ ((((l >> 4) & 1) << 3) | (((l >> 3) & 1) << 2) | (((l >> 2) & 1) << 1) | ((l >> 1) & 1));
        const int64_t row = ((b0 + bm / M) << 4) + r;
        codesT[(b0 * M << 4) + t] = row < n ? codes[row * M + m] : (uint8_t)0;
    }
}
void launch_transpose_codes(const uint8_t* d_codes, int64_t n, int M, int64_t row0, uint8_t* d_codesT, cudaStream_t st) {
    const int64_t total = (((n + 15) >> 4) - (row0 >> 4)) * (int64_t)M * 16;
    if (total <= 0) return;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 32);
    transpose_codes_kernel<<<blocks, 256, 0, st>>>(d_codes, n, M, row0, d_codesT);
}

// ---- LUT image in the gather layout: lutg[q][(m>>6)*16384 + c*64 + (m&63)] (floats) ----
__global__ void lut_to_gather_layout_kernel(const float* __restrict__ lut, int M, int groups, float* __restrict__ lutg) {
    const int q = blockIdx.x;
    const float* src = lut + (size_t)q * M * VDB_NCENT;
    float* dst = lutg + (size_t)q * groups * 16384;
    // thread -> (c, m) with m fastest: coalesced 256 B writes, strided (L2-resident) reads
    for (int t = threadIdx.x; t < groups * 16384; t += blockDim.x) {
        const int g = t >> 14, c = (t >> 6) & 255, ml = t & 63, m = g * 64 + ml;
        dst[t] = m < M ? src[m * VDB_NCENT + c] : 0.0f;
    }
}
void launch_lut_to_gather_layout(const float* d_lut, int nq, int M, float* d_lutg, cudaStream_t st) {
    if (nq <= 0) return;
    lut_to_gather_layout_kernel<<<nq, 512, 0, st>>>(d_lut, M, (M + 63) / 64, d_lutg);
}

// ---- the scan ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {     // codes are read once per CTA: do not pollute L1
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// loads are never predicated: a lane past M reads the row's last chunk (its LUT column is all zero, so whatever it gathers
// adds 0.0), a block past the chunk's end re-reads the last valid block (its rows are not reported)
template <int NJ>
__device__ __forceinline__ void load_block(const uint4* __restrict__ p, int lastoff, uint4 (&v)[NJ]) {
#pragma unroll
    for (int j = 0; j < NJ - 1; j++) v[j] = ld_stream(p + 32 * j);
    v[NJ - 1] = ld_stream(p + lastoff);
}

// gather + transposing butterfly for one 16-row block held in v[]; returns this lane's row total (valid on every lane pair)
template <int NJ>
__device__ __forceinline__ float block_distance(const uint4 (&v)[NJ], const uint32_t (&lbase)[NJ], const unsigned char* __restrict__ lutb) {
    float acc[16];
#pragma unroll
    for (int j = 0; j < NJ; j++) {
        const uint32_t w[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float v0 = *reinterpret_cast<const float*>(lutb + __byte_perm(w[i], lbase[j], 0x7604));
            const float v1 = *reinterpret_cast<const float*>(lutb + __byte_perm(w[i], lbase[j], 0x7614));
            const float v2 = *reinterpret_cast<const float*>(lutb + __byte_perm(w[i], lbase[j], 0x7624));
            const float v3 = *reinterpret_cast<const float*>(lutb + __byte_perm(w[i], lbase[j], 0x7634));
            if (j == 0) { acc[4 * i + 0] = v0; acc[4 * i + 1] = v1; acc[4 * i + 2] = v2; acc[4 * i + 3] = v3; }
            else {
                acc[4 * i + 0] = __fadd_rn(acc[4 * i + 0], v0); acc[4 * i + 1] = __fadd_rn(acc[4 * i + 1], v1);
                acc[4 * i + 2] = __fadd_rn(acc[4 * i + 2], v2); acc[4 * i + 3] = __fadd_rn(acc[4 * i + 3], v3);
            }
        }
    }
    // 16 partial sums x 32 lanes -> one total per row (each on two lanes). The chunk's rows were stored pre-permuted by
    // this lane's bits (transpose_codes_kernel), so every lane keeps acc[r] and sends acc[r + half]: no selects.
#pragma unroll
    for (int r = 0; r < 8; r++) acc[r] = __fadd_rn(acc[r], __shfl_xor_sync(0xffffffffu, acc[r + 8], 16));
#pragma unroll
    for (int r = 0; r < 4; r++) acc[r] = __fadd_rn(acc[r], __shfl_xor_sync(0xffffffffu, acc[r + 4], 8));
#pragma unroll
    for (int r = 0; r < 2; r++) acc[r] = __fadd_rn(acc[r], __shfl_xor_sync(0xffffffffu, acc[r + 2], 4));
    acc[0] = __fadd_rn(acc[0], __shfl_xor_sync(0xffffffffu, acc[1], 2));
    return __fadd_rn(acc[0], __shfl_xor_sync(0xffffffffu, acc[0], 1));
}

// NJ = ceil(M/32) 16-byte loads per lane and block. grid = (chunks, queries).
template <int NJ>
__global__ void __launch_bounds__(AF_NT, 1)
adc_scan_fast_kernel(const uint4* __restrict__ codesT, int64_t n, int M, const float* __restrict__ lutg, int groups, int R2, int cap,
                     int64_t blocks_per_chunk, uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_rows) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* lut = reinterpret_cast<float*>(smem_raw);                                  // [groups][256][64]
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw + (size_t)groups * 65536);   // [cap] candidate keys
    __shared__ int s_cnt;
    __shared__ float s_tau;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
    {
        const float4* src = reinterpret_cast<const float4*>(lutg + (size_t)q * groups * 16384);
        float4* dst = reinterpret_cast<float4*>(lut);
        for (int i = tid; i < groups * 4096; i += AF_NT) dst[i] = __ldg(src + i);
    }
    for (int i = tid; i < cap; i += AF_NT) buf[i] = KEY_MAX;
    if (tid == 0) { s_cnt = 0; s_tau = __int_as_float(0x7f800000); }
    __syncthreads();
    const int64_t nblk = (n + 15) >> 4;
    const int64_t b0 = (int64_t)chunk * blocks_per_chunk, b1 = min(nblk, b0 + blocks_per_chunk);
    // lane constants: byte offset of (group, m & 63) for each j; the code byte is permuted into byte 1 (c * 256)
    uint32_t lbase[NJ];
#pragma unroll
    for (int j = 0; j < NJ; j++) lbase[j] = ((uint32_t)(j >> 1) << 16) | (uint32_t)((lane + 32 * (j & 1)) * 4);
    const unsigned char* lutb = reinterpret_cast<const unsigned char*>(lut);
    const int row_of_lane = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);

    // block indices relative to the chunk start (32-bit); rows of the chunk that exist: nrel
    const int nb = (int)(b1 - b0);
    const int nrel = (int)min((int64_t)nb << 4, n - (b0 << 4));
    const int lastoff = min(lane + 32 * (NJ - 1), M - 1) - lane;
    const int64_t bstride = (int64_t)AF_WARPS * M;
    const uint4* bp = codesT + (b0 + min(warp, nb - 1)) * M + lane;     // this lane's chunk of this warp's current block
    uint4 va[NJ], vb[NJ];
    load_block<NJ>(bp, lastoff, va);
    auto advance = [&](int next_blk) { if (next_blk < nb) bp += bstride; };   // clamp: past the end, re-read the last block
    auto report = [&](float d, int blk) {
        const int row = (blk << 4) + row_of_lane;
        if (!(lane & 1) && row < nrel && d < s_tau) {
            const int slot = atomicAdd(&s_cnt, 1);
            buf[slot] = ((uint64_t)f2ord(d) << 32) | (uint32_t)((b0 << 4) + row);
        }
    };
    for (int sb = 0; sb < nb; sb += AF_WARPS * AF_K) {
#pragma unroll 1
        for (int kk = 0; kk < AF_K; kk += 2) {
            const int blk = sb + kk * AF_WARPS + warp;
            advance(blk + AF_WARPS);
            load_block<NJ>(bp, lastoff, vb);                           // next block in flight while this one is gathered
            report(block_distance<NJ>(va, lbase, lutb), blk);
            advance(blk + 2 * AF_WARPS);
            load_block<NJ>(bp, lastoff, va);
            report(block_distance<NJ>(vb, lbase, lutb), blk + AF_WARPS);
        }
        __syncthreads();
        if (s_cnt > cap - AF_SPAN) {                      // uniform: s_cnt only changes between barriers
            block_bitonic_sort<false>(buf, nullptr, cap);
            if (tid == 0) {
                const int c = min(s_cnt, R2);
                s_cnt = c;
                if (c >= R2) s_tau = ord2f((uint32_t)(buf[R2 - 1] >> 32));
            }
            __syncthreads();
            for (int j = R2 + tid; j < cap; j += AF_NT) buf[j] = KEY_MAX;
            __syncthreads();
        }
    }
    block_bitonic_sort<false>(buf, nullptr, cap);
    uint64_t* ok = out_keys + ((size_t)q * nchunks + chunk) * R2;
    uint32_t* orow = out_rows + ((size_t)q * nchunks + chunk) * R2;
    for (int j = tid; j < R2; j += AF_NT) { const uint64_t k = buf[j]; ok[j] = k; orow[j] = k == KEY_MAX ? 0xffffffffu : (uint32_t)k; }
}

// ---- exact re-evaluation + certificate ---------------------------------------------------------------------------------
// One CTA per query. in: R2 shortlisted rows (ascending fast key; 0xffffffff = none). out: the R best by the EXACT order-R3
// distance and (dist, id); flag[q] = 1 when the shortlist cannot be proven to contain the exact top R.
__global__ void adc_refine_kernel(const uint8_t* __restrict__ codes, int M, const float* __restrict__ luts, const int32_t* __restrict__ ids,
                                  int id_offset, const uint64_t* __restrict__ in_keys, const uint32_t* __restrict__ in_rows, int R2,
                                  int cap, int R, uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_rows, int* __restrict__ flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* bkey = reinterpret_cast<uint64_t*>(smem_raw);
    uint32_t* brow = reinterpret_cast<uint32_t*>(bkey + cap);
    const int q = blockIdx.x;
    const float* lut = luts + (size_t)q * M * VDB_NCENT;
    for (int j = threadIdx.x; j < cap; j += blockDim.x) {
        uint64_t key = KEY_MAX; uint32_t row = 0xffffffffu;
        if (j < R2) {
            row = in_rows[(size_t)q * R2 + j];
            if (row != 0xffffffffu) {
                const uint8_t* c = codes + (size_t)row * M;
                float d = 0.0f;
                for (int m = 0; m < M; m++) d = __fadd_rn(d, __ldg(lut + m * VDB_NCENT + c[m]));     // order R3
                key = make_key(d, ids ? ids[row] : (int32_t)(id_offset + row));
            }
        }
        bkey[j] = key; brow[j] = row;
    }
    block_bitonic_sort<true>(bkey, brow, cap);
    for (int j = threadIdx.x; j < R; j += blockDim.x) { out_keys[(size_t)q * R + j] = bkey[j]; out_rows[(size_t)q * R + j] = brow[j]; }
    if (threadIdx.x == 0) {
        int f = 0;
        const uint64_t last = in_keys[(size_t)q * R2 + R2 - 1];
        if (last != KEY_MAX && bkey[R - 1] != KEY_MAX) {      // shortlist full: rows outside it exist (fast distance >= last)
            const float f_last = ord2f((uint32_t)(last >> 32));
            const float e_r = key_dist(bkey[R - 1]);
            const float delta = 3.0f * (float)M * 5.9604645e-8f;            // 1.5 x (2 * M * 2^-24)
            if (!(e_r < f_last * (1.0f - delta))) f = 1;
        }
        flag[q] = f;
    }
}

int adc_fast_supported(int M, int R) { return M <= 192 && R <= 4096; }
int adc_fast_r2(int R) { int r2 = R + std::max(22, R / 8); return (r2 + 31) & ~31; }

// d_keys_tmp/d_rows_tmp: [nq][nchunks][R2] (+ [nq][R2] after them when nchunks > 1). Returns 0, or non-zero if unsupported.
int launch_adc_scan_fast(const uint8_t* d_codesT, const uint8_t* d_codes, int64_t n, int M, const float* d_luts, const float* d_lutg, int nq,
                         const int32_t* d_ids, int id_offset, int R, int nchunks, uint64_t* d_keys_tmp, uint32_t* d_rows_tmp,
                         uint64_t* d_keys, uint32_t* d_rows, int* d_flag, cudaStream_t st) {
    if (!adc_fast_supported(M, R)) return 1;
    const int R2 = adc_fast_r2(R);
    const int groups = (M + 63) / 64, nj = (M + 31) / 32;
    const int cap = next_pow2(R2 + AF_SPAN);
    const size_t smem = (size_t)groups * 65536 + (size_t)cap * 8;
    if (smem > 226 * 1024) return 1;
    const int64_t nblk = (n + 15) >> 4;
    const int64_t bpc = (nblk + nchunks - 1) / nchunks;
    dim3 grid(nchunks, nq);
    uint64_t* mk = nchunks > 1 ? d_keys_tmp + (size_t)nq * nchunks * R2 : d_keys_tmp;
    uint32_t* mr = nchunks > 1 ? d_rows_tmp + (size_t)nq * nchunks * R2 : d_rows_tmp;
#define AF_LAUNCH(NJ)                                                                                                        \
    do {                                                                                                                     \
        cudaFuncSetAttribute(adc_scan_fast_kernel<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);              \
        adc_scan_fast_kernel<NJ><<<grid, AF_NT, smem, st>>>(reinterpret_cast<const uint4*>(d_codesT), n, M, d_lutg, groups, R2, cap, bpc, \
                                                            d_keys_tmp, d_rows_tmp);                                          \
    } while (0)
    switch (nj) {
        case 1: AF_LAUNCH(1); break;
        case 2: AF_LAUNCH(2); break;
        case 3: AF_LAUNCH(3); break;
        case 4: AF_LAUNCH(4); break;
        case 5: AF_LAUNCH(5); break;
        default: AF_LAUNCH(6); break;
    }
#undef AF_LAUNCH
    if (nchunks > 1) launch_merge_keys(d_keys_tmp, d_rows_tmp, nq, nchunks, R2, mk, mr, st);
    const int rcap = next_pow2(R2);
    const size_t rsmem = (size_t)rcap * 12;
    cudaFuncSetAttribute(adc_refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsmem);
    adc_refine_kernel<<<nq, 256, rsmem, st>>>(d_codes, M, d_luts, d_ids, id_offset, mk, mr, R2, rcap, R, d_keys, d_rows, d_flag);
    return 0;
}
