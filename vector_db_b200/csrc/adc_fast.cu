// adc_fast.cu — the ADC code scan at shared-memory-gather speed.
//
// Reference semantics: PQIndex::search second half (native/index/PQIndex.cpp:173-198): d[i] = sum_m lut[m][codes[i][m]],
// m ascending (order R3), then the k smallest by (dist, id) (:200-205).
//
// Why not one thread per row (pq_kernels.cu adc_scan_topr_kernel, kept as the exact fallback): 32 lanes gathering
// lut[m][c_lane] for the SAME m hit random banks of one 1 KB LUT row — ~3.5-way conflicts — and each lane reads its own
// 125-byte row (uncoalesced). This kernel instead
//   * stores codes TRANSPOSED in blocks of 16 rows: codesT[block][m][16 rows] — a lane loads 16 bytes = one subspace of 16
//     consecutive rows; a HALF-warp works one block (lane l16 takes subspaces l16 + 16 j), the two halves of a warp take two
//     consecutive blocks, so a warp's load is two runs of 256 contiguous bytes (coalesced, 128-bit);
//   * lays the LUT out in shared memory as  offset(m, c) = (m >> 6) * 64 KB + c * 256 B + (m & 63) * 4 B, so lane l's
//     gathers all land in bank l: conflict-free BY CONSTRUCTION for any code values, and the whole address is ONE
//     byte-permute of (code word, lane constant): PRMT + LDS + FADD per gathered value;
//   * reduces the 16 per-lane partial sums across the 16 lanes of the half-warp with a transposing butterfly: 15 shuffles
//     per 32 rows (a whole-warp butterfly needs 17 per 16), every lane ends with one row;
//   * keeps a per-CTA candidate buffer with a lazily tightened threshold; after warm-up almost every row fails one compare.
// The sums are associated differently from R3 (within 2*M*2^-24 relative), so the scan returns a shortlist of R2 > R rows
// and adc_refine_kernel re-evaluates exactly those in order R3, sorts by (dist, id) and CERTIFIES that no row outside the
// shortlist can belong to the exact top R; a query that cannot be certified (massive ties) is redone by the exact kernel.
// Result: ids and distances bit-identical to the reference, at gather speed.
#include <algorithm>

#include "kernels.h"

#define AF_NT 512
#define AF_WARPS (AF_NT / 32)
#ifndef AF_K
#define AF_K 4                                  // pairs of 16-row blocks per warp between barriers (even)
#endif
#define AF_SPAN (32 * AF_WARPS * AF_K)          // rows per super-iteration (= max insertions between barriers)

// ---- codes [n][M] -> codesT [ceil(n/16)][Mp][16], Mp = M rounded up to 8 chunks so that every 16-row block starts on a
//      128-byte line (M = 125: 2048-byte blocks instead of 2000; the half-warp's 256-byte runs then touch 2 lines, not 3);
//      rows past n and chunks past M are code 0 (never reported / gathered from all-zero LUT columns).
//      The 16-byte chunk of subspace m is read by lane (m & 15) of a half-warp; inside the chunk byte a holds row
//      a ^ (m & 15), the order that lane's butterfly wants (see the scan kernel): no selects in the transposing reduction. ----
__global__ void transpose_codes_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, int Mp, int64_t row0, uint8_t* __restrict__ codesT) {
    // one thread per output byte of blocks covering rows [row0, n); chunks m in [M, Mp) are zero padding
    const int64_t b0 = row0 >> 4;
    const int64_t total = (((n + 15) >> 4) - b0) * (int64_t)Mp * 16;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t bm = t >> 4;
        const int m = (int)(bm % Mp);
        const int r = (int)(t & 15) ^ (m & 15);
        const int64_t row = ((b0 + bm / Mp) << 4) + r;
        codesT[(b0 * Mp << 4) + t] = (row < n && m < M) ? codes[row * M + m] : (uint8_t)0;
    }
}
void launch_transpose_codes(const uint8_t* d_codes, int64_t n, int M, int64_t row0, uint8_t* d_codesT, cudaStream_t st) {
    const int Mp = adc_mpad(M);
    const int64_t total = (((n + 15) >> 4) - (row0 >> 4)) * (int64_t)Mp * 16;
    if (total <= 0) return;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 32);
    transpose_codes_kernel<<<blocks, 256, 0, st>>>(d_codes, n, M, Mp, row0, d_codesT);
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

// One 16-row block is worked by a HALF-warp: lane l16 of half h gathers subspaces m = 64 g + l16 + 16 (jj ^ h), jj = 0..3, of
// LUT group g (a "quad" of four 16-byte chunks per lane and group), so at every step the 32 lanes of the warp — two blocks —
// touch 32 different banks. Loads are never predicated: a lane whose subspace is past M reads the row's last chunk (its LUT
// column is all zero, so whatever it gathers adds 0.0); a block past the chunk's end re-reads the last valid block (its rows
// are not reported).
template <int NQ>
__device__ __forceinline__ void load_quad(const uint4* __restrict__ pe, const uint4* __restrict__ po, int g, const int (&offl)[4], uint4 (&v)[4]) {
    if (g < NQ - 1) {
        v[0] = ld_stream(pe + 64 * g); v[1] = ld_stream(po + 64 * g); v[2] = ld_stream(pe + 64 * g + 32); v[3] = ld_stream(po + 64 * g + 32);
    } else {
        v[0] = ld_stream(pe + offl[0]); v[1] = ld_stream(po + offl[1]); v[2] = ld_stream(pe + offl[2]); v[3] = ld_stream(po + offl[3]);
    }
}

// two fp32 additions in one instruction (sm_100 FADD2, add.rn.f32x2): the same round-to-nearest sums as two FADDs, half the
// issue slots — the scan is co-limited by issue (PRMT + LDS + add per gathered value) and the shared-memory crossbar
__device__ __forceinline__ void fadd2(float& a0, float& a1, float b0, float b1) {
    asm("{\n\t.reg .b64 ra, rb;\n\tmov.b64 ra, {%0, %1};\n\tmov.b64 rb, {%2, %3};\n\tadd.rn.f32x2 ra, ra, rb;\n\tmov.b64 {%0, %1}, ra;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}

// 64 gathers of one quad into the 16 per-lane partial sums; the group's 64 KB image is an immediate offset of the LDS
template <bool FIRST>
__device__ __forceinline__ void gather_quad(const uint4 (&v)[4], const uint32_t (&lbase)[4], const unsigned char* __restrict__ lutg, float (&acc)[16]) {
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
        const uint32_t w[4] = {v[jj].x, v[jj].y, v[jj].z, v[jj].w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float v0 = *reinterpret_cast<const float*>(lutg + __byte_perm(w[i], lbase[jj], 0x7604));
            const float v1 = *reinterpret_cast<const float*>(lutg + __byte_perm(w[i], lbase[jj], 0x7614));
            const float v2 = *reinterpret_cast<const float*>(lutg + __byte_perm(w[i], lbase[jj], 0x7624));
            const float v3 = *reinterpret_cast<const float*>(lutg + __byte_perm(w[i], lbase[jj], 0x7634));
            if (FIRST && jj == 0) { acc[4 * i + 0] = v0; acc[4 * i + 1] = v1; acc[4 * i + 2] = v2; acc[4 * i + 3] = v3; }
            else { fadd2(acc[4 * i + 0], acc[4 * i + 1], v0, v1); fadd2(acc[4 * i + 2], acc[4 * i + 3], v2, v3); }
        }
    }
}

// 16 partial sums x 16 lanes -> one row total per lane (row l16 of the half-warp's block). The chunk's rows were stored
// pre-permuted by this lane's bits (transpose_codes_kernel: acc[a] is row a ^ l16), so every lane keeps acc[r] and sends
// acc[r + half]: 15 shuffles per 32 rows, no selects.
__device__ __forceinline__ float half_warp_transpose_sum(float (&acc)[16]) {
#pragma unroll
    for (int r = 0; r < 8; r += 2)
        fadd2(acc[r], acc[r + 1], __shfl_xor_sync(0xffffffffu, acc[r + 8], 8), __shfl_xor_sync(0xffffffffu, acc[r + 9], 8));
#pragma unroll
    for (int r = 0; r < 4; r += 2)
        fadd2(acc[r], acc[r + 1], __shfl_xor_sync(0xffffffffu, acc[r + 4], 4), __shfl_xor_sync(0xffffffffu, acc[r + 5], 4));
    fadd2(acc[0], acc[1], __shfl_xor_sync(0xffffffffu, acc[2], 2), __shfl_xor_sync(0xffffffffu, acc[3], 2));
    return __fadd_rn(acc[0], __shfl_xor_sync(0xffffffffu, acc[1], 1));
}

// NQ = ceil(M/64) LUT groups = quads per lane and block. grid = (chunks, queries).
template <int NQ>
__global__ void __launch_bounds__(AF_NT, 1)
adc_scan_fast_kernel(const uint4* __restrict__ codesT, int64_t n, int M, const float* __restrict__ lutg, int R2, int cap,
                     int64_t blocks_per_chunk, uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_rows) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* lut = reinterpret_cast<float*>(smem_raw);                                  // [NQ][256][64]
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw + (size_t)NQ * 65536);       // [cap] candidate keys
    __shared__ int s_cnt;
    __shared__ float s_tau;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, h = lane >> 4, l16 = lane & 15;
    const int q = blockIdx.y, chunk = blockIdx.x, nchunks = gridDim.x;
    {
        const float4* src = reinterpret_cast<const float4*>(lutg + (size_t)q * NQ * 16384);
        float4* dst = reinterpret_cast<float4*>(lut);
        for (int i = tid; i < NQ * 4096; i += AF_NT) dst[i] = __ldg(src + i);
    }
    for (int i = tid; i < cap; i += AF_NT) buf[i] = KEY_MAX;
    if (tid == 0) { s_cnt = 0; s_tau = __int_as_float(0x7f800000); }
    __syncthreads();
    const int64_t nblk = (n + 15) >> 4;
    const int64_t b0 = (int64_t)chunk * blocks_per_chunk, b1 = min(nblk, b0 + blocks_per_chunk);
    // lane constants: byte offset of (m & 63) inside a group for each jj; the code byte is permuted into byte 1 (c * 256)
    uint32_t lbase[4];
    int offl[4];          // last group: chunk offsets relative to pe (even jj) / po (odd jj), clamped to subspace M - 1
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
        const int ml = l16 + 16 * (jj ^ h);
        lbase[jj] = (uint32_t)(ml * 4);
        offl[jj] = min(64 * (NQ - 1) + ml, M - 1) - (l16 + 16 * ((jj & 1) ^ h));
    }
    const unsigned char* lutb = reinterpret_cast<const unsigned char*>(lut);

    // block indices relative to the chunk start (32-bit); rows of the chunk that exist: nrel
    const int nb = (int)(b1 - b0);
    const int nrel = (int)min((int64_t)nb << 4, n - (b0 << 4));
    const uint4* base_e = codesT + b0 * M + l16 + 16 * h;          // this lane's even-jj chunks of block 0 of the chunk
    const uint4* base_o = codesT + b0 * M + l16 + 16 * (1 - h);
    auto block_off = [&](int pair) { return (int64_t)(min(2 * pair + h, nb - 1) * M); };   // < 2^31: nb * M uint4s per chunk
    uint4 va[4], vb[4];
    float acc[16];
    auto report = [&](float d, int pair) {
        const int row = ((2 * pair + h) << 4) + l16;
        if (row < nrel && d < s_tau) {
            const int slot = atomicAdd(&s_cnt, 1);
            buf[slot] = ((uint64_t)f2ord(d) << 32) | (uint32_t)((b0 << 4) + row);
        }
    };
    const int npairs = (nb + 1) >> 1;
    int64_t o = block_off(warp);
    load_quad<NQ>(base_e + o, base_o + o, 0, offl, va);
    for (int sp = 0; sp < npairs; sp += AF_WARPS * AF_K) {
#pragma unroll 1
        for (int kk = 0; kk < AF_K; kk += 2) {
            // two block pairs = 2 NQ quads, alternating between the two register buffers; the next quad is always in flight
            const int pair = sp + kk * AF_WARPS + warp;
#pragma unroll
            for (int t = 0; t < 2 * NQ; t++) {
                const int g = t % NQ, pr = t / NQ;                        // compile-time after unrolling
                const int gn = (t + 1) % NQ, prn = (t + 1) / NQ;          // the quad to prefetch (prn == 2: first of the next iteration)
                if (gn == 0) o = block_off(pair + prn * AF_WARPS);
                if (t & 1) load_quad<NQ>(base_e + o, base_o + o, gn, offl, va); else load_quad<NQ>(base_e + o, base_o + o, gn, offl, vb);
                const unsigned char* lg = lutb + g * 65536;
                if (t & 1) { if (g == 0) gather_quad<true>(vb, lbase, lg, acc); else gather_quad<false>(vb, lbase, lg, acc); }
                else       { if (g == 0) gather_quad<true>(va, lbase, lg, acc); else gather_quad<false>(va, lbase, lg, acc); }
                if (g == NQ - 1) report(half_warp_transpose_sum(acc), pair + pr * AF_WARPS);
            }
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
    const int groups = (M + 63) / 64;
    const int cap = next_pow2(R2 + AF_SPAN);
    const size_t smem = (size_t)groups * 65536 + (size_t)cap * 8;
    if (smem > 226 * 1024) return 1;
    const int64_t nblk = (n + 15) >> 4;
    const int64_t bpc = (nblk + nchunks - 1) / nchunks;
    dim3 grid(nchunks, nq);
    uint64_t* mk = nchunks > 1 ? d_keys_tmp + (size_t)nq * nchunks * R2 : d_keys_tmp;
    uint32_t* mr = nchunks > 1 ? d_rows_tmp + (size_t)nq * nchunks * R2 : d_rows_tmp;
#define AF_LAUNCH(NQ)                                                                                                        \
    do {                                                                                                                     \
        cudaFuncSetAttribute(adc_scan_fast_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);              \
        adc_scan_fast_kernel<NQ><<<grid, AF_NT, smem, st>>>(reinterpret_cast<const uint4*>(d_codesT), n, adc_mpad(M), d_lutg, R2, cap, bpc, \
                                                            d_keys_tmp, d_rows_tmp);                                          \
    } while (0)
    switch (groups) {
        case 1: AF_LAUNCH(1); break;
        case 2: AF_LAUNCH(2); break;
        default: AF_LAUNCH(3); break;
    }
#undef AF_LAUNCH
    if (nchunks > 1) launch_merge_keys(d_keys_tmp, d_rows_tmp, nq, nchunks, R2, mk, mr, st);
    const int rcap = next_pow2(R2);
    const size_t rsmem = (size_t)rcap * 12;
    cudaFuncSetAttribute(adc_refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsmem);
    adc_refine_kernel<<<nq, 256, rsmem, st>>>(d_codes, M, d_luts, d_ids, id_offset, mk, mr, R2, rcap, R, d_keys, d_rows, d_flag);
    return 0;
}
