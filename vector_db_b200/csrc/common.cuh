// common.cuh — shared device helpers for libvectordb.so (sm_100a only).
// fp32 reduction orders follow the reference implementations named in SURVEY.md §8a; explicit
// __fmul_rn/__fadd_rn/__fmaf_rn keep nvcc from contracting or reassociating them.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <stdint.h>

#define VDB_NCENT 256

enum { ARITH_JAVA_SEQ = 0, ARITH_NATIVE_AVX2 = 1 };
enum { ORDER_SEQ = 0, ORDER_R1 = 1, ORDER_R2 = 2 };

// ---- synthetic data (twin of oracle.cpp synth_value) ----
__host__ __device__ __forceinline__ uint64_t vdb_mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ float vdb_synth(uint64_t seed, uint64_t idx) {
    uint64_t z = vdb_mix64(idx + seed * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull);
    uint32_t k = (uint32_t)(z >> 40);
    return (float)((int32_t)k - (1 << 23)) * (1.0f / (float)(1 << 23));
}

// ---- (distance,id) keys: unsigned 64-bit compare == lexicographic (float distance, signed id) ----
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t b = __float_as_uint(f);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    uint32_t b = u ^ ((u >> 31) ? 0x80000000u : 0xffffffffu);
    return __uint_as_float(b);
}
__device__ __forceinline__ uint64_t make_key(float d, int32_t id) {
    return ((uint64_t)f2ord(d) << 32) | (uint64_t)((uint32_t)id ^ 0x80000000u);
}
__device__ __forceinline__ float key_dist(uint64_t k) { return ord2f((uint32_t)(k >> 32)); }
__device__ __forceinline__ int32_t key_id(uint64_t k) { return (int32_t)((uint32_t)k ^ 0x80000000u); }
#define KEY_MAX 0xffffffffffffffffull

// ---- single-thread distance orders ----
// R6 / Java sequential: sum = sum + round(diff*diff)   (HnswPqIndex.java:741-750, DistanceScalar.cpp:7-14)
__device__ __forceinline__ float l2sq_seq(const float* __restrict__ a, const float* __restrict__ b, int d) {
    float s = 0.0f;
    for (int i = 0; i < d; i++) {
        float df = __fsub_rn(a[i], b[i]);
        s = __fadd_rn(s, __fmul_rn(df, df));
    }
    return s;
}
__device__ __forceinline__ void l2sq_lanes8(const float* __restrict__ a, const float* __restrict__ b, int d, float l[8], int& i) {
#pragma unroll
    for (int j = 0; j < 8; j++) l[j] = 0.0f;
    for (i = 0; i + 8 <= d; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            float df = __fsub_rn(a[i + j], b[i + j]);
            l[j] = __fmaf_rn(df, df, l[j]);
        }
    }
}
// R1: euclideanDistanceAVX2 (DistanceAVX2.cpp:9-33)
__device__ __forceinline__ float l2sq_r1(const float* __restrict__ a, const float* __restrict__ b, int d) {
    float l[8]; int i;
    l2sq_lanes8(a, b, d, l, i);
    float s0 = __fadd_rn(l[0], l[4]), s1 = __fadd_rn(l[1], l[5]), s2 = __fadd_rn(l[2], l[6]), s3 = __fadd_rn(l[3], l[7]);
    float t = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
    for (; i < d; i++) { float df = __fsub_rn(a[i], b[i]); t = __fadd_rn(t, __fmul_rn(df, df)); }
    return t;
}
// R2: batchEuclideanDistanceAVX2 (DistanceAVX2.cpp:59-89)
__device__ __forceinline__ float l2sq_r2(const float* __restrict__ a, const float* __restrict__ b, int d) {
    float l[8]; int i;
    l2sq_lanes8(a, b, d, l, i);
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; j++) s = __fadd_rn(s, l[j]);
    for (; i < d; i++) { float df = __fsub_rn(a[i], b[i]); s = __fadd_rn(s, __fmul_rn(df, df)); }
    return s;
}
__device__ __forceinline__ float l2sq_order(const float* a, const float* b, int d, int order) {
    return order == ORDER_R1 ? l2sq_r1(a, b, d) : (order == ORDER_R2 ? l2sq_r2(a, b, d) : l2sq_seq(a, b, d));
}
__device__ __forceinline__ float dot_seq(const float* __restrict__ a, const float* __restrict__ b, int d) {
    float s = 0.0f;
    for (int i = 0; i < d; i++) s = __fadd_rn(s, __fmul_rn(a[i], b[i]));
    return s;
}
// Java Vector.getNorm (core/Vector.java:93-102)
__device__ __forceinline__ float norm_java(const float* __restrict__ a, int d) {
    float s = 0.0f;
    for (int i = 0; i < d; i++) s = __fadd_rn(s, __fmul_rn(a[i], a[i]));
    return __fsqrt_rn(s);
}
// HnswIndex.calculateDistance (HnswIndex.java:126-133): 1 - dot / (na * nb)
__device__ __forceinline__ float cosine_dist_from_dot(float dot, float na, float nb) {
    return __fsub_rn(1.0f, __fdiv_rn(dot, __fmul_rn(na, nb)));
}

// ---- warp-cooperative order used by the traversal kernel (oracle twin: l2sq_gpu / dot_gpu) ----
__device__ __forceinline__ float warp_tree_sum(float v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
    return v;
}
// a: row in global memory, q: query in shared memory; D % 4 == 0 handled vectorised, ragged tail scalar
__device__ __forceinline__ float warp_l2sq(const float* __restrict__ row, const float* __restrict__ q, int D, int lane) {
    float acc = 0.0f;
    const int nch = D >> 2;
    const float4* r4 = reinterpret_cast<const float4*>(row);
    const float4* q4 = reinterpret_cast<const float4*>(q);
    for (int c = lane; c < nch; c += 32) {
        float4 x = __ldg(r4 + c);
        float4 y = q4[c];
        float d0 = __fsub_rn(x.x, y.x), d1 = __fsub_rn(x.y, y.y), d2 = __fsub_rn(x.z, y.z), d3 = __fsub_rn(x.w, y.w);
        acc = __fmaf_rn(d0, d0, acc); acc = __fmaf_rn(d1, d1, acc); acc = __fmaf_rn(d2, d2, acc); acc = __fmaf_rn(d3, d3, acc);
    }
    int i = (nch << 2) + lane;
    if (i < D) { float df = __fsub_rn(row[i], q[i]); acc = __fmaf_rn(df, df, acc); }
    return warp_tree_sum(acc);
}
__device__ __forceinline__ float warp_dot(const float* __restrict__ row, const float* __restrict__ q, int D, int lane) {
    float acc = 0.0f;
    const int nch = D >> 2;
    const float4* r4 = reinterpret_cast<const float4*>(row);
    const float4* q4 = reinterpret_cast<const float4*>(q);
    for (int c = lane; c < nch; c += 32) {
        float4 x = __ldg(r4 + c);
        float4 y = q4[c];
        acc = __fmaf_rn(x.x, y.x, acc); acc = __fmaf_rn(x.y, y.y, acc); acc = __fmaf_rn(x.z, y.z, acc); acc = __fmaf_rn(x.w, y.w, acc);
    }
    int i = (nch << 2) + lane;
    if (i < D) acc = __fmaf_rn(row[i], q[i], acc);
    return warp_tree_sum(acc);
}

// ---- block-wide bitonic sort of 64-bit keys (+ optional 32-bit payload) in shared memory, n = power of two ----
template <bool PAYLOAD>
__device__ __forceinline__ void block_bitonic_sort(uint64_t* keys, uint32_t* pay, int n) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                int ixj = i ^ j;
                if (ixj > i) {
                    uint64_t a = keys[i], b = keys[ixj];
                    bool up = ((i & k) == 0);
                    if ((a > b) == up) {
                        keys[i] = b; keys[ixj] = a;
                        if (PAYLOAD) { uint32_t t = pay[i]; pay[i] = pay[ixj]; pay[ixj] = t; }
                    }
                }
            }
        }
    }
    __syncthreads();
}

constexpr int RS_TILE = 32 * 33 + 32;      // rescore32: per-warp transpose tile + row numbers (floats)

// Exact distances of 32 rows by one warp, in the reference's summation order, with coalesced row reads: lane l loads
// element (32c + l) of every row, the per-element terms go through a padded shared-memory tile, and lane l then sums row
// l's 32 terms in index order. MODE 0: R6 sum += (q-x)^2 (HnswPqIndex.java:741-750). MODE 1: R1 eight fma lanes + pairwise
// combine + scalar tail (DistanceAVX2.cpp:9-33). MODE 2: dot, sum += x*q (HnswIndex.java:126-133).
template <int MODE>
__device__ __forceinline__ float rescore32(const float* __restrict__ base, int dim, const float* __restrict__ sq, uint32_t myrow,
                                           float* __restrict__ tile, int lane) {
    volatile uint32_t* rows = reinterpret_cast<volatile uint32_t*>(tile + 32 * 33);     // the warp's 32 row numbers, read back as broadcasts
    __syncwarp();
    rows[lane] = myrow;
    __syncwarp();
    float acc = 0.0f, l[8];
#pragma unroll
    for (int j = 0; j < 8; j++) l[j] = 0.0f;
    const int d8 = dim & ~7;
    const int nchunk = (dim + 31) >> 5;
    // rows that are whole float4s (dim % 4 == 0, 16-byte aligned base): lane (g = lane / 8, j = lane % 8) loads columns 4j..4j+3 of rows
    // 4i + g — one 128-bit load instruction covers 4 rows x 32 columns, so a chunk takes 8 loads (and 8 address computations) per lane
    // instead of 32; the kernels around this are issue-bound, and address arithmetic was half of their instructions. The terms land in
    // the same tile slots and are summed in the same order as on the scalar path below.
    const bool vec4 = (dim & 3) == 0 && ((reinterpret_cast<uintptr_t>(base) & 15) == 0) && ((reinterpret_cast<uintptr_t>(sq) & 15) == 0);
    const int g4 = lane >> 3, j4 = lane & 7;
    for (int c = 0; c < nchunk; c++) {
        __syncwarp();
        if (vec4) {
            const int col4 = c * 32 + 4 * j4;
            const bool cok4 = col4 < dim;
            const float4 q4 = cok4 ? *reinterpret_cast<const float4*>(sq + col4) : make_float4(0.f, 0.f, 0.f, 0.f);
            float4 v[8];
#pragma unroll
            for (int i = 0; i < 8; i++)
                v[i] = cok4 ? __ldg(reinterpret_cast<const float4*>(base + (size_t)rows[4 * i + g4] * dim + col4)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                float* t = tile + (4 * i + g4) * 33 + 4 * j4;
                t[0] = MODE == 2 ? __fmul_rn(v[i].x, q4.x) : __fsub_rn(q4.x, v[i].x);
                t[1] = MODE == 2 ? __fmul_rn(v[i].y, q4.y) : __fsub_rn(q4.y, v[i].y);
                t[2] = MODE == 2 ? __fmul_rn(v[i].z, q4.z) : __fsub_rn(q4.z, v[i].z);
                t[3] = MODE == 2 ? __fmul_rn(v[i].w, q4.w) : __fsub_rn(q4.w, v[i].w);
            }
        } else {
            const int col = c * 32 + lane;
            const bool cok = col < dim;
            const float qv = cok ? sq[col] : 0.0f;
#pragma unroll
            for (int r0 = 0; r0 < 32; r0 += 16) {
                float v[16];
#pragma unroll
                for (int r = 0; r < 16; r++) v[r] = cok ? __ldg(base + (size_t)rows[r0 + r] * dim + col) : 0.0f;
#pragma unroll
                for (int r = 0; r < 16; r++) tile[(r0 + r) * 33 + lane] = MODE == 2 ? __fmul_rn(v[r], qv) : __fsub_rn(qv, v[r]);
            }
        }
        __syncwarp();
        const int lim = min(32, (MODE == 1 ? d8 : dim) - c * 32);
#pragma unroll
        for (int i = 0; i < 32; i++) {
            if (i < lim) {
                const float x = tile[lane * 33 + i];
                if (MODE == 0) acc = __fadd_rn(acc, __fmul_rn(x, x));
                else if (MODE == 1) l[i & 7] = __fmaf_rn(x, x, l[i & 7]);
                else acc = __fadd_rn(acc, x);
            }
        }
    }
    if (MODE == 1) {
        const float s0 = __fadd_rn(l[0], l[4]), s1 = __fadd_rn(l[1], l[5]), s2 = __fadd_rn(l[2], l[6]), s3 = __fadd_rn(l[3], l[7]);
        acc = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
        const int c0 = (nchunk - 1) * 32;                    // the tail [d8, dim) lies inside the last chunk, still in the tile
        for (int gi = d8; gi < dim; gi++) { const float x = tile[lane * 33 + (gi - c0)]; acc = __fadd_rn(acc, __fmul_rn(x, x)); }
    }
    return acc;
}

static inline int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }
