// oracle/oracle.cpp — TEST INFRASTRUCTURE ONLY (never linked, imported or called by the product).
//
// CPU restatement of the reference's batched query path (elex-fu/vector_db), written from the
// reference's behaviour, not copied from it. Every function cites the reference file:line it
// follows (paths relative to /root/reference). Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference leg may load the shared object built from this file.
//
// Parity pinning:
//   * native-arithmetic functions (R1/R2/R3 orders, PQ train/encode/LUT/ADC/top-k) are pinned
//     against the UNMODIFIED reference compiled into oracle/_ref (tests/test_oracle_vs_ref.py)
//     and against committed golden vectors generated from it (tests/golden/).
//   * Java-semantics functions (HnswPqIndex / HnswIndex / Vector) cannot be pinned against a JVM
//     here (no JDK): they are pinned against the reference's own known-answer tests
//     (VectorTest.java, HnswIndexTest.java, RecallOptimizationTest.java thresholds) only.
//
// Build: see oracle/Makefile (g++ -O2 -std=c++17 -ffp-contract=off -fopenmp).
// -ffp-contract=off matters: every "mul then add" below must stay two roundings unless fmaf is
// written explicitly (the reference's scalar loops compile to vmulss+vaddss).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <queue>
#include <utility>
#include <vector>

#define ORC_API extern "C" __attribute__((visibility("default")))

// ---------------------------------------------------------------------------------------------
// Synthetic data: counter-based uniform [-1,1) generator (distribution of
// RecallOptimizationTest.java:303-309 `nextFloat()*2-1`, 24-bit mantissa like Random.nextFloat).
// The CUDA fill kernel in vector_db_b200/csrc implements the identical hash.
// ---------------------------------------------------------------------------------------------
static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline float synth_value(uint64_t seed, uint64_t idx) {
    uint64_t z = mix64(idx + seed * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull);
    uint32_t k = (uint32_t)(z >> 40);                 // 24 bits
    return (float)((int32_t)k - (1 << 23)) * (1.0f / (float)(1 << 23));  // exact: 2u-1
}
ORC_API void orc_fill_uniform(uint64_t seed, int64_t start, int64_t count, float* out) {
    for (int64_t i = 0; i < count; i++) out[i] = synth_value(seed, (uint64_t)(start + i));
}

// ---------------------------------------------------------------------------------------------
// fp32 reduction orders
// ---------------------------------------------------------------------------------------------
// R6: Java sequential, mul then add. HnswPqIndex.java:741-750; core/Vector.java:61-72 (pre-sqrt);
// native DistanceScalar.cpp:7-14.
static inline float l2sq_seq(const float* a, const float* b, int d) {
    float sum = 0.0f;
    for (int i = 0; i < d; i++) {
        float diff = a[i] - b[i];
        sum += diff * diff;
    }
    return sum;
}
// Lane accumulators of the 8-wide FMA loop shared by R1 and R2 (DistanceAVX2.cpp:13-18, 68-73).
static inline void l2sq_lanes8(const float* a, const float* b, int d, float lane[8], int* done) {
    for (int j = 0; j < 8; j++) lane[j] = 0.0f;
    int i = 0;
    for (; i + 8 <= d; i += 8)
        for (int j = 0; j < 8; j++) {
            float diff = a[i + j] - b[i + j];
            lane[j] = fmaf(diff, diff, lane[j]);
        }
    *done = i;
}
// R1: euclideanDistanceAVX2, DistanceAVX2.cpp:9-33 — lanes, then (l0+l4)+(l1+l5) + (l2+l6)+(l3+l7),
// tail appended sequentially mul-then-add.
static inline float l2sq_r1(const float* a, const float* b, int d) {
    float l[8]; int i;
    l2sq_lanes8(a, b, d, l, &i);
    float s0 = l[0] + l[4], s1 = l[1] + l[5], s2 = l[2] + l[6], s3 = l[3] + l[7];
    float total = (s0 + s1) + (s2 + s3);
    for (; i < d; i++) { float diff = a[i] - b[i]; total += diff * diff; }
    return total;
}
// R2: batchEuclideanDistanceAVX2, DistanceAVX2.cpp:59-89 — lanes, then sequential sum from 0.
static inline float l2sq_r2(const float* a, const float* b, int d) {
    float l[8]; int i;
    l2sq_lanes8(a, b, d, l, &i);
    float sum = 0.0f;
    for (int k = 0; k < 8; k++) sum += l[k];
    for (; i < d; i++) { float diff = a[i] - b[i]; sum += diff * diff; }
    return sum;
}
// The product's warp-cooperative traversal order (NOT a reference order; restated here so the
// traversal LOGIC can be checked bit-exactly under identical arithmetic — DESIGN.md §Parity):
// lane l owns float4 chunks c = l, l+32, ...; per element acc = fma(diff,diff,acc); then an xor
// butterfly (16,8,4,2,1). Butterfly adds are commutative so every lane ends with the same value.
static inline float warp_tree(float lane[32]) {
    for (int off = 16; off >= 1; off >>= 1) {
        float nxt[32];
        for (int l = 0; l < 32; l++) nxt[l] = lane[l] + lane[l ^ off];
        std::memcpy(lane, nxt, sizeof(nxt));
    }
    return lane[0];
}
static inline float l2sq_gpu(const float* a, const float* b, int d) {
    float lane[32];
    for (int l = 0; l < 32; l++) lane[l] = 0.0f;
    int nch = d / 4;
    for (int c = 0; c < nch; c++) {
        int l = c & 31;
        for (int e = 0; e < 4; e++) {
            float diff = a[4 * c + e] - b[4 * c + e];
            lane[l] = fmaf(diff, diff, lane[l]);
        }
    }
    for (int i = nch * 4; i < d; i++) {           // ragged tail: rows are zero-padded, so it is one more (partial) chunk
        float diff = a[i] - b[i];
        int l = nch & 31;
        lane[l] = fmaf(diff, diff, lane[l]);
    }
    return warp_tree(lane);
}
static inline float dot_seq(const float* a, const float* b, int d) {   // core/Vector.java:82-85
    float s = 0.0f;
    for (int i = 0; i < d; i++) s += a[i] * b[i];
    return s;
}
static inline float dot_gpu(const float* a, const float* b, int d) {
    float lane[32];
    for (int l = 0; l < 32; l++) lane[l] = 0.0f;
    int nch = d / 4;
    for (int c = 0; c < nch; c++) {
        int l = c & 31;
        for (int e = 0; e < 4; e++) lane[l] = fmaf(a[4 * c + e], b[4 * c + e], lane[l]);
    }
    for (int i = nch * 4; i < d; i++) { int l = nch & 31; lane[l] = fmaf(a[i], b[i], lane[l]); }
    return warp_tree(lane);
}
// core/Vector.java:93-102 — sqrt of the sequential sum of squares ((float)Math.sqrt(double) of a
// float is the correctly rounded sqrtf).
static inline float norm_java(const float* a, int d) {
    float s = 0.0f;
    for (int i = 0; i < d; i++) s += a[i] * a[i];
    return sqrtf(s);
}

enum { ARITH_JAVA_SEQ = 0, ARITH_NATIVE_AVX2 = 1, ARITH_GPU_TREE = 2 };

ORC_API float orc_l2sq(const float* a, const float* b, int d, int order) {
    switch (order) {
        case 1: return l2sq_r1(a, b, d);
        case 2: return l2sq_r2(a, b, d);
        case 3: return l2sq_gpu(a, b, d);
        default: return l2sq_seq(a, b, d);
    }
}
// Java known answers (VectorTest.java:39-109): euclidean WITH sqrt, cosine similarity, norm, normalize.
ORC_API float orc_java_euclidean(const float* a, const float* b, int d) { return sqrtf(l2sq_seq(a, b, d)); }
ORC_API float orc_java_norm(const float* a, int d) { return norm_java(a, d); }
ORC_API float orc_java_cosine_similarity(const float* a, const float* b, int d) {
    return dot_seq(a, b, d) / (norm_java(a, d) * norm_java(b, d));       // Vector.java:77-88
}
ORC_API void orc_java_normalize(const float* a, int d, float* out) {      // Vector.java:107-116
    float n = norm_java(a, d);
    for (int i = 0; i < d; i++) out[i] = a[i] / n;
}
ORC_API void orc_java_normalize_rows(const float* a, int64_t n, int d, float* out) {   // row-wise, OpenMP (test sizes up to 1M rows)
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) orc_java_normalize(a + i * d, d, out + i * d);
}
// native "cosine": 1 - dot with the 8-lane FMA order (DistanceAVX2.cpp:35-57). Listed for completeness.
ORC_API float orc_native_cosine(const float* a, const float* b, int d) {
    float l[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i = 0;
    for (; i + 8 <= d; i += 8) for (int j = 0; j < 8; j++) l[j] = fmaf(a[i + j], b[i + j], l[j]);
    float s0 = l[0] + l[4], s1 = l[1] + l[5], s2 = l[2] + l[6], s3 = l[3] + l[7];
    float dot = (s0 + s1) + (s2 + s3);
    for (; i < d; i++) dot += a[i] * b[i];
    return 1.0f - dot;
}

// config/CompressionConfig.java:114-130 recommendedConfig: M = max(8, d/8), decrement until d%M==0.
ORC_API int orc_recommended_pq_m(int dim) {
    int m = dim / 8;
    if (m < 8) m = 8;
    while (dim % m != 0 && m > 1) m--;
    return m;
}

// ---------------------------------------------------------------------------------------------
// RNGs used by the reference's k-means seeding
// ---------------------------------------------------------------------------------------------
// java.util.Random — the published 48-bit LCG (JDK javadoc). Used by HnswPqIndex.java:256.
struct JavaRandom {
    uint64_t seed;
    explicit JavaRandom(int64_t s) { seed = ((uint64_t)s ^ 0x5DEECE66Dull) & ((1ull << 48) - 1); }
    int32_t next(int bits) {
        seed = (seed * 0x5DEECE66Dull + 0xBull) & ((1ull << 48) - 1);
        return (int32_t)((int64_t)seed >> (48 - bits));
    }
    int32_t nextInt(int32_t bound) {
        int32_t r = next(31);
        int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        for (int32_t u = r; (int32_t)((uint32_t)u - (uint32_t)(r = u % bound) + (uint32_t)m) < 0; u = next(31)) {}
        return r;
    }
    float nextFloat() { return (float)next(24) / (float)(1 << 24); }
};
// kind 0: nextInt() = next(32); 1: nextInt(bound); 2: nextFloat() — one stream per call (java.util.Random javadoc)
ORC_API void orc_java_random_stream(int64_t seed, int kind, int32_t bound, int n, int32_t* ints, float* floats) {
    JavaRandom r(seed);
    for (int i = 0; i < n; i++) {
        if (kind == 0) ints[i] = r.next(32);
        else if (kind == 1) ints[i] = r.nextInt(bound);
        else floats[i] = r.nextFloat();
    }
}
ORC_API void orc_java_random_probe(int64_t seed, int32_t bound, int n, int32_t* ints, float* floats) {
    JavaRandom r(seed);
    for (int i = 0; i < n; i++) { ints[i] = r.nextInt(bound); floats[i] = r.nextFloat(); }
}

// std::mt19937 + libstdc++ (GCC 11+) uniform_int_distribution<int> / uniform_real_distribution<float>
// as instantiated by native/index/HNSWPQIndex.cpp:151-192 and PQIndex.cpp:52-60. The reference pins
// neither compiler nor library; this restates GCC 13's algorithms (Lemire's nearly-divisionless
// bounded integers; generate_canonical<float,24>) and is checked against oracle/_ref.
struct Mt19937 {
    uint32_t mt[624]; int idx;
    explicit Mt19937(uint32_t s) {
        mt[0] = s;
        for (int i = 1; i < 624; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
        idx = 624;
    }
    uint32_t operator()() {
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
    int uniform_int(int lo, int hi) {                 // inclusive range
        uint32_t range = (uint32_t)(hi - lo) + 1u;    // callers never pass the full 2^32 range
        uint64_t product = (uint64_t)(*this)() * (uint64_t)range;
        uint32_t low = (uint32_t)product;
        if (low < range) {
            uint32_t threshold = (0u - range) % range;
            while (low < threshold) { product = (uint64_t)(*this)() * (uint64_t)range; low = (uint32_t)product; }
        }
        return lo + (int)(product >> 32);
    }
    float uniform_real(float a, float b) {
        float sum = (float)(*this)() * 1.0f;          // uint32 -> float rounds to nearest
        float ret = sum / 4294967296.0f;
        if (ret >= 1.0f) ret = nextafterf(1.0f, 0.0f);
        return ret * (b - a) + a;
    }
};

// ---------------------------------------------------------------------------------------------
// PQ: train / encode / LUT / ADC / flat search
// ---------------------------------------------------------------------------------------------
static inline float subdist(const float* a, const float* b, int sd, int arith) {
    // Java computeSubspaceDistance (HnswPqIndex.java:343-350) = sequential; native distanceFunc_ = R1.
    return arith == ARITH_NATIVE_AVX2 ? l2sq_r1(a, b, sd) : l2sq_seq(a, b, sd);
}
// findNearestCentroid: HnswPqIndex.java:325-338 / HNSWPQIndex.cpp:246-260 / PQIndex.cpp:107-121 —
// strict '<' from FLT_MAX, lowest index wins ties.
static inline int nearest_centroid(const float* cb_m, const float* x, int sd, int ncent, int arith) {
    int nearest = 0;
    float best = std::numeric_limits<float>::max();
    for (int c = 0; c < ncent; c++) {
        float d = subdist(x, cb_m + (size_t)c * sd, sd, arith);
        if (d < best) { best = d; nearest = c; }
    }
    return nearest;
}

// flavour 0: Java HnswPqIndex.train (HnswPqIndex.java:194-338)
// flavour 1: native HNSWPQIndex::train (HNSWPQIndex.cpp:127-244)
// flavour 2: native PQIndex::train (PQIndex.cpp:29-105)
// codebooks out: [M][256][sd]
ORC_API void orc_pq_train(const float* samples, int n, int dim, int M, int ncent, int iters, int flavour,
                          float* codebooks) {
    const int sd = dim / M;
    #pragma omp parallel for schedule(dynamic)
    for (int m = 0; m < M; m++) {
        float* cb = codebooks + (size_t)m * ncent * sd;
        std::vector<float> sub((size_t)n * sd);
        for (int i = 0; i < n; i++) std::memcpy(&sub[(size_t)i * sd], samples + (size_t)i * dim + (size_t)m * sd, sd * sizeof(float));
        const int arith = flavour == 0 ? ARITH_JAVA_SEQ : ARITH_NATIVE_AVX2;

        if (flavour == 0) {
            // k-means++ with java.util.Random(42+m): HnswPqIndex.java:255-293
            for (size_t z = 0; z < (size_t)ncent * sd; z++) cb[z] = 0.0f;   // Java arrays start zeroed (:103)
            JavaRandom rnd(42 + m);
            int first = rnd.nextInt(n);
            std::memcpy(cb, &sub[(size_t)first * sd], sd * sizeof(float));
            std::vector<float> mind(n, std::numeric_limits<float>::max());
            for (int c = 1; c < ncent && c < n; c++) {
                float total = 0.0f;
                for (int i = 0; i < n; i++) {
                    float d = l2sq_seq(&sub[(size_t)i * sd], cb + (size_t)(c - 1) * sd, sd);
                    if (d < mind[i]) mind[i] = d;
                    total += mind[i];
                }
                float target = rnd.nextFloat() * total;
                float cum = 0.0f; int sel = 0;
                for (int i = 0; i < n; i++) { cum += mind[i]; if (cum >= target) { sel = i; break; } }
                std::memcpy(cb + (size_t)c * sd, &sub[(size_t)sel * sd], sd * sizeof(float));
            }
            // Lloyd: HnswPqIndex.java:231-249; assignments re-zeroed every iteration (:234) so "changed"
            // means "some sample not assigned to centroid 0"; empty clusters keep the old centre (:312-319).
            for (int it = 0; it < iters; it++) {
                std::vector<int> asg(n, 0);
                bool changed = false;
                for (int i = 0; i < n; i++) {
                    int nn = nearest_centroid(cb, &sub[(size_t)i * sd], sd, ncent, arith);
                    if (asg[i] != nn) { asg[i] = nn; changed = true; }
                }
                if (!changed && it > 5) break;
                std::vector<int> cnt(ncent, 0);
                std::vector<float> acc((size_t)ncent * sd, 0.0f);
                for (int i = 0; i < n; i++) {
                    int c = asg[i]; cnt[c]++;
                    for (int d = 0; d < sd; d++) acc[(size_t)c * sd + d] += sub[(size_t)i * sd + d];
                }
                for (int c = 0; c < ncent; c++) if (cnt[c] > 0) {
                    float inv = 1.0f / (float)cnt[c];
                    for (int d = 0; d < sd; d++) cb[(size_t)c * sd + d] = acc[(size_t)c * sd + d] * inv;
                }
            }
        } else {
            Mt19937 rng(42u + (uint32_t)m);
            if (flavour == 1) {
                // HNSWPQIndex.cpp:150-192
                int first = rng.uniform_int(0, n - 1);
                std::memcpy(cb, &sub[(size_t)first * sd], sd * sizeof(float));
                std::vector<float> mind(n, std::numeric_limits<float>::max());
                for (int c = 1; c < ncent; c++) {
                    float total = 0.0f;
                    for (int i = 0; i < n; i++) {
                        float d = l2sq_r1(&sub[(size_t)i * sd], cb + (size_t)(c - 1) * sd, sd);
                        if (d < mind[i]) mind[i] = d;
                        total += mind[i];
                    }
                    float target = rng.uniform_real(0.0f, total);
                    float cum = 0.0f; int sel = 0;
                    for (int i = 0; i < n; i++) { cum += mind[i]; if (cum >= target) { sel = i; break; } }
                    std::memcpy(cb + (size_t)c * sd, &sub[(size_t)sel * sd], sd * sizeof(float));
                }
            } else {
                // PQIndex.cpp:52-60: 256 uniform draws (with replacement)
                for (int c = 0; c < ncent; c++) {
                    int s = rng.uniform_int(0, n - 1);
                    std::memcpy(cb + (size_t)c * sd, &sub[(size_t)s * sd], sd * sizeof(float));
                }
            }
            // Lloyd: HNSWPQIndex.cpp:194-243 / PQIndex.cpp:62-104 — assignments persist (start 0), stop when
            // unchanged, centroids zero-filled then accumulated IN PLACE, empty cluster stays the zero vector.
            std::vector<int> asg(n, 0), cnt(ncent, 0);
            for (int it = 0; it < iters; it++) {
                bool changed = false;
                for (int i = 0; i < n; i++) {
                    int nn = nearest_centroid(cb, &sub[(size_t)i * sd], sd, ncent, arith);
                    if (asg[i] != nn) { asg[i] = nn; changed = true; }
                }
                if (!changed) break;
                std::fill(cnt.begin(), cnt.end(), 0);
                for (size_t z = 0; z < (size_t)ncent * sd; z++) cb[z] = 0.0f;
                for (int i = 0; i < n; i++) {
                    int c = asg[i];
                    for (int d = 0; d < sd; d++) cb[(size_t)c * sd + d] += sub[(size_t)i * sd + d];
                    cnt[c]++;
                }
                for (int c = 0; c < ncent; c++) if (cnt[c] > 0) {
                    float inv = 1.0f / (float)cnt[c];
                    for (int d = 0; d < sd; d++) cb[(size_t)c * sd + d] *= inv;
                }
            }
        }
    }
}

// encode: HnswPqIndex.java:355-366 (arith 0) / HNSWPQIndex.cpp:262-267, PQIndex.cpp:139-144 (arith 1)
ORC_API void orc_pq_encode(const float* x, int64_t n, int dim, int M, int ncent, const float* codebooks,
                           int arith, uint8_t* codes) {
    const int sd = dim / M;
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++)
        for (int m = 0; m < M; m++)
            codes[i * M + m] = (uint8_t)nearest_centroid(codebooks + (size_t)m * ncent * sd,
                                                         x + i * dim + (size_t)m * sd, sd, ncent, arith);
}

// LUT: native PQIndex.cpp:157-163 with batchEuclideanDistanceAVX2 = R2 (arith 1);
// Java PqIndex.java:306-330 sequential (arith 0). lut out: [M][256]
ORC_API void orc_pq_lut(const float* q, int dim, int M, int ncent, const float* codebooks, int arith, float* lut) {
    const int sd = dim / M;
    for (int m = 0; m < M; m++)
        for (int c = 0; c < ncent; c++) {
            const float* cen = codebooks + ((size_t)m * ncent + c) * sd;
            lut[(size_t)m * ncent + c] = arith == ARITH_NATIVE_AVX2 ? l2sq_r2(q + (size_t)m * sd, cen, sd)
                                                                    : l2sq_seq(q + (size_t)m * sd, cen, sd);
        }
}

// ADC scan, order R3: dist = dist + lut[m][code[m]], m ascending (PQIndex.cpp:173-198, ADCUtils.cpp:11-17,
// Java PqIndex.java:279-284).
ORC_API void orc_adc_scan(const float* lut, const uint8_t* codes, int64_t n, int M, int ncent, float* out) {
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float d = 0.0f;
        const uint8_t* c = codes + i * M;
        for (int m = 0; m < M; m++) d += lut[(size_t)m * ncent + c[m]];
        out[i] = d;
    }
}

// direct PQ distance, order R5 (HnswPqIndex.java:371-386; HNSWPQIndex.cpp:269-289)
static inline float pq_direct(const float* q, const uint8_t* code, const float* codebooks, int M, int ncent, int sd) {
    float dist = 0.0f;
    for (int m = 0; m < M; m++) {
        const float* cen = codebooks + ((size_t)m * ncent + code[m]) * sd;
        for (int d = 0; d < sd; d++) { float diff = q[m * sd + d] - cen[d]; dist += diff * diff; }
    }
    return dist;
}
// Twin of the product's CTA-wide upper-layer PQ distance (hnsw_kernels.cu greedy_level_pq): the same quantity summed as
// 8 partial sums of ceil(M/8) consecutive subspaces each (every subspace term summed sequentially from 0), partials added
// in order. Not the Java order (R5) — within 1e-6 relative of it; used only when arith == ARITH_GPU_TREE.
static inline float pq_direct_parts(const float* q, const uint8_t* code, const float* codebooks, int M, int ncent, int sd) {
    const int mp = (M + 7) / 8;
    float total = 0.0f;
    for (int part = 0; part < 8; part++) {
        float ps = 0.0f;
        for (int m = part * mp; m < M && m < (part + 1) * mp; m++) {
            const float* cen = codebooks + ((size_t)m * ncent + code[m]) * sd;
            float t = 0.0f;
            for (int d = 0; d < sd; d++) { float diff = q[m * sd + d] - cen[d]; t += diff * diff; }
            ps += t;
        }
        total = part == 0 ? ps : total + ps;
    }
    return total;
}
ORC_API float orc_pq_direct(const float* q, const uint8_t* code, const float* codebooks, int dim, int M, int ncent) {
    return pq_direct(q, code, codebooks, M, ncent, dim / M);
}

typedef std::pair<float, int> DI;   // ordered by (distance, id): PQIndex.cpp:200-205 partial_sort on pair

static void topk_pairs(std::vector<DI>& v, int k) {
    if ((int)v.size() > k) { std::partial_sort(v.begin(), v.begin() + k, v.end()); v.resize(k); }
    else std::sort(v.begin(), v.end());
}

// Flat PQ search = PQIndex::search (PQIndex.cpp:146-213): LUT (R2) -> ADC (R3) -> (dist,id) top-k.
// Optional exact rerank of the best `rerank` candidates (C5 of BASELINE.json): exact order `exact_order`
// (0 = R6 sequential, 1 = R1), then (dist,id) top-k. ids may be NULL (id = row).
ORC_API void orc_pq_flat_search(const float* queries, int nq, int dim, int M, int ncent, const float* codebooks,
                                const uint8_t* codes, const int32_t* ids, int64_t n, int arith, int k,
                                int rerank, const float* base, int exact_order,
                                int32_t* out_ids, float* out_dists, int nthreads) {
    #pragma omp parallel for schedule(dynamic) num_threads(nthreads > 0 ? nthreads : 1)
    for (int qi = 0; qi < nq; qi++) {
        const float* q = queries + (size_t)qi * dim;
        std::vector<float> lut((size_t)M * ncent);
        orc_pq_lut(q, dim, M, ncent, codebooks, arith, lut.data());
        std::vector<DI> v; v.reserve(n);
        for (int64_t i = 0; i < n; i++) {
            float d = 0.0f;
            const uint8_t* c = codes + i * M;
            for (int m = 0; m < M; m++) d += lut[(size_t)m * ncent + c[m]];
            v.emplace_back(d, ids ? ids[i] : (int)i);
        }
        if (rerank > 0 && base) {
            // shortlist keyed by (adc, row) so it is well defined, then exact re-score
            std::vector<std::pair<DI, int64_t>> sl; sl.reserve(n);
            for (int64_t i = 0; i < n; i++) sl.push_back({v[i], i});
            int R = (int)std::min<int64_t>(rerank, n);
            std::partial_sort(sl.begin(), sl.begin() + R, sl.end());
            std::vector<DI> ex;
            for (int r = 0; r < R; r++) {
                int64_t row = sl[r].second;
                float d = exact_order == 1 ? l2sq_r1(q, base + row * dim, dim) : l2sq_seq(q, base + row * dim, dim);
                ex.emplace_back(d, sl[r].first.second);
            }
            v.swap(ex);
        }
        topk_pairs(v, k);
        for (int j = 0; j < k; j++) {
            out_ids[(size_t)qi * k + j] = j < (int)v.size() ? v[j].second : -1;
            out_dists[(size_t)qi * k + j] = j < (int)v.size() ? v[j].first : std::numeric_limits<float>::infinity();
        }
    }
}

// Flat exact search — the ground truth of RecallOptimizationTest.java:151-164 (sq-L2, R6) and the
// cosine analogue with HnswIndex's distance (HnswIndex.java:126-133). metric 0 = sq-L2, 1 = cosine
// (base/query must already be normalised the Java way; norms recomputed as Java does).
// order: 0 = R6 sequential, 1 = R1.  Ties by id ((dist,id) ordering; Java's heap order is unspecified).
ORC_API void orc_flat_search(const float* base, const int32_t* ids, int64_t n, int dim, const float* queries, int nq,
                             int k, int metric, int order, int32_t* out_ids, float* out_dists, int nthreads) {
    std::vector<float> bn;
    if (metric == 1) { bn.resize(n); for (int64_t i = 0; i < n; i++) bn[i] = norm_java(base + i * dim, dim); }
    #pragma omp parallel for schedule(dynamic) num_threads(nthreads > 0 ? nthreads : 1)
    for (int qi = 0; qi < nq; qi++) {
        const float* q = queries + (size_t)qi * dim;
        float qn = metric == 1 ? norm_java(q, dim) : 0.0f;
        std::priority_queue<DI> heap;   // max-heap of the k best
        for (int64_t i = 0; i < n; i++) {
            float d;
            if (metric == 1) d = 1.0f - dot_seq(base + i * dim, q, dim) / (bn[i] * qn);
            else d = order == 1 ? l2sq_r1(q, base + i * dim, dim) : l2sq_seq(q, base + i * dim, dim);
            DI e(d, ids ? ids[i] : (int)i);
            if ((int)heap.size() < k) heap.push(e);
            else if (e < heap.top()) { heap.pop(); heap.push(e); }
        }
        std::vector<DI> v;
        while (!heap.empty()) { v.push_back(heap.top()); heap.pop(); }
        std::sort(v.begin(), v.end());
        for (int j = 0; j < k; j++) {
            out_ids[(size_t)qi * k + j] = j < (int)v.size() ? v[j].second : -1;
            out_dists[(size_t)qi * k + j] = j < (int)v.size() ? v[j].first : std::numeric_limits<float>::infinity();
        }
    }
}

// ---------------------------------------------------------------------------------------------
// HNSW — restatement of the two Java classes VectorDatabase actually instantiates
//   kind 0: HnswPqIndex  (sq-L2, no normalisation, PQ codes for upper-layer search)  index/HnswPqIndex.java
//   kind 1: HnswIndex    (cosine on normalised vectors)                              index/HnswIndex.java
// Deliberate deviation (documented in DESIGN.md): levels come from a SEEDED generator with the same law
// (P(level>=l) = e^-l, capped at maxLevel-1) because Java uses ThreadLocalRandom (HnswPqIndex.java:752-759).
// Heap ties are broken by id ((dist,id) lexicographic); Java's PriorityQueue leaves them unspecified.
// ---------------------------------------------------------------------------------------------
struct Stats { int64_t dist_evals = 0, expansions = 0; };

struct Hnsw {
    int kind, dim, maxElements, M = 32, efC, ef, maxLevel;
    uint64_t level_seed, level_ctr = 0;
    int arith = 0;                  // 0 = Java sequential (reference), 2 = product's warp-tree order
    int efCap = 5000;               // HnswPqIndex.java:621 hard-codes 5000; settable only for the recall-vs-ef study (bench.py)
    // PQ
    int pqM = 0, ncent = 256, pqIters = 25, sd = 0;
    bool trained = false;
    std::vector<float> codebooks;   // [pqM][256][sd]
    std::vector<uint8_t> codes;     // [index][pqM]
    std::vector<float> training;    // collected samples (row-major), HnswPqIndex.java:170-175
    int ntrain = 0;
    // data (index = insertion order since last clear)
    std::vector<float> vecs;
    std::vector<float> norms;       // Java Vector.norm of the stored (normalised) vector
    std::vector<int32_t> ids;
    std::vector<int> level;
    std::vector<std::vector<std::vector<int>>> graph;   // [level][index] -> neighbour indices
    std::vector<std::vector<uint8_t>> present;           // [level][index] node has a list at that level
    int entry = -1;
    // removeVector (HnswPqIndex.java:542-596): a removed node stays as a dead index (Java: id gone from `vectors`)
    std::vector<uint8_t> dead;
    int nlive = 0;
    // Java `currentSize` / idToIndex: code row = currentSize at insert time, currentSize-- on removal (:393,:467,:591), so an
    // insert after a removal overwrites the code row of an earlier vector. jindex[index] = that row; codes are indexed by it.
    std::vector<int> jindex;
    int javaSize = 0;
    // scratch for the insert path (searches use a per-thread Scratch)
    struct Scratch { std::vector<uint32_t> visit; uint32_t epoch = 0; } own;

    int size() const { return (int)ids.size(); }
    const float* vec(int i) const { return &vecs[(size_t)i * dim]; }

    float dist(const float* q, float qnorm, int i, Stats* st) const {
        if (st) st->dist_evals++;
        if (kind == 0) return arith == ARITH_GPU_TREE ? l2sq_gpu(vec(i), q, dim) : l2sq_seq(vec(i), q, dim);
        float dot = arith == ARITH_GPU_TREE ? dot_gpu(vec(i), q, dim) : dot_seq(vec(i), q, dim);
        return 1.0f - dot / (norms[i] * qnorm);                   // HnswIndex.java:126-133
    }
    float pqdist(const float* q, int i) const {
        const size_t row = (size_t)jindex[i];                       // idToIndex.get(id) (:804,:814)
        if (arith == 2) return pq_direct_parts(q, &codes[row * pqM], codebooks.data(), pqM, ncent, sd);
        return pq_direct(q, &codes[row * pqM], codebooks.data(), pqM, ncent, sd);
    }

    int assignLevel() {                                            // HnswPqIndex.java:752-759 (seeded)
        int lv = 0;
        const double p = 1.0 / M_E;
        while (true) {
            double u = (double)(mix64(level_seed * 0x9E3779B97F4A7C15ull + (level_ctr++)) >> 11) * (1.0 / 9007199254740992.0);
            if (u < p && lv < maxLevel - 1) lv++; else break;
        }
        return lv;
    }
};

struct Cand { float d; int id; int idx; };
struct CandLess { bool operator()(const Cand& a, const Cand& b) const { return a.d < b.d || (a.d == b.d && a.id < b.id); } };
struct CandGreater { bool operator()(const Cand& a, const Cand& b) const { return CandLess()(b, a); } };

// searchLayerClosest: HnswPqIndex.java:761-792 / HnswIndex.java:511-552 — move to the FIRST improving
// neighbour (break), repeat until no neighbour improves. compressed=true: HnswPqIndex.java:794-827.
static int search_closest(const Hnsw& h, const float* q, float qn, int ep, int lvl, bool compressed, Stats* st) {
    int cur = ep;
    float curd = compressed ? h.pqdist(q, cur) : h.dist(q, qn, cur, st);
    bool changed;
    do {
        changed = false;
        if (!h.present[lvl][cur]) return cur;
        const std::vector<int>& nb = h.graph[lvl][cur];
        if (nb.empty()) return cur;
        for (int n : nb) {
            if (h.dead[n]) continue;                                // `!vectors.containsKey(neighborId)` (:778,:811)
            float d = compressed ? h.pqdist(q, n) : h.dist(q, qn, n, st);
            if (d < curd) { curd = d; cur = n; changed = true; break; }
        }
    } while (changed);
    return cur;
}

// variant 0: HnswPqIndex.searchLayerWithSize (HnswPqIndex.java:905-974) — furthest refreshed on every insert,
//            break when cand > furthest AND results full.
// variant 1: HnswPqIndex.searchLayer (:829-895) — furthest refreshed only on eviction, `continue` when
//            cand > furthest AND results full.
// variant 2: HnswIndex.searchLayer (HnswIndex.java:557-644) — furthest refreshed only on eviction, break when
//            cand > furthest (no "full" condition).
static void search_layer(const Hnsw& h, Hnsw::Scratch& sc, const float* q, float qn, int ep, int ef, int lvl, int variant,
                         std::vector<Cand>& out, Stats* st) {
    out.clear();
    std::priority_queue<Cand, std::vector<Cand>, CandLess> results;       // max-heap
    std::priority_queue<Cand, std::vector<Cand>, CandGreater> cands;      // min-heap
    if ((int)sc.visit.size() < h.size()) sc.visit.resize(h.maxElements, 0);
    if (++sc.epoch == 0) { std::fill(sc.visit.begin(), sc.visit.end(), 0); sc.epoch = 1; }
    float d0 = h.dist(q, qn, ep, st);
    Cand e{d0, h.ids[ep], ep};
    results.push(e); cands.push(e);
    sc.visit[ep] = sc.epoch;
    float furthest = d0;
    while (!cands.empty()) {
        Cand c = cands.top(); cands.pop();
        if (variant == 2) { if (c.d > furthest) break; }
        else if (c.d > furthest && (int)results.size() >= ef) { if (variant == 0) break; else continue; }
        if (!h.present[lvl][c.idx]) continue;
        const std::vector<int>& nb = h.graph[lvl][c.idx];
        if (nb.empty()) continue;
        if (st) st->expansions++;
        for (int n : nb) {
            if (sc.visit[n] == sc.epoch || h.dead[n]) continue;     // visited, or no longer stored (:869,:944): not marked
            sc.visit[n] = sc.epoch;
            float d = h.dist(q, qn, n, st);
            if ((int)results.size() < ef || d < furthest) {
                Cand x{d, h.ids[n], n};
                cands.push(x); results.push(x);
                if (variant == 0) {
                    if ((int)results.size() > ef) results.pop();
                    furthest = results.top().d;
                } else if ((int)results.size() > ef) {
                    results.pop();
                    furthest = results.top().d;
                }
            }
        }
    }
    out.resize(results.size());
    for (int i = (int)results.size() - 1; i >= 0; i--) { out[i] = results.top(); results.pop(); }
    // results.sort(by distance) — already ascending by (d,id)
}

// selectNeighbors: HnswPqIndex.java:1043-1064 / HnswIndex.java:637-662 — if <= m keep given order, else
// stable sort by distance (SearchResult.compareTo, distance only) and take the first m.
static void select_neighbors(const std::vector<Cand>& c, int m, std::vector<int>& out) {
    out.clear();
    if ((int)c.size() <= m) { for (auto& x : c) out.push_back(x.idx); return; }
    std::vector<Cand> s(c);
    std::stable_sort(s.begin(), s.end(), [](const Cand& a, const Cand& b) { return a.d < b.d; });
    for (int i = 0; i < m; i++) out.push_back(s[i].idx);
}

// insert: HnswPqIndex.java:391-460 (raw, ef=efConstruction) / :465-539 (compressed, ef=max(efC,200)) /
// HnswIndex.java:139-248 (ef=efConstruction).
static void hnsw_insert(Hnsw& h, int id, const float* v, bool compressed_phase) {
    int index = h.size();
    h.ids.push_back(id);
    h.vecs.insert(h.vecs.end(), v, v + h.dim);
    h.norms.push_back(h.kind == 1 ? norm_java(v, h.dim) : 0.0f);
    h.dead.push_back(0); h.nlive++;
    const int jrow = h.javaSize++;                                    // index = currentSize++ (:393,:467)
    h.jindex.push_back(jrow);
    if (h.kind == 0) {
        if (h.codes.size() < (size_t)(jrow + 1) * h.pqM) h.codes.resize((size_t)(jrow + 1) * h.pqM, 0);   // Java codes[][] start zeroed (:104)
        if (compressed_phase)
            for (int m = 0; m < h.pqM; m++)
                h.codes[(size_t)jrow * h.pqM + m] = (uint8_t)nearest_centroid(&h.codebooks[(size_t)m * h.ncent * h.sd], v + m * h.sd, h.sd, h.ncent, ARITH_JAVA_SEQ);
    }
    int lv = h.assignLevel();
    h.level.push_back(lv);
    for (int l = 0; l < h.maxLevel; l++) { h.graph[l].emplace_back(); h.present[l].push_back(0); }
    if (h.entry == -1) {
        h.entry = index;
        for (int l = 0; l <= lv; l++) h.present[l][index] = 1;
        return;
    }
    const float* q = h.vec(index);
    float qn = h.norms[index];
    int cur = h.entry;
    for (int l = h.maxLevel - 1; l > lv; l--) cur = search_closest(h, q, qn, cur, l, false, nullptr);
    int ef = h.kind == 0 ? (compressed_phase ? std::max(h.efC, 200) : h.efC) : h.efC;
    int variant = h.kind == 0 ? 1 : 2;
    std::vector<Cand> nbrs, tmp; std::vector<int> sel, pruned;
    for (int l = std::min(lv, h.maxLevel - 1); l >= 0; l--) {
        search_layer(h, h.own, q, qn, cur, ef, l, variant, nbrs, nullptr);
        select_neighbors(nbrs, h.M, sel);
        h.graph[l][index] = sel; h.present[l][index] = 1;
        for (int nb : sel) {
            std::vector<int>& conn = h.graph[l][nb];
            h.present[l][nb] = 1;
            conn.push_back(index);
            if ((int)conn.size() > h.M) {
                tmp.clear();
                for (int nn : conn) {
                    if (h.dead[nn]) continue;                       // `nnVector != null` (:435): dangling links drop out here
                    // calculateDistance(neighborVector, nnVector): distance of stored vectors
                    float d;
                    if (h.kind == 0) d = l2sq_seq(h.vec(nb), h.vec(nn), h.dim);
                    else d = 1.0f - dot_seq(h.vec(nb), h.vec(nn), h.dim) / (h.norms[nb] * h.norms[nn]);
                    tmp.push_back(Cand{d, h.ids[nn], nn});
                }
                select_neighbors(tmp, h.M, pruned);
                conn = pruned;
            }
        }
        cur = sel.empty() ? index : sel[0];
    }
    if (lv > h.level[h.entry]) h.entry = index;
}

static void hnsw_train(Hnsw& h) {                                  // HnswPqIndex.java:194-215
    if (h.trained || h.ntrain == 0) return;
    h.codebooks.assign((size_t)h.pqM * h.ncent * h.sd, 0.0f);
    orc_pq_train(h.training.data(), h.ntrain, h.dim, h.pqM, h.ncent, h.pqIters, 0, h.codebooks.data());
    h.trained = true;
}

ORC_API void* orc_hnsw_create(int kind, int dim, int maxElements, int M, int efC, int ef, int pqM, int pqIters,
                              uint64_t level_seed) {
    Hnsw* h = new Hnsw();
    h->kind = kind; h->dim = dim; h->maxElements = maxElements; h->M = M; h->efC = efC; h->ef = ef;
    h->maxLevel = (int)(std::log((double)maxElements) / std::log((double)M)) + 1;   // HnswPqIndex.java:89
    h->level_seed = level_seed;
    if (kind == 0) { h->pqM = pqM; h->sd = dim / pqM; h->pqIters = pqIters; }
    h->graph.resize(h->maxLevel); h->present.resize(h->maxLevel);
    return h;
}
ORC_API void orc_hnsw_destroy(void* p) { delete (Hnsw*)p; }
ORC_API void orc_hnsw_set_arith(void* p, int arith) { ((Hnsw*)p)->arith = arith; }

// addVector: HnswPqIndex.java:153-189 (training trigger at min(10000, maxElements/10) samples) /
// HnswIndex.java:139-248 (normalise on insert :154).
ORC_API int orc_hnsw_add(void* p, int id, const float* v) {
    Hnsw& h = *(Hnsw*)p;
    for (int i = 0; i < h.size(); i++) if (!h.dead[i] && h.ids[i] == id) return 0;    // vectors.containsKey(id) (:157-159)
    if (h.nlive >= h.maxElements) return 0;                                          // vectors.size() >= maxElements (:161-163)
    if (h.kind == 1) {
        std::vector<float> nv(h.dim);
        orc_java_normalize(v, h.dim, nv.data());
        hnsw_insert(h, id, nv.data(), false);
        return 1;
    }
    if (!h.trained && h.ntrain < 10000) {
        h.training.insert(h.training.end(), v, v + h.dim); h.ntrain++;
        if (h.ntrain >= std::min(10000, h.maxElements / 10)) hnsw_train(h);
    }
    hnsw_insert(h, id, v, h.trained);
    return 1;
}
ORC_API void orc_hnsw_add_batch(void* p, const int32_t* ids, const float* v, int64_t n) {
    Hnsw& h = *(Hnsw*)p;
    for (int64_t i = 0; i < n; i++) orc_hnsw_add(p, ids ? ids[i] : (int)i, v + i * h.dim);
}
// buildIndex: HnswPqIndex.java:676-718 / HnswIndex.java:430-478 — train if needed, clear, re-add by id.
ORC_API void orc_hnsw_rebuild(void* p) {
    Hnsw& h = *(Hnsw*)p;
    if (h.kind == 0 && !h.trained && h.ntrain > 0) hnsw_train(h);
    std::vector<float> sv; sv.swap(h.vecs);
    std::vector<int32_t> sid; sid.swap(h.ids);
    std::vector<uint8_t> sdead; sdead.swap(h.dead);
    int n = (int)sid.size();
    std::vector<int> order;
    for (int i = 0; i < n; i++) if (!sdead[i]) order.push_back(i);     // new HashMap<>(vectors) (:687): live vectors only
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return sid[a] < sid[b]; });
    h.norms.clear(); h.level.clear(); h.codes.clear(); h.entry = -1; h.jindex.clear(); h.javaSize = 0; h.nlive = 0;
    for (int l = 0; l < h.maxLevel; l++) { h.graph[l].clear(); h.present[l].clear(); }
    for (int i : order) {
        // HnswIndex re-normalises already-normalised vectors on re-add (addVector :154) — restated as is.
        orc_hnsw_add(p, sid[i], &sv[(size_t)i * h.dim]);
    }
}
// removeVector: HnswPqIndex.java:542-596 / HnswIndex.java:254-317. Returns 1 when the id was stored.
ORC_API int orc_hnsw_remove(void* p, int id) {
    Hnsw& h = *(Hnsw*)p;
    int idx = -1;
    for (int i = 0; i < h.size(); i++) if (!h.dead[i] && h.ids[i] == id) { idx = i; break; }
    if (idx < 0) return 0;
    for (int l = 0; l <= h.level[idx] && l < h.maxLevel; l++) {
        std::vector<int> copy = h.graph[l][idx];                      // neighborsCopy (:562)
        for (int nb : copy) {
            std::vector<int>& conn = h.graph[l][nb];
            auto it = std::find(conn.begin(), conn.end(), idx);       // remove(Integer.valueOf(id)): first occurrence
            if (it != conn.end()) conn.erase(it);
        }
        h.graph[l][idx].clear(); h.present[l][idx] = 0;               // currentGraph.remove(id)
    }
    if (idx == h.entry) {
        if (h.nlive == 1) h.entry = -1;
        else {
            // first key of the ConcurrentHashMap<>(maxElements) in iteration order, skipping id (:577-584): ascending bin
            // ((h ^ h>>>16) & 0x7fffffff) & (table-1), table = tableSizeFor(1.5*maxElements+1); insertion order inside a bin
            uint64_t want = (uint64_t)h.maxElements + ((uint64_t)h.maxElements >> 1) + 1, table = 1;
            while (table < want && table < (1ull << 30)) table <<= 1;
            int best = -1; uint64_t bestbin = ~0ull;
            for (int i = 0; i < h.size(); i++) {
                if (h.dead[i] || i == idx) continue;
                uint32_t hc = (uint32_t)h.ids[i];
                uint64_t bin = ((hc ^ (hc >> 16)) & 0x7fffffffu) & (table - 1);
                if (bin < bestbin) { bestbin = bin; best = i; }
            }
            h.entry = best;
        }
    }
    h.dead[idx] = 1; h.nlive--;
    if (h.kind == 0) h.javaSize--;                                    // currentSize-- (:591)
    return 1;
}
ORC_API int orc_hnsw_live(void* p) { return ((Hnsw*)p)->nlive; }
ORC_API void orc_hnsw_get_dead(void* p, uint8_t* out) { Hnsw& h = *(Hnsw*)p; std::memcpy(out, h.dead.data(), h.dead.size()); }
// state of a snapshot that carries removals: dead[n], code-row table jindex[n] (int64), Java currentSize, level counter
ORC_API void orc_hnsw_set_dead(void* p, const uint8_t* dead, const int64_t* jindex, int64_t javaSize, uint64_t level_ctr) {
    Hnsw& h = *(Hnsw*)p;
    const int n = h.size();
    h.nlive = 0;
    for (int i = 0; i < n; i++) { h.dead[i] = dead[i]; if (!dead[i]) h.nlive++; h.jindex[i] = (int)jindex[i]; }
    h.javaSize = (int)javaSize; h.level_ctr = level_ctr;
    // codes arrive per vector (row i = code of vector i); rows sharing a Java code row hold identical codes by construction
    if (h.kind == 0 && !h.codes.empty()) {
        std::vector<uint8_t> byrow(h.codes);
        for (int i = 0; i < n; i++) std::memcpy(&h.codes[(size_t)h.jindex[i] * h.pqM], &byrow[(size_t)i * h.pqM], (size_t)h.pqM);
    }
}
ORC_API void orc_hnsw_get_jindex(void* p, int64_t* out) { Hnsw& h = *(Hnsw*)p; for (int i = 0; i < h.size(); i++) out[i] = h.jindex[i]; }
ORC_API uint64_t orc_hnsw_level_ctr(void* p) { return ((Hnsw*)p)->level_ctr; }
ORC_API int64_t orc_hnsw_java_size(void* p) { return ((Hnsw*)p)->javaSize; }
ORC_API void orc_hnsw_get_ids(void* p, int32_t* out) { Hnsw& h = *(Hnsw*)p; for (int i = 0; i < h.size(); i++) out[i] = h.ids[i]; }
ORC_API int orc_hnsw_size(void* p) { return ((Hnsw*)p)->size(); }
ORC_API int orc_hnsw_trained(void* p) { return ((Hnsw*)p)->trained ? 1 : 0; }
ORC_API int orc_hnsw_max_level(void* p) { return ((Hnsw*)p)->maxLevel; }
ORC_API int orc_hnsw_entry(void* p) { return ((Hnsw*)p)->entry; }
ORC_API void orc_hnsw_get_codebooks(void* p, float* out) { Hnsw& h = *(Hnsw*)p; std::memcpy(out, h.codebooks.data(), h.codebooks.size() * sizeof(float)); }
ORC_API void orc_hnsw_set_codebooks(void* p, const float* cb) {
    Hnsw& h = *(Hnsw*)p; h.codebooks.assign(cb, cb + (size_t)h.pqM * h.ncent * h.sd); h.trained = true;
}
// codes as the walk sees them: row i = the code row of vector i (codes[idToIndex[id]])
ORC_API void orc_hnsw_get_codes(void* p, uint8_t* out) {
    Hnsw& h = *(Hnsw*)p;
    for (int i = 0; i < h.size(); i++) std::memcpy(out + (size_t)i * h.pqM, &h.codes[(size_t)h.jindex[i] * h.pqM], (size_t)h.pqM);
}
ORC_API void orc_hnsw_get_vectors(void* p, float* out) { Hnsw& h = *(Hnsw*)p; std::memcpy(out, h.vecs.data(), h.vecs.size() * sizeof(float)); }
ORC_API void orc_hnsw_get_norms(void* p, float* out) { Hnsw& h = *(Hnsw*)p; std::memcpy(out, h.norms.data(), h.norms.size() * sizeof(float)); }
ORC_API void orc_hnsw_get_levels(void* p, int32_t* out) { Hnsw& h = *(Hnsw*)p; for (int i = 0; i < h.size(); i++) out[i] = h.level[i]; }
// Export one level as a fixed-degree table adj[n][M] (unused = -1) + present[n]. Lists never exceed M
// after an insert completes (pruned to M when > M).
ORC_API void orc_hnsw_get_level_adj(void* p, int lvl, int32_t* adj, uint8_t* present) {
    Hnsw& h = *(Hnsw*)p;
    for (int i = 0; i < h.size(); i++) {
        present[i] = h.present[lvl][i];
        const std::vector<int>& nb = h.graph[lvl][i];
        for (int j = 0; j < h.M; j++) adj[(size_t)i * h.M + j] = j < (int)nb.size() ? nb[j] : -1;
    }
}

static inline int java_search_ef(int n, int k, int ef, int cap = 5000) {   // HnswPqIndex.java:616-621
    int byRatio = (int)(n * 0.15);
    int byK = k * 100;
    int s = std::max(std::max(byRatio, byK), ef);
    s = std::min(s, n);
    return std::min(s, cap);
}
ORC_API int orc_java_search_ef(int n, int k, int ef) { return java_search_ef(n, k, ef); }
ORC_API void orc_hnsw_set_ef_cap(void* p, int cap) { ((Hnsw*)p)->efCap = cap > 0 ? cap : 5000; }

// searchNearest: HnswPqIndex.java:599-669 / HnswIndex.java:328-365. stats[0]=distance evals, [1]=expansions.
static void hnsw_search_one(const Hnsw& h, Hnsw::Scratch& sc, const float* qin, int k, int32_t* out_ids, float* out_d, int64_t* stats) {
    for (int j = 0; j < k; j++) { out_ids[j] = -1; out_d[j] = std::numeric_limits<float>::infinity(); }
    if (h.nlive == 0 || h.entry < 0) return;
    Stats st;
    std::vector<float> qbuf; const float* q = qin; float qn = 0.0f;
    if (h.kind == 1) { qbuf.resize(h.dim); orc_java_normalize(qin, h.dim, qbuf.data()); q = qbuf.data(); qn = norm_java(q, h.dim); }
    int cur = h.entry;
    std::vector<Cand> cands;
    if (h.kind == 0) {
        int searchEf = java_search_ef(h.nlive, k, h.ef, h.efCap);          // dataSize = vectors.size() (:615)
        for (int l = h.maxLevel - 1; l > 0; l--) cur = search_closest(h, q, qn, cur, l, h.trained, &st);
        if (h.trained) search_layer(h, sc, q, qn, cur, std::max(searchEf, k * 100), 0, 0, cands, &st);
        else search_layer(h, sc, q, qn, cur, searchEf, 0, 1, cands, &st);
        // rerank ALL candidates with the exact distance in Java order (:646-657), stable sort by distance (:660)
        for (auto& c : cands) c.d = l2sq_seq(q, h.vec(c.idx), h.dim);
        std::sort(cands.begin(), cands.end(), CandLess());
    } else {
        int searchEf = std::max(h.ef, k * 4);                       // HnswIndex.java:347
        for (int l = h.maxLevel - 1; l > 0; l--) cur = search_closest(h, q, qn, cur, l, false, &st);
        search_layer(h, sc, q, qn, cur, searchEf, 0, 2, cands, &st);
        if (h.arith != ARITH_JAVA_SEQ) {
            for (auto& c : cands) c.d = 1.0f - dot_seq(h.vec(c.idx), q, h.dim) / (h.norms[c.idx] * qn);
            std::sort(cands.begin(), cands.end(), CandLess());
        }
    }
    for (int j = 0; j < k && j < (int)cands.size(); j++) { out_ids[j] = cands[j].id; out_d[j] = cands[j].d; }
    if (stats) { stats[0] += st.dist_evals; stats[1] += st.expansions; }
}

ORC_API void orc_hnsw_search(void* p, const float* queries, int nq, int k, int32_t* out_ids, float* out_d,
                             int64_t* stats, int nthreads) {
    Hnsw& h0 = *(Hnsw*)p;
    if (stats) { stats[0] = 0; stats[1] = 0; }
    if (nthreads <= 1) {
        for (int i = 0; i < nq; i++) hnsw_search_one(h0, h0.own, queries + (size_t)i * h0.dim, k, out_ids + (size_t)i * k, out_d + (size_t)i * k, stats);
        return;
    }
    // all-cores variant: queries are independent; each thread needs private visited scratch.
    #pragma omp parallel num_threads(nthreads)
    {
        Hnsw::Scratch sc;          // per-thread visited stamps; the index itself is shared read-only
        int64_t ls[2] = {0, 0};
        #pragma omp for schedule(dynamic, 4)
        for (int i = 0; i < nq; i++) hnsw_search_one(h0, sc, queries + (size_t)i * h0.dim, k, out_ids + (size_t)i * k, out_d + (size_t)i * k, ls);
        #pragma omp critical
        if (stats) { stats[0] += ls[0]; stats[1] += ls[1]; }
    }
}

// ---- test hooks: ONE layer search / ONE greedy descent with explicit parameters (the micro-fixtures of
//      tests/golden/micro_fixtures.json drive the termination rules on hand-derived 5..8-node graphs) ----
ORC_API int orc_hnsw_search_layer(void* p, const float* qin, int ep, int ef, int level, int variant, int32_t* out_idx, float* out_d,
                                  int64_t* stats) {
    Hnsw& h = *(Hnsw*)p;
    std::vector<float> qbuf; const float* q = qin; float qn = 0.0f;
    if (h.kind == 1) { qbuf.resize(h.dim); orc_java_normalize(qin, h.dim, qbuf.data()); q = qbuf.data(); qn = norm_java(q, h.dim); }
    std::vector<Cand> out; Stats st;
    search_layer(h, h.own, q, qn, ep, ef, level, variant, out, &st);
    for (size_t i = 0; i < out.size(); i++) { out_idx[i] = out[i].idx; out_d[i] = out[i].d; }
    if (stats) { stats[0] = st.dist_evals; stats[1] = st.expansions; }
    return (int)out.size();
}
ORC_API int orc_hnsw_search_closest(void* p, const float* qin, int ep, int level, int compressed, int64_t* stats) {
    Hnsw& h = *(Hnsw*)p;
    std::vector<float> qbuf; const float* q = qin; float qn = 0.0f;
    if (h.kind == 1) { qbuf.resize(h.dim); orc_java_normalize(qin, h.dim, qbuf.data()); q = qbuf.data(); qn = norm_java(q, h.dim); }
    Stats st;
    int r = search_closest(h, q, qn, ep, level, compressed != 0, &st);
    if (stats) { stats[0] = st.dist_evals; stats[1] = st.expansions; }
    return r;
}

// Import a graph built elsewhere (the product's builder / a snapshot) so the CPU search runs on the SAME graph,
// codebooks and codes as the GPU ("same persisted graph and codebooks", BASELINE.json north_star).
// vecs[n][dim] are stored as is (already normalised for kind 1); adj[max_level][n][M] is -1 padded.
ORC_API void orc_hnsw_import(void* p, int n, const float* vecs, const int32_t* ids, const int32_t* levels,
                             const int32_t* adj, int entry, const float* codebooks, const uint8_t* codes) {
    Hnsw& h = *(Hnsw*)p;
    h.vecs.assign(vecs, vecs + (size_t)n * h.dim);
    h.ids.resize(n);
    for (int i = 0; i < n; i++) h.ids[i] = ids ? ids[i] : i;
    h.norms.resize(n);
    for (int i = 0; i < n; i++) h.norms[i] = h.kind == 1 ? norm_java(h.vec(i), h.dim) : 0.0f;
    h.level.assign(levels, levels + n);
    for (int l = 0; l < h.maxLevel; l++) {
        h.graph[l].assign(n, {});
        h.present[l].assign(n, 0);
        for (int i = 0; i < n; i++) {
            const int32_t* row = adj + ((size_t)l * n + i) * h.M;
            for (int j = 0; j < h.M && row[j] >= 0; j++) h.graph[l][i].push_back(row[j]);
            h.present[l][i] = (levels[i] >= l) ? 1 : 0;
        }
    }
    h.entry = entry;
    h.dead.assign(n, 0); h.nlive = n; h.javaSize = n;
    h.jindex.resize(n);
    for (int i = 0; i < n; i++) h.jindex[i] = i;
    if (h.kind == 0 && codebooks && codes) {
        h.codebooks.assign(codebooks, codebooks + (size_t)h.pqM * h.ncent * h.sd);
        h.codes.assign(codes, codes + (size_t)n * h.pqM);
        h.trained = true;
    }
}
