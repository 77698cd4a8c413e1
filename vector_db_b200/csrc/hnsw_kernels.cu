// hnsw_kernels.cu — HNSW query kernel: greedy upper-layer descent (a4, a3) + layer-0 beam search (a5) + exact
// re-score and top-k (a7, a8) in ONE launch; one CTA per query, persistent CTAs loop over the batch.
//
// Semantics reproduced:
//   variant 0  Java HnswPqIndex.searchNearest, trained   index/HnswPqIndex.java:599-669 + searchLayerWithSize :905-974
//   variant 1  Java HnswPqIndex.searchNearest, untrained                              + searchLayer         :829-895
//   variant 2  Java HnswIndex.searchNearest (cosine)     index/HnswIndex.java:328-365 + searchLayer         :557-644
//
// B200-first formulation (not the reference's two PriorityQueues):
//   * the Java result max-heap and candidate min-heap are ONE array of 64-bit keys (ordered-distance<<32 | expanded<<31 |
//     node) kept SORTED in shared memory and truncated at ef: "results" = the array, "candidates" = its entries without
//     the expanded bit, "furthest" = key[ef-1]. A candidate the Java loop would never pop (worse than furthest once the
//     result set is full) is exactly an entry that falls off the end of the array.
//   * a ROUND pops the `batch` best unexpanded entries at once, fetches their neighbour lists (one coalesced 128 B row
//     per node), test-and-sets a shared-memory visited bitset, computes ALL fresh distances with the whole CTA (a lane
//     group per row, 16-byte loads, several rows in flight per lane, xor-butterfly reduction), sorts the accepted keys
//     and merges them into the array by ranking. batch == 1 is the Java order; batch > 1 is a relaxation that reaches the
//     same fixed point (every entry closer than the final furthest is expanded) — DESIGN.md §4 states the difference.
//   * traversal rows are read either as fp32 (strict: the oracle's ARITH_GPU_TREE twin reproduces every bit) or from a
//     bf16 copy (half the bytes; L2-resident at 100k x 512-d). Either way the answer is re-scored from the fp32 rows in
//     the reference's sequential order (HnswPqIndex.java:646-664), so returned distances are bit-identical to Java's.
#include <cuda_bf16.h>

#include "kernels.h"

#define HN_NT 256
#define HN_WARPS (HN_NT / 32)
#define EXP_BIT 0x80000000ull
#define HN_FMAX 1024      // fresh rows per round (batch * deg <= HN_FMAX)
#define HN_MAXB 64

struct HnswShared {
    int sel[HN_MAXB];
    int wsum[HN_WARPS];
    int nsel;
    int nfresh;
    int nacc;
    int count;        // entries in pool (<= ef)
    int evicted;      // Java: first results.poll() has happened (furthest is live)
    int lb0;
    float dentry;
    float qnorm;
    int cur;
};

__device__ __forceinline__ uint64_t pool_key(float d, int node) { return ((uint64_t)f2ord(d) << 32) | (uint32_t)node; }
__device__ __forceinline__ float pool_dist(uint64_t k) { return ord2f((uint32_t)(k >> 32)); }

// sequential-order exact distance with 16-byte loads when the row allows it (same arithmetic order as l2sq_seq/dot_seq)
__device__ __forceinline__ float l2sq_seq_v(const float* __restrict__ row, const float* __restrict__ q, int d) {
    float s = 0.0f;
    if ((d & 3) == 0) {
        const float4* r4 = reinterpret_cast<const float4*>(row);
        const float4* q4 = reinterpret_cast<const float4*>(q);
        for (int c = 0; c < (d >> 2); c++) {
            const float4 x = __ldg(r4 + c); const float4 y = q4[c];
            float t;
            t = __fsub_rn(x.x, y.x); s = __fadd_rn(s, __fmul_rn(t, t));
            t = __fsub_rn(x.y, y.y); s = __fadd_rn(s, __fmul_rn(t, t));
            t = __fsub_rn(x.z, y.z); s = __fadd_rn(s, __fmul_rn(t, t));
            t = __fsub_rn(x.w, y.w); s = __fadd_rn(s, __fmul_rn(t, t));
        }
        return s;
    }
    return l2sq_seq(row, q, d);
}
__device__ __forceinline__ float dot_seq_v(const float* __restrict__ row, const float* __restrict__ q, int d) {
    float s = 0.0f;
    if ((d & 3) == 0) {
        const float4* r4 = reinterpret_cast<const float4*>(row);
        const float4* q4 = reinterpret_cast<const float4*>(q);
        for (int c = 0; c < (d >> 2); c++) {
            const float4 x = __ldg(r4 + c); const float4 y = q4[c];
            s = __fadd_rn(s, __fmul_rn(x.x, y.x)); s = __fadd_rn(s, __fmul_rn(x.y, y.y));
            s = __fadd_rn(s, __fmul_rn(x.z, y.z)); s = __fadd_rn(s, __fmul_rn(x.w, y.w));
        }
        return s;
    }
    return dot_seq(row, q, d);
}
// exact distance of one node to the query in the reference's sequential order (one lane per node)
__device__ __forceinline__ float exact_seq_dist(const HnswSearchParams& p, const float* sq, float qnorm, int node) {
    const float* row = p.base + (size_t)node * p.dim;
    if (p.metric == 0) return l2sq_seq_v(row, sq, p.dim);
    return cosine_dist_from_dot(dot_seq_v(row, sq, p.dim), p.norms[node], qnorm);
}
// direct PQ distance, order R5 (HnswPqIndex.java:371-386)
__device__ __forceinline__ float pq_direct_dist(const HnswSearchParams& p, const float* sq, int node) {
    const int sd = p.dim / p.pq_m;
    const uint8_t* code = p.codes + (size_t)node * p.pq_m;
    float dist = 0.0f;
    if ((sd & 3) == 0) {
        for (int m = 0; m < p.pq_m; m++) {
            const float4* cen = reinterpret_cast<const float4*>(p.codebooks + ((size_t)m * VDB_NCENT + code[m]) * sd);
            const float4* qq = reinterpret_cast<const float4*>(sq + m * sd);
            for (int c = 0; c < (sd >> 2); c++) {
                const float4 x = __ldg(cen + c); const float4 y = qq[c];
                float t;
                t = __fsub_rn(y.x, x.x); dist = __fadd_rn(dist, __fmul_rn(t, t));
                t = __fsub_rn(y.y, x.y); dist = __fadd_rn(dist, __fmul_rn(t, t));
                t = __fsub_rn(y.z, x.z); dist = __fadd_rn(dist, __fmul_rn(t, t));
                t = __fsub_rn(y.w, x.w); dist = __fadd_rn(dist, __fmul_rn(t, t));
            }
        }
        return dist;
    }
    for (int m = 0; m < p.pq_m; m++) {
        const float* cen = p.codebooks + ((size_t)m * VDB_NCENT + code[m]) * sd;
        for (int d = 0; d < sd; d++) {
            float df = __fsub_rn(sq[m * sd + d], cen[d]);
            dist = __fadd_rn(dist, __fmul_rn(df, df));
        }
    }
    return dist;
}

// searchLayerClosest / searchLayerClosestCompressed (HnswPqIndex.java:761-827, HnswIndex.java:511-552): executed by
// warp 0; lane j evaluates neighbour j, the first improving neighbour in list order wins (ballot + ffs == `break`).
__device__ int greedy_level(const HnswSearchParams& p, const float* sq, float qnorm, int cur, int level, bool compressed,
                            unsigned long long& evals) {
    const int lane = threadIdx.x & 31;
    const int32_t* adjl = p.adj + (size_t)level * p.n * p.deg;
    float curd = 0.0f;
    if (lane == 0) curd = compressed ? pq_direct_dist(p, sq, cur) : exact_seq_dist(p, sq, qnorm, cur);
    curd = __shfl_sync(0xffffffffu, curd, 0);
    if (lane == 0 && !compressed) evals++;
    while (true) {
        bool moved = false;
        for (int j0 = 0; j0 < p.deg && !moved; j0 += 32) {
            int j = j0 + lane;
            int nb = j < p.deg ? adjl[(size_t)cur * p.deg + j] : -1;
            float d = 0.0f;
            bool better = false;
            if (nb >= 0) {
                d = compressed ? pq_direct_dist(p, sq, nb) : exact_seq_dist(p, sq, qnorm, nb);
                better = d < curd;
            }
            const unsigned mask = __ballot_sync(0xffffffffu, better);
            const unsigned valid = __ballot_sync(0xffffffffu, nb >= 0);
            if (mask) {
                int first = __ffs(mask) - 1;
                if (lane == 0 && !compressed) evals += first + 1;   // Java evaluates up to and including the winner
                cur = __shfl_sync(0xffffffffu, nb, first);
                curd = __shfl_sync(0xffffffffu, d, first);
                moved = true;
            } else if (lane == 0 && !compressed) {
                evals += __popc(valid);
            }
        }
        if (!moved) break;
    }
    return cur;
}

// ---- traversal distance: a group of G lanes per row, NCH 16-byte chunks per lane, U rows in flight ----------------
// fp32 rows: chunk = 4 floats; the per-lane fma order and the xor butterfly are mirrored by the oracle (l2sq_gpu/dot_gpu).
// bf16 rows: chunk = 8 elements widened to fp32 (exact), same accumulation.
template <bool BF16> struct Chunk;
template <> struct Chunk<false> { static constexpr int EL = 4; };
template <> struct Chunk<true> { static constexpr int EL = 8; };

template <bool BF16>
__device__ __forceinline__ void unpack(const uint4& v, float* x) {
    if (BF16) {
        x[0] = __uint_as_float(v.x << 16); x[1] = __uint_as_float(v.x & 0xffff0000u);
        x[2] = __uint_as_float(v.y << 16); x[3] = __uint_as_float(v.y & 0xffff0000u);
        x[4] = __uint_as_float(v.z << 16); x[5] = __uint_as_float(v.z & 0xffff0000u);
        x[6] = __uint_as_float(v.w << 16); x[7] = __uint_as_float(v.w & 0xffff0000u);
    } else {
        x[0] = __uint_as_float(v.x); x[1] = __uint_as_float(v.y); x[2] = __uint_as_float(v.z); x[3] = __uint_as_float(v.w);
    }
}

// distances of fresh rows fnode[0..nf) -> accepted keys appended to fkey (slot via shared atomic). All threads call.
template <bool BF16, int G, int NCH, int U>
__device__ __forceinline__ void fresh_distances(const HnswSearchParams& p, const float* __restrict__ sq, float qnorm,
                                                const int* __restrict__ fnode, int nf, uint64_t* __restrict__ fkey,
                                                int* nacc, bool full, float bound) {
    constexpr int EL = Chunk<BF16>::EL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int GPW = 32 / G;
    const int g = lane % G, grp = warp * GPW + lane / G;
    constexpr int NG = HN_WARPS * GPW;
    const int nchunk = p.tstride >> 4;                 // 16-byte chunks per row
    // query slice of this lane, in registers
    float qr[NCH][EL];
#pragma unroll
    for (int c = 0; c < NCH; c++) {
        const int ch = g + c * G;
#pragma unroll
        for (int e = 0; e < EL; e++) qr[c][e] = ch < nchunk ? sq[ch * EL + e] : 0.0f;
    }
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    for (int base = 0; base < nf; base += NG * U) {
        int node[U];
        uint4 v[U][NCH];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int j = base + grp * U + u;
            node[u] = j < nf ? fnode[j] : -1;
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                const int ch = g + c * G;
                v[u][c] = make_uint4(0u, 0u, 0u, 0u);
                if (node[u] >= 0 && ch < nchunk)
                    v[u][c] = __ldg(reinterpret_cast<const uint4*>(tb + (size_t)node[u] * p.tstride) + ch);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            float acc = 0.0f;
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                float x[EL];
                unpack<BF16>(v[u][c], x);
                if (p.metric == 0) {
#pragma unroll
                    for (int e = 0; e < EL; e++) { const float df = __fsub_rn(x[e], qr[c][e]); acc = __fmaf_rn(df, df, acc); }
                } else {
#pragma unroll
                    for (int e = 0; e < EL; e++) acc = __fmaf_rn(x[e], qr[c][e], acc);
                }
            }
#pragma unroll
            for (int off = G / 2; off >= 1; off >>= 1) acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, off));
            if (g == 0 && node[u] >= 0) {
                const float d = p.metric == 0 ? acc : cosine_dist_from_dot(acc, p.norms[node[u]], qnorm);
                if (!full || d < bound) {
                    const int slot = atomicAdd(nacc, 1);
                    fkey[slot] = pool_key(d, node[u]);
                }
            }
        }
    }
}

// rows with more than 4*32 chunks (D > 512 fp32 / 1024 bf16): chunks are walked in passes of 128 per row
template <bool BF16>
__device__ __forceinline__ void fresh_distances_long(const HnswSearchParams& p, const float* __restrict__ sq, float qnorm,
                                                     const int* __restrict__ fnode, int nf, uint64_t* __restrict__ fkey,
                                                     int* nacc, bool full, float bound) {
    constexpr int EL = Chunk<BF16>::EL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nchunk = p.tstride >> 4;
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    for (int base = 0; base < nf; base += HN_WARPS * 2) {
        float acc[2] = {0.0f, 0.0f};
        int node[2];
#pragma unroll
        for (int u = 0; u < 2; u++) { const int j = base + warp * 2 + u; node[u] = j < nf ? fnode[j] : -1; }
        for (int c0 = 0; c0 < nchunk; c0 += 128) {
            uint4 v[2][4];
#pragma unroll
            for (int u = 0; u < 2; u++)
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int ch = c0 + lane + c * 32;
                    v[u][c] = make_uint4(0u, 0u, 0u, 0u);
                    if (node[u] >= 0 && ch < nchunk) v[u][c] = __ldg(reinterpret_cast<const uint4*>(tb + (size_t)node[u] * p.tstride) + ch);
                }
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int ch = c0 + lane + c * 32;
                float qv[EL];
#pragma unroll
                for (int e = 0; e < EL; e++) qv[e] = ch < nchunk ? sq[ch * EL + e] : 0.0f;
#pragma unroll
                for (int u = 0; u < 2; u++) {
                    float x[EL];
                    unpack<BF16>(v[u][c], x);
                    if (p.metric == 0) {
#pragma unroll
                        for (int e = 0; e < EL; e++) { const float df = __fsub_rn(x[e], qv[e]); acc[u] = __fmaf_rn(df, df, acc[u]); }
                    } else {
#pragma unroll
                        for (int e = 0; e < EL; e++) acc[u] = __fmaf_rn(x[e], qv[e], acc[u]);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            float a = acc[u];
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) a = __fadd_rn(a, __shfl_xor_sync(0xffffffffu, a, off));
            if (lane == 0 && node[u] >= 0) {
                const float d = p.metric == 0 ? a : cosine_dist_from_dot(a, p.norms[node[u]], qnorm);
                if (!full || d < bound) { const int slot = atomicAdd(nacc, 1); fkey[slot] = pool_key(d, node[u]); }
            }
        }
    }
}

template <bool BF16>
__device__ __forceinline__ void fresh_dispatch(const HnswSearchParams& p, const float* sq, float qnorm, const int* fnode, int nf,
                                               uint64_t* fkey, int* nacc, bool full, float bound) {
    const int nchunk = p.tstride >> 4;
    if (nchunk <= 4) fresh_distances<BF16, 4, 1, 8>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else if (nchunk <= 8) fresh_distances<BF16, 8, 1, 8>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else if (nchunk <= 16) fresh_distances<BF16, 16, 1, 8>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else if (nchunk <= 32) fresh_distances<BF16, 32, 1, 8>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else if (nchunk <= 64) fresh_distances<BF16, 32, 2, 4>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else if (nchunk <= 96) fresh_distances<BF16, 32, 3, 2>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else if (nchunk <= 128) fresh_distances<BF16, 32, 4, 2>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
    else fresh_distances_long<BF16>(p, sq, qnorm, fnode, nf, fkey, nacc, full, bound);
}

// first index whose key (expanded bit ignored) is >= key. The expanded bit is NOT part of the order, so marking an
// entry never disturbs the sorted invariant (keys stay unique: one entry per node).
__device__ __forceinline__ int lower_bound_u64(const uint64_t* __restrict__ a, int n, uint64_t key) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((a[mid] & ~EXP_BIT) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

template <bool BF16>
__global__ void __launch_bounds__(HN_NT, 2) hnsw_search_kernel(const HnswSearchParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int qpad = (p.dim + 7) & ~7;
    float* sq = reinterpret_cast<float*>(smem_raw);                       // [qpad] query (zero padded)
    uint64_t* pool = reinterpret_cast<uint64_t*>(sq + qpad);              // [ef] sorted
    uint64_t* fkey = pool + p.ef;                                         // [HN_FMAX] accepted fresh keys
    int* fnode = reinterpret_cast<int*>(fkey + HN_FMAX);                  // [HN_FMAX] fresh nodes
    const int vwords = (p.n + 31) >> 5;
    uint32_t* vis = p.visited_in_smem ? reinterpret_cast<uint32_t*>(fnode + HN_FMAX)
                                      : p.visited_scratch + (size_t)blockIdx.x * vwords;
    __shared__ HnswShared S;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ef = p.ef, variant = p.variant, B = p.batch;
    unsigned long long evals = 0, exps = 0;

    for (int q = blockIdx.x; q < p.nq; q += gridDim.x) {
        // ---- A. stage the query (HnswIndex normalises it: HnswIndex.java:344, core/Vector.java:107-116) ----
        __syncthreads();
        for (int i = tid; i < qpad; i += HN_NT) sq[i] = i < p.dim ? p.queries[(size_t)q * p.dim + i] : 0.0f;
        for (int i = tid; i < vwords; i += HN_NT) vis[i] = 0u;
        __syncthreads();
        if (p.metric == 1) {
            if (tid == 0) S.qnorm = norm_java(sq, p.dim);
            __syncthreads();
            const float nrm = S.qnorm;
            __syncthreads();
            for (int i = tid; i < p.dim; i += HN_NT) sq[i] = __fdiv_rn(sq[i], nrm);
            __syncthreads();
            if (tid == 0) S.qnorm = norm_java(sq, p.dim);
        } else if (tid == 0) S.qnorm = 0.0f;
        __syncthreads();
        const float qnorm = S.qnorm;

        // ---- B. greedy descent on levels max_level-1 .. 1 (warp 0), then the entry point's layer-0 distance ----
        if (warp == 0) {
            int cur = p.entry;
            const bool comp = (p.metric == 0) && p.trained;
            for (int l = p.max_level - 1; l > 0; l--) cur = greedy_level(p, sq, qnorm, cur, l, comp, evals);
            if (lane == 0) { S.cur = cur; fnode[0] = cur; S.nacc = 0; vis[cur >> 5] |= 1u << (cur & 31); }
        }
        __syncthreads();
        fresh_dispatch<BF16>(p, sq, qnorm, fnode, 1, fkey, &S.nacc, false, 0.0f);
        __syncthreads();
        if (tid == 0) {
            pool[0] = fkey[0];
            S.dentry = pool_dist(fkey[0]);
            S.count = 1; S.evicted = 0;
            evals++;
        }
        __syncthreads();

        // ---- C. layer-0 search in rounds ----
        while (true) {
            // C1. candidates.poll() x B: the first B unexpanded entries of the eligible prefix (sorted array)
            const int count = S.count;
            const bool full = count >= ef;
            int lim = count;
            if ((variant == 2 && !S.evicted) || (variant == 1 && full && !S.evicted))
                lim = lower_bound_u64(pool, count, ((uint64_t)f2ord(S.dentry) << 32) | 0x7fffffffull);   // entries with d <= dentry (Java: `c.d > furthest` stops)
            const int per = (lim + HN_NT - 1) / HN_NT;
            const int i0 = min(lim, tid * per), i1 = min(lim, i0 + per);
            int c = 0;
            for (int i = i0; i < i1; i++) c += (pool[i] & EXP_BIT) ? 0 : 1;
            int incl = c;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) { const int o = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += o; }
            if (lane == 31) S.wsum[warp] = incl;
            __syncthreads();
            int before = incl - c;
            for (int w = 0; w < warp; w++) before += S.wsum[w];
            int total = 0;
            for (int w = 0; w < HN_WARPS; w++) total += S.wsum[w];
            if (total == 0) break;                                    // no eligible candidate left: search ends
            if (before < B && c > 0) {
                int s = before;
                for (int i = i0; i < i1 && s < B; i++) {
                    const uint64_t kk = pool[i];
                    if (!(kk & EXP_BIT)) { S.sel[s++] = (int)(kk & 0x7fffffffull); pool[i] = kk | EXP_BIT; }
                }
            }
            if (tid == 0) { S.nsel = min(total, B); S.nfresh = 0; S.nacc = 0; }
            __syncthreads();
            const int nsel = S.nsel;
            // C2. neighbour lists + visited test-and-set; fresh nodes appended warp-aggregated
            for (int s = warp; s < nsel; s += HN_WARPS) {
                const int32_t* row = p.adj + (size_t)S.sel[s] * p.deg;
                for (int j0 = 0; j0 < p.deg; j0 += 32) {
                    const int j = j0 + lane;
                    const int nb = j < p.deg ? __ldg(row + j) : -1;
                    bool fresh = false;
                    if (nb >= 0) {
                        const uint32_t bit = 1u << (nb & 31);
                        const uint32_t old = atomicOr(&vis[nb >> 5], bit);
                        fresh = !(old & bit);
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, fresh);
                    int basec = 0;
                    if (lane == 0 && m) basec = atomicAdd(&S.nfresh, __popc(m));
                    basec = __shfl_sync(0xffffffffu, basec, 0);
                    if (fresh) fnode[basec + __popc(m & ((1u << lane) - 1u))] = nb;
                }
            }
            __syncthreads();
            const int nf = S.nfresh;
            if (tid == 0) { exps += nsel; evals += nf; }
            // C3. distances of the fresh rows; acceptance: results.size() < ef || d < furthest
            //     (furthest live: HnswPqIndex.java:953-963; stale == entry distance until the first eviction: :876-884,
            //      HnswIndex.java:611-625)
            float bound = 0.0f;
            if (full) bound = (variant == 0 || S.evicted) ? pool_dist(pool[ef - 1]) : S.dentry;
            fresh_dispatch<BF16>(p, sq, qnorm, fnode, nf, fkey, &S.nacc, full, bound);
            __syncthreads();
            const int nacc = S.nacc;
            if (nacc == 0) continue;
            // C4. sort the accepted keys, merge into the sorted array by ranking, truncate at ef
            int fp2 = 1; while (fp2 < nacc) fp2 <<= 1;
            for (int i = nacc + tid; i < fp2; i += HN_NT) fkey[i] = KEY_MAX;
            if (fp2 > 1) block_bitonic_sort<false>(fkey, nullptr, fp2); else __syncthreads();
            // positions of the fresh keys against the OLD array (<= HN_FMAX / HN_NT per thread)
            uint64_t fk[HN_FMAX / HN_NT]; int fpos[HN_FMAX / HN_NT];
#pragma unroll
            for (int t = 0; t < HN_FMAX / HN_NT; t++) {
                const int j = tid + t * HN_NT;
                fpos[t] = -1;
                if (j < nacc) { fk[t] = fkey[j]; fpos[t] = j + lower_bound_u64(pool, count, fk[t]); }
            }
            if (tid == 0) S.lb0 = lower_bound_u64(pool, count, fkey[0]);
            __syncthreads();
            const int lb0 = S.lb0;
            // move array entries upward, top chunk first (an entry never moves down, so a chunk only overwrites slots of
            // chunks already moved or of itself, which were read before the barrier)
            constexpr int PERC = 4;
            for (int hi = count; hi > lb0; hi -= HN_NT * PERC) {
                const int lo = max(lb0, hi - HN_NT * PERC);
                uint64_t e[PERC]; int pos[PERC];
#pragma unroll
                for (int t = 0; t < PERC; t++) {
                    const int i = lo + tid + t * HN_NT;
                    pos[t] = -1;
                    if (i < hi) { e[t] = pool[i]; pos[t] = i + lower_bound_u64(fkey, nacc, e[t] & ~EXP_BIT); }
                }
                __syncthreads();
#pragma unroll
                for (int t = 0; t < PERC; t++) if (pos[t] >= 0 && pos[t] < ef) pool[pos[t]] = e[t];
                __syncthreads();
            }
#pragma unroll
            for (int t = 0; t < HN_FMAX / HN_NT; t++) if (fpos[t] >= 0 && fpos[t] < ef) pool[fpos[t]] = fk[t];
            if (tid == 0) {
                const int nc = count + nacc;
                if (nc > ef) S.evicted = 1;
                S.count = min(nc, ef);
            }
            __syncthreads();
        }

        // ---- D. final: best entries by traversal distance -> exact re-score in the reference's sequential order
        //         (HnswPqIndex.java:646-664 reranks every candidate; with exact traversal distances the values are
        //         identical, with bf16 rows the shortlist margin covers the rounding) -> (dist,id) sort -> top-k ----
        __syncthreads();
        int rs = 32;
        while (rs < p.k + 22) rs <<= 1;
        if (rs > HN_FMAX) rs = HN_FMAX;
        const int have = min(S.count, rs);
        for (int j = tid; j < rs; j += HN_NT) {
            uint64_t key = KEY_MAX;
            if (j < have) {
                const int node = (int)(pool[j] & 0x7fffffffull);
                key = make_key(exact_seq_dist(p, sq, qnorm, node), p.ids[node]);
            }
            fkey[j] = key;
        }
        block_bitonic_sort<false>(fkey, nullptr, rs);
        for (int j = tid; j < p.k; j += HN_NT) {
            const uint64_t key = j < rs ? fkey[j] : KEY_MAX;
            if (key == KEY_MAX) { p.out_ids[(size_t)q * p.k + j] = -1; p.out_dists[(size_t)q * p.k + j] = __int_as_float(0x7f800000); }
            else { p.out_ids[(size_t)q * p.k + j] = key_id(key); p.out_dists[(size_t)q * p.k + j] = key_dist(key); }
        }
    }
    if (p.counters && tid == 0) {
        if (evals) atomicAdd(&p.counters[0], evals);
        if (exps) atomicAdd(&p.counters[1], exps);
    }
}

size_t hnsw_search_smem_bytes(const HnswSearchParams& p) {
    size_t b = (size_t)((p.dim + 7) & ~7) * 4 + (size_t)p.ef * 8 + (size_t)HN_FMAX * 12;
    if (p.visited_in_smem) b += (size_t)((p.n + 31) / 32) * 4;
    return b;
}
int hnsw_max_batch(int deg) { return std::min(HN_MAXB, HN_FMAX / std::max(deg, 1)); }
int launch_hnsw_search(const HnswSearchParams& p, int grid, cudaStream_t st) {
    if (p.nq <= 0) return 0;
    if (p.k + 22 > HN_FMAX || p.deg > 64 || p.batch < 1 || p.batch * p.deg > HN_FMAX || p.batch > HN_MAXB) return 1;
    if ((p.tstride & 15) != 0) return 1;
    const size_t smem = hnsw_search_smem_bytes(p);
    if (p.tstore == 1) {
        if (cudaFuncSetAttribute(hnsw_search_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 2;
        hnsw_search_kernel<true><<<grid, HN_NT, smem, st>>>(p);
    } else {
        if (cudaFuncSetAttribute(hnsw_search_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 2;
        hnsw_search_kernel<false><<<grid, HN_NT, smem, st>>>(p);
    }
    return 0;
}

// ---- traversal copy of the base rows: zero-padded to a multiple of 8 elements, fp32 or bf16 (round to nearest even) ----
__global__ void make_traversal_rows_kernel(const float* __restrict__ base, int64_t n, int dim, int dpad, int bf16, void* __restrict__ out) {
    const int64_t total = n * dpad;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / dpad; const int c = (int)(i % dpad);
        const float v = c < dim ? base[r * dim + c] : 0.0f;
        if (bf16) reinterpret_cast<__nv_bfloat16*>(out)[i] = __float2bfloat16_rn(v);
        else reinterpret_cast<float*>(out)[i] = v;
    }
}
void launch_make_traversal_rows(const float* d_base, int64_t n, int dim, int dpad, int bf16, void* d_out, cudaStream_t st) {
    if (n <= 0) return;
    const int64_t total = n * dpad;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    make_traversal_rows_kernel<<<blocks, 256, 0, st>>>(d_base, n, dim, dpad, bf16, d_out);
}
