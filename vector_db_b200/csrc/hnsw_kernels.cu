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
#include <cuda_fp8.h>

#include "kernels.h"

#define HN_NT 256
#define HN_WARPS (HN_NT / 32)
#define EXP_BIT 0x80000000ull
#define HN_FMAX 2048      // fresh rows per round (batch * deg <= HN_FMAX)
#define HN_MAXB 64
#ifndef HN_SMALL_SORT
#define HN_SMALL_SORT 128   // rounds with at most this many accepted keys are sorted by counting ranks
#endif


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
// `vis` holds only the removed-node bits at this point: a dead neighbour is skipped (`!vectors.containsKey`, HnswPqIndex.java:778)
__device__ __forceinline__ bool node_dead(const uint32_t* vis, int nb) { return (vis[nb >> 5] >> (nb & 31)) & 1u; }
__device__ int greedy_level(const HnswSearchParams& p, const float* sq, float qnorm, int cur, int level, bool compressed,
                            unsigned long long& evals, const uint32_t* vis) {
    const int lane = threadIdx.x & 31;
    const int32_t* adjl = p.adj + (size_t)level * p.adj_stride * p.deg;
    float curd = 0.0f;
    if (lane == 0) curd = compressed ? pq_direct_dist(p, sq, cur) : exact_seq_dist(p, sq, qnorm, cur);
    curd = __shfl_sync(0xffffffffu, curd, 0);
    if (lane == 0 && !compressed) evals++;
    while (true) {
        bool moved = false;
        for (int j0 = 0; j0 < p.deg && !moved; j0 += 32) {
            int j = j0 + lane;
            int nb = j < p.deg ? adjl[(size_t)cur * p.deg + j] : -1;
            if (nb >= 0 && p.dead && node_dead(vis, nb)) nb = -1;
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

// searchLayerClosestCompressed (HnswPqIndex.java:794-827) with the whole CTA: thread (neighbour = tid/8, part = tid%8)
// sums ceil(M/8) consecutive subspace terms of the direct PQ distance (a3, HnswPqIndex.java:371-386), the 8 partials are
// added in order (oracle twin: pq_direct_parts), the first improving neighbour in list order wins (`break` :821).
// ~1 memory round trip per hop instead of a 512-term serial chain on one lane. Barriers inside; all threads call.
struct GreedyShared { float d[64]; int best; int cur; float curd; };
__device__ __forceinline__ float pq_partial(const HnswSearchParams& p, const float* __restrict__ sq, int node, int part) {
    const int sd = p.dim / p.pq_m, mp = (p.pq_m + 7) >> 3;
    const uint8_t* code = p.codes + (size_t)node * p.pq_m;
    float ps = 0.0f;
    const int m1 = min(p.pq_m, (part + 1) * mp);
    for (int m = part * mp; m < m1; m++) {
        const float* cen = p.codebooks + ((size_t)m * VDB_NCENT + code[m]) * sd;
        float t = 0.0f;
        if ((sd & 3) == 0) {
            for (int c = 0; c < (sd >> 2); c++) {
                const float4 x = __ldg(reinterpret_cast<const float4*>(cen) + c);
                const float4 y = reinterpret_cast<const float4*>(sq + m * sd)[c];
                float df;
                df = __fsub_rn(y.x, x.x); t = __fadd_rn(t, __fmul_rn(df, df));
                df = __fsub_rn(y.y, x.y); t = __fadd_rn(t, __fmul_rn(df, df));
                df = __fsub_rn(y.z, x.z); t = __fadd_rn(t, __fmul_rn(df, df));
                df = __fsub_rn(y.w, x.w); t = __fadd_rn(t, __fmul_rn(df, df));
            }
        } else {
            for (int d = 0; d < sd; d++) { const float df = __fsub_rn(sq[m * sd + d], __ldg(cen + d)); t = __fadd_rn(t, __fmul_rn(df, df)); }
        }
        ps = __fadd_rn(ps, t);
    }
    return ps;
}
__device__ __forceinline__ float pq_combine8(float ps, int lane) {      // partials of lanes 8k..8k+7 added in order
    const int base = lane & ~7;
    float total = __shfl_sync(0xffffffffu, ps, base);
#pragma unroll
    for (int i = 1; i < 8; i++) total = __fadd_rn(total, __shfl_sync(0xffffffffu, ps, base + i));
    return total;
}
__device__ int greedy_level_pq(const HnswSearchParams& p, const float* sq, int cur, int level, GreedyShared& G, const uint32_t* vis) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int32_t* adjl = p.adj + (size_t)level * p.adj_stride * p.deg;
    const int nbr = tid >> 3, part = tid & 7;
    {
        const float d = pq_combine8(pq_partial(p, sq, cur, part), lane);
        if (tid == 0) { G.cur = cur; G.curd = d; }
    }
    __syncthreads();
    while (true) {
        cur = G.cur;
        const float curd = G.curd;
        if (tid == 0) G.best = 0x7fffffff;
        __syncthreads();
        for (int j0 = 0; j0 < p.deg; j0 += HN_NT / 8) {
            const int j = j0 + nbr;
            int nb = j < p.deg ? __ldg(adjl + (size_t)cur * p.deg + j) : -1;
            if (nb >= 0 && p.dead && node_dead(vis, nb)) nb = -1;
            const float ps = nb >= 0 ? pq_partial(p, sq, nb, part) : 0.0f;
            const float d = pq_combine8(ps, lane);
            if (part == 0 && nb >= 0) { G.d[j] = d; if (d < curd) atomicMin(&G.best, j); }
        }
        __syncthreads();
        const int best = G.best;
        if (best == 0x7fffffff) break;
        __syncthreads();
        if (tid == 0) { G.cur = __ldg(adjl + (size_t)cur * p.deg + best); G.curd = G.d[best]; }
        __syncthreads();
    }
    return G.cur;
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


// Distance of one traversal row from its per-lane partial sums.
//   fp32 rows (strict): acc = sum (x-q)^2  (sq-L2)  or  sum x*q (cosine)  — the oracle's l2sq_gpu / dot_gpu order.
//   bf16 rows (fast)  : acc = sum x*q for both metrics; sq-L2 is ||q||^2 + ||x||^2 - 2 x.q with ||x||^2 precomputed from
//                       the bf16 row (2 instructions per element instead of 3; the answer is re-scored exactly anyway).
// rowscal = the row's scalar fetched together with the row: Java norm (cosine) or ||x_bf16||^2 (bf16 sq-L2)
template <bool BF16>
__device__ __forceinline__ float row_scalar(const HnswSearchParams& p, int node) {
    if (p.metric == 1) return __ldg(p.norms + node);
    if (BF16) return __ldg(p.tnorm + node);
    return 0.0f;
}
template <bool BF16>
__device__ __forceinline__ float finish_dist(const HnswSearchParams& p, float acc, float rowscal, float qnorm, float qq) {
    if (p.metric == 1) return cosine_dist_from_dot(acc, rowscal, qnorm);
    if (BF16) return fmaxf(__fmaf_rn(-2.0f, acc, __fadd_rn(qq, rowscal)), 0.0f);
    return acc;
}

struct FreshOut {
    uint64_t* fkey;   // accepted keys, appended
    int* nacc;        // shared counter
    bool full;        // results.size() >= ef
    float bound;      // furthest (live or stale)
    bool ordered;     // keep EVERY key at its fresh-list position j (fkey[j]): the round is replayed in Java's order afterwards
};
__device__ __forceinline__ void accept_key(const FreshOut& o, float d, int node, int j) {
    if (o.ordered) { o.fkey[j] = pool_key(d, node); return; }
    if (!o.full || d < o.bound) {
        const int slot = atomicAdd(o.nacc, 1);
        o.fkey[slot] = pool_key(d, node);
    }
}

// issue the loads of one pass: U rows x NCH chunks per lane (+ the row scalar on the group's first lane)
template <bool BF16, int G, int NCH, int U>
__device__ __forceinline__ void issue_rows(const HnswSearchParams& p, const unsigned char* __restrict__ tb, const int* __restrict__ fnode,
                                           int nf, int base, int grp, int g, int nchunk, int (&nd)[U], float (&sc)[U], uint4 (&vv)[U][NCH]) {
#pragma unroll
    for (int u = 0; u < U; u++) {
        const int j = base + grp * U + u;
        nd[u] = j < nf ? fnode[j] : -1;
        sc[u] = (g == 0 && nd[u] >= 0) ? row_scalar<BF16>(p, nd[u]) : 0.0f;
#pragma unroll
        for (int c = 0; c < NCH; c++) {
            const int ch = g + c * G;
            vv[u][c] = make_uint4(0u, 0u, 0u, 0u);
            if (nd[u] >= 0 && ch < nchunk) vv[u][c] = __ldg(reinterpret_cast<const uint4*>(tb + (size_t)nd[u] * p.tstride) + ch);
        }
    }
}

// distances of fresh rows fnode[0..nf): a group of G lanes per row, NCH 16-byte chunks per lane, U rows per pass and the
// next pass's loads issued before the current pass is reduced (software pipeline). All threads of the CTA call.
template <bool BF16, int G, int NCH, int U, bool PIPE>
__device__ __noinline__ void fresh_distances(const HnswSearchParams& p, const float* __restrict__ sq, float qnorm, float qq,
                                                const int* __restrict__ fnode, int nf, const FreshOut& o) {
    constexpr int EL = Chunk<BF16>::EL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int GPW = 32 / G;
    const int g = lane % G, grp = warp * GPW + lane / G;
    constexpr int NG = HN_WARPS * GPW;
    const int nchunk = p.tstride >> 4;                 // 16-byte chunks per row
    float qr[NCH][EL];                                 // query slice of this lane, in registers
#pragma unroll
    for (int c = 0; c < NCH; c++) {
        const int ch = g + c * G;
#pragma unroll
        for (int e = 0; e < EL; e++) qr[c][e] = ch < nchunk ? sq[ch * EL + e] : 0.0f;
    }
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    int node[U], nnode[U];
    float rs[U], nrs[U];
    uint4 v[U][NCH], nv[U][NCH];
    issue_rows<BF16, G, NCH, U>(p, tb, fnode, nf, 0, grp, g, nchunk, node, rs, v);
    for (int base = 0; base < nf; base += NG * U) {
        const bool more = PIPE && base + NG * U < nf;
        if (more) issue_rows<BF16, G, NCH, U>(p, tb, fnode, nf, base + NG * U, grp, g, nchunk, nnode, nrs, nv);
#pragma unroll
        for (int u = 0; u < U; u++) {
            float acc = 0.0f;
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                float x[EL];
                unpack<BF16>(v[u][c], x);
                if (!BF16 && p.metric == 0) {
#pragma unroll
                    for (int e = 0; e < EL; e++) { const float df = __fsub_rn(x[e], qr[c][e]); acc = __fmaf_rn(df, df, acc); }
                } else {
#pragma unroll
                    for (int e = 0; e < EL; e++) acc = __fmaf_rn(x[e], qr[c][e], acc);
                }
            }
#pragma unroll
            for (int off = G / 2; off >= 1; off >>= 1) acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, off));
            if (g == 0 && node[u] >= 0) accept_key(o, finish_dist<BF16>(p, acc, rs[u], qnorm, qq), node[u], base + grp * U + u);
        }
        if (more) {
#pragma unroll
            for (int u = 0; u < U; u++) {
                node[u] = nnode[u]; rs[u] = nrs[u];
#pragma unroll
                for (int c = 0; c < NCH; c++) v[u][c] = nv[u][c];
            }
        } else if (!PIPE && base + NG * U < nf) {
            issue_rows<BF16, G, NCH, U>(p, tb, fnode, nf, base + NG * U, grp, g, nchunk, node, rs, v);
        }
    }
}

// rows with more than 4*32 chunks (D > 512 fp32 / 1024 bf16): chunks are walked in passes of 128 per row
template <bool BF16>
__device__ __noinline__ void fresh_distances_long(const HnswSearchParams& p, const float* __restrict__ sq, float qnorm, float qq,
                                                     const int* __restrict__ fnode, int nf, const FreshOut& o) {
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
                    if (!BF16 && p.metric == 0) {
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
            if (lane == 0 && node[u] >= 0) accept_key(o, finish_dist<BF16>(p, a, row_scalar<BF16>(p, node[u]), qnorm, qq), node[u], base + warp * 2 + u);
        }
    }
}

template <bool BF16>
__device__ __forceinline__ void fresh_dispatch(const HnswSearchParams& p, const float* sq, float qnorm, float qq, const int* fnode,
                                               int nf, const FreshOut& o) {
    const int nchunk = p.tstride >> 4;
    if (nchunk <= 4) fresh_distances<BF16, 4, 1, 4, true>(p, sq, qnorm, qq, fnode, nf, o);
    else if (nchunk <= 8) fresh_distances<BF16, 8, 1, 4, true>(p, sq, qnorm, qq, fnode, nf, o);
    else if (nchunk <= 16) fresh_distances<BF16, 16, 1, 4, true>(p, sq, qnorm, qq, fnode, nf, o);
    else if (nchunk <= 32) fresh_distances<BF16, 32, 1, 4, true>(p, sq, qnorm, qq, fnode, nf, o);
    else if (nchunk <= 64) fresh_distances<BF16, 32, 2, 4, true>(p, sq, qnorm, qq, fnode, nf, o);
    else if (nchunk <= 96) fresh_distances<BF16, 32, 3, 2, false>(p, sq, qnorm, qq, fnode, nf, o);
    else if (nchunk <= 128) fresh_distances<BF16, 32, 4, 2, false>(p, sq, qnorm, qq, fnode, nf, o);
    else fresh_distances_long<BF16>(p, sq, qnorm, qq, fnode, nf, o);
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
// number of entries of the ascending int array r[0..n) that are <= x
__device__ __forceinline__ int count_le_i32(const int* __restrict__ r, int n, int x) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (r[mid] <= x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// N binary searches in lock-step: the trip count depends on n only (halving form), so the N dependent chains of shared-memory
// loads are independent of each other and overlap — the `while (lo < hi)` forms above have data-dependent trip counts and run
// one after the other. Same results as count_le_i32 / lower_bound_u64.
template <int N>
__device__ __forceinline__ void count_le_i32_multi(const int* __restrict__ r, int n, const int (&x)[N], int (&out)[N]) {
    if (n <= 0) {
#pragma unroll
        for (int t = 0; t < N; t++) out[t] = 0;
        return;
    }
    int base[N];
#pragma unroll
    for (int t = 0; t < N; t++) base[t] = 0;
    int len = n;
    while (len > 1) {
        const int half = len >> 1;
#pragma unroll
        for (int t = 0; t < N; t++) base[t] += (r[base[t] + half - 1] <= x[t]) ? half : 0;
        len -= half;
    }
#pragma unroll
    for (int t = 0; t < N; t++) out[t] = base[t] + ((r[base[t]] <= x[t]) ? 1 : 0);
}
template <int N>
__device__ __forceinline__ void lower_bound_u64_multi(const uint64_t* __restrict__ a, int n, const uint64_t (&key)[N], int (&out)[N]) {
    if (n <= 0) {
#pragma unroll
        for (int t = 0; t < N; t++) out[t] = 0;
        return;
    }
    int base[N];
#pragma unroll
    for (int t = 0; t < N; t++) base[t] = 0;
    int len = n;
    while (len > 1) {
        const int half = len >> 1;
#pragma unroll
        for (int t = 0; t < N; t++) base[t] += ((a[base[t] + half - 1] & ~EXP_BIT) < key[t]) ? half : 0;
        len -= half;
    }
#pragma unroll
    for (int t = 0; t < N; t++) out[t] = base[t] + (((a[base[t]] & ~EXP_BIT) < key[t]) ? 1 : 0);
}

// ---- bf16 traversal distances on the tensor pipe ------------------------------------------------------------------------
// The fresh rows of a round form a [rows x D] bf16 matrix that is multiplied by ONE vector (the query): a gather-bound
// matrix-vector product. It is issued as mma.sync m16n8k16 (16 rows per warp tile; the 8 result columns all hold the same
// query, column 0 is read) because the rows arrive straight from global memory in registers — each lane's 16-byte load
// IS its A fragment for two MMAs (the k index is permuted consistently between A and B, a dot product does not care) —
// so there is no shared-memory staging for a tcgen05 descriptor to point at. ~110 instructions per 16 rows instead of
// ~1000 on the FMA pipe; the kernel is issue-bound, not tensor-bound (profiles/).
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// sqh: the query rounded to bf16, zero padded, in shared memory ([tstride/2] elements). All threads of the CTA call.
// FULL: the row is a whole number of 16-block stages (tstride % 256 == 0, e.g. 512-d bf16): no per-load bound checks
// 32-byte global load (sm_100: LDG.E.256): lanes c = 0..3 of a row group cover one whole 128-byte line of their row, so a
// warp-wide load touches 8 lines for 128 useful bytes each (the 16-byte form: 8 lines for 64 bytes each). Rows are read once
// per query: no L1 allocation (the 2 x 111 KB of shared memory leave ~30 KB of L1). Measured on C3 (1184 queries):
// 16-byte loads + L2 prefetch 11.95 ms, 32-byte loads 11.42, 32-byte loads without the prefetch 11.00, + no_allocate 10.74.
struct U8 { uint4 lo, hi; };
__device__ __forceinline__ U8 ld256_nc(const unsigned char* p) {
    U8 r;
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w) : "l"(p));
    return r;
}
// ---- fp8 (e4m3) traversal rows, opt-in (trav_store = 2): half the row bytes of the bf16 copy. A row is stored as
//      e4m3(x * s_row) with a per-row power-of-two scale (exact to undo); the query as e4m3 with a per-query power-of-two
//      scale, by default in TWO terms (hi + lo / 16: the query side then adds no rounding worth the name; HN_FP8_LO = 0
//      keeps the hi term only and halves the MMAs); m16n8k32 MMAs accumulate in fp32.
#ifndef HN_FP8_LO
#define HN_FP8_LO 1
#endif
__device__ __forceinline__ float fp8_pow2_scale(float maxabs) {       // largest element -> [120, 240] (e4m3 max finite: 448)
    if (!(maxabs > 0.0f)) return 1.0f;
    const int e = min(100, max(-100, (int)floorf(log2f(240.0f / maxabs))));
    return exp2f((float)e);
}
__device__ __forceinline__ float fp8_to_float(__nv_fp8_storage_t h) {
    return __half2float(__half(__nv_cvt_fp8_to_halfraw(h, __NV_E4M3)));
}
__device__ __forceinline__ void mma_e4m3_16832(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// sq8: the query's hi term in bytes [0, dim), its lo term (x 16) in [dim, 2 dim). qinv = 1 / query scale.
// One 32-byte load per lane and row = 32 k-values = four MMAs per query term; NL loads per row in flight.
template <int NL>
__device__ __forceinline__ void fresh_distances_fp8(const HnswSearchParams& p, const unsigned char* __restrict__ sq8, float qnorm, float qq, float qinv,
                                                    const int* __restrict__ fnode, int nf, const FreshOut& o) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, c = lane & 3;
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    const int tstride = p.tstride;
    const int nst = tstride / (NL * 128);
    for (int tile = warp; tile * 16 < nf; tile += HN_WARPS) {
        const int j0 = tile * 16 + g, j1 = j0 + 8;
        const int n0 = fnode[min(j0, nf - 1)], n1 = fnode[min(j1, nf - 1)];
        const unsigned char* r0 = tb + (size_t)n0 * tstride + c * 32;
        const unsigned char* r1 = tb + (size_t)n1 * tstride + c * 32;
        float sc0 = 0.0f, sc1 = 0.0f, is0 = 0.0f, is1 = 0.0f;
        if (c == 0 && j0 < nf) { sc0 = p.metric == 1 ? __ldg(p.norms + n0) : __ldg(p.tnorm + n0); is0 = __ldg(p.tscale + n0); }
        if (c == 0 && j1 < nf) { sc1 = p.metric == 1 ? __ldg(p.norms + n1) : __ldg(p.tnorm + n1); is1 = __ldg(p.tscale + n1); }
        float ah[4] = {0.0f, 0.0f, 0.0f, 0.0f}, al[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        for (int st = 0; st < nst; st++) {
            U8 A0[NL], A1[NL];
#pragma unroll
            for (int i = 0; i < NL; i++) { A0[i] = ld256_nc(r0 + (st * NL + i) * 128); A1[i] = ld256_nc(r1 + (st * NL + i) * 128); }
#pragma unroll
            for (int i = 0; i < NL; i++) {
                const uint4* qh = reinterpret_cast<const uint4*>(sq8 + (st * NL + i) * 128 + c * 32);
                const uint4 h0 = qh[0], h1 = qh[1];
                mma_e4m3_16832(ah, A0[i].lo.x, A1[i].lo.x, A0[i].lo.y, A1[i].lo.y, h0.x, h0.y);
                mma_e4m3_16832(ah, A0[i].lo.z, A1[i].lo.z, A0[i].lo.w, A1[i].lo.w, h0.z, h0.w);
                mma_e4m3_16832(ah, A0[i].hi.x, A1[i].hi.x, A0[i].hi.y, A1[i].hi.y, h1.x, h1.y);
                mma_e4m3_16832(ah, A0[i].hi.z, A1[i].hi.z, A0[i].hi.w, A1[i].hi.w, h1.z, h1.w);
#if HN_FP8_LO
                const uint4* ql = reinterpret_cast<const uint4*>(sq8 + tstride + (st * NL + i) * 128 + c * 32);
                const uint4 l0 = ql[0], l1 = ql[1];
                mma_e4m3_16832(al, A0[i].lo.x, A1[i].lo.x, A0[i].lo.y, A1[i].lo.y, l0.x, l0.y);
                mma_e4m3_16832(al, A0[i].lo.z, A1[i].lo.z, A0[i].lo.w, A1[i].lo.w, l0.z, l0.w);
                mma_e4m3_16832(al, A0[i].hi.x, A1[i].hi.x, A0[i].hi.y, A1[i].hi.y, l1.x, l1.y);
                mma_e4m3_16832(al, A0[i].hi.z, A1[i].hi.z, A0[i].hi.w, A1[i].hi.w, l1.z, l1.w);
#endif
            }
        }
        if (c == 0) {
            if (j0 < nf) {
                const float dot = __fmul_rn(__fmaf_rn(al[0], 0.0625f, ah[0]), __fmul_rn(is0, qinv));
                accept_key(o, p.metric == 1 ? cosine_dist_from_dot(dot, sc0, qnorm) : fmaxf(__fmaf_rn(-2.0f, dot, __fadd_rn(qq, sc0)), 0.0f), n0, j0);
            }
            if (j1 < nf) {
                const float dot = __fmul_rn(__fmaf_rn(al[2], 0.0625f, ah[2]), __fmul_rn(is1, qinv));
                accept_key(o, p.metric == 1 ? cosine_dist_from_dot(dot, sc1, qnorm) : fmaxf(__fmaf_rn(-2.0f, dot, __fadd_rn(qq, sc1)), 0.0f), n1, j1);
            }
        }
    }
}

// ---- int8 traversal rows, opt-in (trav_store = 3): half the row bytes of the bf16 copy, NATIVE tensor op on sm_100a
//      (IMMA.16832.S8.S8; the e4m3 mma.sync above is lowered to F2FP unpacks + two HMMA per MMA, which doubles the ALU work of
//      the phase). Row = round(x * 127 / max|x_row|), query likewise; int32 accumulation is exact, dot = acc * s_row * s_query.
__device__ __forceinline__ void mma_s8_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// NL = 32-byte loads per lane, row and stage. Measured on C3 (B200, 4 096 queries): the whole 512-byte row in one stage (NL = 4,
// twice the bytes in flight per warp) 23.0 ms vs 22.6 ms with NL = 2 — like the bf16 rings, more loads in flight per warp do not help.
template <int NL>
__device__ __forceinline__ void fresh_distances_i8(const HnswSearchParams& p, const unsigned char* __restrict__ sq8, float qnorm, float qq, float qinv,
                                                   const int* __restrict__ fnode, int nf, const FreshOut& o) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, c = lane & 3;
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    const int tstride = p.tstride;
    const int nst = tstride / (NL * 128);
    for (int tile = warp; tile * 16 < nf; tile += HN_WARPS) {
        const int j0 = tile * 16 + g, j1 = j0 + 8;
        const int n0 = fnode[min(j0, nf - 1)], n1 = fnode[min(j1, nf - 1)];
        const unsigned char* r0 = tb + (size_t)n0 * tstride + c * 32;
        const unsigned char* r1 = tb + (size_t)n1 * tstride + c * 32;
        float sc0 = 0.0f, sc1 = 0.0f, is0 = 0.0f, is1 = 0.0f;
        if (c == 0 && j0 < nf) { sc0 = p.metric == 1 ? __ldg(p.norms + n0) : __ldg(p.tnorm + n0); is0 = __ldg(p.tscale + n0); }
        if (c == 0 && j1 < nf) { sc1 = p.metric == 1 ? __ldg(p.norms + n1) : __ldg(p.tnorm + n1); is1 = __ldg(p.tscale + n1); }
        int a0[4] = {0, 0, 0, 0}, a1[4] = {0, 0, 0, 0};
        for (int st = 0; st < nst; st++) {
            U8 A0[NL], A1[NL];
#pragma unroll
            for (int i = 0; i < NL; i++) { A0[i] = ld256_nc(r0 + (st * NL + i) * 128); A1[i] = ld256_nc(r1 + (st * NL + i) * 128); }
#pragma unroll
            for (int i = 0; i < NL; i++) {
                const uint4* qh = reinterpret_cast<const uint4*>(sq8 + (st * NL + i) * 128 + c * 32);
                const uint4 h0 = qh[0], h1 = qh[1];
                mma_s8_16832(a0, A0[i].lo.x, A1[i].lo.x, A0[i].lo.y, A1[i].lo.y, h0.x, h0.y);
                mma_s8_16832(a1, A0[i].lo.z, A1[i].lo.z, A0[i].lo.w, A1[i].lo.w, h0.z, h0.w);
                mma_s8_16832(a0, A0[i].hi.x, A1[i].hi.x, A0[i].hi.y, A1[i].hi.y, h1.x, h1.y);
                mma_s8_16832(a1, A0[i].hi.z, A1[i].hi.z, A0[i].hi.w, A1[i].hi.w, h1.z, h1.w);
            }
        }
        if (c == 0) {
            if (j0 < nf) {
                const float dot = __fmul_rn((float)(a0[0] + a1[0]), __fmul_rn(is0, qinv));
                accept_key(o, p.metric == 1 ? cosine_dist_from_dot(dot, sc0, qnorm) : fmaxf(__fmaf_rn(-2.0f, dot, __fadd_rn(qq, sc0)), 0.0f), n0, j0);
            }
            if (j1 < nf) {
                const float dot = __fmul_rn((float)(a0[2] + a1[2]), __fmul_rn(is1, qinv));
                accept_key(o, p.metric == 1 ? cosine_dist_from_dot(dot, sc1, qnorm) : fmaxf(__fmaf_rn(-2.0f, dot, __fadd_rn(qq, sc1)), 0.0f), n1, j1);
            }
        }
    }
}

#ifndef HN_NL
#define HN_NL 2           // 32-byte loads per lane, row and stage (a stage = HN_NL * 128 bytes of each row)
#endif
// Whole-stage rows (tstride % 256 == 0, e.g. 512-d bf16): 32-byte loads, no bound checks. One stage in flight per warp and
// NO register double buffer: measured on C3 (1184 queries, B200) the double-buffered forms were all slower — 16-byte loads +
// prefetched next stage 11.95 ms, 32-byte loads + next stage 10.74 (the second buffer spills at the 128-register cap), a
// four-deep register ring 14.9, a non-inlined copy with its own registers 12.4 — than load / multiply / load: 9.78 ms.
// The phase is not short of loads in flight (16 warps x 4 KB per SM); it is short of registers.
__device__ __forceinline__ void fresh_distances_mma_full(const HnswSearchParams& p, const uint4* __restrict__ sqh, float qnorm, float qq,
                                                         const int* __restrict__ fnode, int nf, const FreshOut& o) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, c = lane & 3;
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    const int tstride = p.tstride;
    const int nst = tstride / (HN_NL * 128);
    for (int tile = warp; tile * 16 < nf; tile += HN_WARPS) {
        const int j0 = tile * 16 + g, j1 = j0 + 8;
        const int n0 = fnode[min(j0, nf - 1)], n1 = fnode[min(j1, nf - 1)];
        const unsigned char* r0 = tb + (size_t)n0 * tstride + c * 32;
        const unsigned char* r1 = tb + (size_t)n1 * tstride + c * 32;
        const float sc0 = (c == 0 && j0 < nf) ? row_scalar<true>(p, n0) : 0.0f;
        const float sc1 = (c == 0 && j1 < nf) ? row_scalar<true>(p, n1) : 0.0f;
        float acc0[4] = {0.0f, 0.0f, 0.0f, 0.0f}, acc1[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        for (int st = 0; st < nst; st++) {
            U8 A0[HN_NL], A1[HN_NL];
#pragma unroll
            for (int i = 0; i < HN_NL; i++) { A0[i] = ld256_nc(r0 + (st * HN_NL + i) * 128); A1[i] = ld256_nc(r1 + (st * HN_NL + i) * 128); }
#pragma unroll
            for (int i = 0; i < HN_NL; i++) {
                const uint4* qp = sqh + ((st * HN_NL + i) * 8 + c * 2);       // the query's 32 bytes at the same k positions
                const uint4 qa = qp[0], qb = qp[1];
                mma_bf16_16816(acc0, A0[i].lo.x, A1[i].lo.x, A0[i].lo.y, A1[i].lo.y, qa.x, qa.y);
                mma_bf16_16816(acc1, A0[i].lo.z, A1[i].lo.z, A0[i].lo.w, A1[i].lo.w, qa.z, qa.w);
                mma_bf16_16816(acc0, A0[i].hi.x, A1[i].hi.x, A0[i].hi.y, A1[i].hi.y, qb.x, qb.y);
                mma_bf16_16816(acc1, A0[i].hi.z, A1[i].hi.z, A0[i].hi.w, A1[i].hi.w, qb.z, qb.w);
            }
        }
        if (c == 0) {
            if (j0 < nf) accept_key(o, finish_dist<true>(p, acc0[0] + acc1[0], sc0, qnorm, qq), n0, j0);
            if (j1 < nf) accept_key(o, finish_dist<true>(p, acc0[2] + acc1[2], sc1, qnorm, qq), n1, j1);
        }
    }
}
// sqh: the query rounded to bf16, zero padded, in shared memory ([tstride/2] elements). All threads of the CTA call.
template <bool FULL>
__device__ __forceinline__ void fresh_distances_mma(const HnswSearchParams& p, const uint4* __restrict__ sqh, float qnorm, float qq,
                                                    const int* __restrict__ fnode, int nf, const FreshOut& o) {
    if constexpr (FULL) { fresh_distances_mma_full(p, sqh, qnorm, qq, fnode, nf, o); return; }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, c = lane & 3;
    const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
    const int nblk = p.tstride >> 4;                  // 16-byte column blocks per row
    const int nst = (nblk + 15) >> 4;                 // stages of 16 blocks (4 per lane)
    for (int tile = warp; tile * 16 < nf; tile += HN_WARPS) {
        const int j0 = tile * 16 + g, j1 = j0 + 8;
        const int n0 = fnode[min(j0, nf - 1)], n1 = fnode[min(j1, nf - 1)];
        const uint4* r0 = reinterpret_cast<const uint4*>(tb + (size_t)n0 * p.tstride);
        const uint4* r1 = reinterpret_cast<const uint4*>(tb + (size_t)n1 * p.tstride);
        const float sc0 = (c == 0 && j0 < nf) ? row_scalar<true>(p, n0) : 0.0f;
        const float sc1 = (c == 0 && j1 < nf) ? row_scalar<true>(p, n1) : 0.0f;
        float acc0[4] = {0.0f, 0.0f, 0.0f, 0.0f}, acc1[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        uint4 A0[4], A1[4], N0[4], N1[4];
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int i = 0; i < 4; i++) { const int cb = c + 4 * i; A0[i] = (FULL || cb < nblk) ? __ldg(r0 + cb) : z; A1[i] = (FULL || cb < nblk) ? __ldg(r1 + cb) : z; }
        for (int st = 0; st < nst; st++) {
            if (st + 1 < nst) {
#pragma unroll
                for (int i = 0; i < 4; i++) { const int cb = c + 4 * (4 * (st + 1) + i); N0[i] = (FULL || cb < nblk) ? __ldg(r0 + cb) : z; N1[i] = (FULL || cb < nblk) ? __ldg(r1 + cb) : z; }
            }
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int cb = c + 4 * (4 * st + i);
                const uint4 qv = (FULL || cb < nblk) ? sqh[cb] : z;
                mma_bf16_16816(acc0, A0[i].x, A1[i].x, A0[i].y, A1[i].y, qv.x, qv.y);
                mma_bf16_16816(acc1, A0[i].z, A1[i].z, A0[i].w, A1[i].w, qv.z, qv.w);
            }
            if (st + 1 < nst) {
#pragma unroll
                for (int i = 0; i < 4; i++) { A0[i] = N0[i]; A1[i] = N1[i]; }
            }
        }
        if (c == 0) {
            if (j0 < nf) accept_key(o, finish_dist<true>(p, acc0[0] + acc1[0], sc0, qnorm, qq), n0, j0);
            if (j1 < nf) accept_key(o, finish_dist<true>(p, acc0[2] + acc1[2], sc1, qnorm, qq), n1, j1);
        }
    }
}

// Bitonic sort of 32 * PER keys held PER per lane (element PER * lane + r) entirely in registers: strides below PER are
// in-register compare-exchanges, larger strides are shuffles with lane ^ (stride / PER). No barrier, no shared memory.
__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int m) {
    const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, m), hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}
template <int PER>
__device__ __forceinline__ void warp_sort_regs(uint64_t (&e)[PER], int lane) {
#pragma unroll
    for (int k = 2; k <= 32 * PER; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < PER) {
#pragma unroll
                for (int r = 0; r < PER; r++) {
                    if ((r & j) == 0) {
                        const bool up = k < PER ? ((r & k) == 0) : ((lane & (k / PER)) == 0);
                        const uint64_t x = e[r], y = e[r | j];
                        if ((x > y) == up) { e[r] = y; e[r | j] = x; }
                    }
                }
            } else {
                const int m = j / PER;
                const bool keepmin = ((lane & m) == 0) == ((lane & (k / PER)) == 0);
#pragma unroll
                for (int r = 0; r < PER; r++) {
                    const uint64_t q = shfl_xor_u64(e[r], m);
                    e[r] = keepmin ? (e[r] < q ? e[r] : q) : (e[r] > q ? e[r] : q);
                }
            }
        }
    }
}
// Sort keys[0..n) (n <= 8 * 32 * PER) into out[0..n): warp w sorts chunk w of 32 * PER keys in registers, then every key finds
// its slot in the union by adding its lower bounds in the other chunks. Two barriers. out must not alias keys; keys is
// clobbered (padded with KEY_MAX up to the end of the last chunk).
template <int PER>
__device__ __forceinline__ void chunk_sort_merge(uint64_t* __restrict__ keys, int n, uint64_t* __restrict__ out) {
    constexpr int CH = 32 * PER;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = warp * CH;
    if (c0 < n) {
        uint64_t e[PER];
#pragma unroll
        for (int r = 0; r < PER; r++) e[r] = c0 + PER * lane + r < n ? keys[c0 + PER * lane + r] : KEY_MAX;
        warp_sort_regs<PER>(e, lane);
#pragma unroll
        for (int r = 0; r < PER; r++) keys[c0 + PER * lane + r] = e[r];
    }
    __syncthreads();
    const int nch = (n + CH - 1) / CH;
    for (int idx = tid; idx < n; idx += HN_NT) {
        const uint64_t key = keys[idx];
        const int mych = idx / CH;
        int r = idx - mych * CH;
        // the <= 7 searches run in lock-step over whole chunks (the ragged last chunk is padded with KEY_MAX, which no key is below)
        int base[HN_WARPS];
#pragma unroll
        for (int ch = 0; ch < HN_WARPS; ch++) base[ch] = ch * CH;
        int len = CH;
        while (len > 1) {
            const int half = len >> 1;
#pragma unroll
            for (int ch = 0; ch < HN_WARPS; ch++)
                if (ch < nch) base[ch] += (keys[base[ch] + half - 1] < key) ? half : 0;
            len -= half;
        }
#pragma unroll
        for (int ch = 0; ch < HN_WARPS; ch++)
            if (ch < nch && ch != mych) r += base[ch] - ch * CH + ((keys[base[ch]] < key) ? 1 : 0);
        out[r] = key;
    }
    __syncthreads();
}

// Small rounds (n <= HN_NT): every key counts the keys below it (broadcast reads, no sorting network, one barrier). Keys are unique.
__device__ __forceinline__ void count_rank_sort(const uint64_t* __restrict__ keys, int n, uint64_t* __restrict__ out) {
    const int tid = threadIdx.x;
    if (tid < n) {
        const uint64_t key = keys[tid];
        int r = 0;
#pragma unroll 8
        for (int j = 0; j < n; j++) r += (keys[j] < key) ? 1 : 0;
        out[r] = key;
    }
    __syncthreads();
}

#define HN_FCAP 2048      // side buffer capacity (entries); must stay >= HN_FMAX (a round's accepted keys always fit)
#define HN_FTH 1024       // side buffer is folded into the main array when it holds more than this

struct HnswShared {
    int candP[HN_MAXB], candF[HN_MAXB];   // indices of the first unexpanded entries of the main array / side buffer
    uint64_t ckP[HN_MAXB], ckF[HN_MAXB];  // their keys (expanded bit clear): the rank-merge of the two lists reads these
    int sel[HN_MAXB];
    int wsum[HN_WARPS], wsum2[HN_WARPS];
    int nfresh, nacc, takenP, nextq;
    int countP, countF;   // entries in the main array (<= ef) / side buffer
    int frontP, frontF;   // every entry below these indices carries the expanded bit
    int whichF;           // which of the two side-buffer halves is current
    int evicted;          // Java: first results.poll() has happened (furthest is live)
    float dentry, qnorm, qq, qinv;
    long long ph[8];      // profile: per-phase cycles of thread 0
    long long tp;
    long long sub[4];     // profile: cycles of the C0 folds, number of C0 folds, scan cycles, cycles of the folds inside sort / merge
    long long hist[4];    // profile: rounds with > 256 accepted keys, sum of accepted keys, rounds with 129..256, sort cycles of the first
};

// Collect the indices of the first (up to) B unexpanded entries of the sorted array a[from..lim) into cand[] (sorted
// order). Returns the number found (may exceed B: only the first B are recorded); *resume = an index below which every
// entry that was not recorded is expanded (the new front if all recorded ones get taken is computed by the caller).
// Contains barriers: all threads call with uniform arguments.
__device__ __forceinline__ int scan_unexpanded(const uint64_t* __restrict__ a, int from, int lim, int B, int* cand, int* wsum,
                                               int* scanned_to) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int found = 0, pos = from;
    while (found < B && pos < lim) {
        const int i = pos + tid;
        const bool un = i < lim && !(a[i] & EXP_BIT);
        const unsigned bal = __ballot_sync(0xffffffffu, un);
        if (lane == 0) wsum[warp] = __popc(bal);
        __syncthreads();
        int before = found, total = 0;
#pragma unroll
        for (int w = 0; w < HN_WARPS; w++) { const int c = wsum[w]; if (w < warp) before += c; total += c; }
        before += __popc(bal & ((1u << lane) - 1u));
        if (un && before < B) cand[before] = i;
        found += total; pos += HN_NT;
        __syncthreads();
    }
    *scanned_to = min(pos, lim);
    return found;
}

// The same scan over TWO sorted arrays at once (main array and side buffer): one pair of barriers per 256-entry step for both,
// and the candidates' keys are recorded next to their indices. Results as scan_unexpanded's, per array.
__device__ __forceinline__ void scan_unexpanded2(const uint64_t* __restrict__ a, int fromA, int limA, const uint64_t* __restrict__ b, int fromB,
                                                 int limB, int B, HnswShared& S, int* foundA_, int* toA, int* foundB_, int* toB) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int foundA = 0, posA = fromA, foundB = 0, posB = fromB;
    while ((foundA < B && posA < limA) || (foundB < B && posB < limB)) {
        const bool doA = foundA < B && posA < limA, doB = foundB < B && posB < limB;
        const int iA = posA + tid, iB = posB + tid;
        uint64_t kA = 0, kB = 0;
        bool unA = false, unB = false;
        if (doA && iA < limA) { kA = a[iA]; unA = !(kA & EXP_BIT); }
        if (doB && iB < limB) { kB = b[iB]; unB = !(kB & EXP_BIT); }
        const unsigned balA = __ballot_sync(0xffffffffu, unA), balB = __ballot_sync(0xffffffffu, unB);
        if (lane == 0) { S.wsum[warp] = __popc(balA); S.wsum2[warp] = __popc(balB); }
        __syncthreads();
        int beforeA = foundA, totalA = 0, beforeB = foundB, totalB = 0;
#pragma unroll
        for (int w = 0; w < HN_WARPS; w++) {
            const int cA = S.wsum[w], cB = S.wsum2[w];
            if (w < warp) { beforeA += cA; beforeB += cB; }
            totalA += cA; totalB += cB;
        }
        beforeA += __popc(balA & ((1u << lane) - 1u));
        beforeB += __popc(balB & ((1u << lane) - 1u));
        if (unA && beforeA < B) { S.candP[beforeA] = iA; S.ckP[beforeA] = kA; }
        if (unB && beforeB < B) { S.candF[beforeB] = iB; S.ckF[beforeB] = kB; }
        if (doA) { foundA += totalA; posA += HN_NT; }
        if (doB) { foundB += totalB; posB += HN_NT; }
        __syncthreads();
    }
    *foundA_ = foundA; *toA = min(posA, limA); *foundB_ = foundB; *toB = min(posB, limB);
}

// rr[j] = number of entries of the sorted array a[0..na) below s[j] (expanded bits ignored), four searches per thread in lock-step
__device__ __forceinline__ void rank_keys(const uint64_t* __restrict__ a, int na, const uint64_t* __restrict__ s, int ns, int* __restrict__ rr) {
    for (int j0 = threadIdx.x; j0 < ns; j0 += 4 * HN_NT) {
        uint64_t key[4]; int r[4];
#pragma unroll
        for (int t = 0; t < 4; t++) { const int j = j0 + t * HN_NT; key[t] = j < ns ? (s[j] & ~EXP_BIT) : 0ull; }
        lower_bound_u64_multi<4>(a, na, key, r);
#pragma unroll
        for (int t = 0; t < 4; t++) { const int j = j0 + t * HN_NT; if (j < ns) rr[j] = r[t]; }
    }
}

// Merge the sorted keys s[0..ns) into the sorted array a[0..na) writing out[0..min(na+ns,cap)) (out != a): every
// element finds its final slot by ranking (rr[j] = #a-entries below s[j]). Barriers inside; rr is scratch [ns].
__device__ __forceinline__ void merge_ranked(const uint64_t* __restrict__ a, int na, const uint64_t* __restrict__ s, int ns,
                                             uint64_t* __restrict__ out, int cap, int* __restrict__ rr) {
    const int tid = threadIdx.x;
    rank_keys(a, na, s, ns, rr);
    __syncthreads();
    const int r0 = rr[0];
    for (int i0 = tid; i0 < na; i0 += 4 * HN_NT) {
        int x[4], c[4];
#pragma unroll
        for (int t = 0; t < 4; t++) x[t] = i0 + t * HN_NT;
        if (i0 + 3 * HN_NT < r0) {                                   // below the first insertion point nothing moves
#pragma unroll
            for (int t = 0; t < 4; t++) out[x[t]] = a[x[t]];
            continue;
        }
        count_le_i32_multi<4>(rr, ns, x, c);
#pragma unroll
        for (int t = 0; t < 4; t++) {
            if (x[t] < na) { const int np = x[t] + c[t]; if (np < cap) out[np] = a[x[t]]; }
        }
    }
    for (int j = tid; j < ns; j += HN_NT) { const int np = rr[j] + j; if (np < cap) out[np] = s[j]; }
    __syncthreads();
}

// Same merge IN PLACE (out == a, capacity cap): entries only move up, so the array is walked top chunk first; a chunk
// overwrites only slots of chunks already moved or of itself (read before the barrier). s must not alias a.
__device__ __forceinline__ void merge_ranked_inplace(uint64_t* __restrict__ a, int na, const uint64_t* __restrict__ s, int ns, int cap,
                                                     int* __restrict__ rr) {
    const int tid = threadIdx.x;
    rank_keys(a, na, s, ns, rr);
    __syncthreads();
    const int r0 = rr[0];
    constexpr int PERC = 4;
    for (int hi = na; hi > r0; hi -= HN_NT * PERC) {
        const int lo = max(r0, hi - HN_NT * PERC);
        uint64_t e[PERC]; int np[PERC], x[PERC], c[PERC];
#pragma unroll
        for (int t = 0; t < PERC; t++) x[t] = lo + tid + t * HN_NT;
        count_le_i32_multi<PERC>(rr, ns, x, c);
#pragma unroll
        for (int t = 0; t < PERC; t++) {
            np[t] = -1;
            if (x[t] < hi) { e[t] = a[x[t]]; np[t] = x[t] + c[t]; }
        }
        __syncthreads();
#pragma unroll
        for (int t = 0; t < PERC; t++) if (np[t] >= 0 && np[t] < cap) a[np[t]] = e[t];
        __syncthreads();
    }
    for (int j = tid; j < ns; j += HN_NT) { const int np = rr[j] + j; if (np < cap) a[np] = s[j]; }
    __syncthreads();
}

// Fold the side buffer into the main array (in place) and remap the bookkeeping. Barriers inside; uniform call.
__device__ __forceinline__ void fold_side_buffer(uint64_t* P, const uint64_t* F0, HnswShared& S, int ef, int* rr) {
    const int cP = S.countP, cF = S.countF, fP = S.frontP, fF = S.frontF;
    const uint64_t* Fc = F0 + (size_t)S.whichF * HN_FCAP;
    __syncthreads();
    merge_ranked_inplace(P, cP, Fc, cF, ef, rr);
    if (threadIdx.x == 0) {
        // old P[fP] lands after every buffer entry ranked at or below it (rr <= fP)
        int nf_ = fP + count_le_i32(rr, cF, fP);
        if (fF < cF) nf_ = min(nf_, rr[fF] + fF);                          // first possibly-unexpanded buffer entry
        S.frontP = min(nf_, ef);
        if (cP + cF > ef) S.evicted = 1;
        S.countP = min(cP + cF, ef); S.countF = 0; S.frontF = 0;
    }
    __syncthreads();
}

// FULL (bf16 rows only): rows are a whole number of 256-byte MMA stages, the distance loop needs no bound checks
// MINB = CTAs per SM the kernel is compiled for: 2 (128 registers per thread) when the search state of a query needs more than a third
// of the shared memory (ef = 5000: 110 KB), 3 (80 registers, some spills) when three queries fit (ef <= ~1700: C1, C2, the graph
// builder's insert searches) — small-ef walks are bookkeeping-bound and gain more from the third resident CTA than they lose to the
// spills (C1: 2.39 -> 2.17 ms per 1000 queries), large-ef walks would only lose (C3 with the 80-register code: 19.6 -> 21.1 ms).
template <bool BF16, bool FULL, int FP8 = 0, bool I8 = false, int MINB = 2>   // FP8: 0 = off, else 32-byte loads per row and stage (1 or 2) of the 8-bit rows; I8: int8 instead of e4m3
__global__ void __launch_bounds__(HN_NT, MINB) hnsw_search_kernel(const HnswSearchParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int qpad = (p.dim + 7) & ~7;
    float* sq = reinterpret_cast<float*>(smem_raw);                       // [qpad] query, fp32 (zero padded)
    uint4* sqh = reinterpret_cast<uint4*>(sq + qpad);                     // [qpad] query, bf16 (tensor-pipe B operand)
    uint64_t* P = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(sqh) + qpad * 2);   // [ef] main sorted array
    uint64_t* F0 = P + p.ef;                                              // [2][HN_FCAP] side buffer, ping-pong
    uint64_t* fkey = F0 + 2 * HN_FCAP;                                    // [HN_FMAX] accepted keys of the round; rank scratch
    int* fnode = reinterpret_cast<int*>(fkey + HN_FMAX);                  // [HN_FMAX] fresh nodes; rank scratch
    const int vwords = (p.n + 31) >> 5;
    uint32_t* vis = p.visited_in_smem ? reinterpret_cast<uint32_t*>(fnode + HN_FMAX)
                                      : p.visited_scratch + (size_t)blockIdx.x * vwords;
    __shared__ HnswShared S;
    __shared__ GreedyShared GS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ef = p.ef, variant = p.variant, B = p.batch;
    const int32_t* adj0 = p.adj + (size_t)p.beam_level * p.adj_stride * p.deg;      // the level the beam walks
    const int uncertain_ok = B == 1 ? 1 : 128;    // side-buffer entries tolerated while candidates come from the array's tail
    unsigned long long evals = 0, exps = 0;
    if (tid < 8) S.ph[tid] = 0;
    if (tid < 4) { S.hist[tid] = 0; S.sub[tid] = 0; }
#define HN_MARK(slot) do { if (p.profile && tid == 0) { const long long now_ = clock64(); S.ph[slot] += now_ - S.tp; S.tp = now_; } } while (0)

    for (int q = blockIdx.x; q < p.nq;) {
        // ---- A. stage the query (HnswIndex normalises it: HnswIndex.java:344, core/Vector.java:107-116) ----
        __syncthreads();
        if (p.profile && tid == 0) S.tp = clock64();
        const float* qsrc = p.qrows ? p.base + (size_t)__ldg(p.qrows + q) * p.dim : p.queries + (size_t)q * p.dim;
        for (int i = tid; i < qpad; i += HN_NT) sq[i] = i < p.dim ? qsrc[i] : 0.0f;
        for (int i = tid; i < vwords; i += HN_NT) vis[i] = p.dead ? __ldg(p.dead + i) : 0u;
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
        if (FP8 && I8) {
            signed char* q8 = reinterpret_cast<signed char*>(sqh);          // [dim] int8 query
            if (warp == 1) {
                float mx = 0.0f;
                for (int i = lane; i < p.dim; i += 32) mx = fmaxf(mx, fabsf(sq[i]));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                if (lane == 0) S.qinv = mx > 0.0f ? __fdiv_rn(mx, 127.0f) : 1.0f;
            }
            __syncthreads();
            const float qinv_ = S.qinv, qs = __fdiv_rn(1.0f, qinv_);
            for (int i = tid; i < p.dim; i += HN_NT) q8[i] = (signed char)max(-127, min(127, __float2int_rn(__fmul_rn(sq[i], qs))));
            __syncthreads();
            if (warp == 1) {                                           // ||q||^2 of the REPRESENTED query
                float a = 0.0f;
                for (int i = lane; i < p.dim; i += 32) { const float v = __fmul_rn((float)q8[i], qinv_); a = __fmaf_rn(v, v, a); }
                a = warp_tree_sum(a);
                if (lane == 0) S.qq = a;
            }
        } else if (FP8) {
            unsigned char* q8 = reinterpret_cast<unsigned char*>(sqh);      // [dim] hi term, [dim] lo term (x 16)
            if (warp == 1) {
                float mx = 0.0f;
                for (int i = lane; i < p.dim; i += 32) mx = fmaxf(mx, fabsf(sq[i]));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                if (lane == 0) S.qinv = __fdiv_rn(1.0f, fp8_pow2_scale(mx));
            }
            __syncthreads();
            const float qinv_ = S.qinv, qs = __fdiv_rn(1.0f, qinv_);
            for (int i = tid; i < p.dim; i += HN_NT) {
                const float v = __fmul_rn(sq[i], qs);
                const __nv_fp8_storage_t h = __nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
                const __nv_fp8_storage_t l = HN_FP8_LO ? __nv_cvt_float_to_fp8(__fmul_rn(__fsub_rn(v, fp8_to_float(h)), 16.0f), __NV_SATFINITE, __NV_E4M3)
                                                       : (__nv_fp8_storage_t)0;
                q8[i] = h; q8[p.dim + i] = l;
            }
            __syncthreads();
            if (warp == 1) {                                           // ||q||^2 of the REPRESENTED query (hi + lo / 16) / scale
                float a = 0.0f;
                for (int i = lane; i < p.dim; i += 32) {
                    const float v = __fmul_rn(__fmaf_rn(fp8_to_float(q8[p.dim + i]), 0.0625f, fp8_to_float(q8[i])), qinv_);
                    a = __fmaf_rn(v, v, a);
                }
                a = warp_tree_sum(a);
                if (lane == 0) S.qq = a;
            }
        } else if (BF16) {
            __nv_bfloat16* h = reinterpret_cast<__nv_bfloat16*>(sqh);
            for (int i = tid; i < qpad; i += HN_NT) h[i] = __float2bfloat16_rn(sq[i]);
            __syncthreads();
            if (warp == 1) {                                           // ||q||^2 of the ROUNDED query (bf16 sq-L2 form)
                float a = 0.0f;
                for (int i = lane; i < qpad; i += 32) { const float v = __bfloat162float(h[i]); a = __fmaf_rn(v, v, a); }
                a = warp_tree_sum(a);
                if (lane == 0) S.qq = a;
            }
        }
        __syncthreads();
        const float qnorm = S.qnorm, qq = S.qq;

        // ---- B. greedy descent on levels max_level-1 .. 1 (warp 0), then the entry point's layer-0 distance ----
        const int qe = p.qentry ? __ldg(p.qentry + q) : -1;
        if (qe >= 0) {
            if (tid == 0) { fnode[0] = qe; S.nacc = 0; vis[qe >> 5] |= 1u << (qe & 31); }
        } else if ((p.metric == 0) && p.trained) {
            int cur = p.entry;
            for (int l = p.max_level - 1; l > p.beam_level; l--) cur = greedy_level_pq(p, sq, cur, l, GS, vis);
            if (tid == 0) { fnode[0] = cur; S.nacc = 0; vis[cur >> 5] |= 1u << (cur & 31); }
        } else if (warp == 0) {
            int cur = p.entry;
            for (int l = p.max_level - 1; l > p.beam_level; l--) cur = greedy_level(p, sq, qnorm, cur, l, false, evals, vis);
            if (lane == 0) { fnode[0] = cur; S.nacc = 0; vis[cur >> 5] |= 1u << (cur & 31); }
        }
        __syncthreads();
        {
            FreshOut o{fkey, &S.nacc, false, 0.0f, false};
            if constexpr (FP8 != 0 && I8) fresh_distances_i8<FP8 ? FP8 : 1>(p, reinterpret_cast<const unsigned char*>(sqh), qnorm, qq, S.qinv, fnode, 1, o);
            else if constexpr (FP8 != 0) fresh_distances_fp8<FP8 ? FP8 : 1>(p, reinterpret_cast<const unsigned char*>(sqh), qnorm, qq, S.qinv, fnode, 1, o);
            else if (BF16) fresh_distances_mma<FULL>(p, sqh, qnorm, qq, fnode, 1, o);
            else fresh_dispatch<false>(p, sq, qnorm, qq, fnode, 1, o);
        }
        __syncthreads();
        if (tid == 0) {
            P[0] = fkey[0];
            S.dentry = pool_dist(fkey[0]);
            S.countP = 1; S.countF = 0; S.frontP = 0; S.frontF = 0; S.whichF = 0; S.evicted = 0; S.nfresh = 0; S.nacc = 0; S.takenP = 0;
            evals++;
        }
        __syncthreads();

        // ---- C. layer-0 search in rounds ----
        HN_MARK(0);
        while (true) {
            // C0. fold the side buffer into the main array when (a) the array is not full yet (early phase: exact
            //     Java bookkeeping), (b) the buffer is large, (c) candidates are about to come from the array's tail while
            //     the buffer holds entries that may push that tail out of the best ef
            {
                const int cP = S.countP, cF = S.countF;
                if (cF > 0 && (cP < ef || cF > HN_FTH || (S.frontP + B > ef - cF && cF >= uncertain_ok))) {
                    const long long t0_ = (p.profile && tid == 0) ? clock64() : 0;
                    fold_side_buffer(P, F0, S, ef, reinterpret_cast<int*>(fkey));     // fkey is idle here
                    if (p.profile && tid == 0) { S.sub[0] += clock64() - t0_; S.sub[1]++; }
                }
            }
            const long long ts_ = (p.profile && tid == 0) ? clock64() : 0;
            const uint64_t* Fc = F0 + (size_t)S.whichF * HN_FCAP;
            uint64_t* Fc_w = F0 + (size_t)S.whichF * HN_FCAP;
            uint64_t* Fn = F0 + (size_t)(S.whichF ^ 1) * HN_FCAP;
            const int countP = S.countP;
            int countF = S.countF;
            const bool full = countP >= ef;
            const int evicted = S.evicted | ((countP + countF > ef) ? 1 : 0);
            int lim = countP;
            if ((variant == 2 && !evicted) || (variant == 1 && full && !evicted))     // Java: `c.d > furthest` stops; furthest is
                lim = lower_bound_u64(P, countP, ((uint64_t)f2ord(S.dentry) << 32) | 0x7fffffffull);   // the entry distance here
            // C1. candidates.poll() x B: the B best unexpanded entries of array + buffer
            int toP, toF, fndP, fndF;
            scan_unexpanded2(P, S.frontP, lim, Fc, S.frontF, countF, B, S, &fndP, &toP, &fndF, &toF);
            if (p.profile && tid == 0) S.sub[2] += clock64() - ts_;
            const int a = min(fndP, B), b = min(fndF, B);
            if (a + b == 0) break;                                    // no eligible candidate left: the search ends
            const int nsel = min(a + b, B);
            if (tid < a + b) {                                        // rank in the union of the two sorted candidate lists
                const bool fromP = tid < a;
                const int me = fromP ? tid : tid - a;
                const int idx = fromP ? S.candP[me] : S.candF[me];
                const uint64_t key = fromP ? S.ckP[me] : S.ckF[me];   // unexpanded: the expanded bit is clear
                const int r = me + (fromP ? lower_bound_u64(S.ckF, b, key) : lower_bound_u64(S.ckP, a, key));
                if (r < B) {
                    S.sel[r] = (int)(key & 0x7fffffffull);
                    if (fromP) { P[idx] |= EXP_BIT; atomicAdd(&S.takenP, 1); } else Fc_w[idx] |= EXP_BIT;
                }
            }
            __syncthreads();
            // new fronts: the first candidate that was NOT taken, else where the scan stopped
            int newfrontP, newfrontF;
            {
                const int takenP = S.takenP;                        // taken ones form a prefix of each candidate list
                const int takenF = nsel - takenP;
                newfrontP = takenP < a ? S.candP[takenP] : (fndP > B ? S.candP[B - 1] + 1 : toP);
                newfrontF = takenF < b ? S.candF[takenF] : (fndF > B ? S.candF[B - 1] + 1 : toF);
            }
            HN_MARK(1);
            // C2. neighbour lists (loads of eight rows in flight per warp) + visited test-and-set; fresh nodes are appended
            //     and their traversal rows pulled towards L2 right away (the misses overlap in one memory round trip and
            //     are under way while the rest of the lists are still being examined)
            {
                const int nj = (p.deg + 31) >> 5;
                const int nitems = nsel * nj;
                const int lines = (p.tstride + 127) >> 7;
                const unsigned char* tb = reinterpret_cast<const unsigned char*>(p.tbase);
                for (int it0 = warp; it0 < nitems; it0 += HN_WARPS * 8) {
                    int nb[8];
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        const int it = it0 + t * HN_WARPS;
                        nb[t] = -1;
                        if (it < nitems) {
                            const int s = nj == 1 ? it : it / nj, j = (it - s * nj) * 32 + lane;
                            if (j < p.deg) nb[t] = __ldg(adj0 + (size_t)S.sel[s] * p.deg + j);
                        }
                    }
                    // visited test-and-set of all eight lists first (independent shared-memory atomics, pipelined), then ONE
                    // append reservation for the warp instead of one per list
                    unsigned m[8];
                    int tot = 0;
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        bool fresh = false;
                        if (nb[t] >= 0) {
                            const uint32_t bit = 1u << (nb[t] & 31);
                            const uint32_t old = atomicOr(&vis[nb[t] >> 5], bit);
                            fresh = !(old & bit);
                        }
                        m[t] = __ballot_sync(0xffffffffu, fresh);
                        tot += __popc(m[t]);
                    }
                    if (tot == 0) continue;
                    int basec = 0;
                    if (lane == 0) basec = atomicAdd(&S.nfresh, tot);
                    basec = __shfl_sync(0xffffffffu, basec, 0);
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        if ((m[t] >> lane) & 1u) {
                            fnode[basec + __popc(m[t] & ((1u << lane) - 1u))] = nb[t];
                            if (p.prefetch) {
                                const unsigned char* rp = tb + (size_t)nb[t] * p.tstride;
                                for (int l = 0; l < lines; l++) asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + ((size_t)l << 7)));
                            }
                        }
                        basec += __popc(m[t]);
                    }
                }
            }
            __syncthreads();
            const int nf = S.nfresh;
            if (tid == 0) { exps += nsel; evals += nf; S.ph[7]++; }
            // Java's searchLayer (HnswPqIndex.java:876-884, HnswIndex.java:611-625) refreshes `furthest` only when an insert evicts:
            // until the first eviction it is the ENTRY distance, so what is accepted in the round where the result set fills up (and in
            // any later round before the first eviction) depends on the order of the neighbours. Such rounds are replayed in that
            // order by one thread below; they are few (at most until the first eviction) and small.
            const bool replay = variant != 0 && !evicted && countF == 0 && countP + nf > ef && nf <= 512;
            HN_MARK(2);
            // C3. distances of the fresh rows; acceptance: results.size() < ef || d < furthest
            //     (furthest live: HnswPqIndex.java:953-963; stale == entry distance until the first eviction: :876-884,
            //      HnswIndex.java:611-625). With a non-empty side buffer P[ef-1] is an upper bound of the live furthest:
            //      a few more keys are accepted and fall off at the next fold — the kept set is the same.
            {
                float bound = 0.0f;
                if (full) bound = (variant == 0 || evicted) ? pool_dist(P[ef - 1]) : S.dentry;
                FreshOut o{fkey, &S.nacc, full, bound, replay};
                if constexpr (FP8 != 0 && I8) fresh_distances_i8<FP8 ? FP8 : 1>(p, reinterpret_cast<const unsigned char*>(sqh), qnorm, qq, S.qinv, fnode, nf, o);
                else if constexpr (FP8 != 0) fresh_distances_fp8<FP8 ? FP8 : 1>(p, reinterpret_cast<const unsigned char*>(sqh), qnorm, qq, S.qinv, fnode, nf, o);
                else if (BF16) fresh_distances_mma<FULL>(p, sqh, qnorm, qq, fnode, nf, o);
                else fresh_dispatch<false>(p, sq, qnorm, qq, fnode, nf, o);
            }
            __syncthreads();
            if (replay) {
                if (tid == 0) {
                    int size = countP, pe = countP - 1, out = 0, ev = 0;
                    float furthest = S.dentry;
                    for (int j = 0; j < nf; j++) {
                        const uint64_t key = fkey[j];
                        if (size < ef || pool_dist(key) < furthest) {
                            fkey[out++] = key; size++;
                            if (size > ef) {                               // resultSet.poll(): the current maximum leaves
                                int mi = -1; uint64_t mk = 0;
                                for (int t = 0; t < out; t++) if ((fkey[t] & ~EXP_BIT) >= mk) { mk = fkey[t] & ~EXP_BIT; mi = t; }
                                if (pe >= 0 && (P[pe] & ~EXP_BIT) > mk) pe--; else { fkey[mi] = fkey[--out]; }
                                size--; ev = 1;
                                float f = pe >= 0 ? pool_dist(P[pe]) : 0.0f;     // furthest = resultSet.peek()
                                for (int t = 0; t < out; t++) f = fmaxf(f, pool_dist(fkey[t]));
                                furthest = f;
                            }
                        }
                    }
                    S.nacc = out;
                    if (ev) S.evicted = 1;
                }
                __syncthreads();
            }
            const int nacc = S.nacc;
            HN_MARK(3);
            int r0 = newfrontF;
            if (nacc > 0) {
                const uint64_t* skeys = fkey;      // the round's accepted keys, sorted
                int* rr = fnode;                   // rank scratch of the merges
                // C4. sort the accepted keys: every warp sorts one chunk of the append list in registers (bitonic network over
                //     shuffles, no barrier), then each key finds its slot in the union by adding its lower bounds in the other
                //     chunks (<= 7 binary searches); two barriers in all, against 45..66 barrier steps of a block-wide bitonic
                //     sort (which took 80 % of this phase in the 16 % of the rounds with more than 256 keys).
                bool direct = false;
                if (nacc <= 8 * 128) {
                    uint64_t* sk = reinterpret_cast<uint64_t*>(fnode);          // fnode (idle after the distance phase) holds 1024 keys
                    if (nacc <= HN_SMALL_SORT) count_rank_sort(fkey, nacc, sk);
                    else if (nacc <= 8 * 64) chunk_sort_merge<2>(fkey, nacc, sk); else chunk_sort_merge<4>(fkey, nacc, sk);
                    skeys = sk; rr = reinterpret_cast<int*>(fkey);              // the append list is dead from here on
                } else {
                    // a large round (the early rounds accept everything): empty the side buffer first, then sort straight into
                    // its idle half, which becomes the new side buffer without a merge (nacc <= HN_FMAX = HN_FCAP)
                    if (countF > 0) {
                        if (tid == 0) { S.frontP = newfrontP; S.frontF = newfrontF; }
                        fold_side_buffer(P, F0, S, ef, fnode);
                        newfrontP = S.frontP; newfrontF = 0; countF = 0;
                    }
                    chunk_sort_merge<8>(fkey, nacc, Fn);
                    direct = true;
                }
                if (p.profile && tid == 0) {
                    S.hist[1] += nacc;
                    if (nacc > HN_NT) { S.hist[0]++; S.hist[3] += clock64() - S.tp; } else if (nacc > 128) S.hist[2]++;
                }
                HN_MARK(4);
                // C5. merge into the side buffer (other half), by ranking; a buffer that would overflow is folded first
                if (direct) r0 = 0;
                else {
                    if (countF + nacc > HN_FCAP) {
                        if (tid == 0) { S.frontP = newfrontP; S.frontF = newfrontF; }
                        fold_side_buffer(P, F0, S, ef, rr);
                        newfrontP = S.frontP; newfrontF = 0; r0 = 0; countF = 0;
                    }
                    merge_ranked(Fc, countF, skeys, nacc, Fn, HN_FCAP, rr);
                    r0 = rr[0];
                }
            }
            if (tid == 0) {
                if (nacc > 0) { S.countF = countF + nacc; S.whichF ^= 1; }
                S.frontP = newfrontP;
                S.frontF = min(newfrontF, r0);
                S.nfresh = 0; S.nacc = 0; S.takenP = 0;
            }
            __syncthreads();
            HN_MARK(5);
        }

        // ---- D. final: fold the side buffer, then best entries by traversal distance -> exact re-score in the
        //         reference's sequential order (HnswPqIndex.java:646-664 reranks every candidate; with exact traversal
        //         distances the values are identical, with bf16 rows the shortlist margin covers the rounding)
        //         -> (dist,id) sort -> top-k ----
        __syncthreads();
        if (S.countF > 0) {
            const int cP = S.countP, cF = S.countF;
            merge_ranked_inplace(P, cP, F0 + (size_t)S.whichF * HN_FCAP, cF, ef, reinterpret_cast<int*>(fkey));
            if (tid == 0) { S.countP = min(cP + cF, ef); S.countF = 0; }
            __syncthreads();
        }
        {
            int rs = 32;
            while (rs < (FP8 ? 4 * p.k + 24 : p.k + 22)) rs <<= 1;      // fp8 rows: a deeper exact re-score absorbs the coarser ranking
            if (rs > HN_FMAX) rs = HN_FMAX;
            const int have = min(S.countP, rs);
            for (int j = tid; j < rs; j += HN_NT) {
                uint64_t key = KEY_MAX;
                if (j < have) {
                    const int node = (int)(P[j] & 0x7fffffffull);
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
        HN_MARK(6);
        // next query: the first one is blockIdx.x, the following ones are handed out dynamically (queries differ in length)
        if (p.qcounter) {
            __syncthreads();
            if (tid == 0) S.nextq = (int)gridDim.x + atomicAdd(p.qcounter, 1);
            __syncthreads();
            q = S.nextq;
        } else q += gridDim.x;
    }
    if (p.counters && tid == 0) {
        if (evals) atomicAdd(&p.counters[0], evals);
        if (exps) atomicAdd(&p.counters[1], exps);
        if (p.profile) {
            for (int i = 0; i < 8; i++) atomicAdd(&p.counters[2 + i], (unsigned long long)S.ph[i]);
            for (int i = 0; i < 4; i++) atomicAdd(&p.counters[12 + i], (unsigned long long)S.hist[i]);
            for (int i = 0; i < 4; i++) atomicAdd(&p.counters[17 + i], (unsigned long long)S.sub[i]);
        }
    }
}

size_t hnsw_search_smem_bytes(const HnswSearchParams& p) {
    size_t b = (size_t)((p.dim + 7) & ~7) * 6 + (size_t)p.ef * 8 + (size_t)HN_FCAP * 16 + (size_t)HN_FMAX * 12;
    if (p.visited_in_smem) b += (size_t)((p.n + 31) / 32) * 4;
    return b;
}
int hnsw_max_batch(int deg) { return std::min(HN_MAXB, HN_FMAX / std::max(deg, 1)); }
template <bool BF16, bool FULL, int FP8, bool I8>
static int hn_launch(const HnswSearchParams& p, int grid, size_t smem, cudaStream_t st) {
    if (p.ctas_per_sm >= 3) {
        if (cudaFuncSetAttribute(hnsw_search_kernel<BF16, FULL, FP8, I8, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 2;
        hnsw_search_kernel<BF16, FULL, FP8, I8, 3><<<grid, HN_NT, smem, st>>>(p);
    } else {
        if (cudaFuncSetAttribute(hnsw_search_kernel<BF16, FULL, FP8, I8, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 2;
        hnsw_search_kernel<BF16, FULL, FP8, I8, 2><<<grid, HN_NT, smem, st>>>(p);
    }
    return 0;
}
int launch_hnsw_search(const HnswSearchParams& p, int grid, cudaStream_t st) {
    if (p.nq <= 0) return 0;
    if (p.k + 22 > HN_FMAX || p.deg > 64 || p.batch < 1 || p.batch * p.deg > HN_FMAX || p.batch > HN_MAXB) return 1;
    if ((p.tstride & 15) != 0) return 1;
    const size_t smem = hnsw_search_smem_bytes(p);
    if (p.tstore == 3) {
        if ((p.tstride & 127) != 0 || 4 * p.k + 24 > HN_FMAX) return 1;
        return (p.tstride & 255) == 0 ? hn_launch<true, false, 2, true>(p, grid, smem, st) : hn_launch<true, false, 1, true>(p, grid, smem, st);
    }
    if (p.tstore == 2) {
        if ((p.tstride & 127) != 0 || 4 * p.k + 24 > HN_FMAX) return 1;
        return (p.tstride & 255) == 0 ? hn_launch<true, false, 2, false>(p, grid, smem, st) : hn_launch<true, false, 1, false>(p, grid, smem, st);
    }
    if (p.tstore == 1) return (p.tstride & 255) == 0 ? hn_launch<true, true, 0, false>(p, grid, smem, st) : hn_launch<true, false, 0, false>(p, grid, smem, st);
    return hn_launch<false, false, 0, false>(p, grid, smem, st);
}

// ---- traversal copy of the base rows: zero-padded to a multiple of 8 elements, fp32 or bf16 (round to nearest even);
//      tnorm[r] = sum of squares of the STORED (rounded) row, sequential fp32 — used by the bf16 sq-L2 form ----
__global__ void make_traversal_rows_kernel(const float* __restrict__ base, int64_t n, int dim, int dpad, int mode, void* __restrict__ out,
                                           float* __restrict__ tnorm, float* __restrict__ tscale) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        float a = 0.0f;
        if (mode == 3) {
            // int8 with a per-row scale (largest element -> 127); row stride = dim bytes
            float mx = 0.0f;
            for (int c = lane; c < dim; c += 32) mx = fmaxf(mx, fabsf(base[r * dim + c]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            const float inv = mx > 0.0f ? __fdiv_rn(mx, 127.0f) : 1.0f, sc = __fdiv_rn(1.0f, inv);
            for (int c = lane; c < dim; c += 32) {
                const int qv = max(-127, min(127, __float2int_rn(__fmul_rn(base[r * dim + c], sc))));
                reinterpret_cast<signed char*>(out)[r * dim + c] = (signed char)qv;
                const float sv = __fmul_rn((float)qv, inv);
                a = __fmaf_rn(sv, sv, a);
            }
            a = warp_tree_sum(a);
            if (lane == 0) { tnorm[r] = a; tscale[r] = inv; }
            continue;
        }
        if (mode == 2) {
            // fp8 e4m3 with a per-row power-of-two scale (largest element lands in [120, 240]); row stride = dim bytes
            float mx = 0.0f;
            for (int c = lane; c < dim; c += 32) mx = fmaxf(mx, fabsf(base[r * dim + c]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            const float sc = fp8_pow2_scale(mx), inv = __fdiv_rn(1.0f, sc);
            for (int c = lane; c < dim; c += 32) {
                const __nv_fp8_storage_t h = __nv_cvt_float_to_fp8(__fmul_rn(base[r * dim + c], sc), __NV_SATFINITE, __NV_E4M3);
                reinterpret_cast<unsigned char*>(out)[r * dim + c] = h;
                const float sv = __fmul_rn(fp8_to_float(h), inv);
                a = __fmaf_rn(sv, sv, a);
            }
            a = warp_tree_sum(a);
            if (lane == 0) { tnorm[r] = a; tscale[r] = inv; }
            continue;
        }
        for (int c = lane; c < dpad; c += 32) {
            const float v = c < dim ? base[r * dim + c] : 0.0f;
            float sv = v;
            if (mode == 1) { const __nv_bfloat16 h = __float2bfloat16_rn(v); reinterpret_cast<__nv_bfloat16*>(out)[r * dpad + c] = h; sv = __bfloat162float(h); }
            else reinterpret_cast<float*>(out)[r * dpad + c] = v;
            a = __fmaf_rn(sv, sv, a);
        }
        a = warp_tree_sum(a);
        if (lane == 0 && tnorm) tnorm[r] = a;
    }
}
__global__ void scatter_adj_rows_kernel(int32_t* __restrict__ adj, const int32_t* __restrict__ payload, const uint64_t* __restrict__ targets,
                                        int cnt, int deg, int64_t stride) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)cnt * deg) return;
    const int r = (int)(i / deg), j = (int)(i - (int64_t)r * deg);
    const uint64_t t = targets[r];
    adj[((size_t)(t >> 32) * stride + (uint32_t)t) * deg + j] = payload[i];
}
void launch_scatter_adj_rows(int32_t* d_adj, const int32_t* d_payload, const uint64_t* d_targets, int cnt, int deg, int64_t adj_stride,
                             cudaStream_t st) {
    if (cnt <= 0) return;
    const int64_t tot = (int64_t)cnt * deg;
    scatter_adj_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(d_adj, d_payload, d_targets, cnt, deg, adj_stride);
}
__global__ void count_outlier_rows_kernel(const float* __restrict__ base, int64_t n, int dim, float ratio2, unsigned long long* __restrict__ count) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    unsigned long long bad = 0;
    for (int64_t r = warp; r < n; r += nwarps) {
        float mx = 0.0f, ss = 0.0f;
        for (int c = lane; c < dim; c += 32) { const float v = base[r * dim + c]; mx = fmaxf(mx, fabsf(v)); ss = __fmaf_rn(v, v, ss); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o)); ss += __shfl_xor_sync(0xffffffffu, ss, o); }
        if (mx * mx * (float)dim > ratio2 * ss) bad++;            // max^2 > ratio^2 * mean square
    }
    if (lane == 0 && bad) atomicAdd(count, bad);
}
void launch_count_outlier_rows(const float* d_base, int64_t n, int dim, float ratio, unsigned long long* d_count, cudaStream_t st) {
    if (n <= 0) return;
    int blocks = (int)std::min<int64_t>((n + 7) / 8, 148 * 16);
    count_outlier_rows_kernel<<<blocks, 256, 0, st>>>(d_base, n, dim, ratio * ratio, d_count);
}
void launch_make_traversal_rows(const float* d_base, int64_t n, int dim, int dpad, int mode, void* d_out, float* d_tnorm, float* d_tscale,
                                cudaStream_t st) {
    if (n <= 0) return;
    int blocks = (int)std::min<int64_t>((n + 7) / 8, 148 * 16);
    make_traversal_rows_kernel<<<blocks, 256, 0, st>>>(d_base, n, dim, dpad, mode, d_out, d_tnorm, d_tscale);
}
