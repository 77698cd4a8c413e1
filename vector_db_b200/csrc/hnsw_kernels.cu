// hnsw_kernels.cu — HNSW query kernel: greedy upper-layer descent (a4, a3) + layer-0 beam search (a5) + exact
// re-score and top-k (a7, a8) in ONE launch, one CTA per query (persistent CTAs loop over the batch).
//
// Semantics reproduced:
//   metric 0  Java HnswPqIndex.searchNearest  index/HnswPqIndex.java:599-669
//   metric 1  Java HnswIndex.searchNearest    index/HnswIndex.java:328-365
//
// Data structure (B200-first, not the reference's two PriorityQueues): the Java result max-heap and candidate min-heap
// are replaced by ONE unsorted pool of 64-bit keys (ordered-distance<<32 | expanded<<31 | node) in shared memory.
//   * "candidates.poll()"          = smallest key without the expanded bit (block-wide min scan);
//   * "furthest / results.size()"  = a lazy threshold tau >= the ef-th smallest key, refreshed by a block-wide bitonic
//                                    sort when the pool fills (accepting a few extra nodes is harmless: see DESIGN.md);
//   * "cand > furthest -> break"   = the popped node is not among the ef smallest keys of the pool; this needs an exact
//                                    rank count only once #expanded >= ef (all smaller keys are already expanded).
// Distances on layer 0 are computed by a whole warp per neighbour row (coalesced 512 B loads + xor-butterfly), the
// upper-layer greedy and the final re-score use the reference's sequential order per lane, so returned distances are
// bit-identical to the Java arithmetic.
#include "kernels.h"

#define HN_NT 128
#define HN_WARPS (HN_NT / 32)
#define EXP_BIT 0x80000000ull
#define RESCORE 32

struct HnswShared {
    int nbr[64];
    int newn[64];
    float newd[64];
    uint64_t red[HN_WARPS];
    int redslot[HN_WARPS];
    int newcnt;
    int pick_slot;
    uint64_t pick_key;
    int flag;
    int rank;
    int jstop;
    int need_compact;
    // search state (written by thread 0 only)
    int count;        // entries in pool
    int nacc;         // accepted so far (Java results.size() before capping)
    int evicted;      // Java: first results.poll() has happened (furthest is live)
    int nexp;         // nodes expanded
    uint64_t tau;     // lazy upper bound of the ef-th smallest key (masked); KEY_MAX = not established
    float dentry;
    float qnorm;
    int cur;
};

__device__ __forceinline__ uint64_t pool_key(float d, int node) { return ((uint64_t)f2ord(d) << 32) | (uint32_t)node; }
__device__ __forceinline__ uint64_t mask_key(uint64_t k) { return k & ~EXP_BIT; }

// exact distance of one node to the query in the reference's sequential order (one lane per node)
__device__ __forceinline__ float exact_seq_dist(const HnswSearchParams& p, const float* sq, float qnorm, int node) {
    const float* row = p.base + (size_t)node * p.dim;
    if (p.metric == 0) return l2sq_seq(row, sq, p.dim);
    return cosine_dist_from_dot(dot_seq(row, sq, p.dim), p.norms[node], qnorm);
}
// direct PQ distance, order R5 (HnswPqIndex.java:371-386)
__device__ __forceinline__ float pq_direct_dist(const HnswSearchParams& p, const float* sq, int node) {
    const int sd = p.dim / p.pq_m;
    const uint8_t* code = p.codes + (size_t)node * p.pq_m;
    float dist = 0.0f;
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

// pad the unused tail, sort the pool ascending, keep the `keep` smallest and refresh tau (block-wide)
__device__ __forceinline__ void pool_compact(uint64_t* pool, int cap, int keep, HnswShared& S) {
    __syncthreads();
    const int cnt = S.count;
    for (int i = cnt + threadIdx.x; i < cap; i += HN_NT) pool[i] = KEY_MAX;
    block_bitonic_sort<false>(pool, nullptr, cap);
    if (threadIdx.x == 0) {
        int c = min(cnt, keep);
        S.count = c;
        if (c >= keep) S.tau = mask_key(pool[keep - 1]);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(HN_NT) hnsw_search_kernel(const HnswSearchParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int dimpad = (p.dim + 3) & ~3;
    float* sq = reinterpret_cast<float*>(smem_raw);
    uint64_t* pool = reinterpret_cast<uint64_t*>(sq + dimpad);
    const int vwords = (p.n + 31) >> 5;
    uint32_t* vis = p.visited_in_smem ? reinterpret_cast<uint32_t*>(pool + p.cap)
                                      : p.visited_scratch + (size_t)blockIdx.x * vwords;
    __shared__ HnswShared S;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ef = p.ef, cap = p.cap, variant = p.variant;
    unsigned long long evals = 0, exps = 0;

    for (int q = blockIdx.x; q < p.nq; q += gridDim.x) {
        // ---- A. stage the query (HnswIndex normalises it: HnswIndex.java:344, core/Vector.java:107-116) ----
        __syncthreads();
        for (int i = tid; i < p.dim; i += HN_NT) sq[i] = p.queries[(size_t)q * p.dim + i];
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

        // ---- B. greedy descent on levels max_level-1 .. 1 (warp 0) ----
        if (warp == 0) {
            int cur = p.entry;
            const bool comp = (p.metric == 0) && p.trained;
            for (int l = p.max_level - 1; l > 0; l--) cur = greedy_level(p, sq, qnorm, cur, l, comp, evals);
            float d0 = p.metric == 0 ? warp_l2sq(p.base + (size_t)cur * p.dim, sq, p.dim, lane)
                                     : cosine_dist_from_dot(warp_dot(p.base + (size_t)cur * p.dim, sq, p.dim, lane), p.norms[cur], qnorm);
            if (lane == 0) {
                S.cur = cur; S.dentry = d0;
                pool[0] = pool_key(d0, cur);
                S.count = 1; S.nacc = 1; S.evicted = 0; S.nexp = 0; S.tau = KEY_MAX;
                vis[cur >> 5] |= 1u << (cur & 31);
                evals++;
            }
        }
        __syncthreads();

        // ---- C. layer-0 best-first search ----
        while (true) {
            // C1. candidates.poll(): smallest unexpanded key
            uint64_t best = KEY_MAX; int bslot = -1;
            const int cnt = S.count;
            for (int i = tid; i < cnt; i += HN_NT) {
                uint64_t kk = pool[i];
                if (!(kk & EXP_BIT) && kk < best) { best = kk; bslot = i; }
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                uint64_t ob = __shfl_xor_sync(0xffffffffu, best, off);
                int os = __shfl_xor_sync(0xffffffffu, bslot, off);
                if (ob < best) { best = ob; bslot = os; }
            }
            if (lane == 0) { S.red[warp] = best; S.redslot[warp] = bslot; }
            __syncthreads();
            if (tid == 0) {
                uint64_t b = S.red[0]; int sl = S.redslot[0];
                for (int w = 1; w < HN_WARPS; w++) if (S.red[w] < b) { b = S.red[w]; sl = S.redslot[w]; }
                S.pick_key = b; S.pick_slot = sl; S.rank = 0;
            }
            __syncthreads();
            const uint64_t ku = S.pick_key;
            if (ku == KEY_MAX) break;                               // candidate set empty
            // C2. termination test
            const float du = ord2f((uint32_t)(ku >> 32));
            const int nacc = S.nacc, evicted = S.evicted, nexp = S.nexp;
            bool stop = false, need_rank = false;
            if (variant == 0) need_rank = (nacc >= ef) && (nexp >= ef);
            else if (!evicted) { const bool worse = du > S.dentry; stop = (variant == 2) ? worse : (worse && nacc >= ef); }
            else need_rank = nexp >= ef;
            if (need_rank) {
                int r = 0;
                for (int i = tid; i < cnt; i += HN_NT) r += (mask_key(pool[i]) < ku) ? 1 : 0;
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) r += __shfl_xor_sync(0xffffffffu, r, off);
                if (lane == 0 && r) atomicAdd(&S.rank, r);
                __syncthreads();
                if (S.rank >= ef) stop = true;
            }
            if (stop) break;
            // C3. expand: neighbour list, visited test-and-set, ordered compaction of the fresh ones (warp 0)
            const int node = (int)(ku & 0x7fffffffull);
            if (warp == 0) {
                const int32_t* row = p.adj + (size_t)node * p.deg;
                int basec = 0;
                for (int j0 = 0; j0 < p.deg; j0 += 32) {
                    const int j = j0 + lane;
                    const int nb = j < p.deg ? row[j] : -1;
                    bool fresh = false;
                    if (nb >= 0) {
                        const uint32_t bit = 1u << (nb & 31);
                        const uint32_t old = atomicOr(&vis[nb >> 5], bit);
                        fresh = !(old & bit);
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, fresh);
                    if (fresh) S.newn[basec + __popc(m & ((1u << lane) - 1u))] = nb;
                    basec += __popc(m);
                }
                if (lane == 0) {
                    S.newcnt = basec;
                    exps++; evals += basec;
                }
            }
            __syncthreads();
            // C4. distances of the fresh neighbours: one warp per row
            const int newcnt = S.newcnt;
            for (int j = warp; j < newcnt; j += HN_WARPS) {
                const int nb = S.newn[j];
                const float* rowp = p.base + (size_t)nb * p.dim;
                float d = p.metric == 0 ? warp_l2sq(rowp, sq, p.dim, lane)
                                        : cosine_dist_from_dot(warp_dot(rowp, sq, p.dim, lane), p.norms[nb], qnorm);
                if (lane == 0) S.newd[j] = d;
            }
            __syncthreads();
            // C5. sequential acceptance in list order (thread 0), interrupted when the pool must be compacted.
            //   Java: if (results.size() < ef || d < furthest) { candidates.add; results.add; evict/refresh furthest }
            //   (HnswPqIndex.java:953-963 live furthest; :876-884 and HnswIndex.java:611-625 furthest refreshed on eviction)
            int j0 = 0;
            bool first_pass = true;
            while (true) {
                if (tid == 0) {
                    int j = j0, need = 0;
                    if (first_pass) { pool[S.pick_slot] = ku | EXP_BIT; S.nexp = nexp + 1; first_pass = false; }   // slot still valid: no compaction since the pick
                    int count = S.count, na = S.nacc, ev = S.evicted;
                    const uint64_t tau = S.tau;
                    const float dentry = S.dentry;
                    for (; j < newcnt; j++) {
                        const float d = S.newd[j];
                        const uint64_t key = pool_key(d, S.newn[j]);
                        bool acc;
                        if (na < ef) acc = true;
                        else if (variant == 0 || ev) acc = key < tau;
                        else acc = d < dentry;                    // stale furthest == entry distance
                        if (!acc) continue;
                        if (count == cap) { need = 1; break; }    // pool full: compact, then retry this neighbour
                        pool[count++] = key; na++;
                        if (variant == 0) { if (na == ef) { need = 2; j++; break; } }          // furthest becomes live
                        else if (!ev && na == ef + 1) { ev = 1; need = 2; j++; break; }         // first eviction
                    }
                    S.count = count; S.nacc = na; S.evicted = ev; S.jstop = j; S.need_compact = need;
                }
                __syncthreads();
                if (!S.need_compact) break;
                j0 = S.jstop;
                pool_compact(pool, cap, ef, S);
            }
        }

        // ---- D. final: best candidates by traversal distance -> exact re-score in the reference's sequential order
        //         (HnswPqIndex.java:646-664 reranks every candidate; values are identical to its traversal distances)
        //         -> (dist,id) sort -> top-k ----
        pool_compact(pool, cap, cap, S);
        int rs = 32;
        while (rs < p.k + 22) rs <<= 1;
        if (rs > cap) rs = cap;
        const int have = min(S.count, rs);
        uint64_t mykey[4];
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const int j = tid + t * HN_NT;
            mykey[t] = KEY_MAX;
            if (j < have) {
                const int node = (int)(pool[j] & 0x7fffffffull);
                mykey[t] = make_key(exact_seq_dist(p, sq, qnorm, node), p.ids[node]);
            }
        }
        __syncthreads();
#pragma unroll
        for (int t = 0; t < 4; t++) { const int j = tid + t * HN_NT; if (j < rs) pool[j] = mykey[t]; }
        block_bitonic_sort<false>(pool, nullptr, rs);
        for (int j = tid; j < p.k; j += HN_NT) {
            const uint64_t key = j < rs ? pool[j] : KEY_MAX;
            if (key == KEY_MAX) { p.out_ids[(size_t)q * p.k + j] = -1; p.out_dists[(size_t)q * p.k + j] = __int_as_float(0x7f800000); }
            else { p.out_ids[(size_t)q * p.k + j] = key_id(key); p.out_dists[(size_t)q * p.k + j] = key_dist(key); }
        }
    }
    if (p.counters) {
        if (evals) atomicAdd(&p.counters[0], evals);
        if (exps) atomicAdd(&p.counters[1], exps);
    }
}

size_t hnsw_search_smem_bytes(const HnswSearchParams& p) {
    size_t b = (size_t)((p.dim + 3) & ~3) * 4 + (size_t)p.cap * 8;
    if (p.visited_in_smem) b += (size_t)((p.n + 31) / 32) * 4;
    return b;
}
int launch_hnsw_search(const HnswSearchParams& p, int grid, cudaStream_t st) {
    if (p.nq <= 0) return 0;
    if (p.k + 22 > 4 * HN_NT || p.deg > 64) return 1;
    const size_t smem = hnsw_search_smem_bytes(p);
    if (cudaFuncSetAttribute(hnsw_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 2;
    hnsw_search_kernel<<<grid, HN_NT, smem, st>>>(p);
    return 0;
}
