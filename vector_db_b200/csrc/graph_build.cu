// graph_build.cu — GPU graph construction for rebuildIndex (reference: Java addVectorCompressed / addVectorRaw /
// selectNeighbors, index/HnswPqIndex.java:391-539, 1043-1064; HnswIndex.java:139-248).
//
// The reference inserts one vector at a time: beam search with ef = max(efConstruction, 200) on every level <= the node's level,
// M nearest as its links, back-links pruned to the M nearest. Here the rows are inserted in BATCHES against the graph built so far:
//   1. levels come from the same seeded law, drawn on the host in row order; the first GB_SEED rows are inserted by the host builder
//      (host_graph.cpp: the graph is tiny there, and it gives the batches something to search);
//   2. per batch (<= 1/256 of the rows already in the graph, <= GB_MAX_BATCH) and per level L from the top down: the beam kernel of the
//      query path (hnsw_search_kernel: greedy descent above L with exact distances, searchLayer on level L in the Java pop order, exact
//      re-score) finds the M nearest graph nodes of every new node that reaches L. All levels are searched against the graph as it was
//      BEFORE the batch; the nodes of one batch do not see each other (0.4 % of the graph: measured, profiles/r2_graph_build.md);
//   3. gb_link_kernel (one CTA per new node) keeps the M nearest, writes the node's links and cached edge distances, and posts a
//      back-link request to each chosen neighbour. (Opt-in VDB_GB_INTRA=1 also brute-forces the other nodes of the batch — a much
//      BETTER graph than the reference's, whose insert search all but stops when its result set fills: not parity, not the default);
//   4. gb_backlink_kernel (one warp per neighbour with requests) appends them; a list that would exceed M keeps its M nearest
//      (selectNeighbors: sorted by distance), using the cached edge distances.
// Not link-for-link the sequential graph (Java's own graphs are not reproducible either: ThreadLocalRandom levels) — the bar is
// recall parity and degree statistics against the host-built graph (tests/test_gpu_graph_build.py). Strict mode keeps the host builder.
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "kernels.h"
#include "vdb_index.h"

#define GB_SEED 2048          // rows inserted by the host builder first
#define GB_MAX_BATCH 4096
#define GB_DIV 64             // a batch holds at most 1/GB_DIV of the rows already in the graph (its nodes do not see each other)
#define GB_INCAP 64           // back-link requests one node can receive per batch and level (further ones are dropped: hubs only)
#define GB_NT 256

namespace {
// ---- 3. links of one new node: found (beam result) + batch-mates (brute force) -> M nearest ----
__global__ void __launch_bounds__(GB_NT) gb_link_kernel(const float* __restrict__ base, const float* __restrict__ norms, int dim, int metric,
                                                        const int* __restrict__ qrows, int nq, int intra, const int32_t* __restrict__ found_ids,
                                                        const float* __restrict__ found_d, int M, int cap, int32_t* __restrict__ adjL,
                                                        float* __restrict__ adjdL, int* __restrict__ in_cnt, uint64_t* __restrict__ in_key) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sq = reinterpret_cast<float*>(smem_raw);
    uint64_t* keys = reinterpret_cast<uint64_t*>(sq + ((dim + 3) & ~3));
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int u = qrows[q];
    for (int i = tid; i < dim; i += GB_NT) sq[i] = base[(size_t)u * dim + i];
    for (int i = tid; i < cap; i += GB_NT) keys[i] = KEY_MAX;
    __syncthreads();
    for (int j = tid; j < M; j += GB_NT) { const int32_t f = found_ids[(size_t)q * M + j]; keys[j] = f >= 0 ? make_key(found_d[(size_t)q * M + j], f) : KEY_MAX; }
    const float un = metric == 1 ? norms[u] : 0.0f;
    for (int m = warp; m < (intra ? nq : 0); m += GB_NT / 32) {
        const int v = qrows[m];
        if (v == u) continue;
        float d;
        if (metric == 0) d = warp_l2sq(base + (size_t)v * dim, sq, dim, lane);
        else d = cosine_dist_from_dot(warp_dot(base + (size_t)v * dim, sq, dim, lane), norms[v], un);
        if (lane == 0) keys[M + m] = make_key(d, v);
    }
    __syncthreads();
    block_bitonic_sort<false>(keys, nullptr, cap);
    for (int j = tid; j < M; j += GB_NT) {
        const uint64_t k = keys[j];
        int32_t nb = -1; float d = -1.0f;
        if (k != KEY_MAX) {
            nb = key_id(k); d = key_dist(k);
            const int slot = atomicAdd(&in_cnt[nb], 1);
            if (slot < GB_INCAP) in_key[(size_t)nb * GB_INCAP + slot] = make_key(d, u);
        }
        adjL[(size_t)u * M + j] = nb; adjdL[(size_t)u * M + j] = d;
    }
}

// ---- 4. back-links: one warp per row with requests; list = existing order + requests by (dist, id); > M -> M nearest ----
__device__ __forceinline__ uint64_t shfl_u64(uint64_t v, int m) {
    return ((uint64_t)__shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m) << 32) | __shfl_xor_sync(0xffffffffu, (uint32_t)v, m);
}
// bitonic sort of 128 keys, 4 per lane (element 4 * lane + r)
__device__ __forceinline__ void warp_sort128(uint64_t (&e)[4], int lane) {
#pragma unroll
    for (int k = 2; k <= 128; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < 4) {
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    if ((r & j) == 0) {
                        const bool up = k < 4 ? ((r & k) == 0) : ((lane & (k / 4)) == 0);
                        const uint64_t a = e[r], b = e[r | j];
                        if ((a > b) == up) { e[r] = b; e[r | j] = a; }
                    }
                }
            } else {
                const int m = j / 4;
                const bool keepmin = ((lane & m) == 0) == ((lane & (k / 4)) == 0);
#pragma unroll
                for (int r = 0; r < 4; r++) { const uint64_t o = shfl_u64(e[r], m); e[r] = keepmin ? (e[r] < o ? e[r] : o) : (e[r] > o ? e[r] : o); }
            }
        }
    }
}
__global__ void gb_backlink_kernel(int rows, int M, int32_t* __restrict__ adjL, float* __restrict__ adjdL, int* __restrict__ in_cnt,
                                   const uint64_t* __restrict__ in_key) {
    const int lane = threadIdx.x & 31;
    const int t = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (t >= rows) return;
    const int c = min(in_cnt[t], GB_INCAP);
    if (c == 0) return;                                   // warp-uniform: every lane reads the same counter
    __syncwarp();
    if (lane == 0) in_cnt[t] = 0;
    int deg = 0;
    for (int j0 = 0; j0 < M; j0 += 32) { const int j = j0 + lane; deg += __popc(__ballot_sync(0xffffffffu, j < M && adjL[(size_t)t * M + j] >= 0)); }
    // requests sorted by (dist, id); a request from a node that is already in the list is dropped (two nodes of one batch that
    // chose each other: the forward link written by gb_link_kernel is the same edge)
    uint64_t e[4];
    int kept = 0;
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int s = 4 * lane + r;
        e[r] = s < c ? in_key[(size_t)t * GB_INCAP + s] : KEY_MAX;
        if (e[r] != KEY_MAX) {
            const int32_t u = key_id(e[r]);
            for (int j = 0; j < deg; j++) if (adjL[(size_t)t * M + j] == u) { e[r] = KEY_MAX; break; }
        }
        kept += e[r] != KEY_MAX ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(0xffffffffu, kept, o);
    const int c2 = kept;
    if (c2 == 0) return;
    warp_sort128(e, lane);
    if (deg + c2 <= M) {        // room for all: existing order kept, requests appended
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int s = 4 * lane + r;
            if (s < c2) { adjL[(size_t)t * M + deg + s] = key_id(e[r]); adjdL[(size_t)t * M + deg + s] = key_dist(e[r]); }
        }
        return;
    }
    // overflow: existing (cached distances) + requests -> the M nearest, ascending (selectNeighbors sorts by distance)
    uint64_t f[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int s = 4 * lane + r;                       // slots [0, deg): existing; [deg, deg + c2): requests (already sorted in e)
        f[r] = KEY_MAX;
        if (s < deg) f[r] = make_key(adjdL[(size_t)t * M + s], adjL[(size_t)t * M + s]);
    }
    // requests live in e at sorted positions 0..c-1 (element index 4 * lane + r): move them behind the existing ones
    __shared__ uint64_t stage[8][128];
    uint64_t* st = stage[(threadIdx.x >> 5) & 7];
#pragma unroll
    for (int r = 0; r < 4; r++) st[4 * lane + r] = e[r];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 4; r++) { const int s = 4 * lane + r; if (s >= deg && s < deg + c2 && s < 128) f[r] = st[s - deg]; }
    __syncwarp();
    warp_sort128(f, lane);
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int s = 4 * lane + r;
        if (s < M) { adjL[(size_t)t * M + s] = key_id(f[r]); adjdL[(size_t)t * M + s] = key_dist(f[r]); }
    }
}
__global__ void gb_iota_kernel(int32_t* p, int n) { const int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = i; }
// start node of query i on this level = the nearest node its search found on the level above (prevpos = its index there), else -1
__global__ void gb_entries_kernel(const int* __restrict__ prevpos, int nq, const int32_t* __restrict__ found_above, int M, int32_t* __restrict__ qentry) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nq) qentry[i] = prevpos[i] >= 0 ? found_above[(size_t)prevpos[i] * M] : -1;
}
}  // namespace

#define GB_FAIL(code, msg) do { vdb_set_error(msg); return (code); } while (0)

int vdb_gpu_build_graph(vdb_index* x, int ef, int variant) {
    const int64_t n = x->n;
    const int dim = x->cfg.dim, M = x->g.M, L = x->g.max_level;
    cudaStream_t st = x->stream;
    HostGraph& g = x->g;
    g.clear();
    g.vecs = x->h_vecs.data(); g.norms = x->h_norms.data();
    // 1. seed graph on the host (draws the first levels), then the remaining levels from the same stream
    const int64_t n0 = std::min<int64_t>(n, GB_SEED);
    for (int64_t i = 0; i < n0; i++) g.insert((int)i, ef, variant);
    std::vector<int32_t> level(g.level.begin(), g.level.end());
    level.resize(n);
    for (int64_t i = n0; i < n; i++) level[i] = g.draw_level();
    VDB_TRY(vdb_upload_rows_and_traversal_copy(x));
    // device graph: adjacency [L][cap][M] + cached edge distances, back-link request buffers
    x->adj_cap = std::max<int64_t>(x->adj_cap, n);
    const int64_t cap = x->adj_cap;
    VDB_TRY(x->d_adj.ensure((size_t)L * cap * M, 0, st));
    DevBuf<float> d_adjd; DevBuf<int> d_incnt, d_qrows; DevBuf<uint64_t> d_inkey; DevBuf<int32_t> d_iota, d_found; DevBuf<float> d_foundd;
    struct Guard { DevBuf<float>& a; DevBuf<int>& b; DevBuf<int>& c; DevBuf<uint64_t>& d; DevBuf<int32_t>& e; DevBuf<int32_t>& f; DevBuf<float>& h;
                   ~Guard() { a.release(); b.release(); c.release(); d.release(); e.release(); f.release(); h.release(); } } guard{d_adjd, d_incnt, d_qrows, d_inkey, d_iota, d_found, d_foundd};
    VDB_TRY(d_adjd.ensure((size_t)L * cap * M, 0, st));
    VDB_TRY(d_incnt.ensure((size_t)n, 0, st));
    VDB_TRY(d_inkey.ensure((size_t)n * GB_INCAP, 0, st));
    VDB_TRY(d_iota.ensure((size_t)n, 0, st));
    VDB_TRY(d_qrows.ensure((size_t)L * GB_MAX_BATCH, 0, st));
    VDB_TRY(d_found.ensure((size_t)L * GB_MAX_BATCH * M, 0, st));
    VDB_TRY(d_foundd.ensure((size_t)L * GB_MAX_BATCH * M, 0, st));
    VDB_CUDA(cudaMemsetAsync(x->d_adj.p, 0xff, (size_t)L * cap * M * 4, st));
    VDB_CUDA(cudaMemsetAsync(d_incnt.p, 0, (size_t)n * 4, st));
    gb_iota_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_iota.p, (int)n);
    for (int l = 0; l < L; l++) {
        VDB_CUDA(cudaMemcpyAsync(x->d_adj.p + (size_t)l * cap * M, g.adj[l].data(), (size_t)n0 * M * 4, cudaMemcpyHostToDevice, st));
        VDB_CUDA(cudaMemcpyAsync(d_adjd.p + (size_t)l * cap * M, g.adjd[l].data(), (size_t)n0 * M * 4, cudaMemcpyHostToDevice, st));
    }
    int entry = g.entry;
    std::vector<int> rows_at, prevpos;
    DevBuf<int> d_prevpos; DevBuf<int32_t> d_qentry;
    struct Guard2 { DevBuf<int>& a; DevBuf<int32_t>& b; ~Guard2() { a.release(); b.release(); } } guard2{d_prevpos, d_qentry};
    VDB_TRY(d_prevpos.ensure((size_t)GB_MAX_BATCH, 0, st));
    VDB_TRY(d_qentry.ensure((size_t)GB_MAX_BATCH, 0, st));
    const bool pq = x->cfg.index_type == VDB_INDEX_HNSW_PQ;
    // 2.-4. batches
    for (int64_t start = n0; start < n;) {
        const int64_t bsz = std::min<int64_t>(n - start, std::max<int64_t>(16, std::min<int64_t>(GB_MAX_BATCH, start / (getenv("VDB_GB_DIV") ? atoi(getenv("VDB_GB_DIV")) : GB_DIV))));
        int top = 0;
        for (int64_t i = start; i < start + bsz; i++) top = std::max(top, (int)level[i]);
        top = std::min(top, L - 1);
        // phase 1: every level's search runs against the graph as it was BEFORE this batch (links written below would otherwise let
        // the greedy descent step onto batch rows whose lower-level lists do not exist yet)
        std::vector<int> nq_at(L, 0);
        for (int lv = top; lv >= 0; lv--) {
            rows_at.clear();
            for (int64_t i = start; i < start + bsz; i++) if (level[i] >= lv) rows_at.push_back((int)i);
            const int nq = nq_at[lv] = (int)rows_at.size();
            if (nq == 0) continue;
            int* qr = d_qrows.p + (size_t)lv * GB_MAX_BATCH;
            VDB_CUDA(cudaMemcpyAsync(qr, rows_at.data(), (size_t)nq * 4, cudaMemcpyHostToDevice, st));   // pageable source: staged before the call returns
            // rows that also searched the level above start from the nearest node found there (no new descent): Java's
            // currentEntryPoint = selectedNeighbors.get(0) (:526-531)
            prevpos.assign(nq, -1);
            if (lv < top) { int pos = 0; for (int i = 0; i < nq; i++) if (level[rows_at[i]] > lv) prevpos[i] = pos++; }
            VDB_CUDA(cudaMemcpyAsync(d_prevpos.p, prevpos.data(), (size_t)nq * 4, cudaMemcpyHostToDevice, st));
            gb_entries_kernel<<<(nq + 255) / 256, 256, 0, st>>>(d_prevpos.p, nq, d_found.p + (size_t)std::min(lv + 1, L - 1) * GB_MAX_BATCH * M, M, d_qentry.p);
            HnswSearchParams p{};
            p.base = x->d_base.p; p.norms = pq ? nullptr : x->d_norms.p; p.ids = d_iota.p; p.adj = x->d_adj.p; p.adj_stride = cap;
            p.codes = nullptr; p.codebooks = nullptr; p.queries = nullptr; p.qrows = qr; p.beam_level = lv; p.qentry = d_qentry.p;
            p.out_ids = d_found.p + (size_t)lv * GB_MAX_BATCH * M; p.out_dists = d_foundd.p + (size_t)lv * GB_MAX_BATCH * M; p.counters = nullptr;
            p.n = (int)start; p.dim = dim; p.deg = M; p.max_level = L; p.entry = entry;
            p.pq_m = x->pq_m; p.trained = 0; p.metric = pq ? 0 : 1; p.variant = variant;
            p.ef = (int)std::min<int64_t>(ef, start); p.k = M; p.nq = nq;
            // ONE candidate per round = the Java pop order. It matters here: Java's searchLayer keeps `furthest` at the entry
            // distance until the first eviction (HnswPqIndex.java:847-884), so the insert search all but stops once the result set is
            // full — that is what shapes the reference's graphs (forward lists hold ~25 % of the exact 32-NN at 10 k x 512-d). Several
            // candidates per round overshoot the ef boundary inside one round, which lifts that stop: forward lists reach ~88 %,
            // Recall@10 on C3 0.93 instead of the reference's 0.84 (profiles/r2_graph_build.md). A better graph, but not the
            // reference's: opt-in through VDB_GB_WIDE_BEAM=1 only.
            p.batch = getenv("VDB_GB_WIDE_BEAM") ? 0 : 1;
            if (p.ef < M) GB_FAIL(VDB_ERR_INVALID, "graph build: seed graph smaller than the link count");
            VDB_TRY(vdb_graph_search_launch(x, p, st));
        }
        // phase 2: links and back-links, level by level
        for (int lv = top; lv >= 0; lv--) {
            const int nq = nq_at[lv];
            if (nq == 0) continue;
            const int* qr = d_qrows.p + (size_t)lv * GB_MAX_BATCH;
            const int kcap = next_pow2(M + nq);
            const size_t smem = (size_t)((dim + 3) & ~3) * 4 + (size_t)kcap * 8;
            VDB_CUDA(cudaFuncSetAttribute(gb_link_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            gb_link_kernel<<<nq, GB_NT, smem, st>>>(x->d_base.p, pq ? nullptr : x->d_norms.p, dim, pq ? 0 : 1, qr, nq, getenv("VDB_GB_INTRA") ? 1 : 0,
                                                    d_found.p + (size_t)lv * GB_MAX_BATCH * M, d_foundd.p + (size_t)lv * GB_MAX_BATCH * M, M,
                                                    kcap, x->d_adj.p + (size_t)lv * cap * M, d_adjd.p + (size_t)lv * cap * M, d_incnt.p, d_inkey.p);
            const int64_t rows = start + bsz;
            gb_backlink_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, st>>>((int)rows, M, x->d_adj.p + (size_t)lv * cap * M,
                                                                                   d_adjd.p + (size_t)lv * cap * M, d_incnt.p, d_inkey.p);
            VDB_CUDA(cudaGetLastError());
        }
        for (int64_t i = start; i < start + bsz; i++) if (level[i] > level[entry]) entry = (int)i;     // HnswPqIndex.java:534-536
        start += bsz;
    }
    // the host graph adopts the result (incremental adds, removals and snapshots keep working on it)
    std::vector<int32_t> adj((size_t)L * n * M);
    for (int l = 0; l < L; l++)
        VDB_CUDA(cudaMemcpyAsync(adj.data() + (size_t)l * n * M, x->d_adj.p + (size_t)l * cap * M, (size_t)n * M * 4, cudaMemcpyDeviceToHost, st));
    VDB_CUDA(cudaStreamSynchronize(st));
    const uint64_t ctr = g.level_ctr;
    g.import_graph(level.data(), adj.data(), (int)n, entry);
    g.level_ctr = ctr;
    g.clear_dirty();                                            // the device copy IS the source: nothing to upload
    x->graph_uploaded = true;
    return VDB_OK;
}
