// flat_gemm.cu — brute-force flat search (a6, a13; SURVEY.md §8 row "Flat search / batched rerank as GEMM") on the
// 5th-generation tensor cores: a bf16 query x base contraction (tcgen05.mma, operands staged by TMA into 128B-swizzled
// shared memory, fp32 accumulators in TMEM) whose epilogue never writes the Q x N matrix — it turns each accumulator
// into an approximate distance, compares it with a per-query threshold and appends the few survivors to a per-query
// candidate list. The shortlist is then re-scored in fp32 in the reference's own summation order
// (HnswPqIndex.java:741-750 / DistanceAVX2.cpp:9-33 / HnswIndex.java:126-133) and CERTIFIED: a query's answer is
// accepted only if the k-th exact distance is strictly below a rigorous lower bound on the exact distance of every row
// that was not re-scored; otherwise the query is flagged and answered by the order-exact scan in flat_kernels.cu.
// So ids and distances are bit-identical to the CPU reference while >99% of the arithmetic runs on tensor cores.
//
//   pass 0  (once per index change)  bf16 mirror of the base: rows centred on the base mean for sq-L2
//           (||q-x||^2 is translation invariant; centring shrinks ||q'|| ||x'|| and with it the bf16 error bound),
//           rows as stored (already normalised, HnswIndex.java:344) for cosine; per-row (a,b) with score = a + b*acc.
//   pass 1  GEMM over a strided sample of <= 32767 rows, scores stored; per query the s-th smallest -> threshold tau.
//   pass 2  GEMM over all rows, fused filter epilogue: score < tau -> append (score,row).
//   pass 3  per query: sort candidates, exact fp32 re-score of the best k2, certify, emit top-k by (dist,id).
#include "kernels.h"
#include <cuda.h>
#include <cuda_bf16.h>
#include <algorithm>
#include <cmath>
#include <cstdio>

namespace {

constexpr int BM = 128;            // queries per tile  (UMMA M, one TMEM lane per query)
constexpr int BN = 256;            // base rows per tile (UMMA N, one TMEM column per row)
constexpr int BK = 64;             // bf16 per k-block = 128 bytes = one swizzle-128B row
constexpr int UK = 16;             // K of one tcgen05.mma kind::f16
constexpr int A_BYTES = BM * BK * 2;                 // 16 KB
// CG = 1: one CTA per tile (128 queries x 256 rows), B tile 32 KB, 4 stages.
// CG = 2: a CTA pair per tile (256 queries x 256 rows, tcgen05 cta_group::2): each CTA stages its 128 queries and HALF of
//         the base rows (16 KB), the pair's tensor cores read both halves -> 32 KB per k-block per SM instead of 48, 6 stages.
template <int CG> struct Cfg {
    static constexpr int B_BYTES = (BN / CG) * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = CG == 2 ? 6 : 4;
    static constexpr int PIPE_BYTES = STAGES * STAGE_BYTES;
};
constexpr int GEMM_THREADS = 320;                    // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..9 epilogue (two per SM sub-partition)
constexpr int TMEM_COLS = 512;                       // two accumulator stages of BN columns
constexpr int SMEM_AB = 2 * BN * 4;                  // per accumulator stage: the additive term a[] of the tile's base rows
constexpr int WBUF = 64;                             // per-epilogue-warp staging buffer for survivors (entries)
constexpr int SMEM_WB = 8 * WBUF * 8 + 64;           // 8 warps x WBUF x (ordered score, lane<<8 | column) + counters
constexpr int GEMM_SMEM = 4 * (A_BYTES + BN * BK * 2) + SMEM_AB + SMEM_WB + 256 + 1024;   // 192 KB pipeline (both configs) + barriers + alignment slack

// ---------------------------------------------------------------------------------------------------------------
// PTX wrappers (sm_100a)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, both operands K-major, one thread issues for the CTA
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
        "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- cta_group::2 forms (a CTA pair = one 2-CTA cluster; the even CTA issues the MMAs for both) ----
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
    uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank)); return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are credited to the barrier at the same offset in the pair's EVEN CTA (peer bit cleared)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once these MMAs are done) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}

// Shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor layout): K-major tile of 128-byte rows,
// SWIZZLE_128B (layout type 2), 8-row groups 1024 bytes apart (SBO), LBO unused for swizzled K-major (1), version 1.
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
// Instruction descriptor (InstrDescriptor): D=f32 (bit 4), A=B=bf16 (bits 7,10), both K-major, N>>3 at bit 17, M>>4 at bit 24.
template <int CG> struct IDesc {                      // CG = 2: M = 256 across the pair
    static constexpr uint32_t value = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((BM * CG) >> 4) << 24);
};


// v[i] for a run-time i without dynamic register indexing: five levels of selects
__device__ __forceinline__ float pick32(const float (&v)[32], int i) {
    float a[16], b[8], c[4];
#pragma unroll
    for (int j = 0; j < 16; j++) a[j] = (i & 1) ? v[2 * j + 1] : v[2 * j];
#pragma unroll
    for (int j = 0; j < 8; j++) b[j] = (i & 2) ? a[2 * j + 1] : a[2 * j];
#pragma unroll
    for (int j = 0; j < 4; j++) c[j] = (i & 4) ? b[2 * j + 1] : b[2 * j];
    const float d0 = (i & 8) ? c[1] : c[0], d1 = (i & 8) ? c[3] : c[2];
    return (i & 16) ? d1 : d0;
}

// ---------------------------------------------------------------------------------------------------------------
// The contraction. One persistent CTA per SM walks tiles t = blockIdx.x, +gridDim.x, ...; tile t covers base rows
// [nt*256, +256) x queries [mt*128, +128) with nt = t / m_tiles, so CTAs running at the same time share base tiles in L2
// and the base streams from HBM once. Three pipelines: shared-memory stages (TMA -> MMA), two TMEM accumulator stages
// (MMA -> epilogue), and the tile walk.
// ---------------------------------------------------------------------------------------------------------------
struct GemmArgs {
    int m_tiles, n_tiles, k_blocks;
    int nq;                    // valid query rows
    int64_t n_rows;            // valid rows of the B operand (sample rows in pass 1)
    int64_t row_step;          // B row j is base row j*row_step
    const float* arow;         // per base row: score = arow[row] + bconst * acc
    float bconst;
    float* scores;             // STORE: [nq][ld_scores]
    int64_t ld_scores;
    const float* tau;          // FILTER: per-query threshold
    uint32_t* cnt;             // FILTER: per-query candidate count (may run past cap: overflow marker)
    uint64_t* cand;            // FILTER: [nq][cap] (ordered score << 32 | row)
    uint32_t cap;
};

template <bool FILTER, int CG>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
flat_gemm_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_b, const GemmArgs g) {
    constexpr int STAGES = Cfg<CG>::STAGES, STAGE_BYTES = Cfg<CG>::STAGE_BYTES;
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;     // position inside the CTA pair
    const int unit = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, nunits = (int)gridDim.x / CG;
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t sbase = (raw + 1023u) & ~1023u;                    // swizzle-128B tiles need 1024-byte alignment
    unsigned char* sm = smem_raw + (sbase - raw);
    float* s_a = reinterpret_cast<float*>(sm + Cfg<CG>::PIPE_BYTES);
    uint2* s_wbuf = reinterpret_cast<uint2*>(sm + Cfg<CG>::PIPE_BYTES + SMEM_AB);
    uint32_t* s_wcnt = reinterpret_cast<uint32_t*>(sm + Cfg<CG>::PIPE_BYTES + SMEM_AB + 8 * WBUF * 8);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + Cfg<CG>::PIPE_BYTES + SMEM_AB + SMEM_WB);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    const uint32_t bar0 = smem_u32(bars);
#define BAR_FULL(s) (bar0 + 8u * (uint32_t)(s))
#define BAR_EMPTY(s) (bar0 + 8u * (uint32_t)(STAGES + (s)))
#define BAR_TFULL(a) (bar0 + 8u * (uint32_t)(2 * STAGES + (a)))
#define BAR_TEMPTY(a) (bar0 + 8u * (uint32_t)(2 * STAGES + 2 + (a)))
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(BAR_FULL(s), CG); mbar_init(BAR_EMPTY(s), 1); }
        for (int a = 0; a < 2; a++) { mbar_init(BAR_TFULL(a), 1); mbar_init(BAR_TEMPTY(a), 8 * CG); }
        fence_barrier_init();
        tma_prefetch_desc(&map_q);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1) { if (CG == 2) tmem_alloc_pair(smem_u32(tmem_slot), TMEM_COLS); else tmem_alloc(smem_u32(tmem_slot), TMEM_COLS); }
    tc_fence_before();
    if (CG == 2) cluster_sync(); else __syncthreads();      // barriers initialised in both CTAs before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    const int total = g.m_tiles * g.n_tiles;

    if (warp == 0) {
        // ===== TMA producer =====
        int stage = 0; uint32_t phase = 0;
        for (int t = unit; t < total; t += nunits) {
            const int nt = t / g.m_tiles, mt = t - nt * g.m_tiles;
            if (lane == 0) {
                for (int kb = 0; kb < g.k_blocks; kb++) {
                    mbar_wait(BAR_EMPTY(stage), phase ^ 1u);
                    if (CG == 1) {
                        mbar_expect_tx(BAR_FULL(stage), STAGE_BYTES);
                        tma_load_2d(sbase + stage * STAGE_BYTES, &map_q, BAR_FULL(stage), kb * BK, mt * BM);
                        tma_load_2d(sbase + stage * STAGE_BYTES + A_BYTES, &map_b, BAR_FULL(stage), kb * BK, nt * BN);
                    } else {
                        // the even CTA's barrier collects both CTAs' bytes: it expects 2 x STAGE_BYTES, the odd CTA just arrives
                        if (rank == 0) mbar_expect_tx(BAR_FULL(stage), 2 * STAGE_BYTES);
                        else mbar_arrive_cluster(map_to_cta(BAR_FULL(stage), 0));
                        tma_load_2d_pair(sbase + stage * STAGE_BYTES, &map_q, BAR_FULL(stage), kb * BK, (mt * 2 + (int)rank) * BM);
                        tma_load_2d_pair(sbase + stage * STAGE_BYTES + A_BYTES, &map_b, BAR_FULL(stage), kb * BK, nt * BN + (int)rank * (BN / 2));
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        int stage = 0; uint32_t phase = 0; int as = 0; uint32_t aphase = 0;
        for (int t = unit; t < total; t += nunits) {
            if (lane == 0 && rank == 0) {
                mbar_wait(BAR_TEMPTY(as), aphase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
                for (int kb = 0; kb < g.k_blocks; kb++) {
                    mbar_wait(BAR_FULL(stage), phase);
                    tc_fence_after();
                    const uint32_t a0 = sbase + stage * STAGE_BYTES;
                    const uint64_t ad = make_kmajor_sw128_desc(a0), bd = make_kmajor_sw128_desc(a0 + A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UK; k++) {         // +32 bytes along K inside the swizzle atom = +2 in the address field
                        if (CG == 2) umma_bf16_pair(d_tmem, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), IDesc<2>::value, (uint32_t)((kb | k) != 0));
                        else umma_bf16(d_tmem, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), IDesc<1>::value, (uint32_t)((kb | k) != 0));
                    }
                    // frees the stage (in both CTAs of a pair) once these MMAs have read it
                    if (CG == 2) umma_commit_pair(BAR_EMPTY(stage)); else umma_commit(BAR_EMPTY(stage));
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
                if (CG == 2) umma_commit_pair(BAR_TFULL(as)); else umma_commit(BAR_TFULL(as));    // accumulator complete -> epilogue
            }
            __syncwarp();
            as ^= 1; if (as == 0) aphase ^= 1u;
        }
    } else {
        // ===== epilogue: 8 warps; warp w may read TMEM lanes [32*(w&3), +32) = 32 queries, and takes one half of the columns.
        // Two warps per scheduler hide each other's latencies; the filter is branch-free in the common case (a running
        // minimum over 32 scores, one compare), and the next tile's a[] / tau are prefetched while this one is processed.
        const int eq = warp & 3;
        const int half = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;
        int as = 0; uint32_t aphase = 0;
        const float INF = __int_as_float(0x7f800000);
        const float bc = g.bconst;
        uint2* wbuf = s_wbuf + (warp - 2) * WBUF;
        volatile uint32_t* wcntv = s_wcnt + (warp - 2);
        uint32_t* wcnt = const_cast<uint32_t*>(wcntv);
        if (lane == 0) *wcnt = 0u;
        __syncwarp();
        float a_next = INF, tau_next = -INF;
        const uint32_t tempty_remote0 = CG == 2 ? map_to_cta(BAR_TEMPTY(0), 0) : 0u, tempty_remote1 = CG == 2 ? map_to_cta(BAR_TEMPTY(1), 0) : 0u;
        if (unit < total) {
            const int nt = unit / g.m_tiles, mt = unit - nt * g.m_tiles;
            const int64_t j = (int64_t)nt * BN + et;
            if (j < g.n_rows) a_next = __ldg(g.arow + j * g.row_step);
            const int q = (mt * CG + (int)rank) * BM + eq * 32 + lane;
            if (FILTER && q < g.nq) tau_next = g.tau[q];
        }
        for (int t = unit; t < total; t += nunits) {
            const int nt = t / g.m_tiles, mt = t - nt * g.m_tiles;
            const int qbase = (mt * CG + (int)rank) * BM + eq * 32;     // first query of this warp's 32 TMEM lanes
            float* sa = s_a + as * BN;
            sa[et] = a_next;
            const float tau = tau_next;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            {
                const int tn = t + nunits;
                a_next = INF; tau_next = -INF;
                if (tn < total) {
                    const int nt2 = tn / g.m_tiles, mt2 = tn - nt2 * g.m_tiles;
                    const int64_t j = (int64_t)nt2 * BN + et;
                    if (j < g.n_rows) a_next = __ldg(g.arow + j * g.row_step);
                    const int q2 = (mt2 * CG + (int)rank) * BM + eq * 32 + lane;
                    if (FILTER && q2 < g.nq) tau_next = g.tau[q2];
                }
            }
            const int q = qbase + lane;
            const bool qok = q < g.nq;
            mbar_wait(BAR_TFULL(as), aphase);
            tc_fence_after();
            const uint32_t trow = tmem_base + ((uint32_t)(eq * 32) << 16) + (uint32_t)(as * BN + half * 128);
#pragma unroll 1
            for (int c0 = 0; c0 < 128; c0 += 64) {
                uint32_t r[2][32];
                tmem_ld32(trow + c0, r[0]);
                tmem_ld32(trow + c0 + 32, r[1]);
                tmem_ld_wait();
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int cb = half * 128 + c0 + h * 32;           // first column of this group inside the tile
                    const float4* sa4 = reinterpret_cast<const float4*>(sa + cb);
                    if (FILTER) {
                        // survivors are rare per thread but not per warp: keep the common path branch-free. Each score's
                        // "below tau" sign bit is funnel-shifted into a per-thread mask (column i -> bit i); the rare set
                        // bits are then extracted with a select tree (registers cannot be indexed dynamically).
                        float sc[32];
                        uint32_t mask = 0u;
#pragma unroll
                        for (int i4 = 28; i4 >= 0; i4 -= 4) {
                            const float4 av = sa4[i4 >> 2];
                            sc[i4 + 3] = fmaf(bc, __uint_as_float(r[h][i4 + 3]), av.w); mask = __funnelshift_l(__float_as_uint(sc[i4 + 3] - tau), mask, 1);
                            sc[i4 + 2] = fmaf(bc, __uint_as_float(r[h][i4 + 2]), av.z); mask = __funnelshift_l(__float_as_uint(sc[i4 + 2] - tau), mask, 1);
                            sc[i4 + 1] = fmaf(bc, __uint_as_float(r[h][i4 + 1]), av.y); mask = __funnelshift_l(__float_as_uint(sc[i4 + 1] - tau), mask, 1);
                            sc[i4] = fmaf(bc, __uint_as_float(r[h][i4]), av.x);         mask = __funnelshift_l(__float_as_uint(sc[i4] - tau), mask, 1);
                        }
                        if (!qok) mask = 0u;                    // padded query rows never append (NaN scores included)
                        while (mask) {
                            const int i = 31 - __clz(mask);
                            mask &= ~(1u << i);
                            const float v = pick32(sc, i);
                            // stage in the warp's shared buffer; flushed with all global atomics in flight at once
                            const uint32_t pos = atomicAdd(wcnt, 1u);
                            if (pos < WBUF) wbuf[pos] = make_uint2(f2ord(v), ((uint32_t)lane << 8) | (uint32_t)(cb + i));
                            else {
                                const uint32_t gp = atomicAdd(g.cnt + q, 1u);
                                if (gp < g.cap) g.cand[(size_t)q * g.cap + gp] = ((uint64_t)f2ord(v) << 32) | (uint32_t)(nt * BN + cb + i);
                            }
                        }
                    } else if (qok) {
                        float4* dst = reinterpret_cast<float4*>(g.scores + (size_t)q * g.ld_scores + (size_t)nt * BN + cb);
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            const float4 av = sa4[i >> 2];
                            float4 o;
                            o.x = fmaf(bc, __uint_as_float(r[h][i]), av.x);
                            o.y = fmaf(bc, __uint_as_float(r[h][i + 1]), av.y);
                            o.z = fmaf(bc, __uint_as_float(r[h][i + 2]), av.z);
                            o.w = fmaf(bc, __uint_as_float(r[h][i + 3]), av.w);
                            dst[i >> 2] = o;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {                                 // TMEM stage drained: the MMA warp (of the pair's even CTA) may overwrite it
                if (CG == 2) mbar_arrive_cluster(as ? tempty_remote1 : tempty_remote0); else mbar_arrive(BAR_TEMPTY(as));
            }
            if (FILTER) {
                const uint32_t nb = min(*wcnt, (uint32_t)WBUF);
                for (uint32_t e = lane; e < nb; e += 32) {
                    const uint2 w = wbuf[e];
                    const int qq = qbase + (int)(w.y >> 8);
                    const uint32_t gp = atomicAdd(g.cnt + qq, 1u);
                    if (gp < g.cap) g.cand[(size_t)qq * g.cap + gp] = ((uint64_t)w.x << 32) | (uint32_t)(nt * BN + (int)(w.y & 255u));
                }
                __syncwarp();
                if (lane == 0) *wcnt = 0u;
                __syncwarp();
            }
            as ^= 1; if (as == 0) aphase ^= 1u;
        }
    }
    tc_fence_before();
    if (CG == 2) cluster_sync(); else __syncthreads();      // nobody leaves while the peer may still signal its barriers
    if (warp == 1) { if (CG == 2) tmem_dealloc_pair(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS); }
#undef BAR_FULL
#undef BAR_EMPTY
#undef BAR_TFULL
#undef BAR_TEMPTY
}


// ---------------------------------------------------------------------------------------------------------------
// pass 0: bf16 mirrors
// ---------------------------------------------------------------------------------------------------------------
__global__ void column_sum_kernel(const float* __restrict__ base, int64_t n, int dim, int64_t rows_per_block, float* __restrict__ sum) {
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block, r1 = min(n, r0 + rows_per_block);
    for (int d = threadIdx.x; d < dim; d += blockDim.x) {
        float s = 0.0f;
        for (int64_t r = r0; r < r1; r++) s += base[r * dim + d];
        atomicAdd(sum + d, s);
    }
}

// one warp per row: out[row][0..dpad) = bf16(x - mu) (sq-L2) or bf16(x) (cosine, rows already normalised);
// arow[row] = ||x-mu||^2 or 0 (score = arow + b*acc with b = -2 / -1); rows in [n, n_pad) get +inf so padded columns never pass.
__global__ void base_mirror_kernel(const float* __restrict__ base, int64_t n, int64_t n_pad, int dim, int dpad, int metric,
                                   const float* __restrict__ colsum, float inv_n, __nv_bfloat16* __restrict__ out,
                                   float* __restrict__ arow, uint32_t* __restrict__ amax_bits) {
    const int lane = threadIdx.x & 31;
    const int64_t w0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, wstep = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = w0; row < n_pad; row += wstep) {
        if (row >= n) { if (lane == 0) arow[row] = __int_as_float(0x7f800000); continue; }
        const float* x = base + row * dim;
        float acc = 0.0f;
        for (int c = lane * 2; c < dpad; c += 64) {
            float v0 = 0.0f, v1 = 0.0f;
            if (c < dim) v0 = x[c] - (metric == 0 ? colsum[c] * inv_n : 0.0f);
            if (c + 1 < dim) v1 = x[c + 1] - (metric == 0 ? colsum[c + 1] * inv_n : 0.0f);
            acc = fmaf(v0, v0, acc); acc = fmaf(v1, v1, acc);
            *reinterpret_cast<__nv_bfloat162*>(out + row * dpad + c) = __floats2bfloat162_rn(v0, v1);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) {
            arow[row] = metric == 0 ? acc : 0.0f;
            atomicMax(amax_bits, __float_as_uint(acc));
        }
    }
}

// one CTA per query. sq-L2: q' = q - mu. cosine: q^ = q / ||q|| the Java way (HnswIndex.java:344, Vector.java:93-116),
// kept in fp32 for the exact re-score as well. qstat = (||q'||^2, Java norm of q^).
__global__ void query_mirror_kernel(const float* __restrict__ queries, int dim, int dpad, int metric, const float* __restrict__ colsum,
                                    float inv_n, __nv_bfloat16* __restrict__ out, float* __restrict__ qf, float2* __restrict__ qstat) {
    extern __shared__ float sq[];
    __shared__ float s_nrm, s_red[32];
    const int q = blockIdx.x;
    for (int i = threadIdx.x; i < dim; i += blockDim.x) sq[i] = queries[(size_t)q * dim + i];
    __syncthreads();
    float qn = 0.0f;
    if (metric == 1) {
        if (threadIdx.x == 0) s_nrm = norm_java(sq, dim);
        __syncthreads();
        const float nrm = s_nrm;
        __syncthreads();
        for (int i = threadIdx.x; i < dim; i += blockDim.x) sq[i] = __fdiv_rn(sq[i], nrm);
        __syncthreads();
        if (threadIdx.x == 0) s_nrm = norm_java(sq, dim);
        __syncthreads();
        qn = s_nrm;
        for (int i = threadIdx.x; i < dim; i += blockDim.x) qf[(size_t)q * dim + i] = sq[i];
    }
    float acc = 0.0f;
    for (int c = threadIdx.x; c < dpad; c += blockDim.x) {
        float v = 0.0f;
        if (c < dim) v = metric == 0 ? sq[c] - colsum[c] * inv_n : sq[c];
        acc = fmaf(v, v, acc);
        out[(size_t)q * dpad + c] = __float2bfloat16_rn(v);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += s_red[w];
        qstat[q] = make_float2(t, qn);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// pass 1b: per-query threshold = the (s+1)-th smallest sample score (s-th index of the sorted list)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void block_bitonic_sort_u32(uint32_t* keys, int n) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int i = threadIdx.x; i < (n >> 1); i += blockDim.x) {
                const int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1)), hi = lo | j;
                const uint32_t a = keys[lo], b = keys[hi];
                const bool up = ((lo & k) == 0);
                if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
            }
        }
    }
    __syncthreads();
}
constexpr int TS_NT = 512, TS_PER = 32, TS_LCAP = 2048;
// Fast path (s < 511): the s+1 smallest per-thread minima are s+1 distinct sample scores, so the (s+1)-th smallest score
// overall is <= their largest ("pivot"); only the few scores <= pivot are collected and sorted. Otherwise: full sort.
__global__ void __launch_bounds__(TS_NT)
threshold_select_kernel(const float* __restrict__ scores, int64_t ld, int S, int P, int s, float* __restrict__ tau,
                        uint32_t* __restrict__ cnt) {
    extern __shared__ uint32_t skeys[];
    __shared__ uint32_t s_mins[TS_NT];
    __shared__ uint32_t s_n;
    const int q = blockIdx.x;
    if (threadIdx.x == 0) { cnt[q] = 0u; s_n = 0u; }
    if (s >= S) { if (threadIdx.x == 0) tau[q] = __int_as_float(0x7f800000); return; }
    uint32_t key[TS_PER];
    uint32_t mn = 0xffffffffu;
#pragma unroll
    for (int j = 0; j < TS_PER; j++) {
        const int i = threadIdx.x + j * TS_NT;
        key[j] = i < S ? f2ord(__ldg(scores + (size_t)q * ld + i)) : 0xffffffffu;
        mn = min(mn, key[j]);
    }
    if (s < TS_NT - 1) {
        // pivot without a block-wide sort: every warp sorts its 32 per-thread minima in registers (15 shuffle stages); with
        // j = ceil((s + 1) / 16), at least 16 j >= s + 1 scores lie at or below the LARGEST of the 16 warps' j-th smallest minima.
        uint32_t v = mn;
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
            for (int j = k >> 1; j > 0; j >>= 1) {
                const uint32_t o = __shfl_xor_sync(0xffffffffu, v, j);
                const bool keepmin = ((lane & j) == 0) == ((lane & k) == 0);
                v = keepmin ? min(v, o) : max(v, o);
            }
        }
        const int jth = (s + 16) / 16;                                 // ceil((s + 1) / 16), <= 32
        const uint32_t wj = __shfl_sync(0xffffffffu, v, jth - 1);
        if (lane == 0) s_mins[threadIdx.x >> 5] = wj;
        __syncthreads();
        uint32_t pivot = 0u;
#pragma unroll
        for (int w = 0; w < TS_NT / 32; w++) pivot = max(pivot, s_mins[w]);
#pragma unroll
        for (int j = 0; j < TS_PER; j++)
            if (key[j] <= pivot) {
                const uint32_t pos = atomicAdd(&s_n, 1u);
                if (pos < TS_LCAP) skeys[pos] = key[j];
            }
        __syncthreads();
        const int n = (int)s_n;
        if (n <= TS_LCAP) {
            // the s-th (0-based) smallest of the n collected scores: every key counts the keys below it (ties: lower slot first)
            for (int i = threadIdx.x; i < n; i += TS_NT) {
                const uint32_t me = skeys[i];
                int r = 0;
                for (int t = 0; t < n; t++) { const uint32_t o = skeys[t]; r += (o < me || (o == me && t < i)) ? 1 : 0; }
                if (r == s) tau[q] = ord2f(me);
            }
            return;
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < TS_PER; j++) {
        const int i = threadIdx.x + j * TS_NT;
        if (i < P) skeys[i] = key[j];
    }
    block_bitonic_sort_u32(skeys, P);
    if (threadIdx.x == 0) tau[q] = ord2f(skeys[s]);
}

// ---------------------------------------------------------------------------------------------------------------
// pass 3: shortlist, exact re-score in the reference order, certification
// ---------------------------------------------------------------------------------------------------------------
struct ShortlistArgs {
    const uint64_t* cand; const uint32_t* cnt; uint32_t cap; int capP;
    const float* tau; const float2* qstat; const uint32_t* amax_bits;
    const float* base; const float* norms; const int32_t* ids; int id_offset; int64_t n; int dim;
    const float* queries;        // fp32 queries used by the exact distance (normalised copy for cosine)
    int metric, order, k, k2, k2p;
    int tile_off;                // byte offset of sel[] (after the survivor arrays / transpose tiles, which share space)
    float err_rel;               // bound on |acc - q'.x'| / (||q'|| ||x'||)
    int32_t* out_ids; float* out_d; int* flag; unsigned long long* stats;   // stats[0] certified, [1] flagged
    int flag_or;                 // later slabs: OR into flag[] and leave the per-query statistics to the first slab
};

__global__ void __launch_bounds__(256, 3)
flat_shortlist_kernel(const ShortlistArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // [hk | rw] survivor scores / rows (dead after the selection, then reused for the per-warp transpose tiles),
    // sel[k2] selected rows, rkey[k2p] exact keys, sq[dim] the query
    uint32_t* hk = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* rw = hk + a.cap;
    float* tiles = reinterpret_cast<float*>(smem_raw);
    uint32_t* sel = reinterpret_cast<uint32_t*>(smem_raw + a.tile_off);
    uint64_t* rkey = reinterpret_cast<uint64_t*>(sel + a.k2);
    float* sq = reinterpret_cast<float*>(rkey + a.k2p);
    __shared__ uint32_t s_cnt0[32], s_nsel, s_ntie;
    __shared__ int s_ok;
    const int q = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cnt = a.cnt[q];
    const int c = (int)min(cnt, a.cap);
    for (int i = threadIdx.x; i < a.dim; i += blockDim.x) sq[i] = a.queries[(size_t)q * a.dim + i];
    const float qn = a.qstat[q].y;
    float* tile = tiles + warp * RS_TILE;
    // Two attempts: the exact re-score is the expensive part (k2 random rows from HBM per query), and for most queries the
    // certificate already holds with the best k2/2 survivors; only if it does not, the full k2 are re-scored.
    const int k2a = (a.k2 >= 256 && a.k2 / 2 >= a.k + 64) ? a.k2 / 2 : a.k2;
    bool ok = false;
    for (int attempt = 0; attempt < 2 && !ok; attempt++) {
        const int kk = attempt == 0 ? k2a : a.k2;
        if (attempt == 1 && (kk == k2a || c <= k2a)) break;      // nothing more to add
        int kp = 32;
        while (kp < kk) kp <<= 1;
        for (int i = threadIdx.x; i < c; i += blockDim.x) {
            const uint64_t key = a.cand[(size_t)q * a.cap + i];
            hk[i] = (uint32_t)(key >> 32); rw[i] = (uint32_t)key;
        }
        if (threadIdx.x < 32) s_cnt0[threadIdx.x] = 0u;
        if (threadIdx.x == 0) { s_nsel = 0u; s_ntie = 0u; }
        __syncthreads();
        const int R = min(c, kk);
        // ---- the kk best survivors by approximate score: radix bisection for the (kk+1)-th smallest ordered score T (its
        //      value is also the bound s_cut for every row left out), then one compaction pass. No sort of the survivors.
        uint32_t T = 0xffffffffu;
        if (c > kk) {
            uint32_t prefix = 0u;
            int r = kk;                                         // 0-based rank looked for
            for (int b = 31; b >= 0; b--) {
                uint32_t local = 0u;
                for (int i = threadIdx.x; i < c; i += blockDim.x) local += (((hk[i] ^ prefix) >> b) == 0u) ? 1u : 0u;
                local = __reduce_add_sync(0xffffffffu, local);
                if (lane == 0 && local) atomicAdd(&s_cnt0[b], local);
                __syncthreads();
                const int cnt0 = (int)s_cnt0[b];                // elements matching the prefix with bit b clear
                if (r >= cnt0) { prefix |= 1u << b; r -= cnt0; }
            }
            T = prefix;
            const int need_ties = r;                            // how many survivors equal to T still belong to the best kk
            for (int i0 = 0; i0 < c; i0 += blockDim.x) {
                const int i = i0 + threadIdx.x;
                const uint32_t h = i < c ? hk[i] : 0xffffffffu;
                const bool lt = i < c && h < T;
                const uint32_t bal = __ballot_sync(0xffffffffu, lt);
                uint32_t base = 0u;
                if (lane == 0 && bal) base = atomicAdd(&s_nsel, (uint32_t)__popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (lt) sel[base + __popc(bal & ((1u << lane) - 1u))] = rw[i];
                else if (i < c && h == T) {
                    const uint32_t p = atomicAdd(&s_ntie, 1u);
                    if ((int)p < need_ties) sel[kk - 1 - (int)p] = rw[i];
                }
            }
        } else {
            for (int i = threadIdx.x; i < c; i += blockDim.x) sel[i] = rw[i];
        }
        __syncthreads();                                        // hk / rw are dead from here on: the tiles reuse their space
        for (int g = warp; g * 32 < kp; g += 8) {
            const int j = g * 32 + lane;
            if (g * 32 >= R) { rkey[j] = KEY_MAX; continue; }
            const uint32_t row = j < R ? sel[j] : sel[0];
            float d;
            if (a.metric == 1) d = cosine_dist_from_dot(rescore32<2>(a.base, a.dim, sq, row, tile, lane), a.norms[row], qn);
            else if (a.order == ORDER_R1) d = rescore32<1>(a.base, a.dim, sq, row, tile, lane);
            else d = rescore32<0>(a.base, a.dim, sq, row, tile, lane);
            rkey[j] = j < R ? make_key(d, a.ids ? a.ids[row] : (int32_t)(a.id_offset + row)) : KEY_MAX;
        }
        block_bitonic_sort<false>(rkey, nullptr, kp);
        if (threadIdx.x == 0) {
            bool good;
            if ((int64_t)cnt == a.n && c <= kk && cnt <= a.cap) {
                good = true;                                     // every row was re-scored
            } else if (cnt > a.cap || R < a.k) {
                good = false;                                    // list overflow, or too few survivors to fill k
            } else {
                // rows not re-scored have approximate score >= s_cut
                const double s_cut = c > kk ? (double)ord2f(T) : (double)a.tau[q];
                const double dk = (double)key_dist(rkey[a.k - 1]);
                const double amax = (double)__uint_as_float(*a.amax_bits), qn2 = (double)a.qstat[q].x;
                const double gamma = (double)a.dim * 2.384185791015625e-07;          // dim * 2^-22: fp32 summation slack
                const double ecs = (double)a.err_rel * sqrt(qn2 * amax);               // |acc - q'.x'| for any row
                if (a.metric == 0) {
                    const double lb = s_cut + qn2 - 2.0 * ecs - (amax + qn2 + fabs(s_cut)) * gamma;
                    good = lb > 0.0 && dk * (1.0 + gamma) < lb * (1.0 - gamma);
                } else {
                    const double u = -s_cut + ecs;                                     // upper bound on dot(x, q^) of an outsider
                    const double lb = 1.0 - (u >= 0.0 ? u * (1.0 + 1e-5) : u * (1.0 - 1e-5)) - gamma - 1e-5;
                    good = dk + gamma + 1e-5 < lb;
                }
            }
            s_ok = good ? 1 : 0;
        }
        __syncthreads();
        ok = s_ok != 0;
        if (ok || attempt == 1 || kk == a.k2 || c <= k2a) {      // final attempt for this query: emit its list
            for (int j = threadIdx.x; j < a.k; j += blockDim.x) {
                const uint64_t key = j < kp ? rkey[j] : KEY_MAX;
                if (key == KEY_MAX) { a.out_ids[(size_t)q * a.k + j] = -1; a.out_d[(size_t)q * a.k + j] = __int_as_float(0x7f800000); }
                else { a.out_ids[(size_t)q * a.k + j] = key_id(key); a.out_d[(size_t)q * a.k + j] = key_dist(key); }
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (a.flag_or) {
            if (!ok && !a.flag[q]) { a.flag[q] = 1; atomicAdd(a.stats + 1, 1ull); atomicAdd(a.stats, ~0ull); }   // certified -> fallback
        } else {
            a.flag[q] = ok ? 0 : 1;
            atomicAdd(a.stats + (ok ? 0 : 1), 1ull);
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
// 2-D bf16 tensor [rows][dpad], row pitch `pitch_bytes`, box = 64 elements (128 bytes) x box_rows, 128-byte swizzle
static int make_map(CUtensorMap* m, const void* ptr, uint64_t rows, int dpad, uint64_t pitch_bytes, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return 1;
    cuuint64_t dims[2] = {(cuuint64_t)dpad, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : 2;
}

static int round_up_i(int64_t v, int m) { return (int)(((v + m - 1) / m) * m); }

FlatGemmPlan flat_gemm_plan(int64_t n, int dim, int k) {
    FlatGemmPlan p;
    p.dpad = round_up_i(dim, BK);
    p.n_pad = ((n + BN - 1) / BN) * BN;
    p.k2 = std::min(2048, round_up_i(std::max(4 * k, k + 128), 128));
    p.k2p = next_pow2(p.k2);
    // sample of the threshold pass: 1 row in 32, between 4 096 and 16 384 rows — a fixed 16 384-row sample is 13 % of a 125 k-row
    // shard (1 M rows over 8 GPUs), and the sample GEMM + threshold selection do not shrink with the shard otherwise
    const int64_t s_target = std::min<int64_t>(16384, std::max<int64_t>(4096, n / 32));
    p.step = std::max<int64_t>(1, (n + s_target - 1) / s_target);
    p.S = (int)((n + p.step - 1) / p.step);
    p.S_pad = round_up_i(p.S, BN);
    p.P = next_pow2(p.S);
    // expected survivors s*step; keep >= 3*k2 so the shortlist is full with overwhelming probability
    p.s = (int)std::min<int64_t>(p.S, std::max<int64_t>(32, (3 * (int64_t)p.k2 + p.step - 1) / p.step));
    const int64_t expect = (int64_t)p.s * p.step;
    p.cap = (uint32_t)std::min<int64_t>(n, (int64_t)round_up_i(3 * expect + 1024, 1024));
    p.capP = next_pow2((int)p.cap);
    return p;
}

// The survivor list of one query must fit one CTA's shared memory in pass 3; with a 16 384-row sample that holds up to
// ~2.6 M rows at k <= 256. Larger indexes are searched slab by slab (each slab its own thresholds, shortlist and certificate)
// and the per-slab top-k lists are merged by (dist,id).
static bool plan_fits(int64_t n, int dim, int k) {
    const FlatGemmPlan p = flat_gemm_plan(n, dim, k);
    return p.capP <= 16384 && p.k2 <= (int)p.cap;
}
int flat_gemm_supported(int64_t n, int dim, int k) {
    return n >= 4096 && n < ((int64_t)1 << 31) && dim >= 16 && dim <= 8192 && k >= 1 && k <= 256;
}
int64_t flat_gemm_slab_rows(int64_t n, int dim, int k) {
    int64_t limit = 2000000;
    if (const char* e = getenv("VDB_FLAT_SLAB_ROWS")) limit = std::max<int64_t>(4096, atoll(e));     // tests: force several slabs
    if (n <= limit && plan_fits(n, dim, k)) return n;
    int64_t slabs = (n + limit - 1) / limit;
    for (;; slabs++) {
        const int64_t rows = (((n + slabs - 1) / slabs) + 255) / 256 * 256;
        if (rows <= 4096 || plan_fits(rows, dim, k)) return rows;
    }
}

int launch_flat_mirror(const float* d_base, int64_t n, int dim, int metric, void* d_mirror, float* d_arow, float* d_colsum, cudaStream_t st) {
    const FlatGemmPlan p = flat_gemm_plan(n, dim, 1);
    cudaMemsetAsync(d_colsum, 0, (size_t)(dim + 4) * 4, st);
    if (metric == 0) {
        const int64_t rpb = 512;
        column_sum_kernel<<<(unsigned)((n + rpb - 1) / rpb), 256, 0, st>>>(d_base, n, dim, rpb, d_colsum);
    }
    base_mirror_kernel<<<148 * 8, 256, 0, st>>>(d_base, n, p.n_pad, dim, p.dpad, metric, d_colsum, 1.0f / (float)n,
                                                reinterpret_cast<__nv_bfloat16*>(d_mirror), d_arow, reinterpret_cast<uint32_t*>(d_colsum + dim));
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

template <bool FILTER, int CG>
static int run_gemm(const CUtensorMap& mq, const CUtensorMap& mb, const GemmArgs& g, cudaStream_t st) {
    // per launch: the attribute is per device and one process may drive several
    if (cudaFuncSetAttribute(flat_gemm_kernel<FILTER, CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM) != cudaSuccess) return 3;
    int sms = 0, dev = 0;
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
    const int total = g.m_tiles * g.n_tiles;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(CG * std::min(total, sms / CG)));
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = GEMM_SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, flat_gemm_kernel<FILTER, CG>, mq, mb, g) == cudaSuccess ? 0 : 4;
}

int launch_flat_gemm_search(const FlatGemmArgs& a, cudaStream_t st) {
    const FlatGemmPlan p = flat_gemm_plan(a.n, a.dim, a.k);
    const float inv_n = 1.0f / (float)(a.n_total > 0 ? a.n_total : a.n);     // the centring mean is over the whole index
    __nv_bfloat16* qm = reinterpret_cast<__nv_bfloat16*>(a.d_qmirror);
    query_mirror_kernel<<<a.nq, 128, (size_t)a.dim * 4, st>>>(a.d_queries, a.dim, p.dpad, a.metric, a.d_colsum, inv_n, qm, a.d_qf, a.d_qstat);
    // CTA pairs (cta_group::2) whenever there is more than one 128-query tile; VDB_FLAT_CG=1 forces single-CTA tiles
    const char* cg_env = getenv("VDB_FLAT_CG");
    const int CG = (a.nq > BM && !(cg_env && cg_env[0] == '1')) ? 2 : 1;
    CUtensorMap mq, mb_sample, mb_all;
    if (make_map(&mq, qm, (uint64_t)a.nq, p.dpad, (uint64_t)p.dpad * 2, BM)) return 10;
    if (make_map(&mb_sample, a.d_mirror, (uint64_t)p.S, p.dpad, (uint64_t)p.dpad * 2 * (uint64_t)p.step, BN / CG)) return 11;
    if (make_map(&mb_all, a.d_mirror, (uint64_t)a.n, p.dpad, (uint64_t)p.dpad * 2, BN / CG)) return 12;
    GemmArgs g{};
    g.m_tiles = (a.nq + BM * CG - 1) / (BM * CG); g.k_blocks = p.dpad / BK; g.nq = a.nq; g.arow = a.d_arow; g.bconst = a.metric == 0 ? -2.0f : -1.0f;
    // pass 1: sample scores -> tau
    g.n_tiles = p.S_pad / BN; g.n_rows = p.S; g.row_step = p.step; g.scores = a.d_scores; g.ld_scores = p.S_pad;
    int rc = CG == 2 ? run_gemm<false, 2>(mq, mb_sample, g, st) : run_gemm<false, 1>(mq, mb_sample, g, st);
    if (rc) return rc;
    const size_t tsm = (size_t)std::max(p.P, TS_LCAP) * 4;
    if (cudaFuncSetAttribute(threshold_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm) != cudaSuccess) return 15;
    threshold_select_kernel<<<a.nq, TS_NT, tsm, st>>>(a.d_scores, p.S_pad, p.S, p.P, p.s, a.d_tau, a.d_cnt);
    // pass 2: all rows, filter epilogue
    g.n_tiles = (int)(p.n_pad / BN); g.n_rows = a.n; g.row_step = 1; g.scores = nullptr;
    g.tau = a.d_tau; g.cnt = a.d_cnt; g.cand = a.d_cand; g.cap = p.cap;
    rc = CG == 2 ? run_gemm<true, 2>(mq, mb_all, g, st) : run_gemm<true, 1>(mq, mb_all, g, st);
    if (rc) return rc;
    // pass 3
    ShortlistArgs s{};
    s.cand = a.d_cand; s.cnt = a.d_cnt; s.cap = p.cap; s.capP = p.capP; s.tau = a.d_tau; s.qstat = a.d_qstat;
    s.amax_bits = reinterpret_cast<const uint32_t*>(a.d_colsum + a.dim);
    s.base = a.d_base; s.norms = a.d_norms; s.ids = a.d_ids; s.id_offset = a.id_offset; s.n = a.n; s.dim = a.dim;
    s.queries = a.metric == 1 ? a.d_qf : a.d_queries;
    s.metric = a.metric; s.order = a.order; s.k = a.k; s.k2 = p.k2; s.k2p = p.k2p;
    s.err_rel = (float)((0.00390625 * (1.0 + 0.001953125) + (double)a.dim * 2.384185791015625e-07) * 1.02);
    s.out_ids = a.d_out_ids; s.out_d = a.d_out_d; s.flag = a.d_flag; s.stats = a.d_stats; s.flag_or = a.flag_or;
    const size_t region = std::max((size_t)p.cap * 8, (size_t)8 * RS_TILE * 4);
    s.tile_off = (int)((region + 15) & ~(size_t)15);              // byte offset of sel[] (tiles alias the survivor arrays)
    const size_t smem = (size_t)s.tile_off + (size_t)p.k2 * 4 + (size_t)p.k2p * 8 + (size_t)((a.dim + 3) & ~3) * 4 + 16;
    if (cudaFuncSetAttribute(flat_shortlist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 13;
    flat_shortlist_kernel<<<a.nq, 256, smem, st>>>(s);
    return cudaGetLastError() == cudaSuccess ? 0 : 14;
}
