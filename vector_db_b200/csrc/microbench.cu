// microbench.cu — MEASURED ceilings for the two gather-bound kernels of the path, so that their rooflines are quoted against
// what the hardware path they actually use can deliver, not against the HBM copy peak (the working sets live in L2 / shared
// memory: see DESIGN.md §4.1, §4.2 and the round-1 ncu summaries). Both are stripped copies of the kernels' inner loads with
// all bookkeeping removed; bench.py runs them live and reports roofline.frac against them with `peak_source`.
//
//   kind 0  row gather   : random rows of `row_bytes` (1024 = 512-d bf16) from a table of `rows` rows (100 k rows = 102 MB, mostly
//                          L2-resident like the traversal copy), read exactly like hnsw_search_kernel's distance phase reads them:
//                          a warp takes 16 rows, lane (g = lane / 4, c = lane % 4) issues 32-byte ld.global.nc.L1::no_allocate
//                          loads (LDG.E.256) of rows g and g + 8 at byte c * 32 of every 128-byte line, two lines in flight per
//                          row, 256 threads per CTA, 2 CTAs per SM (= 16 warps per SM). No MMA, no candidate lists: the loaded
//                          words are XOR-folded so the loads cannot be dropped.  units = bytes gathered.
//   kind 1  row gather, deep: same addresses with 8 lines per row in flight per lane (the load-level parallelism a register ring
//                          would give if registers were free) — the upper bound of the L2 -> SM path for this access pattern.
//   kind 4  row gather, one CTA per SM: kind 0 with 8 warps per SM instead of 16 — what ONE query's distance phase can have in flight.
//   kind 2  LDS gather   : the ADC scan's shared-memory gather at its floor: 512 threads, one CTA per SM, a 128 KB LUT image in
//                          the scan's layout offset(m, c) = (m >> 6) * 64 KB + c * 256 + (m & 63) * 4 (conflict-free for any
//                          codes), address = one PRMT of (code word, lane constant), one LDS and one FADD per gathered value; code
//                          words come from registers (an LCG), so no global traffic at all.  units = gathers.
//   kind 3  LDS gather, two queries per word: the ceiling of a T = 2 scan (bf16 pairs unpacked to fp32) — measured to decide
//                          against building it.  units = query-gathers.
#include <algorithm>
#include <vector>

#include "kernels.h"
#include "vdb_index.h"

namespace {
struct U8v { uint4 lo, hi; };
__device__ __forceinline__ U8v ld256(const unsigned char* p) {
    U8v r;
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t fold(const U8v& a) { return a.lo.x ^ a.lo.y ^ a.lo.z ^ a.lo.w ^ a.hi.x ^ a.hi.y ^ a.hi.z ^ a.hi.w; }

template <int NL>     // 128-byte lines of each row in flight per lane
__global__ void __launch_bounds__(256, 2) row_gather_kernel(const unsigned char* __restrict__ table, const int* __restrict__ rows, int nrows_per_cta,
                                                           int row_bytes, uint32_t* __restrict__ sink) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, c = lane & 3;
    const int* my = rows + (size_t)blockIdx.x * nrows_per_cta;
    const int nst = row_bytes / (NL * 128);
    uint32_t acc = 0;
    for (int tile = warp; tile * 16 < nrows_per_cta; tile += 8) {
        const int n0 = my[tile * 16 + g], n1 = my[tile * 16 + g + 8];
        const unsigned char* r0 = table + (size_t)n0 * row_bytes + c * 32;
        const unsigned char* r1 = table + (size_t)n1 * row_bytes + c * 32;
        for (int st = 0; st < nst; st++) {
            U8v A0[NL], A1[NL];
#pragma unroll
            for (int i = 0; i < NL; i++) { A0[i] = ld256(r0 + (st * NL + i) * 128); A1[i] = ld256(r1 + (st * NL + i) * 128); }
#pragma unroll
            for (int i = 0; i < NL; i++) acc ^= fold(A0[i]) ^ fold(A1[i]);
        }
    }
    if (acc == 0x9e3779b9u) sink[0] = acc;      // practically never: keeps the loads alive
}

__global__ void __launch_bounds__(512, 1) lds_gather_kernel(int iters, float* __restrict__ sink) {
    extern __shared__ float lut[];                       // 2 groups x 256 centroids x 64 subspaces = 128 KB
    for (int i = threadIdx.x; i < 2 * 16384; i += 512) lut[i] = (float)(i & 1023) * 1e-3f;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    // lane constant of the scan: byte offsets (m & 63) * 4 with m & 31 == lane -> bank == lane, conflict-free for any code
    const uint32_t lane_off = (uint32_t)lane * 4u;
    uint32_t w = 0x12345u * (threadIdx.x + 1) + blockIdx.x;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    const char* base = reinterpret_cast<const char*>(lut);
    for (int it = 0; it < iters; it++) {
        w = w * 1664525u + 1013904223u;                  // four code bytes
        // address = code * 256 + lane_off (+ group offset as an immediate): one PRMT builds (code << 8 | lane_off)
        const uint32_t o0 = __byte_perm(w, lane_off, 0x6504);                // byte j of w -> bits 8..15, lane_off -> bits 0..7, zeros above
        const uint32_t o1 = __byte_perm(w, lane_off, 0x6514);
        const uint32_t o2 = __byte_perm(w, lane_off, 0x6524);
        const uint32_t o3 = __byte_perm(w, lane_off, 0x6534);
        a0 += *reinterpret_cast<const float*>(base + o0);
        a1 += *reinterpret_cast<const float*>(base + o1 + 65536);
        a2 += *reinterpret_cast<const float*>(base + o2 + 128);
        a3 += *reinterpret_cast<const float*>(base + o3 + 65536 + 128);
    }
    if (a0 + a1 + a2 + a3 == -1.0f) sink[0] = a0;
}
// kind 3: the same gather where one 32-bit LUT word carries the bf16 entries of TWO queries (lo half = query A, hi half = query B):
// PRMT + LDS + 2 unpack ops + 2 FADD per word = two query-gathers. units = query-gathers.
__global__ void __launch_bounds__(512, 1) lds_gather_pair_kernel(int iters, float* __restrict__ sink) {
    extern __shared__ float lut[];
    for (int i = threadIdx.x; i < 2 * 16384; i += 512) lut[i] = __uint_as_float(0x3f803f80u + (uint32_t)(i & 63));
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t lane_off = (uint32_t)lane * 4u;
    uint32_t w = 0x12345u * (threadIdx.x + 1) + blockIdx.x;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
    const char* base = reinterpret_cast<const char*>(lut);
    for (int it = 0; it < iters; it++) {
        w = w * 1664525u + 1013904223u;
        const uint32_t w0 = *reinterpret_cast<const uint32_t*>(base + __byte_perm(w, lane_off, 0x6504));
        const uint32_t w1 = *reinterpret_cast<const uint32_t*>(base + __byte_perm(w, lane_off, 0x6514) + 65536);
        const uint32_t w2 = *reinterpret_cast<const uint32_t*>(base + __byte_perm(w, lane_off, 0x6524) + 128);
        const uint32_t w3 = *reinterpret_cast<const uint32_t*>(base + __byte_perm(w, lane_off, 0x6534) + 65536 + 128);
        a0 += __uint_as_float(w0 << 16); b0 += __uint_as_float(w0 & 0xffff0000u);
        a1 += __uint_as_float(w1 << 16); b1 += __uint_as_float(w1 & 0xffff0000u);
        a2 += __uint_as_float(w2 << 16); b2 += __uint_as_float(w2 & 0xffff0000u);
        a3 += __uint_as_float(w3 << 16); b3 += __uint_as_float(w3 & 0xffff0000u);
    }
    if (a0 + a1 + a2 + a3 + b0 + b1 + b2 + b3 == -1.0f) sink[0] = a0;
}
}  // namespace

// out[0] = milliseconds per launch (median of `reps`), out[1] = units per launch (bytes gathered / gathers)
extern "C" int vdb_microbench_device(int32_t kind, int64_t rows, int32_t row_bytes, int32_t work, int32_t reps, double* out) try {
    if (!out || reps <= 0 || reps > 100) { vdb_set_error("vdb_microbench_device: bad arguments"); return VDB_ERR_INVALID; }
    cudaDeviceProp prop; int dev = 0;
    VDB_CUDA(cudaGetDevice(&dev));
    VDB_CUDA(cudaGetDeviceProperties(&prop, dev));
    const int sms = prop.multiProcessorCount;
    cudaEvent_t e0, e1;
    VDB_CUDA(cudaEventCreate(&e0)); VDB_CUDA(cudaEventCreate(&e1));
    std::vector<float> ms;
    double units = 0;
    if (kind == 0 || kind == 1 || kind == 4) {            // kind 4 = kind 0 with ONE 8-warp CTA per SM (what one query's distance phase has)
        // kind 0 keeps 2 lines per row in flight (row_bytes % 256 == 0: 512-byte int8 rows included), kind 1 keeps 8 (or 4 for 512-byte rows)
        if (rows <= 0 || row_bytes <= 0 || (row_bytes % (kind != 1 ? 256 : 512)) != 0 || work <= 0) { vdb_set_error("row gather: rows > 0, row_bytes % 256 (kind 0) / % 512 (kind 1) == 0, work = rows gathered per CTA"); return VDB_ERR_INVALID; }
        const int grid = kind == 4 ? sms : sms * 2, per_cta = (work + 127) / 128 * 128;
        unsigned char* table = nullptr; int* idx = nullptr; uint32_t* sink = nullptr;
        VDB_CUDA(cudaMalloc(&table, (size_t)rows * row_bytes));
        VDB_CUDA(cudaMemset(table, 1, (size_t)rows * row_bytes));
        std::vector<int> h((size_t)grid * per_cta);
        uint64_t s = 88172645463325252ull;
        for (auto& v : h) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; v = (int)(s % (uint64_t)rows); }
        VDB_CUDA(cudaMalloc(&idx, h.size() * 4)); VDB_CUDA(cudaMalloc(&sink, 4));
        VDB_CUDA(cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
        for (int r = 0; r < reps + 2; r++) {
            cudaEventRecord(e0);
            if (kind == 0 || kind == 4) row_gather_kernel<2><<<grid, 256>>>(table, idx, per_cta, row_bytes, sink);
            else if (row_bytes % 1024 == 0) row_gather_kernel<8><<<grid, 256>>>(table, idx, per_cta, row_bytes, sink);
            else row_gather_kernel<4><<<grid, 256>>>(table, idx, per_cta, row_bytes, sink);
            cudaEventRecord(e1);
            VDB_CUDA(cudaEventSynchronize(e1));
            float t; cudaEventElapsedTime(&t, e0, e1);
            if (r >= 2) ms.push_back(t);
        }
        units = (double)grid * per_cta * row_bytes;
        cudaFree(table); cudaFree(idx); cudaFree(sink);
    } else if (kind == 3) {
        if (work <= 0) { vdb_set_error("lds pair gather: work = iterations per thread"); return VDB_ERR_INVALID; }
        float* sink = nullptr;
        VDB_CUDA(cudaMalloc(&sink, 4));
        VDB_CUDA(cudaFuncSetAttribute(lds_gather_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
        for (int r = 0; r < reps + 2; r++) {
            cudaEventRecord(e0);
            lds_gather_pair_kernel<<<sms, 512, 131072>>>(work, sink);
            cudaEventRecord(e1);
            VDB_CUDA(cudaEventSynchronize(e1));
            float t; cudaEventElapsedTime(&t, e0, e1);
            if (r >= 2) ms.push_back(t);
        }
        units = (double)sms * 512 * (double)work * 8.0;
        cudaFree(sink);
    } else if (kind == 2) {
        if (work <= 0) { vdb_set_error("lds gather: work = iterations per thread"); return VDB_ERR_INVALID; }
        float* sink = nullptr;
        VDB_CUDA(cudaMalloc(&sink, 4));
        VDB_CUDA(cudaFuncSetAttribute(lds_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
        for (int r = 0; r < reps + 2; r++) {
            cudaEventRecord(e0);
            lds_gather_kernel<<<sms, 512, 131072>>>(work, sink);
            cudaEventRecord(e1);
            VDB_CUDA(cudaEventSynchronize(e1));
            float t; cudaEventElapsedTime(&t, e0, e1);
            if (r >= 2) ms.push_back(t);
        }
        units = (double)sms * 512 * (double)work * 4.0;
        cudaFree(sink);
    } else { vdb_set_error("vdb_microbench_device: unknown kind"); return VDB_ERR_INVALID; }
    VDB_CUDA(cudaGetLastError());
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    std::sort(ms.begin(), ms.end());
    out[0] = ms[ms.size() / 2]; out[1] = units;
    return VDB_OK;
} catch (...) { return vdb_caught("vdb_microbench_device"); }
