// noalloc.cu — vdb_flat_search_noalloc_device: the order-exact flat scan (flat_kernels.cu) for
// callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). It has NO
// counterpart in the reference.
//
// The chunked scan needs scratch: per (query, chunk) a sorted list of k (dist,id) keys + row numbers, and the merged list per
// query. The index-bound entry points keep that in growable cudaMalloc'd buffers (vdb_index.h DevBuf); this one keeps it in ONE
// file-scope device array of fixed size, VDB_NOALLOC_SCRATCH_BYTES (8 MiB), which every GPU that loads the library holds until
// the process exits. Consequences, all documented in include/vdb.h:
//   * requests whose scratch would not fit are rejected with VDB_ERR_INVALID and a message that states the limit;
//   * every call shares the array and finds the previous call's contents in it. Nothing in it has to start at zero: each
//     (query, chunk) list is written in full by flat_exact_kernel before merge_keys reads it, and the merged list is written in
//     full before keys_to_results reads it — so nothing is cleared;
//   * two calls that overlap on the GPU (different streams, or two host threads) race on the array. One call at a time per GPU.
// The call does no cudaMalloc / cudaFree / cudaMallocHost and calls no allocating library routine.
#include <mutex>
#include <string>

#include "kernels.h"
#include "vdb_index.h"

#define VDB_NOALLOC_SCRATCH_BYTES (8u << 20)
#define VDB_NOALLOC_MAX_DEVICES 64

// In host code this name is only a host-side stand-in for the symbol, NOT a device address: its address must never be handed
// to a kernel. The device address is fetched once per device with cudaGetSymbolAddress (below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[VDB_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_mu;
unsigned char* g_scratch_addr[VDB_NOALLOC_MAX_DEVICES] = {nullptr};   // device address of g_noalloc_scratch, per device ordinal

// set-up, once per device, with that device current
int scratch_for_current_device(unsigned char** out) {
    int dev = -1;
    VDB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= VDB_NOALLOC_MAX_DEVICES) { vdb_set_error("vdb_flat_search_noalloc_device: device ordinal out of range"); return VDB_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        VDB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) { vdb_set_error("cudaGetSymbolAddress returned a null address for the scratch array"); return VDB_ERR_CUDA; }
        g_scratch_addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_addr[dev];
    return VDB_OK;
}
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
}  // namespace

extern "C" int vdb_flat_search_noalloc_device(const float* d_base, int64_t n, int32_t dim, int32_t order, const float* d_queries,
                                              int64_t nq, int32_t k, int32_t* d_out_ids, float* d_out_dists, void* stream) {
    if (!d_base || !d_queries || !d_out_ids || !d_out_dists || n <= 0 || dim <= 0 || nq < 0 || k <= 0) {
        vdb_set_error("vdb_flat_search_noalloc_device: bad arguments");
        return VDB_ERR_INVALID;
    }
    if (order != VDB_ARITH_JAVA_SEQ && order != VDB_ARITH_NATIVE_AVX2) { vdb_set_error("vdb_flat_search_noalloc_device: unknown reduction order"); return VDB_ERR_INVALID; }
    if (k > 1024 || dim > 8192 || n > 0x7fffffffLL || nq > 65535) {
        vdb_set_error("vdb_flat_search_noalloc_device: limits are k <= 1024, dim <= 8192, n < 2^31, nq <= 65535");
        return VDB_ERR_INVALID;
    }
    if (nq == 0) return VDB_OK;
    // chunks per query: about two CTAs per SM in flight, at least 1024 rows per chunk
    int nch = (int)((2 * 148 + nq - 1) / nq);
    if (nch > 64) nch = 64;
    while (nch > 1 && (n / nch < 1024 || (int64_t)nch * k > 8192)) nch--;       // merge_keys sorts nch * k keys in shared memory
    // scratch: keys_tmp[nq][nch][k] u64 | keys[nq][k] u64 | rows_tmp[nq][nch][k] u32 | rows[nq][k] u32, shrinking the chunk count
    // before giving up
    size_t o_keys = 0, o_rows_tmp = 0, o_rows = 0, total = 0;
    for (;; nch--) {
        const size_t lists = (size_t)nq * k;
        o_keys = align256(lists * nch * 8);
        o_rows_tmp = o_keys + align256(lists * 8);
        o_rows = o_rows_tmp + align256(lists * nch * 4);
        total = o_rows + align256(lists * 4);
        if (total <= VDB_NOALLOC_SCRATCH_BYTES || nch == 1) break;
    }
    if (total > VDB_NOALLOC_SCRATCH_BYTES) {
        vdb_set_error("vdb_flat_search_noalloc_device: nq * k = " + std::to_string((long long)nq * k) + " needs " + std::to_string(total) +
                      " bytes of scratch; the fixed scratch array holds " + std::to_string(VDB_NOALLOC_SCRATCH_BYTES) +
                      " bytes (nq * k <= ~340 000): split the batch or use vdb_search_batch_device");
        return VDB_ERR_INVALID;
    }
    unsigned char* s = nullptr;
    VDB_TRY(scratch_for_current_device(&s));
    cudaStream_t st = (cudaStream_t)stream;
    uint64_t* keys_tmp = reinterpret_cast<uint64_t*>(s);
    uint64_t* keys = reinterpret_cast<uint64_t*>(s + o_keys);
    uint32_t* rows_tmp = reinterpret_cast<uint32_t*>(s + o_rows_tmp);
    uint32_t* rows = reinterpret_cast<uint32_t*>(s + o_rows);
    launch_flat_search_exact(d_base, nullptr, nullptr, 0, n, dim, d_queries, (int)nq, k, VDB_METRIC_SQL2,
                             order == VDB_ARITH_NATIVE_AVX2 ? ORDER_R1 : ORDER_SEQ, nch, keys_tmp, rows_tmp, keys, rows, st, nullptr);
    launch_keys_to_results(keys, k, k, nq, d_out_ids, d_out_dists, st, nullptr);
    VDB_CUDA(cudaGetLastError());
    return VDB_OK;
}
