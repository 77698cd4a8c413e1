// multi_gpu.cu — one index over several GPUs of the box FROM ONE PROCESS (vdb_config.n_devices > 1). A JVM cannot
// run torchrun: the Java caller's single nativeSearchBatch (native/bridge/VectorDBJNI.cpp:184-200) has to reach all GPUs through one
// vdb_search_batch. The reference has no multi-device path at all; the rule that makes the answer well defined is its (dist,id)
// order (PQIndex.cpp:200-205).
//
// Layout: FLAT / PQ rows are split by contiguous id range, one single-device vdb_index ("shard") per GPU; codebooks replicated
// (trained on shard 0, copied: bit-identical). HNSW graphs do not shard by id range (a walk touches arbitrary nodes): rejected here,
// run replicas instead.
// Search (per sub-batch of <= 1024 queries), all copies device-to-device over NVLink (cudaMemcpyPeerAsync), ordering by events:
//   FLAT:  every shard searches its rows -> [qb][k] (dist,id) lists -> gathered on the front device -> merge_topk_kernel.
//   PQ:    phase 1 on every shard (LUT + ADC scan -> local top-R keys) -> keys gathered -> select_kth_key_kernel gives the GLOBAL
//          R-th key per query -> sent back -> phase 2 on every shard re-scores only the rows it owns among the global top-R ->
//          top-k lists gathered -> merge. The answer is therefore the single-GPU answer bit for bit, for any number of GPUs, and the
//          rerank row gathers (R x dim x 4 B per query) are divided among the GPUs instead of repeated on each.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "kernels.h"
#include "vdb_index.h"

#define MG_FAIL(code, msg) do { vdb_set_error(msg); return (code); } while (0)

int vdb_multi_create(vdb_index* f) {
    const int g = f->cfg.n_devices;
    int ndev = 0;
    VDB_CUDA(cudaGetDeviceCount(&ndev));
    if (f->cfg.index_type != VDB_INDEX_FLAT && f->cfg.index_type != VDB_INDEX_PQ)
        MG_FAIL(VDB_ERR_INVALID, "n_devices > 1: only FLAT and PQ indexes shard by id range (HNSW graphs: run one replica per GPU)");
    // VDB_MULTI_SAME_DEVICE=1 puts every shard on cfg.device: the same code path (peer copies degenerate to device copies), so the
    // sharded logic is testable on a one-GPU box
    const bool same = getenv("VDB_MULTI_SAME_DEVICE") != nullptr;
    if (!same && f->cfg.device + g > ndev) MG_FAIL(VDB_ERR_INVALID, "n_devices: not enough CUDA devices on this box");
    for (int s = 0; s < g; s++) {
        vdb_config c = f->cfg;
        c.n_devices = 1; c.device = same ? f->cfg.device : f->cfg.device + s;
        // capacity is a limit, not an allocation: every shard may take the whole index (rows added after the first search all go
        // to the last shard until the next vdb_build); the front index enforces max_elements
        vdb_index* sh = nullptr;
        VDB_TRY(vdb_create(&c, &sh));
        f->shards.push_back(sh);
        cudaEvent_t e;
        VDB_CUDA(cudaSetDevice(c.device));
        VDB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        f->shard_ev.push_back(e);
        for (int t = 0; t < g && !same; t++) if (t != s) { cudaDeviceEnablePeerAccess(f->cfg.device + t, 0); cudaGetLastError(); }   // already enabled / unsupported: staged copies still work
    }
    VDB_CUDA(cudaSetDevice(f->cfg.device));
    VDB_CUDA(cudaEventCreateWithFlags(&f->ev_front, cudaEventDisableTiming));
    return VDB_OK;
}
void vdb_multi_destroy(vdb_index* f) {
    for (size_t s = 0; s < f->shards.size(); s++) {
        cudaSetDevice(f->shards[s]->cfg.device);
        if (s < f->shard_ev.size()) cudaEventDestroy(f->shard_ev[s]);
        vdb_destroy(f->shards[s]);
    }
    f->shards.clear(); f->shard_ev.clear();
    cudaSetDevice(f->cfg.device);
    if (f->ev_front) cudaEventDestroy(f->ev_front);
    f->g_keys.release(); f->g_kth.release(); f->g_d.release(); f->g_i.release();
}

int vdb_multi_train(vdb_index* f, const float* samples, int64_t ns) {
    VDB_TRY(vdb_train(f->shards[0], samples, ns));
    std::vector<float> cb((size_t)f->pq_m * VDB_NCENT * f->sd);
    VDB_TRY(vdb_get_codebooks(f->shards[0], cb.data()));
    for (size_t s = 1; s < f->shards.size(); s++) VDB_TRY(vdb_set_codebooks(f->shards[s], cb.data()));
    f->trained = true;
    return VDB_OK;
}

// host rows added since the last call go to the shards. Rows are split by contiguous range over the CURRENT total, so a bulk load
// followed by searches is balanced; rows added later are appended to the last shard (rebuild = vdb_build re-splits).
int vdb_multi_sync(vdb_index* f) {
    const int g = (int)f->shards.size();
    if (f->device_only || f->n_distributed >= f->n) return VDB_OK;
    const int dim = f->cfg.dim;
    if (f->n_distributed == 0) {
        for (int s = 0; s < g; s++) {
            const int64_t a = f->n * s / g, b = f->n * (s + 1) / g;
            if (b > a) VDB_TRY(vdb_add_batch(f->shards[s], f->h_ids.data() + a, f->h_vecs.data() + (size_t)a * dim, b - a));
        }
    } else {
        const int64_t a = f->n_distributed;
        VDB_TRY(vdb_add_batch(f->shards[g - 1], f->h_ids.data() + a, f->h_vecs.data() + (size_t)a * dim, f->n - a));
    }
    f->n_distributed = f->n;
    for (int s = 0; s < g; s++) { std::lock_guard<std::mutex> lk(f->shards[s]->mu); VDB_TRY(vdb_sync_index_nolock(f->shards[s])); }
    return VDB_OK;
}
// vdb_build (rebuildIndex) of a host-backed multi-device index: the rows are split evenly again (rows that arrived after the first
// search had all gone to the last shard). The shards are recreated and refilled; trained codebooks carry over.
int vdb_multi_rebuild(vdb_index* f) {
    if (f->device_only) return vdb_multi_sync(f);
    std::vector<float> cb;
    const bool pq = f->cfg.index_type == VDB_INDEX_PQ;
    if (pq && f->trained) {
        cb.resize((size_t)f->pq_m * VDB_NCENT * f->sd);
        VDB_TRY(vdb_get_codebooks(f->shards[0], cb.data()));
    }
    vdb_multi_destroy(f);
    f->ev_front = nullptr;
    VDB_TRY(vdb_multi_create(f));
    if (!cb.empty()) for (auto* sh : f->shards) VDB_TRY(vdb_set_codebooks(sh, cb.data()));
    f->n_distributed = 0;
    return vdb_multi_sync(f);
}
int vdb_multi_fill_synthetic(vdb_index* f, uint64_t seed, int64_t cnt) {
    const int g = (int)f->shards.size();
    if (f->n != 0) MG_FAIL(VDB_ERR_INVALID, "multi-device synthetic fill: the index must be empty");
    for (int s = 0; s < g; s++) {
        const int64_t a = cnt * s / g, b = cnt * (s + 1) / g;
        f->shards[s]->cfg.id_offset = f->cfg.id_offset + (int32_t)a;      // global row = id: the generator is keyed on it
        if (b > a) VDB_TRY(vdb_fill_synthetic(f->shards[s], seed, b - a));
    }
    f->device_only = true; f->n = cnt; f->n_distributed = cnt;
    return VDB_OK;
}
int vdb_multi_encode_all(vdb_index* f) {
    for (auto* sh : f->shards) if (vdb_size(sh) > 0) VDB_TRY(vdb_encode_all(sh));
    return VDB_OK;
}
int vdb_multi_stats(vdb_index* f, vdb_stats* out) {
    vdb_stats t{};
    for (auto* sh : f->shards) {
        vdb_stats s{};
        VDB_TRY(vdb_get_stats(sh, &s));
        t.adc_rows += s.adc_rows; t.flat_certified += s.flat_certified; t.flat_fallback += s.flat_fallback;
        t.last_kernel_ms = std::max(t.last_kernel_ms, s.last_kernel_ms); t.last_total_ms = std::max(t.last_total_ms, s.last_total_ms);
    }
    t.queries = f->stats.queries; t.shortlist_retries = f->stats.shortlist_retries;
    *out = t;
    return VDB_OK;
}

// Run fn(shard) for every shard, each on its own host thread: the shards' search calls block the calling thread in places
// (certification read-backs), so issuing them from one thread would serialise the GPUs. Errors travel back with their message
// (vdb_last_error is per thread).
template <class F>
static int for_each_shard(vdb_index* f, F fn) {
    const int g = (int)f->shards.size();
    std::vector<int> rc(g, VDB_OK);
    std::vector<std::string> msg(g);
    auto run = [&](int s) {
        if (cudaSetDevice(f->shards[s]->cfg.device) != cudaSuccess) { rc[s] = VDB_ERR_CUDA; msg[s] = "cudaSetDevice failed"; return; }
        rc[s] = fn(s);
        if (rc[s] != VDB_OK) msg[s] = vdb_last_error();
    };
    std::vector<std::thread> th;
    for (int s = 1; s < g; s++) {
        try { th.emplace_back(run, s); } catch (...) { run(s); }       // no thread to be had: do it here (nothing may throw across the C ABI)
    }
    run(0);
    for (auto& t : th) t.join();
    cudaSetDevice(f->cfg.device);
    for (int s = 0; s < g; s++) if (rc[s] != VDB_OK) { vdb_set_error(msg[s]); return rc[s]; }
    return VDB_OK;
}

int vdb_multi_search(vdb_index* f, const float* q, bool q_on_device, int64_t nq, int k, int32_t* out_ids, float* out_d, cudaStream_t user_stream) {
    const int g = (int)f->shards.size(), dim = f->cfg.dim, d0 = f->cfg.device;
    VDB_TRY(vdb_multi_sync(f));
    if (nq == 0) return VDB_OK;
    const bool pq = f->cfg.index_type == VDB_INDEX_PQ;
    cudaStream_t fs = f->stream;
    VDB_CUDA(cudaSetDevice(d0));
    int32_t* o_ids = out_ids; float* o_d = out_d;
    if (!q_on_device) {
        VDB_TRY(f->p_q.ensure((size_t)nq * dim)); VDB_TRY(f->p_d.ensure((size_t)nq * k)); VDB_TRY(f->p_ids.ensure((size_t)nq * k));
        VDB_TRY(f->d_out_d.ensure((size_t)nq * k, 0, fs)); VDB_TRY(f->d_out_ids.ensure((size_t)nq * k, 0, fs));
        std::memcpy(f->p_q.p, q, (size_t)nq * dim * 4);
        o_ids = f->d_out_ids.p; o_d = f->d_out_d.p;
    } else {
        VDB_CUDA(cudaEventRecord(f->ev_front, user_stream));              // the caller's queries are ready on its stream
    }
    if (f->n == 0 || (pq && !f->trained)) {                                // empty / untrained: empty lists (PQIndex.cpp:147-150)
        launch_keys_to_results(nullptr, 0, k, nq, o_ids, o_d, fs);
    } else {
        const int R = pq ? (f->cfg.rerank > 0 ? std::max(f->cfg.rerank, k) : k) : 0;
        if (pq && (int64_t)g * R > 16384) MG_FAIL(VDB_ERR_INVALID, "multi-device PQ search: n_devices x shortlist > 16384 not supported");
        if ((int64_t)g * k > 8192) MG_FAIL(VDB_ERR_INVALID, "multi-device search: n_devices x k > 8192 not supported");
        const int64_t QB = pq ? 1024 : 8192;        // flat: whole batches (the GEMM wants many query tiles); PQ: the LUT scratch is per query
        VDB_TRY(f->g_d.ensure((size_t)g * nq * k, 0, fs)); VDB_TRY(f->g_i.ensure((size_t)g * nq * k, 0, fs));
        if (pq) { VDB_TRY(f->g_keys.ensure((size_t)g * QB * R, 0, fs)); VDB_TRY(f->g_kth.ensure((size_t)QB, 0, fs)); VDB_TRY(f->g_unc.ensure(1, 0, fs)); }
        // Each shard keeps a LOCAL shortlist of about twice its expected share of the global top-R (the scan's candidate bookkeeping
        // grows with the list length and does not shrink with the shard): with rows spread evenly a shard holds Binomial(R, 1/g) of
        // the global shortlist, 2 R / g + 64 is more than ten standard deviations up. select_kth_key_kernel certifies the boundary per
        // query (no shard may run out of keys below it); a sub-batch with an uncertain query is repeated with full-length lists.
        const int Rloc = (pq && g > 2 && !getenv("VDB_MULTI_FULL_SHORTLIST")) ? std::min(R, std::max(k, 2 * ((R + g - 1) / g) + 64)) : R;
        // queries to every shard
        for (int s = 0; s < g; s++) {
            vdb_index* sh = f->shards[s];
            VDB_CUDA(cudaSetDevice(sh->cfg.device));
            VDB_TRY(sh->d_q.ensure((size_t)nq * dim, 0, sh->stream));
            VDB_TRY(sh->d_out_d.ensure((size_t)QB * k, 0, sh->stream)); VDB_TRY(sh->d_out_ids.ensure((size_t)QB * k, 0, sh->stream));
            if (pq) VDB_TRY(sh->d_kth.ensure((size_t)QB, 0, sh->stream));
            if (q_on_device) {
                VDB_CUDA(cudaStreamWaitEvent(sh->stream, f->ev_front, 0));
                VDB_CUDA(cudaMemcpyPeerAsync(sh->d_q.p, sh->cfg.device, q, d0, (size_t)nq * dim * 4, sh->stream));
            } else {
                VDB_CUDA(cudaMemcpyAsync(sh->d_q.p, f->p_q.p, (size_t)nq * dim * 4, cudaMemcpyHostToDevice, sh->stream));
            }
        }
        VDB_CUDA(cudaSetDevice(d0));
        for (int64_t q0 = 0; q0 < nq; q0 += QB) {
            const int64_t qb = std::min(QB, nq - q0);
            float* gd = f->g_d.p + (size_t)g * q0 * k; int32_t* gi = f->g_i.p + (size_t)g * q0 * k;
            if (pq) {
                for (int attempt = 0; attempt < 2; attempt++) {
                    const int L = attempt == 0 ? Rloc : R;              // list length of this attempt
                    for (int s = 0; s < g; s++)
                        if (vdb_size(f->shards[s]) == 0) VDB_CUDA(cudaMemsetAsync(f->g_keys.p + (size_t)s * qb * L, 0xff, (size_t)qb * L * 8, fs));
                    VDB_CUDA(cudaMemsetAsync(f->g_unc.p, 0, sizeof(int), fs));
                    VDB_CUDA(cudaEventRecord(f->ev_front, fs));
                    VDB_TRY(for_each_shard(f, [&](int s) -> int {
                        vdb_index* sh = f->shards[s];
                        VDB_CUDA(cudaStreamWaitEvent(sh->stream, f->ev_front, 0));       // g_keys of the previous sub-batch / attempt consumed
                        if (vdb_size(sh) > 0) {
                            sh->shortlist_override = L < R ? L : 0;
                            VDB_TRY(vdb_pq_phase1(sh, sh->d_q.p + q0 * dim, qb, k, sh->stream));
                            VDB_CUDA(cudaMemcpyPeerAsync(f->g_keys.p + (size_t)s * qb * L, d0, sh->d_keys.p, sh->cfg.device, (size_t)qb * L * 8, sh->stream));
                        }
                        VDB_CUDA(cudaEventRecord(f->shard_ev[s], sh->stream));
                        return VDB_OK;
                    }));
                    for (int s = 0; s < g; s++) VDB_CUDA(cudaStreamWaitEvent(fs, f->shard_ev[s], 0));
                    launch_select_kth_key(f->g_keys.p, g, (int)qb, L, R, f->g_kth.p, L < R ? f->g_unc.p : nullptr, fs);
                    VDB_CUDA(cudaEventRecord(f->ev_front, fs));
                    if (L == R) break;
                    int unc = 0;
                    VDB_CUDA(cudaMemcpyAsync(&unc, f->g_unc.p, sizeof(int), cudaMemcpyDeviceToHost, fs));
                    VDB_CUDA(cudaStreamSynchronize(fs));
                    if (unc == 0) break;
                    f->stats.shortlist_retries += unc;                  // reported: queries that needed full-length lists
                }
            }
            VDB_TRY(for_each_shard(f, [&](int s) -> int {
                vdb_index* sh = f->shards[s];
                if (vdb_size(sh) > 0) {
                    if (pq) {
                        VDB_CUDA(cudaStreamWaitEvent(sh->stream, f->ev_front, 0));
                        VDB_CUDA(cudaMemcpyPeerAsync(sh->d_kth.p, sh->cfg.device, f->g_kth.p, d0, (size_t)qb * 8, sh->stream));
                        VDB_TRY(vdb_pq_phase2(sh, sh->d_q.p + q0 * dim, qb, k, sh->d_kth.p, sh->d_out_ids.p, sh->d_out_d.p, sh->stream));
                    } else {
                        std::lock_guard<std::mutex> lk(sh->mu);
                        VDB_TRY(vdb_search_device_nolock(sh, sh->d_q.p + q0 * dim, qb, k, sh->d_out_ids.p, sh->d_out_d.p, sh->stream));
                    }
                } else {
                    launch_keys_to_results(nullptr, 0, k, qb, sh->d_out_ids.p, sh->d_out_d.p, sh->stream);
                }
                VDB_CUDA(cudaMemcpyPeerAsync(gd + (size_t)s * qb * k, d0, sh->d_out_d.p, sh->cfg.device, (size_t)qb * k * 4, sh->stream));
                VDB_CUDA(cudaMemcpyPeerAsync(gi + (size_t)s * qb * k, d0, sh->d_out_ids.p, sh->cfg.device, (size_t)qb * k * 4, sh->stream));
                VDB_CUDA(cudaEventRecord(f->shard_ev[s], sh->stream));
                return VDB_OK;
            }));
            VDB_CUDA(cudaSetDevice(d0));
            for (int s = 0; s < g; s++) VDB_CUDA(cudaStreamWaitEvent(fs, f->shard_ev[s], 0));
            launch_merge_topk(gd, gi, g, qb, k, o_d + q0 * k, o_ids + q0 * k, fs);
            VDB_CUDA(cudaGetLastError());
            // the shards' result buffers are reused by the next sub-batch: its phase 2 comes after this sub-batch's copies in stream order
        }
    }
    VDB_CUDA(cudaSetDevice(d0));
    if (!q_on_device) {
        VDB_CUDA(cudaMemcpyAsync(f->p_ids.p, o_ids, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, fs));
        VDB_CUDA(cudaMemcpyAsync(f->p_d.p, o_d, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, fs));
        VDB_CUDA(cudaStreamSynchronize(fs));
        std::memcpy(out_ids, f->p_ids.p, (size_t)nq * k * 4);
        std::memcpy(out_d, f->p_d.p, (size_t)nq * k * 4);
    } else if (user_stream != fs) {
        VDB_CUDA(cudaEventRecord(f->ev_front, fs));
        VDB_CUDA(cudaStreamWaitEvent(user_stream, f->ev_front, 0));
    }
    f->stats.queries += nq;
    return VDB_OK;
}
