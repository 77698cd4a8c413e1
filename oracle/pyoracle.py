"""ctypes bindings for the CPU oracle (oracle/liboracle.so) and the compiled reference (oracle/_ref/libvdbref.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference leg. The product package (vector_db_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")

ARITH_JAVA_SEQ, ARITH_NATIVE_AVX2, ARITH_GPU_TREE = 0, 1, 2
TRAIN_JAVA, TRAIN_NATIVE_HNSWPQ, TRAIN_NATIVE_PQ = 0, 1, 2
KIND_HNSWPQ, KIND_HNSW = 0, 1


def build(force=False):
    """Compile liboracle.so (always possible) and oracle/_ref (only where /root/reference exists)."""
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"])
    ref = os.path.join(_HERE, "_ref", "libvdbref.so")
    if os.path.isdir("/root/reference/native") and (force or not os.path.exists(ref)):
        subprocess.check_call(["make", "-C", _HERE, "ref"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(os.path.join(_HERE, "liboracle.so"))
        L.orc_fill_uniform.argtypes = [C.c_uint64, C.c_int64, C.c_int64, _f32p]
        L.orc_l2sq.argtypes = [_f32p, _f32p, C.c_int, C.c_int]; L.orc_l2sq.restype = C.c_float
        for n in ("orc_java_euclidean", "orc_java_cosine_similarity", "orc_native_cosine"):
            getattr(L, n).argtypes = [_f32p, _f32p, C.c_int]; getattr(L, n).restype = C.c_float
        L.orc_java_norm.argtypes = [_f32p, C.c_int]; L.orc_java_norm.restype = C.c_float
        L.orc_java_normalize.argtypes = [_f32p, C.c_int, _f32p]
        L.orc_java_normalize_rows.argtypes = [_f32p, C.c_int64, C.c_int, _f32p]
        L.orc_recommended_pq_m.argtypes = [C.c_int]; L.orc_recommended_pq_m.restype = C.c_int
        L.orc_java_random_probe.argtypes = [C.c_int64, C.c_int32, C.c_int, _i32p, _f32p]
        L.orc_pq_train.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.orc_pq_encode.argtypes = [_f32p, C.c_int64, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, _u8p]
        L.orc_pq_lut.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, _f32p]
        L.orc_adc_scan.argtypes = [_f32p, _u8p, C.c_int64, C.c_int, C.c_int, _f32p]
        L.orc_pq_direct.argtypes = [_f32p, _u8p, _f32p, C.c_int, C.c_int, C.c_int]; L.orc_pq_direct.restype = C.c_float
        L.orc_pq_flat_search.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _u8p, C.c_void_p, C.c_int64,
                                         C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, _i32p, _f32p, C.c_int]
        L.orc_flat_search.argtypes = [_f32p, C.c_void_p, C.c_int64, C.c_int, _f32p, C.c_int, C.c_int, C.c_int, C.c_int,
                                      _i32p, _f32p, C.c_int]
        L.orc_hnsw_create.argtypes = [C.c_int] * 8 + [C.c_uint64]; L.orc_hnsw_create.restype = C.c_void_p
        L.orc_hnsw_destroy.argtypes = [C.c_void_p]
        L.orc_hnsw_set_arith.argtypes = [C.c_void_p, C.c_int]
        L.orc_hnsw_set_ef_cap.argtypes = [C.c_void_p, C.c_int]
        L.orc_hnsw_add.argtypes = [C.c_void_p, C.c_int, _f32p]; L.orc_hnsw_add.restype = C.c_int
        L.orc_hnsw_add_batch.argtypes = [C.c_void_p, C.c_void_p, _f32p, C.c_int64]
        L.orc_hnsw_rebuild.argtypes = [C.c_void_p]
        for n in ("orc_hnsw_size", "orc_hnsw_trained", "orc_hnsw_max_level", "orc_hnsw_entry"):
            getattr(L, n).argtypes = [C.c_void_p]; getattr(L, n).restype = C.c_int
        L.orc_hnsw_get_codebooks.argtypes = [C.c_void_p, _f32p]
        L.orc_hnsw_set_codebooks.argtypes = [C.c_void_p, _f32p]
        L.orc_hnsw_get_codes.argtypes = [C.c_void_p, _u8p]
        L.orc_hnsw_get_vectors.argtypes = [C.c_void_p, _f32p]
        L.orc_hnsw_get_norms.argtypes = [C.c_void_p, _f32p]
        L.orc_hnsw_get_levels.argtypes = [C.c_void_p, _i32p]
        L.orc_hnsw_get_level_adj.argtypes = [C.c_void_p, C.c_int, _i32p, _u8p]
        L.orc_hnsw_import.argtypes = [C.c_void_p, C.c_int, _f32p, C.c_void_p, _i32p, _i32p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_java_search_ef.argtypes = [C.c_int] * 3; L.orc_java_search_ef.restype = C.c_int
        L.orc_java_random_stream.argtypes = [C.c_int64, C.c_int, C.c_int32, C.c_int, _i32p, _f32p]
        L.orc_hnsw_search_layer.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, _f32p, _i64p]
        L.orc_hnsw_search_layer.restype = C.c_int
        L.orc_hnsw_search_closest.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_int, _i64p]
        L.orc_hnsw_search_closest.restype = C.c_int
        L.orc_hnsw_remove.argtypes = [C.c_void_p, C.c_int]; L.orc_hnsw_remove.restype = C.c_int
        L.orc_hnsw_live.argtypes = [C.c_void_p]; L.orc_hnsw_live.restype = C.c_int
        L.orc_hnsw_get_dead.argtypes = [C.c_void_p, _u8p]
        L.orc_hnsw_set_dead.argtypes = [C.c_void_p, _u8p, _i64p, C.c_int64, C.c_uint64]
        L.orc_hnsw_get_jindex.argtypes = [C.c_void_p, _i64p]
        L.orc_hnsw_get_ids.argtypes = [C.c_void_p, _i32p]
        L.orc_hnsw_level_ctr.argtypes = [C.c_void_p]; L.orc_hnsw_level_ctr.restype = C.c_uint64
        L.orc_hnsw_java_size.argtypes = [C.c_void_p]; L.orc_hnsw_java_size.restype = C.c_int64
        L.orc_hnsw_search.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p, _f32p, _i64p, C.c_int]
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def fill_uniform(seed, start, count):
    out = np.empty(count, dtype=np.float32)
    lib().orc_fill_uniform(seed, start, count, out)
    return out


def synth(seed, n, dim):
    return fill_uniform(seed, 0, n * dim).reshape(n, dim)


def l2sq(a, b, order=0):
    a, b = _f32(a), _f32(b)
    return float(lib().orc_l2sq(a, b, a.size, order))


def pq_train(samples, M, iters=25, flavour=TRAIN_JAVA, ncent=256):
    s = _f32(samples)
    n, dim = s.shape
    cb = np.zeros((M, ncent, dim // M), dtype=np.float32)
    lib().orc_pq_train(s, n, dim, M, ncent, iters, flavour, cb)
    return cb


def pq_encode(x, codebooks, arith):
    x = _f32(x)
    n, dim = x.shape
    M, ncent, _ = codebooks.shape
    codes = np.empty((n, M), dtype=np.uint8)
    lib().orc_pq_encode(x, n, dim, M, ncent, _f32(codebooks), arith, codes)
    return codes


def pq_lut(q, codebooks, arith):
    q = _f32(q)
    M, ncent, _ = codebooks.shape
    lut = np.empty((M, ncent), dtype=np.float32)
    lib().orc_pq_lut(q, q.size, M, ncent, _f32(codebooks), arith, lut)
    return lut


def adc_scan(lut, codes):
    n, M = codes.shape
    out = np.empty(n, dtype=np.float32)
    lib().orc_adc_scan(_f32(lut), np.ascontiguousarray(codes), n, M, lut.shape[1], out)
    return out


def pq_direct(q, code, codebooks):
    M, ncent, sd = codebooks.shape
    return float(lib().orc_pq_direct(_f32(q), np.ascontiguousarray(code), _f32(codebooks), M * sd, M, ncent))


def pq_flat_search(queries, codebooks, codes, k, arith=ARITH_NATIVE_AVX2, ids=None, rerank=0, base=None,
                   exact_order=0, nthreads=1):
    q = _f32(queries)
    nq, dim = q.shape
    M, ncent, _ = codebooks.shape
    n = codes.shape[0]
    out_ids = np.empty((nq, k), dtype=np.int32)
    out_d = np.empty((nq, k), dtype=np.float32)
    idp = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32).ctypes.data
    if base is not None:
        base = _f32(base)
    bp = None if base is None else base.ctypes.data
    lib().orc_pq_flat_search(q, nq, dim, M, ncent, _f32(codebooks), np.ascontiguousarray(codes), idp, n, arith, k,
                             rerank, bp, exact_order, out_ids, out_d, nthreads)
    return out_ids, out_d


def flat_search(base, queries, k, metric=0, order=0, ids=None, nthreads=1):
    base, q = _f32(base), _f32(queries)
    n, dim = base.shape
    nq = q.shape[0]
    out_ids = np.empty((nq, k), dtype=np.int32)
    out_d = np.empty((nq, k), dtype=np.float32)
    idp = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32).ctypes.data
    lib().orc_flat_search(base, idp, n, dim, q, nq, k, metric, order, out_ids, out_d, nthreads)
    return out_ids, out_d


def java_normalize(v):
    v = _f32(v)
    out = np.empty_like(v)
    if v.ndim == 1:
        lib().orc_java_normalize(v, v.size, out)
    else:
        for i in range(v.shape[0]):
            lib().orc_java_normalize(v[i], v.shape[1], out[i])
    return out


def java_normalize_rows(v):
    """row-wise Vector.normalize (core/Vector.java:107-116) in C with OpenMP: for the 1M-row cosine checks"""
    v = _f32(v)
    out = np.empty_like(v)
    lib().orc_java_normalize_rows(v, v.shape[0], v.shape[1], out)
    return out


class OracleHnsw:
    """Restatement of Java HnswPqIndex (kind 0) / HnswIndex (kind 1)."""

    def __init__(self, kind, dim, max_elements, M=32, efC=None, ef=None, pq_m=0, pq_iters=25, level_seed=1234):
        if efC is None:
            efC = 64 if kind == KIND_HNSWPQ else 400      # HnswPqIndex.java:28 / HnswIndex.java:22
        if ef is None:
            ef = 64 if kind == KIND_HNSWPQ else 400       # HnswPqIndex.java:29 / HnswIndex.java:23
        self.kind, self.dim, self.M, self.pq_m = kind, dim, M, pq_m
        self.h = lib().orc_hnsw_create(kind, dim, max_elements, M, efC, ef, pq_m, pq_iters, level_seed)

    def close(self):
        if self.h:
            lib().orc_hnsw_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_arith(self, arith):
        lib().orc_hnsw_set_arith(self.h, arith)

    def set_ef_cap(self, cap):
        """searchEf cap (HnswPqIndex.java:621 hard-codes 5000); other values are a documented extension"""
        lib().orc_hnsw_set_ef_cap(self.h, int(cap))

    def add_batch(self, vecs, ids=None):
        v = _f32(vecs)
        idp = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32).ctypes.data
        lib().orc_hnsw_add_batch(self.h, idp, v, v.shape[0])

    def search_layer(self, q, ep, ef, level, variant):
        """test hook: one searchLayer* call (variant 0 searchLayerWithSize, 1 HnswPqIndex.searchLayer, 2 HnswIndex.searchLayer)
        -> (node indices ascending by distance, distances, (dist_evals, expansions))"""
        idx = np.empty(self.size, dtype=np.int32); d = np.empty(self.size, dtype=np.float32); st = np.zeros(2, dtype=np.int64)
        c = lib().orc_hnsw_search_layer(self.h, _f32(q), int(ep), int(ef), int(level), int(variant), idx, d, st)
        return idx[:c], d[:c], st

    def search_closest(self, q, ep, level, compressed=False):
        st = np.zeros(2, dtype=np.int64)
        return lib().orc_hnsw_search_closest(self.h, _f32(q), int(ep), int(level), int(bool(compressed)), st), st

    def add(self, id_, vec):
        """addVector: False for a duplicate id or a full index (HnswPqIndex.java:157-163)"""
        return bool(lib().orc_hnsw_add(self.h, int(id_), _f32(vec)))

    def remove(self, id_):
        """removeVector (HnswPqIndex.java:542-596 / HnswIndex.java:254-317)"""
        return bool(lib().orc_hnsw_remove(self.h, int(id_)))

    @property
    def live(self):
        return lib().orc_hnsw_live(self.h)

    def dead(self):
        out = np.zeros(self.size, dtype=np.uint8)
        lib().orc_hnsw_get_dead(self.h, out)
        return out

    def ext_ids(self):
        out = np.empty(self.size, dtype=np.int32)
        lib().orc_hnsw_get_ids(self.h, out)
        return out

    def rebuild(self):
        lib().orc_hnsw_rebuild(self.h)

    def set_codebooks(self, cb):
        lib().orc_hnsw_set_codebooks(self.h, _f32(cb))

    @property
    def size(self):
        return lib().orc_hnsw_size(self.h)

    @property
    def trained(self):
        return bool(lib().orc_hnsw_trained(self.h))

    @property
    def max_level(self):
        return lib().orc_hnsw_max_level(self.h)

    @property
    def entry(self):
        return lib().orc_hnsw_entry(self.h)

    def codebooks(self):
        cb = np.empty((self.pq_m, 256, self.dim // self.pq_m), dtype=np.float32)
        lib().orc_hnsw_get_codebooks(self.h, cb)
        return cb

    def codes(self):
        c = np.empty((self.size, self.pq_m), dtype=np.uint8)
        lib().orc_hnsw_get_codes(self.h, c)
        return c

    def vectors(self):
        v = np.empty((self.size, self.dim), dtype=np.float32)
        lib().orc_hnsw_get_vectors(self.h, v)
        return v

    def norms(self):
        v = np.empty(self.size, dtype=np.float32)
        lib().orc_hnsw_get_norms(self.h, v)
        return v

    def levels(self):
        v = np.empty(self.size, dtype=np.int32)
        lib().orc_hnsw_get_levels(self.h, v)
        return v

    def level_adj(self, lvl):
        adj = np.empty((self.size, self.M), dtype=np.int32)
        present = np.empty(self.size, dtype=np.uint8)
        lib().orc_hnsw_get_level_adj(self.h, lvl, adj, present)
        return adj, present

    def import_state(self, vecs, levels, adj, entry, ids=None, codebooks=None, codes=None):
        """load vectors + graph (+ PQ state) built elsewhere; adj is [max_level][n][M]"""
        v = _f32(vecs)
        lv = np.ascontiguousarray(levels, dtype=np.int32)
        a = np.ascontiguousarray(adj, dtype=np.int32)
        assert a.shape == (self.max_level, v.shape[0], self.M), a.shape
        idp = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32).ctypes.data
        cb = None if codebooks is None else _f32(codebooks)
        cd = None if codes is None else np.ascontiguousarray(codes, dtype=np.uint8)
        lib().orc_hnsw_import(self.h, v.shape[0], v, idp, lv, a, int(entry), None if cb is None else cb.ctypes.data,
                              None if cd is None else cd.ctypes.data)

    def export_graph(self):
        """(levels[n], [adj_l[n][M] for l in range(max_level)], entry) in the layout vdb_set_graph takes."""
        return self.levels(), [self.level_adj(l)[0] for l in range(self.max_level)], self.entry

    def search(self, queries, k, nthreads=1):
        q = _f32(queries)
        nq = q.shape[0]
        ids = np.empty((nq, k), dtype=np.int32)
        d = np.empty((nq, k), dtype=np.float32)
        stats = np.zeros(2, dtype=np.int64)
        lib().orc_hnsw_search(self.h, q, nq, k, ids, d, stats, nthreads)
        return ids, d, stats


# ------------------------------------------------------------------------------------------------
# VDBSNAP2 — the snapshot format of vdb_save / vdb_load (vector_db_b200/csrc/snapshot.cu), read and written here so that
# the CPU oracle and the GPU library load "the same persisted graph and codebooks" from ONE file (SURVEY.md §8f rank 2;
# the reference's own save()/load() are empty: native/index/HNSWPQIndex.cpp:774-780, VectorIndex.h:21-25).
# ------------------------------------------------------------------------------------------------
_CFG_FIELDS = ["dim", "max_elements", "index_type", "metric", "arith", "hnsw_m", "ef_construction", "ef_search", "pq_m", "pq_bits",
               "pq_iters", "train_flavour", "rerank", "device", "id_offset", "search_ef_cap"]
_CFG_DTYPE = np.dtype([(n, "<i4") for n in _CFG_FIELDS] + [("level_seed", "<u8")] +
                      [(n, "<i4") for n in ("trav_store", "trav_batch", "n_devices", "reserved2")])
_HDR_DTYPE = np.dtype([("n", "<i8"), ("trained", "<i4"), ("max_level", "<i4"), ("entry", "<i4"), ("any_dead", "<i4"), ("raw_rows", "<i8"),
                       ("level_ctr", "<u8"), ("n_train", "<i8"), ("java_size", "<i8"), ("n_coded", "<i8")])


def snapshot_read(path):
    """-> dict(cfg, hdr, ids, vectors, norms, codebooks, codes, samples, levels, adj[max_level][n][deg], dead, jindex)"""
    with open(path, "rb") as f:
        if f.read(8) != b"VDBSNAP2":
            raise ValueError("not a VDBSNAP2 snapshot")
        cfg_bytes = int(np.frombuffer(f.read(4), "<u4")[0])
        if cfg_bytes != _CFG_DTYPE.itemsize:
            raise ValueError("vdb_config size mismatch: %d != %d" % (cfg_bytes, _CFG_DTYPE.itemsize))
        cfg = np.frombuffer(f.read(cfg_bytes), _CFG_DTYPE)[0]
        hdr = np.frombuffer(f.read(_HDR_DTYPE.itemsize), _HDR_DTYPE)[0]
        n, dim, M = int(hdr["n"]), int(cfg["dim"]), int(cfg["pq_m"])
        rd = lambda dt, cnt: np.frombuffer(f.read(np.dtype(dt).itemsize * cnt), dt, cnt).copy()
        out = {"cfg": cfg, "hdr": hdr, "ids": rd("<i4", n), "vectors": rd("<f4", n * dim).reshape(n, dim)}
        out["norms"] = rd("<f4", n) if int(cfg["metric"]) == 1 else None
        uses_pq = int(cfg["index_type"]) in (1, 3) and M > 0
        out["codebooks"] = out["codes"] = out["samples"] = None
        if uses_pq and int(hdr["trained"]):
            out["codebooks"] = rd("<f4", M * 256 * (dim // M)).reshape(M, 256, dim // M)
            out["codes"] = rd("u1", n * M).reshape(n, M)
        elif uses_pq:
            out["samples"] = rd("<f4", int(hdr["n_train"]) * dim).reshape(-1, dim)
        if int(cfg["index_type"]) in (2, 3):
            L, deg = int(hdr["max_level"]), int(cfg["hnsw_m"])
            out["levels"] = rd("<i4", n)
            out["adj"] = rd("<i4", L * n * deg).reshape(L, n, deg)
            out["dead"] = rd("u1", n)
            out["jindex"] = rd("<i8", n)
        if f.read(1):
            raise ValueError("trailing bytes in snapshot")
    return out


def snapshot_write(path, snap):
    with open(path, "wb") as f:
        f.write(b"VDBSNAP2")
        f.write(np.array([_CFG_DTYPE.itemsize], "<u4").tobytes())
        f.write(np.array([snap["cfg"]], _CFG_DTYPE).tobytes())
        f.write(np.array([snap["hdr"]], _HDR_DTYPE).tobytes())
        for key, dt in (("ids", "<i4"), ("vectors", "<f4"), ("norms", "<f4"), ("codebooks", "<f4"), ("codes", "u1"), ("samples", "<f4"),
                        ("levels", "<i4"), ("adj", "<i4"), ("dead", "u1"), ("jindex", "<i8")):
            if snap.get(key) is not None:
                f.write(np.ascontiguousarray(snap[key], dtype=dt).tobytes())


def oracle_from_snapshot(path):
    """build an OracleHnsw holding exactly the state a vdb_save file describes (graph indexes)"""
    s = snapshot_read(path)
    cfg, hdr = s["cfg"], s["hdr"]
    kind = KIND_HNSWPQ if int(cfg["index_type"]) == 3 else KIND_HNSW
    o = OracleHnsw(kind, int(cfg["dim"]), int(cfg["max_elements"]), M=int(cfg["hnsw_m"]), efC=int(cfg["ef_construction"]),
                   ef=int(cfg["ef_search"]), pq_m=int(cfg["pq_m"]) if kind == KIND_HNSWPQ else 0, pq_iters=int(cfg["pq_iters"]),
                   level_seed=int(cfg["level_seed"]))
    o.import_state(s["vectors"], s["levels"], s["adj"], int(hdr["entry"]), ids=s["ids"], codebooks=s["codebooks"], codes=s["codes"])
    lib().orc_hnsw_set_dead(o.h, np.ascontiguousarray(s["dead"]), np.ascontiguousarray(s["jindex"]), int(hdr["java_size"]), int(hdr["level_ctr"]))
    return o


def oracle_to_snapshot(o, path, max_elements, ef_construction, ef_search, pq_iters=25, level_seed=1234):
    """write the oracle's state as a VDBSNAP2 file vdb_load accepts (graph indexes, trained or compression off)"""
    n = o.size
    hnswpq = o.kind == KIND_HNSWPQ
    cfg = np.zeros((), _CFG_DTYPE)
    cfg["dim"], cfg["max_elements"], cfg["index_type"], cfg["metric"] = o.dim, max_elements, 3 if hnswpq else 2, 0 if hnswpq else 1
    cfg["hnsw_m"], cfg["ef_construction"], cfg["ef_search"] = o.M, ef_construction, ef_search
    cfg["pq_m"], cfg["pq_bits"], cfg["pq_iters"], cfg["level_seed"] = (o.pq_m if hnswpq else 0), 8, pq_iters, level_seed
    dead = o.dead()
    jindex = np.zeros(n, dtype=np.int64)
    lib().orc_hnsw_get_jindex(o.h, jindex)
    hdr = np.zeros((), _HDR_DTYPE)
    trained = hnswpq and o.trained
    hdr["n"], hdr["trained"], hdr["max_level"], hdr["entry"], hdr["any_dead"] = n, int(trained), o.max_level, o.entry, int(dead.any())
    hdr["level_ctr"], hdr["java_size"], hdr["n_coded"] = lib().orc_hnsw_level_ctr(o.h), lib().orc_hnsw_java_size(o.h), n if trained else 0
    levels, adj, _ = o.export_graph()
    snap = {"cfg": cfg, "hdr": hdr, "ids": o.ext_ids(), "vectors": o.vectors(), "norms": o.norms() if not hnswpq else None,
            "codebooks": o.codebooks() if trained else None, "codes": o.codes() if trained else None,
            "samples": np.zeros((0, o.dim), np.float32) if (hnswpq and not trained) else None,
            "levels": levels, "adj": np.stack(adj, 0), "dead": dead, "jindex": jindex}
    snapshot_write(path, snap)


# ------------------------------------------------------------------------------------------------
# compiled reference (oracle/_ref) — present only where `make -C oracle ref` has run
# ------------------------------------------------------------------------------------------------
_ref = None


def ref_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libvdbref.so"))


def ref():
    global _ref
    if _ref is None:
        R = C.CDLL(os.path.join(_HERE, "_ref", "libvdbref.so"))
        R.ref_isa_name.restype = C.c_char_p
        for n in ("ref_l2sq_dispatch", "ref_l2sq_avx2", "ref_l2sq_scalar", "ref_cosine_avx2"):
            getattr(R, n).argtypes = [_f32p, _f32p, C.c_int]; getattr(R, n).restype = C.c_float
        R.ref_batch_l2.argtypes = [_f32p, _f32p, C.c_int, C.c_int, _f32p]
        for n in ("ref_adc_scalar", "ref_adc_avx2"):
            getattr(R, n).argtypes = [_f32p, _u8p, C.c_int, C.c_int]; getattr(R, n).restype = C.c_float
        R.ref_pq_create.argtypes = [C.c_int] * 5; R.ref_pq_create.restype = C.c_void_p
        R.ref_pq_destroy.argtypes = [C.c_void_p]
        R.ref_pq_train.argtypes = [C.c_void_p, C.c_int, _f32p]; R.ref_pq_train.restype = C.c_int
        R.ref_pq_add_batch.argtypes = [C.c_void_p, _f32p, _i32p, C.c_int]; R.ref_pq_add_batch.restype = C.c_int
        R.ref_pq_search.argtypes = [C.c_void_p, _f32p, C.c_int, _i32p, _f32p]; R.ref_pq_search.restype = C.c_int
        R.ref_pq_search_batch.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p, _f32p]
        R.ref_pq_size.argtypes = [C.c_void_p]; R.ref_pq_size.restype = C.c_int
        R.ref_pq_get_codebooks.argtypes = [C.c_void_p, _f32p]
        R.ref_pq_get_codes.argtypes = [C.c_void_p, _u8p]
        R.ref_hnswpq_create.argtypes = [C.c_int] * 8; R.ref_hnswpq_create.restype = C.c_void_p
        R.ref_hnswpq_destroy.argtypes = [C.c_void_p]
        R.ref_hnswpq_train.argtypes = [C.c_void_p, C.c_int, _f32p]; R.ref_hnswpq_train.restype = C.c_int
        R.ref_hnswpq_get_codebooks.argtypes = [C.c_void_p, _f32p]
        R.ref_hnswpq_encode.argtypes = [C.c_void_p, _f32p, C.c_int, _u8p]
        R.ref_hnswpq_add.argtypes = [C.c_void_p, C.c_int, _f32p]; R.ref_hnswpq_add.restype = C.c_int
        R.ref_hnswpq_search.argtypes = [C.c_void_p, _f32p, C.c_int, _i32p, _f32p]; R.ref_hnswpq_search.restype = C.c_int
        R.ref_hnswpq_search_batch.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p, _f32p]
        R.ref_hnswpq_pq_distance.argtypes = [C.c_void_p, _f32p, C.c_int]; R.ref_hnswpq_pq_distance.restype = C.c_float
        if hasattr(R, "ref_hnsw_create"):
            R.ref_hnsw_create.argtypes = [C.c_int] * 5; R.ref_hnsw_create.restype = C.c_void_p
            R.ref_hnsw_destroy.argtypes = [C.c_void_p]
            R.ref_hnsw_add.argtypes = [C.c_void_p, C.c_int, _f32p]; R.ref_hnsw_add.restype = C.c_int
            R.ref_hnsw_add_batch.argtypes = [C.c_void_p, _f32p, _i32p, C.c_int]; R.ref_hnsw_add_batch.restype = C.c_int
            R.ref_hnsw_size.argtypes = [C.c_void_p]; R.ref_hnsw_size.restype = C.c_int
            R.ref_hnsw_search.argtypes = [C.c_void_p, _f32p, C.c_int, _i32p, _f32p]; R.ref_hnsw_search.restype = C.c_int
            R.ref_hnsw_search_batch.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p, _f32p]
            R.ref_hnsw_entry.argtypes = [C.c_void_p]; R.ref_hnsw_entry.restype = C.c_int
            R.ref_hnsw_node_level.argtypes = [C.c_void_p, C.c_int]; R.ref_hnsw_node_level.restype = C.c_int
            R.ref_hnsw_neighbors.argtypes = [C.c_void_p, C.c_int, C.c_int, _i32p, C.c_int]; R.ref_hnsw_neighbors.restype = C.c_int
        _ref = R
    return _ref


class RefPQ:
    """The reference's native PQIndex, unmodified (native/index/PQIndex.cpp)."""

    def __init__(self, dim, max_elements, M, nbits=8, max_iter=25):
        self.dim, self.M = dim, M
        self.h = ref().ref_pq_create(dim, max_elements, M, nbits, max_iter)
        assert self.h

    def close(self):
        if self.h:
            ref().ref_pq_destroy(self.h)
            self.h = None

    def train(self, samples):
        s = _f32(samples)
        assert ref().ref_pq_train(self.h, s.shape[0], s) == 0

    def add_batch(self, vecs, ids):
        v = _f32(vecs)
        assert ref().ref_pq_add_batch(self.h, v, np.ascontiguousarray(ids, dtype=np.int32), v.shape[0]) == 0

    def codebooks(self):
        cb = np.empty((self.M, 256, self.dim // self.M), dtype=np.float32)
        ref().ref_pq_get_codebooks(self.h, cb)
        return cb

    def codes(self):
        c = np.empty((ref().ref_pq_size(self.h), self.M), dtype=np.uint8)
        ref().ref_pq_get_codes(self.h, c)
        return c

    def search(self, q, k):
        ids = np.full(k, -1, dtype=np.int32)
        d = np.full(k, np.inf, dtype=np.float32)
        cnt = ref().ref_pq_search(self.h, _f32(q), k, ids, d)
        return ids, d, cnt

    def search_batch(self, queries, k):
        q = _f32(queries)
        ids = np.full((q.shape[0], k), -1, dtype=np.int32)
        d = np.full((q.shape[0], k), np.inf, dtype=np.float32)
        ref().ref_pq_search_batch(self.h, q, q.shape[0], k, ids, d)
        return ids, d


class RefHnswPq:
    """The reference's native HNSWPQIndex, unmodified (native/index/HNSWPQIndex.cpp)."""

    def __init__(self, dim, max_elements, pq_m, M=32, efC=64, ef=64, pq_bits=8, pq_iters=25):
        self.dim, self.pq_m = dim, pq_m
        self.h = ref().ref_hnswpq_create(dim, max_elements, M, efC, ef, pq_m, pq_bits, pq_iters)
        assert self.h

    def close(self):
        if self.h:
            ref().ref_hnswpq_destroy(self.h)
            self.h = None

    def train(self, samples):
        s = _f32(samples)
        assert ref().ref_hnswpq_train(self.h, s.shape[0], s) == 0

    def codebooks(self):
        cb = np.empty((self.pq_m, 256, self.dim // self.pq_m), dtype=np.float32)
        ref().ref_hnswpq_get_codebooks(self.h, cb)
        return cb

    def encode(self, vecs):
        v = _f32(vecs)
        codes = np.empty((v.shape[0], self.pq_m), dtype=np.uint8)
        ref().ref_hnswpq_encode(self.h, v, v.shape[0], codes)
        return codes

    def add(self, id_, v):
        assert ref().ref_hnswpq_add(self.h, int(id_), _f32(v)) == 0

    def search_batch(self, queries, k):
        q = _f32(queries)
        ids = np.full((q.shape[0], k), -1, dtype=np.int32)
        d = np.full((q.shape[0], k), np.inf, dtype=np.float32)
        ref().ref_hnswpq_search_batch(self.h, q, q.shape[0], k, ids, d)
        return ids, d


class RefHnsw:
    """The reference's native HNSWIndex, unmodified (native/index/HNSWIndex.cpp; BLAS calls through oracle/shim/cblas.h)."""

    def __init__(self, dim, max_elements, M=32, efC=64, ef=64):
        self.dim, self.M = dim, M
        self.h = ref().ref_hnsw_create(dim, max_elements, M, efC, ef)
        assert self.h

    def close(self):
        if self.h:
            ref().ref_hnsw_destroy(self.h)
            self.h = None

    def add(self, id_, v):
        return ref().ref_hnsw_add(self.h, int(id_), _f32(v)) == 0

    def add_batch(self, vecs, ids):
        v = _f32(vecs)
        assert ref().ref_hnsw_add_batch(self.h, v, np.ascontiguousarray(ids, dtype=np.int32), v.shape[0]) == 0

    def size(self):
        return ref().ref_hnsw_size(self.h)

    def search(self, q, k):
        ids = np.full(k, -1, dtype=np.int32)
        d = np.full(k, np.inf, dtype=np.float32)
        cnt = ref().ref_hnsw_search(self.h, _f32(q), k, ids, d)
        return ids, d, cnt

    def search_batch(self, queries, k):
        q = _f32(queries)
        ids = np.full((q.shape[0], k), -1, dtype=np.int32)
        d = np.full((q.shape[0], k), np.inf, dtype=np.float32)
        ref().ref_hnsw_search_batch(self.h, q, q.shape[0], k, ids, d)
        return ids, d

    def graph(self):
        """(levels[n], {level: adj[n][2M] padded with -1}, entry)"""
        n = self.size()
        levels = np.array([ref().ref_hnsw_node_level(self.h, i) for i in range(n)], dtype=np.int32)
        adj = {}
        buf = np.empty(4 * self.M, dtype=np.int32)
        for lv in range(int(levels.max()) + 1 if n else 0):
            a = np.full((n, 2 * self.M), -1, dtype=np.int32)
            for i in np.nonzero(levels >= lv)[0]:
                c = ref().ref_hnsw_neighbors(self.h, int(i), lv, buf, buf.size)
                a[i, :min(c, 2 * self.M)] = buf[:min(c, 2 * self.M)]
            adj[lv] = a
        return levels, adj, ref().ref_hnsw_entry(self.h)
