"""ctypes binding of libvectordb.so (include/vdb.h). Fails loudly when the CUDA library is missing: there is no
CPU fallback anywhere in this package."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VECTORDB_NATIVE_PATH", os.path.join(_HERE, "libvectordb.so"))   # NativeLoader.java:61-64

VDB_OK = 0
INDEX_FLAT, INDEX_PQ, INDEX_HNSW, INDEX_HNSW_PQ = 0, 1, 2, 3
METRIC_SQL2, METRIC_COSINE = 0, 1
ARITH_JAVA_SEQ, ARITH_NATIVE_AVX2 = 0, 1
TRAIN_JAVA, TRAIN_NATIVE_HNSWPQ, TRAIN_NATIVE_PQ = 0, 1, 2
TRAV_AUTO, TRAV_F32, TRAV_FP8, TRAV_INT8, TRAV_BF16 = 0, 1, 2, 3, 4

# every symbol include/vdb.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "vdb_last_error", "vdb_version", "vdb_default_config", "vdb_create", "vdb_destroy", "vdb_add_batch", "vdb_train",
    "vdb_build", "vdb_search_batch", "vdb_search_batch_device", "vdb_size", "vdb_is_trained", "vdb_get_stats",
    "vdb_set_codebooks", "vdb_get_codebooks", "vdb_get_codes", "vdb_set_codes", "vdb_set_graph", "vdb_get_graph",
    "vdb_max_level", "vdb_save", "vdb_load", "vdb_fill_synthetic", "vdb_encode_all", "vdb_pq_lut_device",
    "vdb_adc_scan_device", "vdb_merge_topk_device", "vdb_distance_device", "vdb_flat_search_noalloc_device", "vdb_remove", "vdb_contains", "vdb_rows", "vdb_microbench_device", "vdb_select_kth_key_ex_device",
    "vdb_pq_shortlist_device", "vdb_pq_rerank_device", "vdb_select_kth_key_device",
]


class VdbConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "dim", "max_elements", "index_type", "metric", "arith", "hnsw_m", "ef_construction", "ef_search", "pq_m", "pq_bits",
        "pq_iters", "train_flavour", "rerank", "device", "id_offset", "search_ef_cap")] + [("level_seed", C.c_uint64)] + [
        (n, C.c_int32) for n in ("trav_store", "trav_batch", "n_devices", "reserved2")]


class VdbStats(C.Structure):
    _fields_ = [("queries", C.c_int64), ("dist_evals", C.c_int64), ("expansions", C.c_int64), ("adc_rows", C.c_int64),
                ("last_kernel_ms", C.c_float), ("last_total_ms", C.c_float), ("flat_certified", C.c_int64), ("flat_fallback", C.c_int64),
                ("trav_format", C.c_int32), ("shortlist_retries", C.c_int32)]


class VdbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("vdb error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libvectordb.so not found at %s — build it with `make -C vector_db_b200/csrc` "
                           "(__graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
    L.vdb_last_error.restype = C.c_char_p
    L.vdb_version.restype = C.c_char_p
    L.vdb_default_config.argtypes = [i32, i32, i32, C.POINTER(VdbConfig)]
    L.vdb_create.argtypes = [C.POINTER(VdbConfig), C.POINTER(vp)]
    L.vdb_destroy.argtypes = [vp]
    L.vdb_add_batch.argtypes = [vp, vp, vp, i64]
    L.vdb_train.argtypes = [vp, vp, i64]
    L.vdb_build.argtypes = [vp]
    L.vdb_search_batch.argtypes = [vp, vp, i64, i32, vp, vp]
    L.vdb_search_batch_device.argtypes = [vp, vp, i64, i32, vp, vp, vp]
    L.vdb_size.argtypes = [vp]; L.vdb_size.restype = i64
    L.vdb_rows.argtypes = [vp]; L.vdb_rows.restype = i64
    L.vdb_microbench_device.argtypes = [i32, i64, i32, i32, i32, C.POINTER(C.c_double)]
    L.vdb_remove.argtypes = [vp, i32, C.POINTER(i32)]
    L.vdb_contains.argtypes = [vp, i32]
    L.vdb_is_trained.argtypes = [vp]
    L.vdb_get_stats.argtypes = [vp, C.POINTER(VdbStats)]
    L.vdb_set_codebooks.argtypes = [vp, vp]
    L.vdb_get_codebooks.argtypes = [vp, vp]
    L.vdb_get_codes.argtypes = [vp, vp]
    L.vdb_set_codes.argtypes = [vp, vp]
    L.vdb_set_graph.argtypes = [vp, vp, vp, i32, i32]
    L.vdb_get_graph.argtypes = [vp, vp, vp, C.POINTER(i32)]
    L.vdb_max_level.argtypes = [vp]; L.vdb_max_level.restype = i32
    L.vdb_save.argtypes = [vp, C.c_char_p]
    L.vdb_load.argtypes = [vp, C.c_char_p]
    L.vdb_fill_synthetic.argtypes = [vp, u64, i64]
    L.vdb_encode_all.argtypes = [vp]
    L.vdb_pq_lut_device.argtypes = [vp, vp, i64, vp, vp]
    L.vdb_adc_scan_device.argtypes = [vp, vp, vp, vp]
    L.vdb_merge_topk_device.argtypes = [vp, vp, i32, i64, i32, vp, vp, vp]
    L.vdb_distance_device.argtypes = [vp, vp, i64, i32, i32, i32, vp, vp]
    L.vdb_pq_shortlist_device.argtypes = [vp, vp, i64, i32, vp, C.POINTER(i32), vp]
    L.vdb_pq_rerank_device.argtypes = [vp, vp, i64, i32, vp, vp, vp, vp]
    L.vdb_select_kth_key_device.argtypes = [vp, i32, i64, i32, vp, vp]
    L.vdb_select_kth_key_ex_device.argtypes = [vp, i32, i64, i32, i32, vp, vp, vp]
    L.vdb_flat_search_noalloc_device.argtypes = [vp, i64, i32, i32, vp, i64, i32, vp, vp, vp]
    _lib = L
    return L


def check(rc):
    if rc != VDB_OK:
        raise VdbError(rc, lib().vdb_last_error().decode("utf-8", "replace"))
