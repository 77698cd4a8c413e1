"""vector_db_b200 — B200 (sm_100a) implementation of elex-fu/vector_db's batched query path behind the reference's own
index plug-in interface. All compute is in libvectordb.so (vector_db_b200/csrc, C ABI in include/vdb.h)."""
from ._lib import (ARITH_JAVA_SEQ, ARITH_NATIVE_AVX2, INDEX_FLAT, INDEX_HNSW, INDEX_HNSW_PQ, INDEX_PQ, LIB_PATH, METRIC_COSINE,
                   METRIC_SQL2, TRAIN_JAVA, TRAIN_NATIVE_HNSWPQ, TRAIN_NATIVE_PQ, VdbError)
from .database import CompressionConfig, CompressionType, IndexType, VectorDatabase
from .index import GpuIndex, NativeFlatIndex, NativeHnswIndex, NativeHnswPqIndex, NativePqIndex, SearchResult

__all__ = [n for n in dir() if not n.startswith("_")]
