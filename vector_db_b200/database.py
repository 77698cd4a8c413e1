"""Python mirror of the reference's Java façade for the query path, so parity tests read like the reference's own
tests (src/test/java/com/vectordb/**). Same names, argument meaning and error behaviour:

  * `CompressionConfig`              config/CompressionConfig.java:12-204
  * `VectorDatabase` / `Builder`     core/VectorDatabase.java:151-221, 284-286, 326-410
  * `IndexType`                      core/VectorDatabase.java createIndex switch (:71-103)

Only what the hot path touches is mirrored: storage (vectors.json), text vectorisation, IVF/LSH/Annoy are out of scope
(SURVEY.md §2 rows 6, 10-12). Every search runs on the GPU index (`vector_db_b200.index`)."""
from dataclasses import dataclass

import numpy as np

from .index import GpuIndex, NativeHnswIndex, NativeHnswPqIndex, NativePqIndex, SearchResult


class IndexType:
    HNSW, ANNOY, LSH, IVF, PQ, HNSWPQ = "HNSW", "ANNOY", "LSH", "IVF", "PQ", "HNSWPQ"


class CompressionType:
    NONE, PQ, HNSWPQ = "NONE", "PQ", "HNSWPQ"


@dataclass
class CompressionConfig:
    enabled: bool = False
    type: str = CompressionType.NONE
    pqSubspaces: int = 64
    pqBits: int = 8
    pqIterations: int = 25

    @staticmethod
    def defaultConfig():
        return CompressionConfig()

    @staticmethod
    def hnswPqConfig(pqSubspaces, pqBits):                       # CompressionConfig.java:93-104
        return CompressionConfig(True, CompressionType.HNSWPQ, pqSubspaces, pqBits, 25)

    @staticmethod
    def pqConfig(pqSubspaces, pqBits):
        return CompressionConfig(True, CompressionType.PQ, pqSubspaces, pqBits, 25)

    @staticmethod
    def _fit(dimension, target_sub_dim, floor):
        m = max(floor, dimension // target_sub_dim)
        while dimension % m != 0 and m > 1:
            m -= 1
        return m

    @staticmethod
    def recommendedConfig(dimension):                            # CompressionConfig.java:114-130 (8 dims / subspace)
        return CompressionConfig.hnswPqConfig(CompressionConfig._fit(dimension, 8, 8), 8)

    @staticmethod
    def highRecallConfig(dimension):                             # 4 dims / subspace
        return CompressionConfig.hnswPqConfig(CompressionConfig._fit(dimension, 4, 8), 8)

    @staticmethod
    def highCompressionConfig(dimension):                        # 16 dims / subspace
        return CompressionConfig.hnswPqConfig(CompressionConfig._fit(dimension, 16, 4), 8)

    def getCompressionRatio(self, dimension):                    # CompressionConfig.java:180-192: 4*d bytes -> M bytes
        if not self.enabled:
            return 1.0
        return (dimension * 4.0) / (self.pqSubspaces * self.pqBits / 8.0)

    def getMemorySavings(self, dimension):
        return (1.0 - 1.0 / self.getCompressionRatio(dimension)) * 100.0


class VectorDatabase:
    """core/VectorDatabase.java — addVector / search / rebuildIndex / getVector / deleteVector / size / close, with the
    index created exactly as createIndex does (:71-103) but landing on the GPU classes."""

    class Builder:
        def __init__(self):
            self.dimension = 0
            self.maxElements = 10000
            self.storagePath = None
            self.indexType = IndexType.HNSW
            self.compressionConfig = CompressionConfig.defaultConfig()
            self.index = None
            self.indexOptions = {}

        def withDimension(self, d):
            self.dimension = d
            return self

        def withMaxElements(self, n):
            self.maxElements = n
            return self

        def withStoragePath(self, p):
            self.storagePath = p
            return self

        def withIndexType(self, t):
            self.indexType = t
            return self

        def withCompression(self, cfg):
            self.compressionConfig = cfg
            return self

        def withCompressionEnabled(self, enabled):
            if enabled and not self.compressionConfig.enabled:
                self.compressionConfig = CompressionConfig.recommendedConfig(self.dimension)
            elif not enabled:
                self.compressionConfig = CompressionConfig.defaultConfig()
            return self

        def withIndex(self, index):                                # VectorDatabase.java:402-405
            self.index = index
            return self

        def withIndexOptions(self, **kw):                          # extra vdb_config fields (device, level_seed, ...)
            self.indexOptions.update(kw)
            return self

        def build(self):
            if self.dimension <= 0:
                raise ValueError("dimension must be positive")
            return VectorDatabase(self)

    def __init__(self, b):
        self.dimension, self.maxElements = b.dimension, b.maxElements
        self.compressionConfig = b.compressionConfig
        self.index = b.index if b.index is not None else self._createIndex(b.indexType, b.compressionConfig, b.indexOptions)
        self._vectors = {}

    def _createIndex(self, indexType, c, opts):
        if c.enabled and c.type == CompressionType.HNSWPQ:
            return NativeHnswPqIndex(self.dimension, self.maxElements, pqM=c.pqSubspaces, pqBits=c.pqBits, pqIters=c.pqIterations, **opts)
        if c.enabled and c.type == CompressionType.PQ:
            return NativePqIndex(self.dimension, self.maxElements, M=c.pqSubspaces, nBits=c.pqBits, **opts)
        if indexType == IndexType.PQ:
            return NativePqIndex(self.dimension, self.maxElements, **opts)
        if indexType == IndexType.HNSWPQ:
            r = CompressionConfig.recommendedConfig(self.dimension)
            return NativeHnswPqIndex(self.dimension, self.maxElements, pqM=r.pqSubspaces, **opts)
        if indexType == IndexType.HNSW:
            return NativeHnswIndex(self.dimension, self.maxElements, **opts)
        raise NotImplementedError("index type %s is outside the GPU query path (SURVEY.md §2 row 6)" % indexType)

    # ---- core/VectorDatabase.java:151-221 ----
    def addVector(self, id_, values):
        v = np.asarray(values, dtype=np.float32)
        if v.ndim != 1 or v.size != self.dimension:
            raise ValueError("向量维度不匹配: expected %d, got %d" % (self.dimension, v.size))   # IllegalArgumentException :152-154
        if id_ in self._vectors:
            return False
        ok = self.index.addVector(int(id_), v)
        if ok:
            self._vectors[id_] = v.copy()
        return ok

    def addVectors(self, ids, matrix):
        """batched form of addVector (one PCIe transfer); same per-row semantics"""
        m = np.ascontiguousarray(matrix, dtype=np.float32)
        if m.ndim != 2 or m.shape[1] != self.dimension:
            raise ValueError("向量维度不匹配")
        self.index.addBatch(np.asarray(ids, dtype=np.int32), m)
        for i, row in zip(ids, m):
            self._vectors[int(i)] = row
        return True

    def getVector(self, id_):
        return self._vectors.get(id_)

    def deleteVector(self, id_):
        """core/VectorDatabase.java:192-205: remove from the storage first (false when absent), then from the index."""
        if id_ not in self._vectors:
            return False
        del self._vectors[id_]
        return self.index.removeVector(int(id_))

    def search(self, query, k):
        q = np.asarray(query, dtype=np.float32)
        if q.ndim != 1 or q.size != self.dimension:
            raise ValueError("查询向量维度不匹配")                                                  # :215-217
        return self.index.searchNearest(q, k)

    def searchBatch(self, queries, k):
        return self.index.searchBatch(queries, k)

    def rebuildIndex(self):                                          # :284-286
        return self.index.buildIndex()

    def size(self):
        return self.index.size()

    def getCompressionRatio(self):
        return self.compressionConfig.getCompressionRatio(self.dimension)

    def close(self):
        self.index.close()
