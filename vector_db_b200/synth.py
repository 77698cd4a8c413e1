"""Synthetic uniform[-1,1) vectors — the host twin of `fill_synth_kernel` / `vdb_synth` (csrc/common.cuh): element `idx`
of stream `seed` is a pure function of (seed, idx), so a shard generated on a GPU (`vdb_fill_synthetic`, offset by
`id_offset`), the same rows generated here on the host, and the oracle's generator all agree bit for bit. Used by
bench.py and tools/ to make inputs without touching the oracle."""
import numpy as np

_M1, _M2 = np.uint64(0xBF58476D1CE4E5B9), np.uint64(0x94D049BB133111EB)
_G, _C = np.uint64(0x9E3779B97F4A7C15), np.uint64(0x632BE59BD9B4E019)


def fill_uniform(seed, start, count, chunk=1 << 24):
    out = np.empty(count, dtype=np.float32)
    with np.errstate(over="ignore"):
        base = np.uint64(seed) * _G + _C
        for c0 in range(0, count, chunk):
            c1 = min(count, c0 + chunk)
            z = np.arange(start + c0, start + c1, dtype=np.uint64) + base
            z = (z ^ (z >> np.uint64(30))) * _M1
            z = (z ^ (z >> np.uint64(27))) * _M2
            z = z ^ (z >> np.uint64(31))
            k = (z >> np.uint64(40)).astype(np.int64)
            out[c0:c1] = ((k - (1 << 23)).astype(np.float32)) * np.float32(1.0 / (1 << 23))
    return out


def synth(seed, n, dim):
    return fill_uniform(seed, 0, n * dim).reshape(n, dim)
