#!/usr/bin/env python
"""tools/sass_table.py [libvectordb.so] — per-kernel SASS opcode table of the shipped library (cuobjdump -sass, sm_100a):
which kernels use the Blackwell paths (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA load, UTCBAR), which use the
legacy mma.sync (HMMA / IMMA), 256-bit global loads, shared-memory gathers, and local-memory traffic (STL / LDL = spills).
Output is markdown for profiles/."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vector_db_b200", "libvectordb.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
COLS = ["UTCHMMA", "LDTM", "UTMALDG", "UTCBAR", "HMMA", "IMMA", "LDG.256", "LDG", "LDS", "STS", "PRMT", "STL", "LDL", "BAR", "total"]
tab, cur, arch = collections.OrderedDict(), None, "?"
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        name = re.sub(r"^void ", "", name)
        name = re.sub(r"\(.*$", "", name)
        cur = tab.setdefault(name, collections.Counter())
        continue
    m = re.search(r"arch = (sm_\w+)", line)
    if m:
        arch = m.group(1)
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur is not None:
        op = m.group(1)
        cur["total"] += 1
        base = op.split(".")[0]
        if base in ("UTCHMMA", "LDTM", "UTMALDG", "UTCBAR", "HMMA", "IMMA", "LDS", "STS", "PRMT", "STL", "LDL", "BAR"):
            cur[base] += 1
        if base == "LDG":
            cur["LDG"] += 1
            if ".256" in op:
                cur["LDG.256"] += 1
print("# SASS opcode counts per kernel — %s (%s), `cuobjdump -sass`, static instruction counts\n" % (os.path.basename(lib), arch))
print("| kernel | " + " | ".join(COLS) + " |")
print("|---|" + "---|" * len(COLS))
for k, c in sorted(tab.items(), key=lambda kv: -kv[1]["total"]):
    print("| `%s` | " % k[:90] + " | ".join(str(c[x]) if c[x] else "·" for x in COLS) + " |")
print("""
Legend. `UTCHMMA` = tcgen05.mma (5th-gen tensor core, operands in shared memory, accumulator in TMEM), `LDTM` = tcgen05.ld (TMEM -> registers),
`UTMALDG` = TMA tile load (cp.async.bulk.tensor), `UTCBAR` = tcgen05.commit -> mbarrier; `HMMA` / `IMMA` = legacy warp-level mma.sync
(bf16 / int8), used by the beam kernel for 16 x 8 query-row products only; `LDG.256` = 32-byte global loads (ld.global.nc.L1::no_allocate.v8.b32);
`STL` / `LDL` = local-memory stores / loads (register spills and indexed local arrays). hnsw_search_kernel template arguments:
<BF16 rows (false = fp32 strict), FULL, 8-bit stage count (0 = off), int8>. The shipped default on C3 is `<true, false, 2, true>` (int8 rows);
the fp32 strict variant `<false, false, 0, false>` is the one with heavy local traffic (it is the bit-parity mode, not the fast path).
Regenerate with `python tools/sass_table.py > profiles/r2_sass_opcodes.md`.""")
