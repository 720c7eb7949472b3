"""Opcode histogram per kernel of librpc_b200.so (cuobjdump -sass): the evidence for bulk copies (UBLKCP), mbarriers
(SYNCS), global atomics / reductions (ATOMG / REDG), wide loads (LDG.*.128 / .256) and, for the MLP kernels, the
tensor-core path (UTCHMMA / UTCMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st).  Usage: python tools/sass_summary.py [lib] > profiles/r02_sass_summary.txt"""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "robustpointclouds_b200/librpc_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
WATCH = ("UBLKCP", "SYNCS", "ATOMG", "REDG", "ATOM", "RED", "LDG", "STG", "LDS", "STS", "UTCHMMA", "UTCQMMA", "UTCMMA", "UTCBAR", "LDTM", "STTM",
         "UTMALDG", "UTMASTG", "BAR", "SHFL", "MATCH", "VOTE", "FFMA", "FMUL", "FADD", "DADD", "MUFU", "CCTL", "ERRBAR", "MEMBAR", "FENCE", "HMMA")
print(f"# {lib}: opcode counts per kernel (static SASS, sm_100a)")
cur, counts, wide = None, None, None
def flush():
    if cur is None: return
    tot = sum(counts.values())
    keys = [k for k in WATCH if counts.get(k)]
    print(f"{cur}\n    instructions {tot}; " + ", ".join(f"{k} {counts[k]}" for k in keys))
    if wide: print("    wide / special forms: " + ", ".join(f"{k} x{v}" for k, v in sorted(wide.items())))
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        flush()
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur, counts, wide = re.sub(r"\(.*", "", name), collections.Counter(), collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
    if m and cur is not None:
        op, mods = m.group(1), m.group(2)
        counts[op] += 1
        if op in ("LDG", "STG", "ATOMG", "REDG", "LDS", "STS") and re.search(r"\.(128|256|64)\b", mods): wide[op + mods] += 1
        if op.startswith(("UTC", "LDTM", "STTM", "UBLKCP", "SYNCS", "UTMA")): wide[op + mods] += 1
flush()
