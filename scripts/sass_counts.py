"""profiles/r02_sass.txt: per-kernel counts of the SASS instructions that prove tcgen05 / TMEM / TMA use
(B200_PROFILING.md: UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor load / store,
UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier), from `cuobjdump -sass` of the built library.
Usage: python scripts/sass_counts.py > profiles/r02_sass.txt"""
import re, subprocess, sys, collections
lib = "image_compression_b200/libhpcodec_sm100a.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "UTCBAR", "SYNCS", "HMMA", "FFMA", "MUFU"]
cur, counts, total = None, collections.OrderedDict(), {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1); counts[cur] = collections.Counter(); total[cur] = 0; continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1); total[cur] += 1
        for k in KEYS:
            if op.startswith(k):
                counts[cur][k] += 1
def demangle(n):
    r = subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip()
    r = re.sub(r">\(.*$", ">", r) if ">(" in r else re.sub(r"\(.*$", "", r)
    return r.replace("(anonymous namespace)::", "")
print("# SASS instruction counts per kernel of", lib, "(cuobjdump -sass; sm_100a)")
print("# kernel | instructions | " + " | ".join(KEYS))
for k, c in counts.items():
    if not any(c[x] for x in ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP")) and total[k] < 400:
        continue
    print(demangle(k), "|", total[k], "|", " | ".join(str(c[x]) for x in KEYS))
