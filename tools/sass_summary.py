"""Per-kernel counts of the SASS mnemonics that show which hardware paths a kernel uses (B200_PROFILING.md):
UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor load / store, UBLKCP = bulk copy,
STAS = st.async (DSMEM), SYNCS = mbarrier, UCGABAR = cluster barrier, HMMA = legacy mma.sync, LDGSTS = cp.async.
Usage: python tools/sass_summary.py [> profiles/rNN_sass.txt]"""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "transformertts_b200", "libtts_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "STAS", "SYNCS", "UCGABAR", "HMMA", "LDGSTS", "MUFU", "REDG", "ATOMG"]
cur, counts, sizes = None, collections.defaultdict(collections.Counter), collections.Counter()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("tts::", "")
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if m and cur:
        sizes[cur] += 1
        op = m.group(1)
        for k in KEYS:
            if op.startswith(k):
                counts[cur][k] += 1
print(f"# cuobjdump -sass {os.path.relpath(so, ROOT)}: instructions per kernel and counts of the marker mnemonics")
print(f"{'kernel':58s} {'instr':>6s}  " + " ".join(f"{k:>8s}" for k in KEYS))
for k in sorted(sizes):
    print(f"{k[:58]:58s} {sizes[k]:6d}  " + " ".join(f"{counts[k][x]:8d}" for x in KEYS))
