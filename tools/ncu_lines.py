"""Stall samples of one kernel launch in an .ncu-rep, aggregated per CUDA source line.
The SASS page of the report is joined, instruction by instruction, with `nvdisasm --print-line-info` of the kernel's
cubin (extracted from libtts_b200.so; build with -lineinfo).  Usage: ncu_lines.py rep kernel_substr [launch] [top]"""
import csv, io, os, re, subprocess, sys, tempfile, collections
rep, ksub = sys.argv[1], sys.argv[2]
k = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "transformertts_b200", "libtts_b200.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
lines = None
for f in sorted(os.listdir(tmp)):
    if not f.endswith(".cubin"):
        continue
    syms = subprocess.run(["cuobjdump", "-elf", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    m = re.findall(r"\.text\.(\S*%s\S*)" % re.escape(ksub), syms)
    if not m:
        continue
    fun = m[0]
    dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    lines, cur, inside = [], None, False
    for ln in dis.splitlines():
        if ln.startswith(".text."):
            inside = ln.startswith(".text." + fun + ":")
            continue
        if not inside:
            continue
        mm = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
        if mm:
            cur = (os.path.basename(mm.group(1)), int(mm.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            lines.append(cur)
    print(f"kernel {fun}: {len(lines)} instructions")
    break
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(k), "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rr) if r and r[0] == "Address"][0]
h = rr[hi]
body = [r for r in rr[hi + 1:] if len(r) == len(h)]
si = h.index("# Samples")
stalls = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
assert lines is not None and len(lines) == len(body), (None if lines is None else len(lines), len(body))
agg = collections.defaultdict(lambda: [0, collections.Counter()])
for r, loc in zip(body, lines):
    n = int(r[si] or 0)
    agg[loc][0] += n
    for i in stalls:
        agg[loc][1][h[i][6:]] += int(r[i] or 0)
tot = sum(v[0] for v in agg.values())
srcs = {}
print(f"{tot} samples")
for loc, (n, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    text = ""
    if loc:
        fn = os.path.join(ROOT, "transformertts_b200", "csrc", loc[0])
        if os.path.exists(fn):
            srcs.setdefault(fn, open(fn).read().splitlines())
            text = srcs[fn][loc[1] - 1].strip()[:90] if loc[1] - 1 < len(srcs[fn]) else ""
    print(f"{n:>6} {100.0 * n / max(tot, 1):5.1f}%  {str(loc):24s} {text:90s} {c.most_common(2)}")
