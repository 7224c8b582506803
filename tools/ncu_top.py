"""Summarise an .ncu-rep: key raw metrics per launch and the top stalled source lines. Usage: ncu_top.py rep [launch]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_bytes.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'sm__cycles_elapsed.max',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active']
for w in want:
    for i, h in enumerate(hdr):
        if h == w:
            print(f"{h} [{units[i]}]:", [r[i][:40] for r in rows[2:]])
n = len(rows) - 2
sel = [int(sys.argv[2])] if len(sys.argv) > 2 else range(n)
for k in sel:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(k), "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(src)))
    hi = [i for i, r in enumerate(rr) if r and r[0] == 'Address'][0]
    h2 = rr[hi]
    body = [r for r in rr[hi + 1:] if len(r) == len(h2)]
    si, sc = h2.index('# Samples'), h2.index('Source')
    stalls = [i for i, h in enumerate(h2) if h.startswith('stall_') and 'Not Issued' not in h]
    tot = sum(int(r[si] or 0) for r in body)
    print(f"--- launch {k}: {tot} samples, {len(body)} instr")
    agg = {h2[i]: sum(int(r[i] or 0) for r in body) for i in stalls}
    print("   ", sorted(agg.items(), key=lambda x: -x[1])[:8])
    for r in sorted(body, key=lambda r: -int(r[si] or 0))[:18]:
        st = sorted(((int(r[i] or 0), h2[i][6:]) for i in stalls), reverse=True)[:2]
        print(f"    {r[si]:>6} {r[sc][:80]:80s} {st}")
