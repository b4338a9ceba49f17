"""Instruction / stall-sample share of the straight-line regions of one kernel, from
`ncu -i REPORT --page source --csv --kernel-name regex:NAME` on stdin."""
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[1]
ia, ie, isamp, it = (hdr.index(k) for k in ("Source", "Instructions Executed", "# Samples", "Avg. Threads Executed"))
data = []
for r in rows[2:]:
    if len(r) <= ie or not r[ie].isdigit():
        break
    data.append(r)
tot = sum(int(r[ie]) for r in data)
ts = sum(int(r[isamp]) for r in data) or 1
print(f"# {rows[0][1][:90]}")
print(f"# {tot} warp instructions executed, {ts} stall samples, {len(data)} SASS lines")
print("#  lines        executions  SASS  inst%  samples%  avg.threads  first instruction")
seg, prev = [], None
for i, r in enumerate(data):
    n = int(r[ie])
    if prev is None or abs(n - prev) > 0.02 * max(n, prev, 1):
        seg.append([i, i, n, 0, 0])
    prev = n
    seg[-1][1] = i
    seg[-1][3] += n
    seg[-1][4] += int(r[isamp])
for a, b, n, ins, sm in seg:
    if ins > tot * 0.01:
        print(f"{a:5d}-{b:<5d} {n:12d} {b - a + 1:5d} {ins / tot * 100:6.1f} {sm / ts * 100:8.1f} {data[a][it]:>10s}    {data[a][ia].strip()[:60]}")
