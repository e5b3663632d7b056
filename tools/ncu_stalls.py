"""Summarise the source page of an .ncu-rep: stall totals and the hottest SASS lines of one kernel launch.
usage: python tools/ncu_stalls.py report.ncu-rep kernel_regex [launch_index] [top_n]"""
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx, "--launch-skip", str(idx),
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[2:] if len(r) >= len(hdr)]
tot = sum(int(r[isamp]) for r in data)
print("instructions", len(data), "samples", tot)
for i in stalls:
    s = sum(int(r[i] or 0) for r in data)
    if s > tot * 0.01:
        print("  %-24s %6d %5.1f%%" % (hdr[i], s, 100.0 * s / tot))
top = sorted(range(len(data)), key=lambda j: -int(data[j][isamp]))[:topn]
for j in sorted(top):
    r = data[j]
    st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stalls), reverse=True)[:2]
    print("%5d %6s %9s %-80s %s" % (j, r[isamp], r[iex], r[isrc][:80], st))
