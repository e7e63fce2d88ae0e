#!/usr/bin/env python
"""Split an ncu source-page capture of one kernel at its barriers: per segment, share of stall samples, share of
executed warp-instructions and the dominant stall reasons.   python tools/ncu_phases.py REP KERNEL_REGEX"""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
import re
pick = next(i for i in secs if re.search(kern, rows[i][1]))
end = next((j for j in secs if j > pick), len(rows))
rows = rows[pick:end]
hdr, data = rows[1], [r for r in rows[2:] if len(r) == len(rows[1])]
iS, iW, iI = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
reasons = [(h, i) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[iW] or 0) for r in data); toti = sum(int(r[iI] or 0) for r in data)
print(rows[0][1] if len(rows[0]) > 1 else kern, "| SASS", len(data), "| samples", tot, "| warp-inst", toti)
acc = 0; acci = 0; start = 0; rs = {h: 0 for h, _ in reasons}
def flush(k, label):
    global acc, acci, rs, start
    top = sorted(rs.items(), key=lambda kv: -kv[1])[:4]
    print("%5d-%5d %-34s samples %5.1f%%  inst %5.1f%%  %s" % (start, k, label[:34], 100.0 * acc / tot, 100.0 * acci / toti,
          " ".join("%s %.0f%%" % (h[6:], 100.0 * v / max(acc, 1)) for h, v in top)))
    acc = 0; acci = 0; rs = {h: 0 for h, _ in reasons}; start = k + 1
for k, r in enumerate(data):
    acc += int(r[iW] or 0); acci += int(r[iI] or 0)
    for h, i in reasons: rs[h] += int(r[i] or 0)
    src = r[iS]
    if 'BAR.' in src or 'RET' in src or 'EXIT' in src: flush(k, src.strip())
flush(len(data) - 1, "end")
