#!/usr/bin/env python3
"""Hot SASS instructions of one profiled launch: python tools/ncu_hot.py <report.ncu-rep> <kernel-regex> [launch-skip] [min-pct]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
minpct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.7
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass', '--kernel-name', 'regex:' + pat,
                      '--launch-skip', skip, '--launch-count', '1'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = rows[1]
iS, iE, iT, iN = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('# Samples')
stalls = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
body = [r for r in rows[2:] if len(r) > iN and r[iN].isdigit()]
tot = sum(int(r[iN]) for r in body); texec = sum(int(r[iE]) for r in body); tthr = sum(int(r[iT]) for r in body)
print(f"# {rows[0][1][:90]}: {len(body)} SASS instructions, {texec} warp-instructions executed, {tthr / max(texec, 1):.2f} active lanes, {tot} samples")
agg = {}
for i, h in stalls:
    agg[h] = sum(int(r[i]) for r in body if r[i].isdigit())
print("# stall mix: " + ", ".join(f"{h[6:]} {100 * v / max(tot, 1):.1f}%" for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
for k, r in enumerate(body):
    s = int(r[iN])
    if 100 * s / max(tot, 1) < minpct:
        continue
    e, t = int(r[iE]), int(r[iT])
    top = sorted(((int(r[i]) if r[i].isdigit() else 0, h[6:]) for i, h in stalls), reverse=True)[:2]
    print(f"{k:5d} {r[iS][:64]:64s} exec {e:9d} lanes {t / max(e, 1):5.1f} samp {100 * s / tot:5.2f}%  " + " ".join(f"{h}:{v}" for v, h in top if v))
