#!/usr/bin/env python
"""Executed-instruction mix by SASS opcode from an ncu report's source page: ncu_opmix.py report.ncu-rep"""
import csv, io, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]; body = rows[hi + 1:]
si, ci, ti = h.index("Source"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
agg = collections.defaultdict(lambda: [0, 0]); tot = [0, 0]
for r in body:
    if len(r) <= ti: continue
    ops = r[si].split()
    op = ops[1] if ops and ops[0].startswith("@") else (ops[0] if ops else "?")
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("I2F", "HADD2", "F2F", "MUFU")) and "." in op else "")
    v = (int(r[ci] or 0), int(r[ti] or 0))
    agg[op][0] += v[0]; agg[op][1] += v[1]; tot[0] += v[0]; tot[1] += v[1]
print("total warp-inst %d, thread-inst %d, eff %.1f" % (tot[0], tot[1], tot[1] / max(1, tot[0])))
for op, v in sorted(agg.items(), key=lambda x: -x[1][0])[:40]:
    print("  %-14s %6.2f%%  eff %4.1f" % (op, 100.0 * v[0] / tot[0], v[1] / max(1, v[0])))
