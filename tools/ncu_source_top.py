"""Condense `ncu --page source --csv` (SASS view) into the lines that matter: per opcode class and per source line,
the share of warp-stall samples and of executed instructions.    python tools/ncu_source_top.py in.csv out.txt"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], newline="")))
hdr_i = next(i for i, r in enumerate(rows) if any("Sampling" in c for c in r))
hdr = rows[hdr_i]
out = open(sys.argv[2], "w")
out.write("columns: %s\n" % hdr)
col = {c: i for i, c in enumerate(hdr)}
samp = next(c for c in hdr if "Sampling" in c and "All" in c)
execd = next((c for c in hdr if c.strip() == "Instructions Executed"), None)
src = next((c for c in hdr if c.strip() in ("Source", "SASS")), hdr[1])
stall_cols = [c for c in hdr if c.startswith("stall_") or c.lower().startswith("warp stall") and c != samp]
tot = collections.Counter()
by_op = collections.Counter()
by_op_exec = collections.Counter()
lines = []
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        continue
    try:
        s = float(r[col[samp]] or 0)
        e = float(r[col[execd]] or 0) if execd else 0
    except ValueError:
        continue
    text = r[col[src]]
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", text)
    op = m.group(1) if m else "?"
    by_op[op] += s
    by_op_exec[op] += e
    tot["samples"] += s
    tot["exec"] += e
    lines.append((s, e, text, r))
out.write("total samples %d, instructions executed %d\n" % (tot["samples"], tot["exec"]))
out.write("\nby opcode: samples%%  exec%%\n")
for op, s in by_op.most_common(25):
    out.write("%-14s %6.2f %6.2f\n" % (op, 100 * s / max(1, tot["samples"]), 100 * by_op_exec[op] / max(1, tot["exec"])))
out.write("\ntop instructions by samples\n")
for s, e, text, r in sorted(lines, key=lambda x: -x[0])[:60]:
    extra = {c: r[col[c]] for c in stall_cols if r[col[c]] not in ("", "0")}
    out.write("%7d %9d  %-60s %s\n" % (s, e, text[:60], extra))
