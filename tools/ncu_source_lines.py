#!/usr/bin/env python3
"""Which source lines issue the instructions / the local-memory traffic / collect the stall samples of a captured kernel
(development tool):  ncu -i x.ncu-rep --page source --csv --print-source sass,cuda > x.csv;  python tools/ncu_source_lines.py x.csv [N]"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
inst, local, samples, long_sb, no_inst = (collections.Counter() for _ in range(5)); text = {}
path, line, hdr = "?", "?", None
for r in rows:
    if not r: continue
    if r[0] == "File Path": path = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; sass_col = [i for i, n in enumerate(r) if n == "Source"][1]; continue
    if hdr is None or len(r) <= hdr["Instructions Executed"]: continue
    if r[0] != "":
        line = "%s:%s" % (path, r[0]); text[line] = r[1].strip()[:110]; continue
    try: n = int(r[hdr["Instructions Executed"]])
    except ValueError: continue
    inst[line] += n
    samples[line] += int(r[hdr["# Samples"]] or 0)
    long_sb[line] += int(r[hdr["stall_long_sb"]] or 0)
    no_inst[line] += int(r[hdr["stall_no_inst"]] or 0)
    m = re.match(r"\s*(@!?U?P\d+\s+)?(\w+)", r[sass_col])
    if m and m.group(2) in ("LDL", "STL"): local[line] += n
for name, c in (("warp instructions", inst), ("local ld/st warp instructions", local), ("stall samples", samples), ("long_scoreboard samples", long_sb), ("no_instruction samples", no_inst)):
    total = sum(c.values())
    print("--- %s: %d" % (name, total))
    for k, v in c.most_common(top): print("%6.2f %%  %-22s %s" % (100.0 * v / max(total, 1), k, text.get(k, "")))
