#!/usr/bin/env python
"""Warp-stall samples of one kernel of an .ncu-rep, by SASS region and by reason (read here, no GPU needed):
python profiles/ncu_stalls.py gpurun_out/x.ncu-rep [kernel-regex] [launch-index]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
rx = sys.argv[2] if len(sys.argv) > 2 else "."
idx = sys.argv[3] if len(sys.argv) > 3 else "1"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f"::regex:{rx}:{idx}"],
                     capture_output=True, text=True).stdout
lines = out.splitlines()
print(lines[0][:200])
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
total = 0
per = []
for r in rows[1:]:
    if len(r) < len(hdr) or not r[col["# Samples"]].isdigit():
        continue
    n = int(r[col["# Samples"]] or 0)
    total += n
    for q in reasons:
        tot[q] += int(r[col[q]] or 0)
    per.append((n, r[col["Address"]], r[col["Source"]], {q: int(r[col[q]] or 0) for q in reasons if int(r[col[q]] or 0)},
                int(r[col["Instructions Executed"]] or 0)))
print("instructions:", len(per), "samples:", total)
print("by reason:", ", ".join(f"{q[6:]} {100 * c / max(total, 1):.1f}%" for q, c in tot.most_common(12)))
# cumulative profile over the instruction stream in 40 chunks
n = len(per)
chunk = max(n // 40, 1)
print("\nsamples by position in the SASS stream (chunks of %d instructions): first opcode of the chunk, share, top reasons" % chunk)
for a in range(0, n, chunk):
    seg = per[a:a + chunk]
    s = sum(p[0] for p in seg)
    c = collections.Counter()
    for p in seg:
        c.update(p[3])
    ex = max(p[4] for p in seg)
    print(f"{a:5d} {100 * s / max(total, 1):5.1f}%  exec/instr {ex:9d}  " + ", ".join(f"{q[6:]} {100 * v / max(s, 1):.0f}%" for q, v in c.most_common(3)))
print("\ntop 40 instructions")
for p in sorted(per, key=lambda p: -p[0])[:40]:
    print(f"{100 * p[0] / max(total, 1):5.2f}%  {p[1]}  {p[2][:90]}  " + ", ".join(f"{q[6:]} {v}" for q, v in sorted(p[3].items(), key=lambda kv: -kv[1])[:3]))
