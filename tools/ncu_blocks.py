"""Heaviest basic blocks of a kernel from `ncu -i <rep> --page source --csv --print-source sass` (development aid):
consecutive SASS instructions with the same execution count are one block; prints length, executions, share of the
kernel's warp instructions, active lanes and stall-sample share.   python tools/ncu_blocks.py <sass.csv> [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 15
H = {k: i for i, k in enumerate(rows[1])}
data = []
for r in rows[2:]:
    if len(r) < 10:
        continue
    try:
        a = int(r[H["Address"]], 16) if r[H["Address"]].startswith("0x") else int(r[H["Address"]])
    except ValueError:
        continue
    data.append((a, r[H["Source"]].strip(), float(r[H["Instructions Executed"]] or 0), float(r[H["Thread Instructions Executed"]] or 0), float(r[H["# Samples"]] or 0)))
ti, ts, base = sum(d[2] for d in data), sum(d[4] for d in data), data[0][0]
blocks, cur = [], None
for a, src, i, t, s in data:
    if cur and abs(cur["i"] - i) <= 1e-9 * max(1.0, i):
        cur["n"] += 1; cur["s"] += s; cur["t"] += t; cur["end"] = a
    else:
        if cur: blocks.append(cur)
        cur = {"start": a, "end": a, "i": i, "n": 1, "s": s, "t": t, "first": src}
blocks.append(cur)
blocks = sorted((b for b in blocks if b["i"] > 0), key=lambda b: -b["i"] * b["n"])
print("%d SASS instructions, %.3e warp instructions, %d stall samples" % (len(data), ti, ts))
for b in blocks[:top]:
    w = b["i"] * b["n"]
    print("%6x-%6x  %3d instr x %.2e = %.2e (%4.1f %%)  lanes %4.1f  samples %4.1f %%  | %s"
          % (b["start"] - base, b["end"] - base, b["n"], b["i"], w, 100 * w / ti, b["t"] / max(1.0, w), 100 * b["s"] / ts, b["first"][:48]))
