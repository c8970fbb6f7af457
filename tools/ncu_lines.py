"""Aggregates `ncu --page source --csv --print-source cuda,sass` per CUDA source line (development aid)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file, hdr, out = None, None, []
for r in rows:
    if len(r) >= 2 and r[0] == "File Name":
        cur_file = r[1].split("/")[-1]; continue
    if len(r) > 4 and r[0] == "Line No":
        hdr = {k: i for i, k in enumerate(r)}; continue
    if hdr and len(r) > 8 and r[0] not in ("", "Line No"):
        def g(k):
            try: return float(r[hdr[k]] or 0)
            except (ValueError, IndexError): return 0.0
        out.append((cur_file, int(r[0]), r[1].strip()[:90], g("Instructions Executed"), g("# Samples"), g("stall_long_sb"), g("stall_no_inst"), g("Thread Instructions Executed")))
ti = sum(o[3] for o in out); ts = sum(o[4] for o in out)
print("total warp-instr %.4g, samples %d" % (ti, ts))
byfile = collections.defaultdict(lambda: [0, 0])
for o in out:
    byfile[o[0]][0] += o[3]; byfile[o[0]][1] += o[4]
for f, (i, s) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("  %-22s instr %5.1f%%  samples %5.1f%%" % (f, 100 * i / ti, 100 * s / ts))
print("top lines by samples:")
for o in sorted(out, key=lambda o: -o[4])[:top]:
    print("%-18s %4d  instr %5.2f%% samp %5.2f%% lsb %5.2f%% noinst %5.2f%% thr/inst %4.1f | %s" % (o[0], o[1], 100 * o[3] / ti, 100 * o[4] / ts, 100 * o[5] / ts, 100 * o[6] / ts, o[7] / max(1, o[3]), o[2]))
