"""Aggregates `ncu --page source --csv --print-source cuda,sass` per CUDA source line (development aid)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file, hdr, out = None, None, []
for r in rows:
    if len(r) >= 2 and r[0] in ("File Name", "File Path"):
        cur_file = r[1].split("/")[-1]; continue
    if len(r) > 4 and r[0] == "Line No":
        hdr = {k: i for i, k in enumerate(r)}; continue
    if hdr and len(r) > 8 and r[0] not in ("", "Line No"):
        def g(k):
            try: return float(r[hdr[k]] or 0)
            except (ValueError, IndexError): return 0.0
        out.append((cur_file, int(r[0]), r[1].strip()[:90], g("Instructions Executed"), g("# Samples"), g("stall_long_sb"), g("stall_no_inst"), g("Thread Instructions Executed")))
# exclusive counts: SASS rows (empty "Line No") listed under the CUDA line they are attributed to
excl = collections.defaultdict(lambda: [0.0, 0.0, 0])
cur_file, hdr, cur_line = None, None, None
for r in rows:
    if len(r) >= 2 and r[0] in ("File Name", "File Path"):
        cur_file = r[1].split("/")[-1]; continue
    if len(r) > 4 and r[0] == "Line No":
        hdr = {k: i for i, k in enumerate(r)}; continue
    if not hdr or len(r) < 9:
        continue
    if r[0] != "":
        cur_line = (cur_file, int(r[0]), r[1].strip()[:80]); continue
    try:
        e = excl[cur_line]; e[0] += float(r[hdr["Instructions Executed"]] or 0); e[1] += float(r[hdr["# Samples"]] or 0); e[2] += 1
    except (ValueError, IndexError):
        pass
te = sum(v[0] for v in excl.values()); tse = sum(v[1] for v in excl.values())
if "--excl" in sys.argv:
    print("EXCLUSIVE (each SASS instruction counted once): total warp-instr %.4g, %d SASS instructions" % (te, sum(v[2] for v in excl.values())))
    byf = collections.defaultdict(lambda: [0.0, 0.0, 0])
    for k, v in excl.items():
        byf[k[0]][0] += v[0]; byf[k[0]][1] += v[1]; byf[k[0]][2] += v[2]
    for f, v in sorted(byf.items(), key=lambda kv: -kv[1][0]):
        print("  %-28s instr %5.1f%%  samples %5.1f%%  sass %d" % (f, 100 * v[0] / te, 100 * v[1] / tse, v[2]))
    for k, v in sorted(excl.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%-16s %4d instr %5.2f%% samp %5.2f%% sass %3d | %s" % (k[0], k[1], 100 * v[0] / te, 100 * v[1] / tse, v[2], k[2]))
    sys.exit(0)
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
