"""Regenerates profiles/ from the files one profiling gpurun call leaves in gpurun_out/ (development aid).

    python tools/make_profiles.py [tag]        (tag defaults to r01)

Reads   gpurun_out/<tag>_bench_short.json       un-profiled `bench.py --steps 2 --warmup 1 --no-cpu-baseline`
        gpurun_out/<tag>_launches.csv           ncu --metrics gpu__time_duration.sum launch list of the same command
        gpurun_out/<tag>_render_persistent_raw.csv   ncu --set full capture, --page raw --csv
        gpurun_out/<tag>_src_cs.csv             the same capture, --page source --csv --print-source cuda,sass
Writes  profiles/<tag>_launches.csv / .md, profiles/<tag>_render_persistent.md, profiles/traffic.json
"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
BENCH_CMD = "python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-other-workloads"
ISSUE_PEAK = 148 * 4 * 1.965e9

bench = None
for line in open(os.path.join(G, tag + "_bench_short.json")):
    if line.startswith("{"):
        bench = json.loads(line)
cfg = bench["config"]
spp = cfg["spp_per_step"]

# ---- launch list -------------------------------------------------------------------------------------------------------
rows = [r for r in csv.reader(open(os.path.join(G, tag + "_launches.csv"))) if len(r) > 14]
hdr = {k: i for i, k in enumerate(rows[0])}
agg = collections.OrderedDict()
for r in rows[1:]:
    name = r[hdr["Kernel Name"]].split("(")[0]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[hdr["Metric Value"]]) / 1e6
total = sum(a[1] for a in agg.values())
with open(os.path.join(P, tag + "_launches.csv"), "w") as f:
    f.write(open(os.path.join(G, tag + "_launches.csv")).read())
with open(os.path.join(P, tag + "_launches.md"), "w") as f:
    f.write("# %s — ncu launch list of `%s` (%s, %s pipeline, %d spp per step)\n\n" % (tag, BENCH_CMD, cfg["workload"], cfg["pipeline"], spp))
    f.write("Command: `ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/%s_launches.csv %s`\n" % (tag, BENCH_CMD))
    f.write("(raw list: `profiles/%s_launches.csv`; per-launch times are cold-cache and serialised — compare shares).\n\n" % tag)
    f.write("| kernel | launches | total ms | share |\n|---|---|---|---|\n")
    for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("| `%s` | %d | %.3f | %.1f %% |\n" % (name[-70:], n, ms, 100 * ms / total))
    rp = sum(ms for name, (n, ms) in agg.items() if "render_persistent" in name)
    f.write("\nThe render kernel (`render_persistent<CfgFast,2>`, one launch per %d-frame step; the list holds warm-up, timed, e2e and the\n"
            "1-frame counting launch `CfgFastCount`) is %.1f %% of device time; the rest is bench.py's own L2-flush fills and the\n"
            "upload-time repack kernels, which run once.\n" % (spp, 100 * rp / total))
    f.write("Un-profiled run of the same command: value %.0f Mrays/s, %.2f ms per step (`gpurun_out/%s_bench_short.json`).\n"
            % (bench["value"], bench["ms_per_step"], tag))

# ---- full capture ------------------------------------------------------------------------------------------------------
raw = list(csv.reader(open(os.path.join(G, tag + "_render_persistent_raw.csv"))))
m = dict(zip(raw[0], raw[2]))
u = dict(zip(raw[0], raw[1]))


def val(k):
    return float(m[k].replace(",", ""))


def scaled(k):          # to base units
    v, unit = val(k), u[k]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1}.get(unit, 1)


dur = scaled("gpu__time_duration.sum")
rd, wr = scaled("dram__bytes_read.sum"), scaled("dram__bytes_write.sum")
inst = val("smsp__inst_executed.sum")
samples = spp * cfg["width"] * cfg["height"]
rays = samples * bench["rays_per_sample"]
hbm = bench["roofline"].get("hbm", bench["roofline"])          # the byte figure: a sub-object when the issue roofline is primary
alg = hbm["bytes_per_ray"] * rays
peak = hbm["peak"]
stalls = sorted(((val(k), k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for k in m
                 if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio") and k.find("selected") < 0), reverse=True)
summary = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), os.path.join(G, tag + "_render_persistent_raw.csv")],
                         capture_output=True, text=True).stdout
lines = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), os.path.join(G, tag + "_src_cs.csv"), "30"],
                       capture_output=True, text=True).stdout
with open(os.path.join(P, tag + "_render_persistent.md"), "w") as f:
    f.write("# %s — `render_persistent<CfgFast,2>` (%s, %d frames per launch), ncu --set full\n\n" % (tag, cfg["workload"], spp))
    f.write("Command: `ncu --set full --clock-control none --import-source on -k regex:render_persistent -s 3 -c 1 -o gpurun_out/%s_render_persistent %s`\n" % (tag, BENCH_CMD))
    f.write("(taken after the same command exited 0 without ncu: %.0f Mrays/s, %.2f ms per step). One launch = %d frames x %dx%d\n= %.3e samples = %.3e rays.\n\n"
            % (bench["value"], bench["ms_per_step"], spp, cfg["width"], cfg["height"], samples, rays))
    f.write("## Headline numbers\n\n| quantity | value |\n|---|---|\n")
    f.write("| duration | %.2f ms |\n" % (dur * 1e3))
    f.write("| warp instructions | %.2fe9 -> %.2e /s = %.0f %% of the %.3e /s issue peak (148 SM x 4 x 1.965 GHz) |\n" % (inst / 1e9, inst / dur, 100 * inst / dur / ISSUE_PEAK, ISSUE_PEAK))
    f.write("| `smsp__issue_active` | %.1f %% ; eligible warps per cycle %.2f |\n" % (val("smsp__issue_active.avg.pct_of_peak_sustained_active"), val("smsp__warps_eligible.avg.per_cycle_active")))
    f.write("| resident warps per SM | %.1f (%d CTAs x 8 warps allowed by the register cap: %d registers/thread) |\n"
            % (val("sm__warps_active.avg.per_cycle_active"), int(val("launch__occupancy_limit_registers")), int(val("launch__registers_per_thread"))))
    f.write("| warp execution efficiency | %.1f / 32 lanes |\n" % val("smsp__thread_inst_executed_per_inst_executed.ratio"))
    f.write("| SM busy | %.2f M of %.2f M cycles (min %.2f M, max %.2f M): no straggler tail |\n"
            % (val("sm__cycles_active.avg") / 1e6, val("sm__cycles_elapsed.avg") / 1e6, val("sm__cycles_active.min") / 1e6, val("sm__cycles_active.max") / 1e6))
    f.write("| DRAM | %.2f GB read + %.2f GB write = %.1f GB per launch = %.2f TB/s (%.0f %% of the measured %.2f TB/s) |\n"
            % (rd / 1e9, wr / 1e9, (rd + wr) / 1e9, (rd + wr) / dur / 1e12, 100 * (rd + wr) / dur / 1e9 / peak, peak / 1e3))
    f.write("| algorithmic bytes (bench.py, %.0f B/ray) | %.1f GB per launch -> %.2f TB/s = %.2f of HBM peak; measured DRAM = %.2fx algorithmic |\n"
            % (hbm["bytes_per_ray"], alg / 1e9, alg / dur / 1e12, alg / dur / 1e9 / peak, (rd + wr) / alg))
    f.write("| L1 hit rate / L2 hit rate | %.1f %% / %.1f %% ; L2 throughput %.0f %% of peak |\n"
            % (val("l1tex__t_sector_hit_rate.pct"), val("lts__t_sector_hit_rate.pct"), val("lts__throughput.avg.pct_of_peak_sustained_elapsed")))
    ll, ls = val("sass__inst_executed_local_loads"), val("sass__inst_executed_local_stores")
    f.write("| local memory (register spills + traversal stack) | %.2fe9 loads + %.2fe9 stores = %.1f %% of instructions |\n" % (ll / 1e9, ls / 1e9, 100 * (ll + ls) / inst))
    f.write("| stalls per issue | %s |\n" % ", ".join("%s %.1f" % (k.replace("_", " "), v) for v, k in stalls[:7]))
    f.write("\n## Raw metrics (tools/ncu_summary.py)\n\n```\n%s```\n" % summary)
    f.write("\n## Hot source lines (tools/ncu_lines.py over `--page source --print-source cuda,sass`; inlined lines are counted in every caller)\n\n```\n%s```\n" % lines)

tj_path = os.path.join(P, "traffic.json")
tj = json.load(open(tj_path)) if os.path.exists(tj_path) else {}
tj[cfg["workload"]] = {
    # what bench.py quotes (only when `sass_hash` equals pt_kernel_sass_hash() of the library it runs)
    "sass_hash": bench.get("kernel_sass_hash"),
    "frames_per_launch": spp,
    "dram_bytes_per_launch": rd + wr,
    "warp_instructions_per_launch": int(inst),
    "threads_per_warp_instruction": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "l2_tex_read_sectors_per_launch": int(val("lts__t_sectors_srcunit_tex_op_read.sum")),
    "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "warps_per_sm": val("sm__warps_active.avg.per_cycle_active"),
    "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"),
    "l1tex_throughput_pct": val("l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
    "kernel_ms_under_ncu": dur * 1e3,
    "launch": "render_persistent<CfgFast,2>, %d frames of %dx%d" % (spp, cfg["width"], cfg["height"]),
    "source": "profiles/%s_render_persistent.md (ncu --set full of `%s`)" % (tag, BENCH_CMD),
}
json.dump(tj, open(tj_path, "w"), indent=1)
print(open(os.path.join(P, tag + "_render_persistent.md")).read()[:2600])
