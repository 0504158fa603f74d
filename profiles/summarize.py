#!/usr/bin/env python
"""Round summary from the ncu artefacts a gpurun call brought back:
  summarize.py TAG launches.csv full.ncu-rep bench.json ["bench.py arguments of the profiled command"]
  ->  profiles/TAG_launches.csv, TAG_traffic.json, TAG_summary.md"""
import csv, io, json, os, re, subprocess, sys
tag, launches, rep, bench = sys.argv[1:5]
cmd_args = sys.argv[5] if len(sys.argv) > 5 else "--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-workloads"
here = os.path.dirname(os.path.abspath(__file__))
rows = [r for r in csv.reader(open(launches, errors="replace")) if len(r) > 10 and r[0].isdigit()]
agg = {}
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "").replace("brb::", "").strip()
    us = float(r[-1].replace(",", "")) / (1.0 if "us" in (r[-3] + r[-2]).lower() else 1.0)
    unit = r[-2]
    if unit in ("ns", "nsecond"): us /= 1e3
    elif unit in ("ms", "msecond"): us *= 1e3
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += us
open(os.path.join(here, tag + "_launches.csv"), "w").write(open(launches, errors="replace").read())
generic = ([k for k in agg if "k_front<1>" in k] if "k_front_fast" in agg else []) + [k for k in agg if k.startswith(("k_index_", "k_validate_batch", "k_pack_genome", "at::"))]    # once per index / upload: not part of a step
step_total = sum(v[1] for k, v in agg.items() if k not in generic)
md = [f"# {tag}: ncu summary\n"]
md.append(f"Launch list: `ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv python bench.py {cmd_args}`")
md.append(f"(raw list: {tag}_launches.csv; per-launch times are cold-cache + serialised, compare shares; `--no-e2e` keeps the list to the resident leg the bench line's `value` is timed on).  The generic front kernel `k_front<1>` (short-read workloads) runs ONCE, untimed,")
md.append("to obtain the reference-defined algorithmic byte count; shares exclude it and the once-per-index / once-per-upload kernels (`k_index_*`, `k_validate_batch`).\n")
md.append("| kernel | launches | total us | share of step | avg us |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    md.append(f"| {k} | {v[0]} | {v[1]:.1f} | {'-' if k in generic else f'{v[1] / step_total:.3f}'} | {v[1] / v[0]:.1f} |")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(out)))
h = rr[0]
want = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "launch__registers_per_thread", "lts__t_sector_hit_rate.pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"]
units = dict(zip(h, rr[1]))
md.append(f"\n`ncu --set full --clock-control none --import-source on` of one step at the same size ({tag}_full.ncu-rep is scratch, not committed):\n")
md.append("| kernel | " + " | ".join(f"{w} [{units.get(w, '')}]" for w in want) + " |\n|" + "---|" * (len(want) + 1))
traffic = {}
def to_bytes(v, u):
    v = float(v.replace(",", "")); return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
for r in rr[2:]:
    d = dict(zip(h, r)); name = re.sub(r"\(.*", "", d["Kernel Name"]).replace("void ", "").replace("brb::", "").strip()
    md.append(f"| {name} | " + " | ".join(d.get(w, "") for w in want) + " |")
    traffic[name] = to_bytes(d["dram__bytes_read.sum"], units["dram__bytes_read.sum"]) + to_bytes(d["dram__bytes_write.sum"], units["dram__bytes_write.sum"])
b = json.load(open(bench))
# provenance: bench.py only quotes this capture while the kernel sources, the workload and its size are the ones it was taken on
json.dump({"engine_src_sha": b["roofline"].get("engine_src_sha"), "workload": b["config"].get("id"), "names": b["config"].get("names_per_gpu"),
           "kernels": traffic}, open(os.path.join(here, tag + "_traffic.json"), "w"), indent=1)
md.append("\nDRAM traffic per launch (read+write): " + ", ".join(f"{k}: {v / 1e9:.2f} GB" for k, v in traffic.items()) +
          f"; sum {sum(traffic.values()) / 1e9:.2f} GB vs {b['roofline']['alg_bytes_per_step'] / 1e9:.2f} GB algorithmic.\n")
md.append("Live bench line of the same build (no profiler):\n\n```\n" + json.dumps(b, indent=1) + "\n```")
open(os.path.join(here, tag + "_summary.md"), "w").write("\n".join(md) + "\n")
print("\n".join(md[:14])); print(traffic)
