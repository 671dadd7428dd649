#!/usr/bin/env python
"""Turn the artefacts of scripts/profile.sh into the committed summaries under profiles/.
usage: scripts/ncu_summarise.py <tag>      (reads gpurun_out/launches_<tag>.csv and gpurun_out/prof_<tag>.ncu-rep)"""
import collections, csv, io, os, subprocess, sys

tag = sys.argv[1]
out_dir = "profiles"
os.makedirs(out_dir, exist_ok=True)

# ---- launch list -> per-kernel totals and shares ------------------------------------------------------
lines = [l for l in open(f"gpurun_out/launches_{tag}.csv") if l.startswith('"')]
agg = collections.OrderedDict()
for x in csv.DictReader(io.StringIO("".join(lines))):
    v = float(x["Metric Value"].replace(",", ""))
    u = x["Metric Unit"]
    v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    k = x["Kernel Name"].split("(")[0]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(v[1] for v in agg.values())
with open(f"{out_dir}/launches_{tag}.md", "w") as f:
    f.write(f"# ncu launch list, {tag}\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` over `python bench.py --steps 2 --warmup 3 "
            "--no-e2e --no-cpu-baseline` (cold-cache, serialised: compare shares).\n\n| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{k}` | {v[0]} | {v[1]:.1f} | {v[1]/v[0]:.1f} | {v[1]/tot:.3f} |\n")

# ---- the step-only launch list (no auxiliary measurements), when present --------------------------------------------
step_csv = f"gpurun_out/launches_{tag}_step.csv"
if os.path.exists(step_csv):
    lines2 = [l for l in open(step_csv) if l.startswith('"')]
    agg2 = collections.OrderedDict()
    for x in csv.DictReader(io.StringIO("".join(lines2))):
        v = float(x["Metric Value"].replace(",", ""))
        u = x["Metric Unit"]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        a = agg2.setdefault(x["Kernel Name"].split("(")[0], [0, 0.0])
        a[0] += 1; a[1] += v
    tot2 = sum(v[1] for v in agg2.values())
    with open(f"{out_dir}/launches_{tag}.md", "a") as f:
        f.write(f"\n## The timed steps only (`--no-aux --no-sweep`: no forward sweeps, k-buffer, model-side or whole-step measurements)\n\n"
                "| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
        for k, v in sorted(agg2.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {v[0]} | {v[1]:.1f} | {v[1]/v[0]:.1f} | {v[1]/tot2:.3f} |\n")

# ---- full capture -> selected raw metrics per kernel ---------------------------------------------------
rep = f"gpurun_out/prof_{tag}.ncu-rep"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__grid_size", "launch__block_size",
        "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
        # issue / pipe utilisation and the warp-state mix (what a statement like "the FMA pipe is N % busy" rests on)
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.per_cycle_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.per_cycle_active",
        "smsp__warps_active.avg.per_cycle_active", "smsp__average_warps_active_per_inst_executed.ratio",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__thread_inst_executed.sum"]
want += [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
want += [h for h in hdr if h.startswith("smsp__sass_thread_inst_executed_op_f") and h.endswith(".sum") and h not in want]
with open(f"{out_dir}/ncu_full_{tag}.md", "w") as f:
    f.write(f"# ncu --set full, {tag}\n\n`ncu --set full --clock-control none --import-source on -k regex:gut_` over one view of "
            "`python bench.py --steps 1 --warmup 3 --views 1 --no-e2e --no-cpu-baseline` (3M Gaussians, 1237x822).\n")
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        f.write(f"\n## `{name}`\n\n| metric | value | unit |\n|---|---|---|\n")
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                f.write(f"| {w} | {r[i]} | {units[i]} |\n")
    # auxiliary captures (k-buffer kernels of the K = 16 configuration, model-side passes): first launch of each kernel
    for extra, what in (("kbuffer", "the K = 16 forward + backward of the auxiliary `sorted_k16` measurement"),
                        ("model", "the model-side passes of the auxiliary `model_side` measurement (N = 3M)")):
        rep2 = f"gpurun_out/prof_{tag}_{extra}.ncu-rep"
        if not os.path.exists(rep2):
            continue
        raw2 = subprocess.run(["ncu", "-i", rep2, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows2 = list(csv.reader(io.StringIO(raw2)))
        if len(rows2) < 3:
            continue
        hdr2, units2, seen = rows2[0], rows2[1], set()
        f.write(f"\n# auxiliary capture: {what}\n")
        for r in rows2[2:]:
            name = r[hdr2.index("Kernel Name")].split("(")[0]
            if name in seen:
                continue
            seen.add(name)
            f.write(f"\n## `{name}`\n\n| metric | value | unit |\n|---|---|---|\n")
            for w in want:
                if w in hdr2:
                    i = hdr2.index(w)
                    f.write(f"| {w} | {r[i]} | {units2[i]} |\n")
print("wrote", f"{out_dir}/launches_{tag}.md", f"{out_dir}/ncu_full_{tag}.md")

# ---- DRAM traffic per launch of each stage's kernel -> profiles/traffic.json (bench.py's roofline.traffic) ----------
import json
STAGE_OF = {"gut_project_kernel": "render::project", "gut_emit_kernel": "render::expand", "gut_tile_bin_sort_kernel": "render::sort",
            "gut_render_kernel": "render::render", "gut_render_backward_kernel": "render-backward::render",
            "gut_view_fold_kernel": "render-backward::project", "gut_project_backward_kernel": "render-backward::project"}
traffic = {"source": f"profiles/ncu_full_{tag}.md (ncu --set full, one view of the bench workload)", "stages": {}}
ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")].split("(")[0].split("::")[-1].split("<")[0]
    for k, st in STAGE_OF.items():
        if name == k:
            b = float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]]
            # a kernel launched per variant (the backward: uniform-origin / per-ray origin): the instance that did the work
            if st not in traffic["stages"] or b > traffic["stages"][st]["dram_bytes_per_launch"]:
                traffic["stages"][st] = {"kernel": k, "dram_bytes_per_launch": b}
json.dump(traffic, open(f"{out_dir}/traffic.json", "w"), indent=1)
print("wrote", f"{out_dir}/traffic.json")
