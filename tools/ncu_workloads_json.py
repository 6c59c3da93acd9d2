#!/usr/bin/env python3
"""Rebuilds profiles/ncu_render_kernel.json — the ncu figures bench.py's roofline record quotes — from the per-capture
summaries profiles/r02_ncu_full_*.csv (tools/ncu_summary.py output). Development tool: python tools/ncu_workloads_json.py"""
import csv, json, os, re
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNITS = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1, "ms": 1, "us": 1e-3}


def load(path):
    d = {}
    for r in csv.reader(open(os.path.join(REPO, path))):
        if len(r) >= 3 and not r[0].startswith("#") and r[0] != "metric":
            d[r[0]] = (r[1], ",".join(r[2:]) if r[0] == "Kernel Name" else r[2])   # ncu_summary.py does not quote the name's commas
    return d


def entry(path, what):
    d = load(path)
    num = lambda k: float(d[k][1]) * UNITS.get(d[k][0], 1)
    stalls = {k.split("stalled_")[1].split("_per_issue")[0]: round(float(d[k][1]), 3) for k in d if "issue_stalled" in k and float(d[k][1]) >= 0.05}
    name = re.sub(r"\((int|bool)\)", "", d["Kernel Name"][1]).replace("void ", "")
    return {
        "kernel": name[:name.index(">(") + 1] if ">(" in name else name.split("(")[0], "capture": what, "source": path,
        "gpu_time_ms_under_ncu": num("gpu__time_duration.sum"),
        "dram_bytes_per_launch": int(num("dram__bytes_read.sum") + num("dram__bytes_write.sum")),
        "dram_read_bytes": int(num("dram__bytes_read.sum")), "dram_write_bytes": int(num("dram__bytes_write.sum")),
        "l2_bytes_per_launch": int(num("lts__t_bytes.sum")), "l1_bytes_per_launch": int(num("l1tex__t_bytes.sum")),
        "warp_instructions_per_launch": int(num("smsp__inst_executed.sum")), "thread_instructions_per_launch": int(num("smsp__thread_inst_executed.sum")),
        "threads_per_instruction": float(d["smsp__thread_inst_executed_per_inst_executed.ratio"][1]),
        "issue_active_pct": round(float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"][1]), 2),
        "l1_hit_pct": round(float(d["l1tex__t_sector_hit_rate.pct"][1]), 1), "l2_hit_pct": round(float(d["lts__t_sector_hit_rate.pct"][1]), 1),
        "l1_throughput_pct": round(float(d["l1tex__throughput.avg.pct_of_peak_sustained_elapsed"][1]), 1),
        "local_ld_st_warp_instructions": int(num("smsp__sass_inst_executed_op_local_ld.sum") + num("smsp__sass_inst_executed_op_local_st.sum")),
        "global_ld_warp_instructions": int(num("smsp__sass_inst_executed_op_global_ld.sum")),
        "registers_per_thread": int(float(d["launch__registers_per_thread"][1])), "warps_eligible_per_cycle": round(float(d["smsp__warps_eligible.avg.per_cycle_active"][1]), 2),
        "stalls_per_issue": dict(sorted(stalls.items(), key=lambda kv: -kv[1])),
    }


CAP = "ncu --set full --metrics lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__thread_inst_executed.sum --clock-control none --import-source on -k regex:%s -s 3 -c 1 on python tools/ncu_frame.py %s (B200RT_DEEP_QUEUE=%s), %s"
out = {
    "what": "ncu figures of the render kernel per bench workload; bench.py's roofline record reads the entry of the workload it times (the '/queued' entry when the frame went through the deep-pixel queue). Instruction and byte counts of a launch do not depend on the box; times under ncu are not bench values.",
    "launch_list": "profiles/r02_launches_bench.csv (bench.py --steps 2 --warmup 3 under ncu --metrics gpu__time_duration.sum --clock-control none, tools/gpu_call28.sh: render_kernel<0,3,32> 1.054 ms avg x 28, fxaa_kernel 0.0356 ms x 31, tile_schedule_kernel 0.0046 ms x 33 -> the render kernel is 96.3 % of a frame's kernel time; the live bench of the same call measures 1.028 of 1.072 ms = 95.9 %; the 3 render_queue_kernel launches are AUTO's comparison frames, the 2 render_kernel<1,8,64> the counting launches)",
    "workloads": {
        "sponza1080": entry("profiles/r02_ncu_full_render_sponza1080.csv", CAP % ("render_kernel", "sponza1080", "off", "tools/gpu_call14.sh")),
        "dielectrics": entry("profiles/r02_ncu_full_render_dielectrics.csv", CAP % ("render_kernel", "dielectrics", "off", "tools/gpu_call14.sh")),
        "dielectrics/queued": entry("profiles/r02_ncu_full_render_queue_dielectrics.csv", CAP % ("render_queue_kernel", "dielectrics", "all: AUTO's queued candidate", "tools/gpu_call31.sh, -s 5")),
        "sibenik4k": entry("profiles/r02_ncu_full_render_sibenik4k.csv", CAP % ("render_kernel", "sibenik4k", "off", "tools/gpu_call14.sh")),
    },
    "fxaa": entry("profiles/r02_ncu_full_fxaa_sponza1080.csv", CAP % ("fxaa_kernel", "sponza1080", "off", "tools/gpu_call14.sh")),
}
json.dump(out, open(os.path.join(REPO, "profiles", "ncu_render_kernel.json"), "w"), indent=1)
for k, v in out["workloads"].items():
    print(k, v["kernel"], v["thread_instructions_per_launch"], v["dram_bytes_per_launch"])
