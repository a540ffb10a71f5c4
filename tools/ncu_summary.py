#!/usr/bin/env python
"""prints the metrics of an .ncu-rep (ncu -i ... --page raw --csv) that the round's profile notes quote:
tools/ncu_summary.py gpurun_out/x.ncu-rep [--json out.json --frames F]"""
import csv, io, json, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg.per_second"]
out = {}
for k in want:
    if k in d:
        print(f"{k:75s} {d[k][0]:>20s} {d[k][1]}")
        out[k] = d[k]
stalls = sorted(((float(v[0].replace(",", "")), k) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio") or k.startswith("smsp__average_warp_latency_issue_stalled")), reverse=True)[:6]
for v, k in stalls:
    print(f"  stall {k.split('stalled_')[-1]:60s} {v:.3f}")
if "--json" in sys.argv:
    frames = int(sys.argv[sys.argv.index("--frames") + 1])
    f = lambda k: float(d[k][0].replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(d[k][1], 1)
    json.dump({"kernel": d["Kernel Name"][0], "frames_per_launch": frames, "dram_bytes_per_launch": f("dram__bytes_read.sum") + f("dram__bytes_write.sum"),
               "dram_bytes_read": f("dram__bytes_read.sum"), "dram_bytes_write": f("dram__bytes_write.sum"), "source": rep},
              open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
