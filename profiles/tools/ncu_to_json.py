"""ncu report of the step kernel (ncu --set full ... -k regex:go2_step_kernel -c 1) -> profiles/step_kernel_ncu.json, the per-launch
figures bench.py quotes (roofline.traffic, issue_frac) so that they come from a capture of the CURRENT build, not a constant.
usage: python profiles/tools/ncu_to_json.py gpurun_out/X.ncu-rep "command the capture profiled" """
import csv
import json
import os
import subprocess
import sys

rep, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
h, u, v = r[0], r[1], r[-1]


def get(k):
    i = h.index(k)
    x = float(v[i].replace(",", ""))
    unit = u[i]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "msecond": 1e-3, "usecond": 1e-6, "second": 1.0, "nsecond": 1e-9, "ms": 1e-3, "us": 1e-6, "s": 1.0, "ns": 1e-9}.get(unit, 1.0)
    return x * scale


d = {
    "source": f"ncu --set full --clock-control none, {os.path.basename(rep)}; {cmd}",
    "kernel": v[h.index("Kernel Name")].split("(")[0],
    "duration_s_under_ncu": get("gpu__time_duration.sum"),
    "dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
    "dram_bytes_read": get("dram__bytes_read.sum"), "dram_bytes_write": get("dram__bytes_write.sum"),
    "warp_instr_per_launch": get("smsp__inst_issued.sum"),
    "l1_hit_pct": get("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": get("lts__t_sector_hit_rate.pct"),
    "long_scoreboard_per_issue": get("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
    "active_lanes": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "local_load_sectors": get("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum"), "local_store_sectors": get("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum"),
    "registers": get("launch__registers_per_thread"),
}
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "step_kernel_ncu.json")
json.dump(d, open(path, "w"), indent=1)
print(json.dumps(d, indent=1))
