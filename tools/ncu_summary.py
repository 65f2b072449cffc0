"""tools/ncu_summary.py <raw.csv> <out.csv> [traffic.json] — condense `ncu --page raw --csv` output to the columns the
roofline discussion uses (one row per captured launch) and refresh profiles/traffic.json."""
import csv
import os
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
K = {
    "time_us": "gpu__time_duration.sum",
    "dram_read_B": "dram__bytes_read.sum",
    "dram_write_B": "dram__bytes_write.sum",
    "regs": "launch__registers_per_thread",
    "grid": "launch__grid_size",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "fp64_pipe_pct": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "issue_active_pct": "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "warp_inst": "smsp__inst_executed.sum",
    "local_ld_inst": "smsp__sass_inst_executed_op_local_ld.sum",
    "local_st_inst": "smsp__sass_inst_executed_op_local_st.sum",
}
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = [["kernel"] + list(K) + ["fp64_flops", "fp64_tflops", "dram_GBps"]]
traffic = {}
for r in data:
    name = r[col["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "").strip()  # "void k_eval_sum<2>(...)" -> k_eval_sum
    row = [name]
    vals = {}
    for k, m in K.items():
        if m not in col:
            row.append("")
            continue
        v = float(r[col[m]].replace(",", ""))
        v *= scale.get(units[col[m]], 1.0)
        vals[k] = v
        row.append("%.6g" % v)
    cyc = float(r[col["smsp__cycles_elapsed.max"]])
    fl = (float(r[col["derived__smsp__sass_thread_inst_executed_op_dfma_pred_on_x2"]])
          + float(r[col["smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed"]])
          + float(r[col["smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed"]])) * cyc
    row += ["%.6g" % fl, "%.4g" % (fl / vals["time_us"] / 1e6), "%.5g" % ((vals["dram_read_B"] + vals["dram_write_B"]) / vals["time_us"] / 1e3)]
    out.append(row)
    traffic.setdefault(name, []).append((vals["dram_read_B"] + vals["dram_write_B"], fl))
csv.writer(open(sys.argv[2], "w")).writerows(out)
if len(sys.argv) > 3:
    t = {}
    if "k_eval_sum" in traffic:
        t["k_eval_sum_dram_bytes_per_launch"] = int(traffic["k_eval_sum"][0][0])  # first k_eval_sum = the Dalitz one
    for k in ("k_amp3_waves", "k_amp4_cache", "k_amp4_sf", "k_amp4_update", "k_eval_fast_amp4", "k_eval_logleaf"):
        if k in traffic:
            t[k + "_fp64_flops_per_launch"] = traffic[k][0][1]
            t[k + "_dram_bytes_per_launch"] = int(traffic[k][0][0])
    import subprocess
    t["source"] = "ncu --set full --clock-control none, tools/profile_once.py at 10^7 events (profiles/r02_ncu_full_summary.csv)"
    head = ""
    try:
        head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    except Exception:
        pass
    if not head:  # the GPU box gets a snapshot without .git: tools/final_capture.sh is sent with the hash in .git_head
        here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        if os.path.exists(os.path.join(here, ".git_head")):
            head = open(os.path.join(here, ".git_head")).read().strip()
    t["git_head"] = head
    json.dump(t, open(sys.argv[3], "w"), indent=1)
print("rows", len(out) - 1)
