"""Selected metrics of the first kernel in an ncu report -> JSON (profiles/*.json).  usage: ncu_metrics.py rep [note]"""
import csv, json, subprocess, sys
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h, r = rows[0], rows[2]
d = dict(zip(h, r))
def f(k):
    try: return float(d[k].replace(",", ""))
    except Exception: return None
unit = dict(zip(h, rows[1]))
def bytes_of(k):
    v = f(k); u = unit.get(k, "")
    if v is None: return None
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
out = {
    "kernel": d.get("Kernel Name"), "report": rep,
    "gpu_time_ms_under_ncu": (f("gpu__time_duration.sum") or 0) / (1e6 if unit.get("gpu__time_duration.sum") in ("ns", "nsecond") else 1e3 if unit.get("gpu__time_duration.sum", "").startswith("us") else 1),
    "dram_bytes_read": bytes_of("dram__bytes_read.sum"), "dram_bytes_write": bytes_of("dram__bytes_write.sum"),
    "registers": f("launch__registers_per_thread"),
    "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "issue_active_pct": f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "inst_executed": f("smsp__inst_executed.sum"),
    "thread_inst_per_inst": f("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "l2_hit_pct": f("lts__t_sector_hit_rate.pct"), "l1_hit_pct": f("l1tex__t_sector_hit_rate.pct"),
    "pipe_utilisation_pct": {k.split("pipe_")[1].split(".")[0]: f(k) for k in h if k.startswith("sm__inst_executed_pipe_") and k.endswith("pct_of_peak_sustained_active") and f(k) and f(k) > 1},
    "stall_per_issue": {k.split("issue_stalled_")[1].split("_per_issue")[0]: f(k) for k in h if "issue_stalled_" in k and k.endswith("per_issue_active.ratio") and f(k) and f(k) > 0.2},
}
if out["dram_bytes_read"] is not None:
    out["dram_bytes_per_launch"] = out["dram_bytes_read"] + out["dram_bytes_write"]
    out["dram_gbs"] = out["dram_bytes_per_launch"] / out["gpu_time_ms_under_ncu"] / 1e6
if len(sys.argv) > 2: out["note"] = sys.argv[2]
print(json.dumps(out, indent=1))
