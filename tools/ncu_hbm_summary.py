"""Summarise the ncu --page raw CSV of tools/hbm_stages.py: per kernel (second repetition = warm), duration, DRAM bytes,
achieved DRAM GB/s and its fraction of MEASURED_PEAKS.json's hbm_gbs, L2 / L1 bytes.
  python tools/ncu_hbm_summary.py gpurun_out/r2_ncu_hbm_stages_raw.csv > profiles/r2_ncu_hbm_stages.json"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
rows = [r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith("==")]
H = rows[0]
col = {n: i for i, n in enumerate(H)}


def num(r, name):
    v = r[col[name]].replace(",", "")
    return float(v) if v not in ("", "n/a") else None


out = []
for r in rows[2:]:
    if len(r) < len(H):
        continue
    dur = num(r, "gpu__time_duration.sum")
    rd, wr = num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
    out.append({"id": int(r[col["ID"]]), "kernel": r[col["Kernel Name"]][:90], "grid": r[col["Grid Size"]], "block": r[col["Block Size"]],
                "us": dur / 1e3, "dram_read_B": rd, "dram_write_B": wr,
                "dram_GBps": (rd + wr) / dur, "frac_of_measured_hbm_peak": (rd + wr) / dur / peak,
                "l2_bytes": num(r, "lts__t_bytes.sum"), "l1_bytes": num(r, "l1tex__t_bytes.sum"),
                "dram_pct_of_peak_ncu": num(r, "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
                "sm_pct_of_peak_ncu": num(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
                "regs": num(r, "launch__registers_per_thread")})
half = len(out) // 2
print(json.dumps({"source": os.path.basename(sys.argv[1]), "command": "ncu --metrics ... --clock-control none python tools/hbm_stages.py "
                  "(second of two repetitions; per-launch, cold-ish caches, serialised)", "hbm_peak_GBps_measured": peak,
                  "launches": out[half:]}, indent=1))
