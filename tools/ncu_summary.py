"""Summarise one `ncu --set full --import-source on` capture of the head kernel for profiles/:
  python tools/ncu_summary.py gpurun_out/head_v10.ncu-rep profiles/r1_ncu_head_full_summary.json [profiles/..._roles.txt]
Writes the metrics DESIGN.md quotes (duration, DRAM bytes = roofline.traffic, tensor-pipe activity, ...) as JSON and,
optionally, the warp-stall samples per role (the source page split at each setmaxnreg) as text."""
import collections
import csv
import os
import io
import json
import subprocess
import sys

KEEP = [
    "Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_reads.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_writes.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
    "launch__cluster_dim_x", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def num(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


def main():
    rep, out_json = sys.argv[1], sys.argv[2]
    rows = page(rep, "raw")
    hdr, unit, val = rows[0], rows[1], rows[2]
    m = {h: {"unit": u, "value": v} for h, u, v in zip(hdr, unit, val) if h in KEEP}
    traffic = sum(float(m[k]["value"]) * UNIT[m[k]["unit"]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    doc = {
        "command": os.environ.get("NCU_SUMMARY_COMMAND", "ncu --set full --clock-control none --import-source on -k regex:head_kernel "
                                                         "-s 2 -c 2 python tools/refine_step.py (first of the two captured launches)"),
        "workload": "configs[1]: 1,000,000 pairs, 512-d fp32 table (M=200,000), K=5, one launch of the FP32_REFINE filter "
                    "(tc::head_kernel, pair mode, IEEE-half operands, staged feature tiles)",
        "report": rep,
        "traffic_bytes_per_launch": traffic,
        "algorithmic_bytes_per_launch": 4108000000,
        "note": "DRAM traffic stays below the algorithmic gather bytes (every row gather counted); the writes are the "
                "4-byte utilities plus evictions of the L2-resident staging scratch. Tensor-bound kernel.",
        "metrics": m,
    }
    with open(out_json, "w") as f:
        json.dump(doc, f, indent=1)
    print("traffic", traffic, "duration", m.get("gpu__time_duration.sum"))
    if len(sys.argv) > 3:
        rows = page(rep, "source")
        hdr = rows[1]
        data = [r for r in rows[2:] if len(r) >= len(hdr)]  # a report with several kernels repeats the header block: first kernel only
        cut = next((i for i, r in enumerate(rows[2:]) if len(r) < len(hdr)), None)
        if cut is not None:
            data = [r for r in rows[2:2 + cut]]
        isrc, isamp = hdr.index("Source"), hdr.index("# Samples")
        stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        marks = [i for i, r in enumerate(data) if "USETMAXREG" in r[isrc]] + [len(data)]
        tot = sum(num(r[isamp]) for r in data)
        with open(sys.argv[3], "w") as f:
            f.write(f"{rows[0][1]}\nwarp-stall samples per role (source page split at each setmaxnreg; roles in code order: "
                    "weight producer, MMA issuer, peer relay, tile loader, epilogue, gather producers, tail)\n")
            for a, b in zip(marks[:-1], marks[1:]):
                s = sum(num(r[isamp]) for r in data[a:b])
                st = collections.Counter()
                for r in data[a:b]:
                    for i in stall:
                        st[hdr[i][6:]] += num(r[i])
                f.write(f"\nSASS lines {a}-{b}: {int(s)} samples ({100 * s / tot:.1f} %)  "
                        + ", ".join(f"{k} {int(v)}" for k, v in st.most_common(5)) + "\n")
                for k in range(a, b):
                    r = data[k]
                    if s and num(r[isamp]) >= 0.03 * s:
                        top = max(((hdr[i][6:], num(r[i])) for i in stall), key=lambda kv: kv[1])
                        f.write(f"    {int(num(r[isamp])):6d}  {r[isrc][:70]:70s} {top[0]}\n")


if __name__ == "__main__":
    main()
