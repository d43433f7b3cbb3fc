#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i ... --page raw --csv`) into a small tracked CSV under profiles/
and update profiles/traffic.json (dram bytes per launch of the sample kernel, used by bench.py's roofline.traffic).
    python scripts/ncu_summary.py gpurun_out/r01h_kernels.ncu-rep profiles/r01_ncu_kernels.csv"""
import csv, io, json, os, subprocess, sys

KEYS = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]


def main(rep, out_csv):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [(k, hdr.index(k)) for k in KEYS if k in hdr]
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([k + (f" [{units[i]}]" if units[i] else "") for k, i in cols])
        for r in data:
            w.writerow([r[i] for _, i in cols])
    # traffic of the sample kernel (first capture of the NB=0 or 256 variant)
    tp = os.path.join(os.path.dirname(out_csv), "traffic.json")
    traffic = json.load(open(tp)) if os.path.isfile(tp) else {}
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    vals = [float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]] for r in data if "ring_sample_band" in r[ik]]
    if vals:
        traffic["ring_sample_b128"] = sum(vals) / len(vals)
        traffic["source"] = os.path.basename(rep) + ": mean dram__bytes_read.sum + dram__bytes_write.sum over %d launches" % len(vals)
        json.dump(traffic, open(tp, "w"), indent=1)
    print("wrote", out_csv, "and", tp)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
