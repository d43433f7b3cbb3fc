#!/usr/bin/env python
"""Brief text summary of an .ncu-rep: headline metrics, stall reasons and the opcode mix of each captured kernel.
    python scripts/ncu_brief.py gpurun_out/x.ncu-rep [--ops]"""
import collections, csv, io, subprocess, sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'smsp__inst_executed.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum']


def main(rep, ops):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u, d = rows[0], rows[1], rows[2:]
    print("kernels:", [r[h.index("Kernel Name")][:60] for r in d])
    for k in h:
        if k in WANT or ('issue_stalled' in k and 'per_issue_active' in k):
            i = h.index(k)
            vals = [r[i] for r in d]
            if 'issue_stalled' in k and all(float(v) < 0.15 for v in vals):
                continue
            print(f"{k.replace('smsp__average_warps_issue_stalled_', 'stall:').replace('_per_issue_active.ratio', '')} [{u[i]}] {vals}")
    if ops:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
        for n, hi in enumerate(hdr):
            h = rows[hi]
            end = hdr[n + 1] - 1 if n + 1 < len(hdr) else len(rows)
            iS, iE, iM = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
            tot, cnt, smp = 0, collections.Counter(), collections.Counter()
            for r in rows[hi + 1:end]:
                try:
                    e = int(r[iE])
                except Exception:
                    continue
                t = r[iS].split()
                op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
                cnt[op] += e; tot += e; smp[op] += int(r[iM])
            print(f"-- kernel {n}: {tot} warp-instr, {end - hi - 1} SASS lines")
            print("   " + "  ".join(f"{op} {100 * c / tot:.1f}%({smp[op]})" for op, c in cnt.most_common(22)))


if __name__ == "__main__":
    main(sys.argv[1], "--ops" in sys.argv)
