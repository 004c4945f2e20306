"""Condense `ncu -i X.ncu-rep --page raw --csv` (one row per launch, one column per metric) into the handful of
metrics DESIGN.md argues from.  Missing metrics are skipped, not fatal.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv
    python tools/summarize_ncu.py gpurun_out/prof_raw.csv "header line" > profiles/rNN_ncu_full_summary.txt
"""
import csv
import sys

KEYS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum",
    "l1tex__m_xbar2l1tex_read_bytes.sum",
    "l1tex__m_l1tex2xbar_write_bytes.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed.sum",
    "smsp__inst_executed.avg.per_cycle_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__inst_executed_op_shared_ld.sum",
    "smsp__inst_executed_op_shared_st.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "launch__registers_per_thread",
    "launch__grid_size",
    "launch__block_size",
    "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
]
STALLS = "smsp__pcsamp_warps_issue_stalled_"


def main(path, header):
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = list(csv.reader(lines))
    names, units, rows = rd[0], rd[1], rd[2:]
    col = {n: i for i, n in enumerate(names)}
    print(header)
    for li, r in enumerate(rows):
        kn = r[col["Kernel Name"]]
        print(f"\n== launch {li}: {kn[:110]}")
        for k in KEYS:
            if k in col and r[col[k]] != "":
                print(f"   {k:82s} {units[col[k]]:10s} {r[col[k]]}")
        stalls = []
        for n, i in col.items():
            if n.startswith(STALLS) and not n.endswith("_not_issued") and r[i] not in ("", "0"):
                try:
                    stalls.append((float(r[i].replace(",", "")), n[len(STALLS):]))
                except ValueError:
                    pass
        tot = sum(v for v, _ in stalls) or 1.0
        top = ", ".join(f"{n} {100 * v / tot:.0f}%" for v, n in sorted(stalls, reverse=True)[:7])
        print(f"   pc-sampling stall mix: {top}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
