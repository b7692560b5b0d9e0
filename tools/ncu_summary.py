"""Turn an ncu report into the small per-launch table kept under profiles/.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep [-o profiles/x_metrics.csv] [--metric NAME ...]

Reads the report with `ncu -i ... --page raw --csv` (runs here, no GPU needed) and writes one row per metric,
one column per captured launch: `metric,unit,launch0,launch1,...`. Without --metric the list below is used: the
figures DESIGN.md and bench.py quote (duration, DRAM bytes, tensor / XU / FMA / ALU pipe activity, issue
slots, occupancy limits, registers). Section prefixes ncu puts before a metric name are dropped.
"""
import argparse
import csv
import io
import subprocess
import sys

DEFAULT = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers",
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("-o", "--out", default=None)
    ap.add_argument("--metric", action="append", default=None)
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], check=True, capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    header, units, launches = rows[start], rows[start + 1], rows[start + 2:]
    want = args.metric or DEFAULT
    col = {}
    for i, h in enumerate(header):          # "SECTION.Sub.metric__name.sum" -> "metric__name.sum"
        for m in want:
            if h == m or h.endswith("." + m):
                col.setdefault(m, i)
    out = io.StringIO()
    wr = csv.writer(out, lineterminator="\n")
    wr.writerow(["metric", "unit"] + [f"launch{k}" for k in range(len(launches))])
    k_name = header.index("Kernel Name")
    wr.writerow(["Kernel Name", ""] + [r[k_name] for r in launches])
    for m in want:
        if m in col:
            wr.writerow([m, units[col[m]]] + [r[col[m]] for r in launches])
    text = out.getvalue()
    if args.out:
        with open(args.out, "w") as f:
            f.write(text)
    sys.stdout.write(text)


if __name__ == "__main__":
    main()
