"""Summaries of an `ncu --set full` report for profiles/: a text table of the metrics the roofline argument uses
and the small JSON bench.py reads for `roofline.traffic` (no literal in bench.py).

    python tools/ncu_summary.py gpurun_out/r2_prof_n1.ncu-rep profiles/r02_ncu_n1
writes profiles/r02_ncu_n1.txt and profiles/r02_ncu_n1.json (the .ncu-rep itself stays in gpurun_out/).
"""
import csv
import io
import json
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.per_cycle_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
]
TO_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(rep, out_base):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out = io.StringIO()
    out.write(f"ncu --set full --clock-control none, report {rep}\n")
    summary = {"source": rep, "kernels": []}
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        out.write(f"\n== {name}\n")
        k = {"name": name}
        for m in METRICS:
            if m in idx:
                out.write(f"  {m:86s} {r[idx[m]]} {units[idx[m]]}\n")
                k[m] = (r[idx[m]], units[idx[m]])
        rd = float(r[idx["dram__bytes_read.sum"]]) * TO_BYTES.get(units[idx["dram__bytes_read.sum"]], 1)
        wr = float(r[idx["dram__bytes_write.sum"]]) * TO_BYTES.get(units[idx["dram__bytes_write.sum"]], 1)
        k["dram_bytes"] = rd + wr
        out.write(f"  dram bytes read + written per launch: {rd + wr:.0f}\n")
        summary["kernels"].append(k)
        short = name.split("(")[0]
        if "dispatch_local" in short:
            summary["dispatch_dram_bytes"] = rd + wr
        elif "combine_local" in short:
            summary["combine_dram_bytes"] = rd + wr
        elif "layout_kernel" in short:
            summary["layout_dram_bytes"] = rd + wr
    open(out_base + ".txt", "w").write(out.getvalue())
    json.dump(summary, open(out_base + ".json", "w"), indent=1)
    print(out.getvalue())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
