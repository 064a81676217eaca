"""Summarise an .ncu-rep (one or more launches of one kernel) into the JSON kept under profiles/.
usage: python scripts/ncu_summary.py gpurun_out/prof_k4.ncu-rep profiles/r02_k4_ncu_summary.json [points_per_launch] [git]
The optional arguments stamp the summary with a `_meta` record (points one profiled launch processed, git hash of the
build that was profiled): bench.py scales `roofline.traffic` from it and reports where the figure comes from."""
import csv
import io
import json
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "smsp__average_warp_latency_per_inst_issued.ratio",
]


def main(rep, out, points_per_launch=None, git=None):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    result = []
    for r in rows[2:]:
        d = dict(zip(head, r))
        u = dict(zip(head, units))
        s = {"Kernel Name": d.get("Kernel Name")}
        for k in KEEP:
            if k in d and d[k] != "":
                s[k] = "%s %s" % (d[k], u.get(k, ""))
        for k, v in d.items():
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k and v:
                if float(v) >= 0.05:
                    s[k] = v
        result.append(s)
    if points_per_launch:
        result.append({"_meta": {"points_per_launch": float(points_per_launch), "git": git, "report": rep}})
    with open(out, "w") as f:
        json.dump(result, f, indent=1)
    print(out, len(result), "launch(es)")


if __name__ == "__main__":
    main(*sys.argv[1:5])
