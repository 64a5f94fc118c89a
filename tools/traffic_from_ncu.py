"""profiles/traffic.json from an ncu raw-page CSV of the enumeration's level-3 launches.

    ncu -i gpurun_out/r3_level3.ncu-rep --page raw --csv > gpurun_out/r3_level3_raw.csv      (here, no GPU needed)
    python tools/traffic_from_ncu.py gpurun_out/r3_level3_raw.csv profiles/r3_level3_ncu_full_summary.csv

Sums dram__bytes_read.sum + dram__bytes_write.sum over the LAST count_unit / scan_groups / fill_unit launch set in
the capture (tools/profile_varlen.py runs the plan a few times; the last run is warm), writes the summary rows that
are worth keeping under profiles/, and stamps traffic.json with the digest of the kernel sources it was captured
from - bench.py quotes the figure only while that digest still matches (bench.csrc_digest)."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sectors.sum", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "smsp__issue_active.avg.pct", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "launch__grid_size", "launch__block_size")


def main():
    src, out_csv = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(src)))
    head = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    cols, units = rows[head], rows[head + 1]
    data = [dict(zip(cols, r)) for r in rows[head + 2:] if len(r) == len(cols)]
    lvl3 = [d for d in data if any(k in d["Kernel Name"] for k in ("count_unit", "fill_unit", "scan_groups"))]
    # the last fill_unit, the count_unit before it and the scan_groups between them
    fi = max(i for i, d in enumerate(lvl3) if "fill_unit" in d["Kernel Name"])
    ci = max(i for i, d in enumerate(lvl3[:fi]) if "count_unit" in d["Kernel Name"])
    pick = lvl3[ci:fi + 1]

    def num(x):
        try:
            return float(x.replace(",", ""))
        except ValueError:
            return 0.0
    unit_of = dict(zip(cols, units))

    def to_bytes(d, name):
        v, u = num(d.get(name, "0")), unit_of.get(name, "byte").lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
    traffic = sum(to_bytes(d, "dram__bytes_read.sum") + to_bytes(d, "dram__bytes_write.sum") for d in pick)
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [k for k in KEEP if k in cols])
        w.writerow(["(unit)"] + [unit_of[k] for k in KEEP if k in cols])
        for d in pick:
            w.writerow([d["Kernel Name"]] + [d[k] for k in KEEP if k in cols])
    import bench
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    dirty = subprocess.run(["git", "-C", ROOT, "status", "--porcelain", "--", "k8s-llm-rca_b200/csrc", "include"], capture_output=True, text=True).stdout.strip()
    tj = {"varlen/level3": int(traffic), "csrc_sha": bench.csrc_digest(), "commit": commit + ("+uncommitted csrc" if dirty else ""),
          "source": os.path.relpath(out_csv, ROOT),
          "note": "dram__bytes_read.sum + dram__bytes_write.sum of the last count_unit + scan_groups + fill_unit launch set "
                  "(enumeration level 3 of config 3) in an ncu --set full capture of tools/profile_varlen.py"}
    json.dump(tj, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(json.dumps(tj, indent=1))
    for d in pick:
        print(d["Kernel Name"][:60], d.get("gpu__time_duration.sum"), unit_of.get("gpu__time_duration.sum"))


if __name__ == "__main__":
    main()
