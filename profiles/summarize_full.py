"""Summarise an `ncu --set full` report exported with `ncu -i X.ncu-rep --page raw --csv`: one block per captured
kernel with the counters the design discussion uses (duration, DRAM traffic, pipe / issue utilisation, launch shape).
usage: python profiles/summarize_full.py gpurun_out/prof_full_raw.csv > profiles/rNN_full_summary.md"""
import csv
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe cycles active %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots active %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("sm__cycles_elapsed.max", "SM cycles elapsed"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "registers / thread"),
]


def main(path):
    rows = list(csv.reader(open(path)))
    H, U = rows[0], rows[1]
    col = {h: i for i, h in enumerate(H)}
    for r in rows[2:]:
        print("### `%s`\n" % r[col["Kernel Name"]][:100])
        print("| counter | value | unit |\n|---|---|---|")
        for key, label in WANT:
            if key in col:
                print("| %s (`%s`) | %s | %s |" % (label, key, r[col[key]], U[col[key]]))
        print()


if __name__ == "__main__":
    main(sys.argv[1])
