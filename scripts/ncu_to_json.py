"""profiles/rNN_traffic.json from an `ncu --set full` capture of `python bench.py --steps 2 --warmup 3 --quick`:
per stage of the step the kernel's dram bytes per launch, duration under ncu and pipe activity - what bench.py's
`roofline.traffic` / `tensor_pipe_active_pct` quote.   python scripts/ncu_to_json.py <report.ncu-rep> <frames> > out.json"""
import csv
import io
import json
import subprocess
import sys

STAGE = {"frame_grid_build": "points_in_boxes(grid)", "box_mask": "points_in_boxes", "pool_from_mask": "roipoint_pool",
         "fps_bq_dup": "fps_ball_query", "fps_bq_cta": "fps_ball_query(long clouds)", "fps_bq": "fps_ball_query",
         "order_clouds": "fps_ball_query(order)", "build_live_tiles": "sa_mlp(tile list)",
         "build_live_units": "sa_mlp(unit list)", "sa_mlp_tc16": "sa_mlp"}
WANT = {"gpu__time_duration.sum": "ncu_us", "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_active_pct",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "fma_pipe_active_pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
        "launch__registers_per_thread": "registers", "smsp__inst_executed.sum": "warp_instructions",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct"}
UNIT = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}


def main():
    rep, frames = sys.argv[1], int(sys.argv[2])
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    kernels = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        stage = next((v for k, v in STAGE.items() if k in name), None)
        if stage is None or stage in kernels:
            continue
        d = dict(kernel=name.split("(")[0][-60:])
        for k, out in WANT.items():
            if k in col and r[col[k]] not in ("", "n/a"):
                d[out] = float(r[col[k]].replace(",", "")) * UNIT.get(units[col[k]], 1.0)
        d["dram_bytes"] = int(d.pop("dram_read", 0) + d.pop("dram_write", 0))
        kernels[stage] = d
    json.dump(dict(source="%s (ncu --set full --clock-control none, one launch per kernel of one step)" % rep.split("/")[-1],
                   frames_per_step=frames, kernels=kernels), sys.stdout, indent=1)


if __name__ == "__main__":
    main()
