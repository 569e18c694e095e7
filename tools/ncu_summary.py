#!/usr/bin/env python3
"""Condense `ncu --set full` captures (.ncu-rep, one launch each) into the JSON bench.py reads
`roofline.traffic` from.   usage: ncu_summary.py out.json class=capture.ncu-rep [class=capture.ncu-rep ...]"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__grid_size", "launch__block_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}


def main():
    out = {}
    for spec in sys.argv[2:]:
        cls, rep = spec.split("=", 1)
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        hdr, units, vals = rows[0], rows[1], rows[2]
        get = lambda k: (float(vals[hdr.index(k)].replace(",", "")), units[hdr.index(k)])
        rd, wr, t = get("dram__bytes_read.sum"), get("dram__bytes_write.sum"), get("gpu__time_duration.sum")
        out[cls] = {"kernel": vals[hdr.index("Kernel Name")].split("(")[0].replace("void ", ""),
                    "capture": "ncu --set full --clock-control none, one launch (%s)" % rep.split("/")[-1],
                    "traffic_bytes": rd[0] * SCALE[rd[1]] + wr[0] * SCALE[wr[1]],
                    "duration_s_under_ncu": t[0] * SCALE[t[1]],
                    "metrics": {k: "%s %s" % (vals[hdr.index(k)], units[hdr.index(k)]) for k in KEEP if k in hdr}}
    with open(sys.argv[1], "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({k: (v["kernel"], v["traffic_bytes"], v["duration_s_under_ncu"]) for k, v in out.items()}))


if __name__ == "__main__":
    main()
