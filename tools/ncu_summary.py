#!/usr/bin/env python
"""Summarise an `ncu --set full` report (.ncu-rep) per kernel: duration, DRAM traffic, tensor-pipe and DRAM
utilisation, occupancy facts.  `ncu -i <rep> --page raw --csv` is read here (no GPU needed).

  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [out.json]

The JSON (kernel name -> averages over the captured launches) is what bench.py reads for `roofline.traffic`."""
import csv
import io
import json
import re
import subprocess
import sys

METRICS = {
    "gpu__time_duration.sum": "duration_us",
    "dram__bytes_read.sum": "dram_read_bytes",
    "dram__bytes_write.sum": "dram_write_bytes",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed": "tensor_pipe_pct_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "xu_pipe_pct_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "launch__registers_per_thread": "registers",
    "launch__shared_mem_per_block": "smem_per_block",
    "launch__grid_size": "grid_size",
    "launch__block_size": "block_size",
    "launch__waves_per_multiprocessor": "waves_per_sm",
    "sm__cycles_active.avg": "sm_cycles_active",
    "sm__cycles_elapsed.max": "sm_cycles_elapsed",
}
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0, "ms": 1e3, "msecond": 1e3,
         "usecond": 1.0, "nsecond": 1e-3}


def main(rep, out=None):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    acc = {}
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).strip()
        d = acc.setdefault(name, {"launches": 0})
        d["launches"] += 1
        for m, key in METRICS.items():
            if m not in col:
                continue
            v = r[col[m]].replace(",", "")
            try:
                v = float(v) * SCALE.get(units[col[m]], 1.0)
            except ValueError:
                continue
            d[key] = d.get(key, 0.0) + v
    for name, d in acc.items():
        n = d["launches"]
        for k in list(d):
            if k != "launches":
                d[k] = round(d[k] / n, 3)
        if "dram_read_bytes" in d:
            d["dram_traffic_bytes"] = d["dram_read_bytes"] + d["dram_write_bytes"]
    js = json.dumps(acc, indent=1)
    print(js)
    if out:
        with open(out, "w") as f:
            f.write(js + "\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
