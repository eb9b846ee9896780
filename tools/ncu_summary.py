#!/usr/bin/env python
"""Summarise an .ncu-rep (`--set full`) into a small text table + JSON of DRAM bytes per instance.

    python tools/ncu_summary.py prof.ncu-rep <instances-per-launch> [out.json]
"""
import csv
import io
import json
import re
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_blocks", "sm__cycles_elapsed.avg"]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def main():
    rep, inst = sys.argv[1], int(sys.argv[2])
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    traffic = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        print("==", name)
        for w in WANT:
            if w in col:
                print(f"   {w:68s} {r[col[w]]:>18s} {units[col[w]]}")
        rd = float(r[col["dram__bytes_read.sum"]]) * SCALE[units[col["dram__bytes_read.sum"]]]
        wr = float(r[col["dram__bytes_write.sum"]]) * SCALE[units[col["dram__bytes_write.sum"]]]
        ins = float(r[col["smsp__inst_executed.sum"]])
        print(f"   {'-> dram bytes per instance':68s} {(rd + wr) / inst:18.0f}")
        print(f"   {'-> warp instructions per instance':68s} {ins / inst:18.0f}")
        short = name.split("(")[0].split("::")[-1]
        # template arguments as ncu prints them -> the names dls_kernel_name() reports
        m = re.match(r"(sched_(?:stream|simple)_kernel)<(?:\(int\))?([0-3])(?:,\s*(?:\(int\))?\d+)?>", short)
        if m:
            short = f"{m.group(1)}<{('DFS', 'Greedy', 'Critical', 'MRU')[int(m.group(2))]}>"
        short = re.sub(r"^sched_stream_mru_kernel<.*>$", "sched_stream_mru_kernel", short)
        short = short.replace("stream_build_staged_kernel", "stream_build_kernel")
        traffic[short] = {"dram_bytes_per_instance": (rd + wr) / inst, "warp_inst_per_instance": ins / inst,
                          "instances_in_capture": inst}
    if len(sys.argv) > 3:
        json.dump(traffic, open(sys.argv[3], "w"), indent=1)


if __name__ == "__main__":
    main()
