#!/usr/bin/env python
"""Condenses an .ncu-rep (ncu --set full) into the handful of per-kernel numbers DESIGN.md and bench.py cite:
    python scripts/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_ncu.json"""
import csv, json, subprocess, sys

KEYS = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed": "issue_active_pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed": "lsu_shared_wavefront_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "shared_bank_conflicts",
    "smsp__inst_executed.sum": "warp_instructions",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid", "launch__block_size": "block",
    "launch__shared_mem_per_block_dynamic": "dyn_smem",
    "smsp__warps_active.avg.per_cycle_active": "warps_active_per_smsp",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
}
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units = rows[0], rows[1]
    res = []
    for row in rows[2:]:
        d = {"kernel": row[head.index("Kernel Name")]}
        for k, name in KEYS.items():
            if k in head:
                i = head.index(k); v = float(row[i].replace(",", "")); u = units[i].split("/")[0]
                if name == "duration": d["duration_ms"] = v * SCALE.get(u, 1.0)
                elif name in ("dram_read", "dram_write", "dyn_smem"): d[name + "_bytes"] = v * SCALE.get(u, 1.0)
                else: d[name] = v
        if "dram_read_bytes" in d:
            d["dram_bytes"] = d["dram_read_bytes"] + d["dram_write_bytes"]
        res.append(d)
    json.dump({"report": sys.argv[1], "note": "ncu --set full --clock-control none; durations are cold-cache, serialised launches", "kernels": res},
              sys.stdout, indent=1)


if __name__ == "__main__":
    main()
