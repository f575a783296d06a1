#!/usr/bin/env python
"""profiles/kernel_traffic.json from the `ncu --set full` summaries tools/ncu_summary.py wrote (bench.py reads it for roofline.traffic)."""
import csv
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CAPTURES = {"imagenet_raw": ("k_toast_gather", "r2_gather"), "imagenet_pglz": ("k_pglz_decode", "r2_pglz_img_final"),
            "criteo": ("k_pglz_decode", "r2_pglz_criteo_final")}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-3, "ms": 1.0, "s": 1e3}


def main():
    out = {}
    for workload, (kernel, name) in CAPTURES.items():
        path = os.path.join(ROOT, "profiles", name + "_ncu_full_summary.csv")
        rows = list(csv.reader(open(path)))
        head, units = rows[0], rows[1]
        row = next(r for r in rows[2:] if kernel in r[head.index("Kernel Name")])

        def get(key):
            i = head.index(key)
            return float(row[i]) * SCALE.get(units[i], 1.0)
        rd, wr = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
        out[workload] = {"kernel": kernel,
                         "source": "profiles/%s_ncu_full_summary.csv (ncu --set full --clock-control none, tools/r2_evidence.sh)" % name,
                         "dram_bytes_read_per_launch": rd, "dram_bytes_write_per_launch": wr, "dram_bytes_per_launch": rd + wr,
                         "duration_ms_under_ncu": get("gpu__time_duration.sum"), "warp_instructions": get("smsp__inst_executed.sum"),
                         "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                         "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active"),
                         "note": "writes still dirty in the 126 MB L2 at kernel end are not counted by dram__bytes_write"}
    with open(os.path.join(ROOT, "profiles", "kernel_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(json.dumps({k: v["dram_bytes_per_launch"] for k, v in out.items()}))


if __name__ == "__main__":
    main()
