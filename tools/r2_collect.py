#!/usr/bin/env python
"""After `bash tools/r2_evidence.sh` under gpurun: summaries, kernel_traffic.json and copies into profiles/, and the numbers the
tables of profiles/README.md quote (printed)."""
import csv
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT, PROF = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
WORKLOADS = ["imagenet_raw", "imagenet_pglz", "criteo", "imagenet_packed"]


def last_json(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


def launches(path):
    rows = [x for x in csv.reader(l for l in open(path) if not l.startswith("=="))]
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    seq = []
    for x in rows[1:]:
        m = re.search(r"(k_\w+(<[^>]*>)?)\(", x[ki]) or re.search(r"(k_\w+)", x[ki])
        seq.append((m.group(1) if m else x[ki][:40], float(x[vi].replace(",", ""))))
    idx = [i for i, (k, _) in enumerate(seq) if k.startswith("k_front")]
    return seq[idx[-2]:idx[-1]]


def main():
    for n in ("r2_gather", "r2_gather_pglz", "r2_pglz_img_final", "r2_pglz_criteo_final", "r2_wide_final"):
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), os.path.join(OUT, n + ".ncu-rep"),
                        os.path.join(PROF, n + "_ncu_full_summary.csv")], check=True)
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "kernel_traffic.py")], check=True)
    for f in ["r2_bench_default.json", "r2_bench_reference_arm.json"] + ["r2_launches_%s.csv" % w for w in WORKLOADS]:
        shutil.copy(os.path.join(OUT, f), os.path.join(PROF, f))
    d = last_json(os.path.join(PROF, "r2_bench_default.json"))

    def show(name, v):
        r, c = v["roofline"], v["cpu_baseline"]
        print("%-16s %.3f ms %.4g rows/s | e2e %.2f ms %.4g rows/s | dominant %.3f ms %.0f GB/s frac %.4f | stages %s | cpu %.4g (%s cores), C port %.4g"
              % (name, v["ms_per_step"], v["value"], v["e2e"]["ms_per_step"], v["e2e"]["value"], r["launch_ms"], r["achieved"], r["frac"],
                 {k: (round(x, 3) if x else x) for k, x in r["stages_ms"].items()}, c["value"], c["cores"], c.get("c_port_openmp_rows_per_s", 0)))
    show("imagenet_raw", d)
    print("whole step / HBM:", d.get("hbm_frac_whole_step"), "launches:", d["gpu_launches"], "clocks:", d["clocks"])
    for k, v in d["configs"].items():
        show(k, v)
    r = last_json(os.path.join(PROF, "r2_bench_reference_arm.json"))
    print("reference arm:", round(r["value"]), {k: round(v["value"]) for k, v in r["configs"].items()})
    for w in WORKLOADS:
        one = launches(os.path.join(PROF, "r2_launches_%s.csv" % w))
        print(w, " · ".join("%s %d" % (k, v) for k, v in one), "total", sum(v for _, v in one))


if __name__ == "__main__":
    main()
