"""Writes profiles/traffic.json from an `ncu --page raw --csv` export of ONE workload's capture:
per kernel, dram__bytes_read.sum + dram__bytes_write.sum and smsp__inst_executed.sum per launch (mean over the captured
launches).  bench.py puts the dominant kernel's figures into roofline.traffic / roofline.issue_slots -- but only when the
capture was taken with the kernel sources the librc.so it loaded was built from: the digest of kernels.cu + kernels.h +
free_flight.h (what csrc/Makefile bakes into rc_build_id()) is recorded here per workload, at capture time.
    python tools/ncu_traffic.py <workload> <raw.csv> [prefix]
prefix: key prefix for captures of an A/B schedule (e.g. "two_kernels/": bench.py's roofline_kernels uses those names)."""
import hashlib
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    workload, path = sys.argv[1], sys.argv[2]
    prefix = sys.argv[3] if len(sys.argv) > 3 else ""
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    acc = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "")
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[col[m]].replace(",", "")) * UNIT[units[col[m]]]
        t_us = float(r[col["gpu__time_duration.sum"]].replace(",", ""))
        t_us *= {"us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(units[col["gpu__time_duration.sum"]], 1.0)
        a = acc.setdefault(name, {"bytes": [], "us": [], "inst": [], "issue": []})
        a["bytes"].append(tot)
        a["us"].append(t_us)
        if "smsp__inst_executed.sum" in col:
            a["inst"].append(float(r[col["smsp__inst_executed.sum"]].replace(",", "")))
        if "smsp__issue_active.avg.pct_of_peak_sustained_active" in col:
            a["issue"].append(float(r[col["smsp__issue_active.avg.pct_of_peak_sustained_active"]].replace(",", "")))
    out_path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        out = json.load(open(out_path))
    except Exception:
        out = {}
    # rc_build_id() of the librc.so that was profiled, written next to the export by tools/ncu_capture.sh; without it, the
    # sources as they stand now (right only if the kernels were not edited since the capture)
    id_path = path.replace("_raw.csv", "_build_id.txt")
    if os.path.exists(id_path):
        build_id = open(id_path).read().strip()
    else:
        digest = hashlib.sha256()
        for name in ("kernels.cu", "kernels.h", "free_flight.h"):
            with open(os.path.join(ROOT, "raytracing_test_b200", "csrc", name), "rb") as f:
                digest.update(f.read())
        build_id = digest.hexdigest()
    ids = out.setdefault("_kernel_sources_sha256", {})
    if ids.get(workload) != build_id:
        out[workload] = {}  # records of other kernel sources do not mix with this capture's
    ids[workload] = build_id
    out.setdefault(workload, {})
    out[workload].update({prefix + k: {"dram_bytes_per_launch": sum(v["bytes"]) / len(v["bytes"]), "launches": len(v["bytes"]),
                         "ncu_duration_us": sum(v["us"]) / len(v["us"]),
                         "warp_instructions_per_launch": sum(v["inst"]) / len(v["inst"]) if v["inst"] else None,
                         "ncu_issue_active_pct": sum(v["issue"]) / len(v["issue"]) if v["issue"] else None,
                         "source": os.path.basename(path)}
                     for k, v in acc.items()})
    json.dump(out, open(out_path, "w"), indent=1)
    print(json.dumps(out[workload], indent=1))


if __name__ == "__main__":
    main()
