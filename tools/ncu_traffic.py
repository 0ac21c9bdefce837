#!/usr/bin/env python
"""`ncu -i X.ncu-rep --page raw --csv` -> profiles/ncu_traffic.json: per-launch DRAM traffic (dram__bytes_read.sum +
dram__bytes_write.sum) of the kernels bench.py quotes, looked up there by KEY instead of being typed into the source.

    python tools/ncu_traffic.py profiles/r02_ncu_full_roofline.csv conv_fprop_3x3_256_14=conv_tc_gemm_pair_kernel \
                                                                bn_bwd_apply_256_56=bn_bwd_apply_kernel

Each KEY=REGEX picks the LAST launch whose kernel name matches (the earlier ones warm the caches)."""
import csv
import json
import os
import re
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path, specs = sys.argv[1], sys.argv[2:]
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    out_path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    out = json.load(open(out_path)) if os.path.exists(out_path) else {}
    for spec in specs:
        key, rx = spec.split("=", 1)
        hits = [r for r in data if re.search(rx, r[col["Kernel Name"]])]
        if not hits:
            print("no launch matches", rx)
            continue
        r = hits[-1]
        rd = float(r[col["dram__bytes_read.sum"]]) * UNIT[units[col["dram__bytes_read.sum"]]]
        wr = float(r[col["dram__bytes_write.sum"]]) * UNIT[units[col["dram__bytes_write.sum"]]]
        us = float(r[col["gpu__time_duration.sum"]]) * {"us": 1.0, "ms": 1e3, "ns": 1e-3}[units[col["gpu__time_duration.sum"]]]
        out[key] = {"dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr, "duration_us_under_ncu": us, "kernel": r[col["Kernel Name"]][:120],
                    "source": os.path.relpath(os.path.abspath(path), ROOT), "launches_matched": len(hits)}
        print(key, out[key])
    json.dump(out, open(out_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
