"""ncu CSV (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv) of `bench.py` ->
profiles/r2_traffic.json: per kernel, DRAM bytes per launch (largest-grid launches only, i.e. the full-size ones) and the
kernel's share of the summed kernel time.  bench.py reads roofline.traffic from that file.
  python tools/traffic_from_ncu.py gpurun_out/dram.csv profiles/r2_traffic.json "<command that was profiled>" """
import csv
import json
import re
import subprocess
import sys
from collections import defaultdict


def main(src, dst, source):
    rows = [r for r in csv.reader(open(src, errors="replace")) if len(r) >= 15 and r[0].isdigit()]
    per = defaultdict(lambda: defaultdict(float))   # launch id -> metric -> value
    name, grid = {}, {}
    for r in rows:
        i = int(r[0])
        name[i] = re.sub(r"^void ", "", r[4]).split("<")[0].split("(")[0]
        grid[i] = r[8]
        try:
            per[i][r[12]] = float(r[14].replace(",", ""))
        except ValueError:
            pass
    unit_ns = {}
    for r in rows:
        if r[12] == "gpu__time_duration.sum":
            unit_ns[int(r[0])] = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6,
                                  "second": 1e9}.get(r[13], 1.0)
    by = defaultdict(list)
    for i in per:
        by[name[i]].append(i)
    total_ns = sum(per[i].get("gpu__time_duration.sum", 0.0) * unit_ns.get(i, 1.0) for i in per)
    out = {"source": source, "head": subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip(),
           "kernels": {}}
    for k, ids in by.items():
        def gsz(i):
            m = re.findall(r"\d+", grid[i])
            return int(m[0]) if m else 0
        big = max(gsz(i) for i in ids)
        full = [i for i in ids if gsz(i) == big]
        rd = sum(per[i].get("dram__bytes_read.sum", 0.0) for i in full) / len(full)
        wr = sum(per[i].get("dram__bytes_write.sum", 0.0) for i in full) / len(full)
        ns = sum(per[i].get("gpu__time_duration.sum", 0.0) * unit_ns.get(i, 1.0) for i in ids)
        out["kernels"][k] = {"bytes_per_launch": rd + wr, "read": rd, "write": wr, "launches": len(ids), "full_size_launches": len(full),
                             "grid": big, "time_share": ns / total_ns if total_ns else None,
                             "note": f"{rd / 1e6:.1f} MB read + {wr / 1e6:.1f} MB written, grid {big}"}
    json.dump(out, open(dst, "w"), indent=1, sort_keys=True)
    for k, v in sorted(out["kernels"].items(), key=lambda kv: -(kv[1]["time_share"] or 0)):
        print(f"{k:40s} share {100 * (v['time_share'] or 0):5.1f}%  {v['note']}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
