"""ncu csv (dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum of the conv_* kernels of one
step) -> profiles/rNN_conv_traffic.json, the file bench.py quotes as roofline.traffic.
usage: python tools/conv_traffic_from_ncu.py gpurun_out/conv_traffic.csv profiles/r02_conv_traffic.json"""
import collections
import csv
import json
import sys


def main(src, dst):
    lines = [l for l in open(src) if not l.startswith("==")]
    rows = collections.defaultdict(dict)
    for r in csv.DictReader(lines):
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        if r["Metric Name"].startswith("dram__bytes"):
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        else:
            v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(unit, 1e-6)
        rows[r["ID"]][r["Metric Name"]] = v
        rows[r["ID"]]["name"] = r["Kernel Name"]
    by = {"fprop": dict(launches=0, dram_read=0.0, dram_write=0.0, ms_serialised=0.0),
          "wgrad": dict(launches=0, dram_read=0.0, dram_write=0.0, ms_serialised=0.0)}
    for r in rows.values():
        kind = "wgrad" if "conv_wgrad" in r["name"] else ("fprop" if "conv_fprop" in r["name"] else None)
        if kind is None:
            continue
        b = by[kind]
        b["launches"] += 1
        b["dram_read"] += r.get("dram__bytes_read.sum", 0.0)
        b["dram_write"] += r.get("dram__bytes_write.sum", 0.0)
        b["ms_serialised"] += r.get("gpu__time_duration.sum", 0.0)
    n = sum(b["launches"] for b in by.values())
    total = sum(b["dram_read"] + b["dram_write"] for b in by.values())
    out = {"source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none "
                     "--profile-from-start off -k regex:conv_ python bench.py --steps 1 --warmup 3 --profile --no-graph "
                     "(1xB200, batch 8, one G+D step; cold cache per launch)",
           "kernel": "conv_fprop_kernel/conv_wgrad_kernel", "launches_per_step": n, "dram_bytes_per_step": total,
           "dram_bytes_per_launch": total / max(n, 1), "by_kind": by}
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps({k: v for k, v in out.items() if k != "source"}))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
