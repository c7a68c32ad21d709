"""Summarise an `ncu --csv --metrics ...` launch list (tools/gpu_ncu_final.sh) per kernel.

    python tools/ncu_summary.py gpurun_out/launches_ResNet50.csv > profiles/rNN_launches_final_summary.csv
"""
import collections
import csv
import re
import sys


def main():
    rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
    rd = csv.DictReader(rows)
    per = collections.OrderedDict()   # launch id -> {metric: value}, kernel name
    for r in rd:
        d = per.setdefault(r["ID"], {"kernel": r["Kernel Name"]})
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1.0)
        d[r["Metric Name"]] = v * scale
    agg = collections.OrderedDict()
    for d in per.values():
        name = d["kernel"].split("(")[0]
        name = re.sub(r"<[^<>]*>$", "", name)          # template arguments
        name = name.split("::")[-1].replace("void ", "")
        a = agg.setdefault(name, [0, 0.0, 0.0, 0.0, 0.0, 0.0])
        t = d.get("gpu__time_duration.sum", 0.0)
        a[0] += 1
        a[1] += t
        a[2] += d.get("dram__bytes_read.sum", 0.0)
        a[3] += d.get("dram__bytes_write.sum", 0.0)
        a[4] += d.get("lts__t_bytes.sum", 0.0)
        a[5] += d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0) * t
    total = sum(a[1] for a in agg.values())
    print("kernel,launches,total_us,share,dram_read_MB,dram_write_MB,l2_bytes_MB,tensor_pipe_active_pct_time_weighted")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name},{a[0]},{a[1]:.1f},{a[1] / total:.3f},{a[2]:.0f},{a[3]:.0f},{a[4]:.0f},{a[5] / a[1] if a[1] else 0:.1f}")
    print(f"# total,{sum(a[0] for a in agg.values())},{total:.1f}")


if __name__ == "__main__":
    main()
