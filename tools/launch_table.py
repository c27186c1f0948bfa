"""Per-kernel averages of an `ncu --metrics ... --csv` launch list.
    python tools/launch_table.py gpurun_out/launches.csv [min-grid-z]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = None, collections.OrderedDict()
for r in rows:
    if len(r) > 5 and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        key = (d["Kernel Name"].split("(")[0].replace("<unnamed>::", "").replace("void ", "")[:40], d["Grid Size"])
        data.setdefault(key, collections.defaultdict(list))[d["Metric Name"]].append(float(d["Metric Value"].replace(",", "")))
tot = 0.0
for (k, g), m in data.items():
    t = sum(m["gpu__time_duration.sum"]) / len(m["gpu__time_duration.sum"]) / 1e3
    line = f"{k:40s} {g:18s} n={len(m['gpu__time_duration.sum']):3d}  {t:9.1f} us"
    if "smsp__inst_executed.sum" in m:
        line += f"  {sum(m['smsp__inst_executed.sum']) / len(m['smsp__inst_executed.sum']) / 1e6:8.2f} M inst"
    if "dram__bytes_read.sum" in m:
        line += f"  rd {sum(m['dram__bytes_read.sum']) / len(m['dram__bytes_read.sum']) / 1e6:8.1f} MB"
        line += f"  wr {sum(m['dram__bytes_write.sum']) / len(m['dram__bytes_write.sum']) / 1e6:8.1f} MB"
    print(line)
