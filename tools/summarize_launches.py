"""Summarise an `ncu --csv --metrics gpu__time_duration.sum[,dram__bytes_*]` launch list:
per kernel name -> launches, total device time, share of all launches, DRAM bytes."""
import collections
import csv
import sys

path = sys.argv[1]
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.defaultdict(lambda: collections.defaultdict(float))
cnt = collections.Counter()
for row in csv.DictReader(lines):
    name = row["Kernel Name"]
    short = name.split("(")[0].replace("void ", "")[-70:]
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    m = row["Metric Name"]
    if m == "gpu__time_duration.sum":
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)   # -> ms
        cnt[short] += 1
    else:
        v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    agg[short][m] += v
tot = sum(a["gpu__time_duration.sum"] for a in agg.values())
print(f"{path}: {sum(cnt.values())} launches, {tot:.2f} ms of device time (cold-cache, serialised under ncu)")
print(f"{'ms':>9s} {'share':>6s} {'n':>5s} {'DRAM rd GB':>10s} {'DRAM wr GB':>10s}  kernel")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"])[:int(sys.argv[2]) if len(sys.argv) > 2 else 16]:
    print(f"{a['gpu__time_duration.sum']:9.2f} {100 * a['gpu__time_duration.sum'] / tot:5.1f}% {cnt[k]:5d} "
          f"{a.get('dram__bytes_read.sum', 0) / 1e9:10.2f} {a.get('dram__bytes_write.sum', 0) / 1e9:10.2f}  {k}")

# per-launch DRAM traffic of the dominant kernel (bench.py's roofline.traffic reads this file)
if len(sys.argv) > 3:
    import json
    keys = [k for k in agg if "tc2_fwd_kernel" in k]
    n = sum(cnt[k] for k in keys)
    rd = sum(agg[k].get("dram__bytes_read.sum", 0) for k in keys)
    wr = sum(agg[k].get("dram__bytes_write.sum", 0) for k in keys)
    if n and (rd + wr) > 0:
        json.dump({"kernel": "tc2_fwd_kernel", "launches": n, "dram_bytes_per_launch": (rd + wr) / n,
                   "dram_read_bytes": rd, "dram_write_bytes": wr, "source": path,
                   "note": "ncu dram__bytes_read.sum + dram__bytes_write.sum over every tc2_fwd_kernel launch of one timed bench step"},
                  open(sys.argv[3], "w"), indent=1)
        print("wrote", sys.argv[3])
